#!/bin/bash
# Round 2, 1-GPU call: the 96-register megakernel with its own occupancy (five blocks per SM), chunk cap sweep,
# per-config bench lines, ncu capture of the megakernel on C3.
TAG=${1:-r2u}; OUT=gpurun_out; mkdir -p $OUT
timeout 900 python -m pytest tests -m gpu -q > $OUT/${TAG}_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -n 4 $OUT/${TAG}_pytest_gpu.log
AB=$OUT/${TAG}_ab_mega_chunk.txt; : > $AB
for rep in 1 2; do
  for cm in 16 32 64 128 256; do
    echo "== RTW_MEGA_CHUNK_MAX=$cm (repeat $rep)" | tee -a $AB
    RTW_MEGA_CHUNK_MAX=$cm timeout 300 python tools/crossover.py C1,C3 100,1000 3 mega 2>&1 | grep -v config | tee -a $AB
    RTW_MEGA_CHUNK_MAX=$cm timeout 300 python tools/crossover.py C2,C4,C5 16,64 3 mega 2>&1 | grep -v config | tee -a $AB
  done
done
for c in C1 C2 C3 C4; do timeout 300 python bench.py --config $c --steps 3 --warmup 3 --cpu-seconds 3 > $OUT/${TAG}_bench_$c.json 2>/dev/null; printf "$c "; python tools/bench_line.py < $OUT/${TAG}_bench_$c.json; done
timeout 600 ncu --set full --clock-control none --import-source on -k regex:mega_kernel -s 3 -c 1 -o $OUT/${TAG}_mega_C3 \
    python bench.py --config C3 --spp 200 --steps 2 --warmup 3 --cpu-seconds 1 > $OUT/${TAG}_ncu_mega.log 2>&1; echo "ncu mega rc=$?"
timeout 300 python tools/light_timing.py 2>&1 | tee $OUT/${TAG}_light_timing.txt
