#!/bin/bash
# Round 2, 1-GPU call: the whole gpu suite on the build with the 2^24 rule and the 96-register megakernel; megakernel
# knobs (mid-walk refill, chunk cap); per-config bench lines; ncu capture of the megakernel.
#   gpurun --timeout 1500 -- 'bash tools/r2t.sh r2t'
TAG=${1:-r2t}; OUT=gpurun_out; mkdir -p $OUT
timeout 900 python -m pytest tests -m gpu -q > $OUT/${TAG}_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -n 8 $OUT/${TAG}_pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $OUT/${TAG}_smoke.log 2>&1; echo "smoke rc=$?"; tail -n 2 $OUT/${TAG}_smoke.log
AB=$OUT/${TAG}_ab_mega_knobs.txt; : > $AB
for rep in 1 2; do
  for rb in 1 8 16 24; do
    echo "== RTW_REFILL_BELOW=$rb (repeat $rep)" | tee -a $AB
    RTW_REFILL_BELOW=$rb timeout 300 python tools/crossover.py C1,C3 100 3 mega 2>&1 | grep -v config | tee -a $AB
    RTW_REFILL_BELOW=$rb timeout 300 python tools/crossover.py C2,C4,C5 16 3 mega 2>&1 | grep -v config | tee -a $AB
  done
  for cm in 64 256 1024; do
    echo "== RTW_MEGA_CHUNK_MAX=$cm (repeat $rep)" | tee -a $AB
    RTW_MEGA_CHUNK_MAX=$cm timeout 300 python tools/crossover.py C1,C3 1000 3 mega 2>&1 | grep -v config | tee -a $AB
    RTW_MEGA_CHUNK_MAX=$cm timeout 300 python tools/crossover.py C2,C4 64 3 mega 2>&1 | grep -v config | tee -a $AB
  done
done
for c in C1 C2 C3 C4; do timeout 300 python bench.py --config $c --steps 3 --warmup 3 --cpu-seconds 3 > $OUT/${TAG}_bench_$c.json 2>/dev/null; printf "$c "; python tools/bench_line.py < $OUT/${TAG}_bench_$c.json; done
timeout 600 ncu --set full --clock-control none --import-source on -k regex:mega_kernel -s 3 -c 1 -o $OUT/${TAG}_mega_C3 \
    python bench.py --config C3 --spp 200 --steps 2 --warmup 3 --cpu-seconds 1 > $OUT/${TAG}_ncu_mega.log 2>&1; echo "ncu mega rc=$?"
