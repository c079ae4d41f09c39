#!/bin/bash
# Round 2, validation of the final build (pipeline rule, 96-register megakernel, one-tail Hit in wf_shade, streaming records):
# Hit in wf_shade: smoke, the headline bench, per-config lines, ncu launch list.
TAG=${1:-r3b}; OUT=gpurun_out; mkdir -p $OUT
nvidia-smi --query-gpu=name,clocks.max.sm,memory.total --format=csv,noheader > $OUT/${TAG}_gpu.txt; nproc >> $OUT/${TAG}_gpu.txt
timeout 900 python -m pytest tests -m gpu -q > $OUT/${TAG}_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -n 3 $OUT/${TAG}_pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $OUT/${TAG}_smoke.log 2>&1; echo "smoke rc=$?"; tail -n 2 $OUT/${TAG}_smoke.log
timeout 600 python bench.py --steps 3 --warmup 3 > $OUT/${TAG}_bench_C5_n1.json 2> $OUT/${TAG}_bench_C5_n1.err; echo "bench rc=$?"; python tools/bench_line.py < $OUT/${TAG}_bench_C5_n1.json
for c in C1 C2 C3 C4; do timeout 300 python bench.py --config $c --steps 3 --warmup 3 --cpu-seconds 3 > $OUT/${TAG}_bench_$c.json 2>/dev/null; printf "$c "; python tools/bench_line.py < $OUT/${TAG}_bench_$c.json; done
timeout 300 python bench.py --spp 1250 --steps 3 --warmup 3 --cpu-seconds 1 > $OUT/${TAG}_bench_C5_spp1250.json 2>/dev/null; printf "C5 1250spp "; python tools/bench_line.py < $OUT/${TAG}_bench_C5_spp1250.json
timeout 300 python bench.py --spp 400 --steps 2 --warmup 3 --cpu-seconds 1 > $OUT/${TAG}_bench_spp400.json 2>/dev/null &&
  timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 1400 --csv --log-file $OUT/${TAG}_launches_bench_spp400.csv \
    python bench.py --spp 400 --steps 1 --warmup 1 --cpu-seconds 1 > $OUT/${TAG}_ncu_launches.log 2>&1; echo "ncu launches rc=$?"
timeout 120 python bench.py --impl reference --steps 1 --warmup 0 > $OUT/${TAG}_bench_reference.json 2>/dev/null; tail -c 400 $OUT/${TAG}_bench_reference.json
