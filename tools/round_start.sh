#!/bin/bash
# First GPU call of a round (run under gpurun from the repo root): the whole `-m gpu` suite, smoke, the
# default bench line, then — each only after its own command exited 0 without ncu — the ncu launch list
# and one full capture of the dominant kernel, all into gpurun_out/ under the given tag. Usage:
#   gpurun --timeout 1500 -- 'bash tools/round_start.sh r2a'
TAG=${1:-rNx}; OUT=gpurun_out; mkdir -p $OUT
timeout 900 python -m pytest tests -m gpu -q -x > $OUT/${TAG}_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 $OUT/${TAG}_pytest_gpu.log
timeout 120 python -c "import __graft_entry__ as g; g.smoke()" > $OUT/${TAG}_smoke.log 2>&1; echo "smoke rc=$?"; tail -1 $OUT/${TAG}_smoke.log
timeout 300 python bench.py > $OUT/${TAG}_bench_C5_n1.json 2> $OUT/${TAG}_bench_C5_n1.err; echo "bench rc=$?"; python tools/bench_line.py < $OUT/${TAG}_bench_C5_n1.json
timeout 300 python bench.py --spp 400 --steps 2 --warmup 3 --cpu-seconds 1 > $OUT/${TAG}_bench_spp400.json 2>/dev/null &&
  timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $OUT/${TAG}_launches_bench_spp400.csv \
    python bench.py --spp 400 --steps 2 --warmup 3 --cpu-seconds 1 > $OUT/${TAG}_ncu_launches.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:wf_extend_kernel -s 150 -c 1 -o $OUT/${TAG}_wf_extend \
    python bench.py --spp 400 --steps 2 --warmup 3 --cpu-seconds 1 > $OUT/${TAG}_ncu_full.log 2>&1; echo "ncu full rc=$?"
