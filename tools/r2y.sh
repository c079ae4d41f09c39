#!/bin/bash
# Round 2, 1-GPU call: guided sample supply at the end of the frame (RTW_WF_GUIDED_SUPPLY=0/1), short frames.
TAG=${1:-r2y}; OUT=gpurun_out; mkdir -p $OUT
timeout 900 python -m pytest tests -m gpu -q -x > $OUT/${TAG}_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -n 3 $OUT/${TAG}_pytest_gpu.log
AB=$OUT/${TAG}_ab_guided_supply.txt; : > $AB
run() { printf "%-34s " "$1" | tee -a $AB; shift; env "$@" 2>&1 | python tools/bench_line.py | tee -a $AB; }
B="timeout 300 python bench.py --steps 3 --warmup 3 --cpu-seconds 1"
for rep in 1 2; do
  for g in 0 1; do
    run "C5 400spp guided=$g"   RTW_WF_GUIDED_SUPPLY=$g $B --spp 400
    run "C5 1250spp guided=$g"  RTW_WF_GUIDED_SUPPLY=$g $B --spp 1250
    run "C4 625spp guided=$g"   RTW_WF_GUIDED_SUPPLY=$g $B --config C4 --spp 625
    run "C2 guided=$g"          RTW_WF_GUIDED_SUPPLY=$g $B --config C2
  done
done
