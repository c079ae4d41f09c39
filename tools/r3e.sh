#!/bin/bash
# Round 2, 1-GPU call: extend's tail zone (the pool's last slots reserved 32 at a time), RTW_WF_TAIL_ZONE = 0 / 1 / 2 / 4
# chunks per warp; the pipeline-rule test.
TAG=${1:-r3e}; OUT=gpurun_out; mkdir -p $OUT
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -q -x -k "pipeline or agree or memory" > $OUT/${TAG}_pytest_subset.log 2>&1; echo "pytest rc=$?"; tail -n 3 $OUT/${TAG}_pytest_subset.log
AB=$OUT/${TAG}_ab_extend_tail_zone.txt; : > $AB
run() { printf "%-34s " "$1" | tee -a $AB; shift; env "$@" 2>&1 | python tools/bench_line.py | tee -a $AB; }
B="timeout 300 python bench.py --steps 3 --warmup 3 --cpu-seconds 1"
for rep in 1 2; do
  for z in 0 1 2 4; do
    run "C5 400spp zone=$z"   RTW_WF_TAIL_ZONE=$z $B --spp 400
    run "C5 1250spp zone=$z"  RTW_WF_TAIL_ZONE=$z $B --spp 1250
    run "C4 625spp zone=$z"   RTW_WF_TAIL_ZONE=$z $B --config C4 --spp 625
  done
done
for z in 0 1; do run "C5 full zone=$z" RTW_WF_TAIL_ZONE=$z $B; done
