#!/bin/bash
# Round 2, third GPU call (1 GPU): the bench's other partitions and the single-process mode on one device (a dry run of what the
# 8-GPU call does), drain-mode A/B with the pool-sized chunk, two half-frame renders side by side on one device.
TAG=${1:-r2c}; OUT=gpurun_out; mkdir -p $OUT
AB=$OUT/${TAG}_ab.txt; : > $AB
run() { printf "%-52s " "$1" | tee -a $AB; shift; env "$@" 2>&1 | python tools/bench_line.py | tee -a $AB; }
B="timeout 300 python bench.py --steps 3 --warmup 3 --cpu-seconds 1"
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_multi_device.py -m gpu -q -x > $OUT/${TAG}_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 $OUT/${TAG}_pytest.log
run "C4 500spp partition tiles (1 GPU)"                 X=1 $B --config C4 --spp 500 --partition tiles
run "C5 200spp partition frames (1 GPU)"                X=1 $B --spp 200 --partition frames
run "C5 1000spp single-process devices 0"               X=1 $B --spp 1000 --single-process --devices 0
run "C5 1000spp single-process devices 0,0"             X=1 $B --spp 1000 --single-process --devices 0,0
run "C5 1000spp single-process devices 0,0,0"           X=1 $B --spp 1000 --single-process --devices 0,0,0
run "C4 500spp single-process tiles devices 0,0"        X=1 $B --config C4 --spp 500 --single-process --devices 0,0 --partition tiles
for rep in 1 2; do
  run "C5 1250spp drain mode (default)"        X=1 $B --spp 1250
  run "C5 1250spp RTW_WF_NO_DRAIN=1"           RTW_WF_NO_DRAIN=1 $B --spp 1250
  run "C5 200spp drain mode (default)"         X=1 $B --spp 200
  run "C5 200spp RTW_WF_NO_DRAIN=1"            RTW_WF_NO_DRAIN=1 $B --spp 200
  run "C5 1250spp greedy collapse"             RTW_BVH_COLLAPSE=greedy $B --spp 1250
done
timeout 600 python bench.py --steps 3 --warmup 3 --cpu-seconds 2 > $OUT/${TAG}_bench_C5_n1.json 2> $OUT/${TAG}_bench_C5_n1.err; echo "bench rc=$?"; python tools/bench_line.py < $OUT/${TAG}_bench_C5_n1.json
