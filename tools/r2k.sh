#!/bin/bash
# Round 2, 1-GPU call: the REDUX vote (one packed reduction per loop iteration instead of ballots) against the default.
TAG=${1:-r2k}; OUT=gpurun_out; mkdir -p $OUT
L=$PWD/raytracing-in-a-weekend_b200/csrc
RTW_B200_LIB=$L/librtw_redux.so timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -q -x -k "path_parity or pipelines_agree or statistically" > $OUT/${TAG}_pytest_redux.log 2>&1; echo "pytest(redux) rc=$?"; tail -2 $OUT/${TAG}_pytest_redux.log
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -q -x > $OUT/${TAG}_pytest_parity.log 2>&1; echo "pytest(parity, default) rc=$?"; tail -2 $OUT/${TAG}_pytest_parity.log
AB=$OUT/${TAG}_ab.txt; : > $AB
run() { printf "%-46s " "$1" | tee -a $AB; shift; env "$@" 2>&1 | python tools/bench_line.py | tee -a $AB; }
B="timeout 300 python bench.py --steps 3 --warmup 3 --cpu-seconds 1"
for rep in 1 2; do
  run "C5 1250spp ballots (default)"   X=1 $B --spp 1250
  run "C5 1250spp REDUX"               RTW_B200_LIB=$L/librtw_redux.so $B --spp 1250
  run "C2 ballots (default)"           X=1 $B --config C2
  run "C2 REDUX"                       RTW_B200_LIB=$L/librtw_redux.so $B --config C2
  run "C4 625spp ballots (default)"    X=1 $B --config C4 --spp 625
  run "C4 625spp REDUX"                RTW_B200_LIB=$L/librtw_redux.so $B --config C4 --spp 625
done
timeout 600 python bench.py --steps 3 --warmup 3 --cpu-seconds 2 > $OUT/${TAG}_bench_C5_n1.json 2>/dev/null; printf "headline: "; python tools/bench_line.py < $OUT/${TAG}_bench_C5_n1.json
