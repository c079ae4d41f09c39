#!/bin/bash
# Round 2, 1-GPU call: eval_prim / eval_hit as one convergent flow (one texture lookup and material load per warp) against
# the build before it (librtw_head.so), every config; the gpu suite on the new build.
TAG=${1:-r2v}; OUT=gpurun_out; mkdir -p $OUT
L=$PWD/raytracing-in-a-weekend_b200/csrc
timeout 900 python -m pytest tests -m gpu -q > $OUT/${TAG}_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -n 4 $OUT/${TAG}_pytest_gpu.log
AB=$OUT/${TAG}_ab_convergent_hit.txt; : > $AB
run() { printf "%-34s " "$1" | tee -a $AB; shift; env "$@" 2>&1 | python tools/bench_line.py | tee -a $AB; }
B="timeout 300 python bench.py --steps 3 --warmup 3 --cpu-seconds 1"
for rep in 1 2; do
  for lib in head b200; do
    run "C5 625spp $lib"  RTW_B200_LIB=$L/librtw_$lib.so $B --spp 625
    run "C4 625spp $lib"  RTW_B200_LIB=$L/librtw_$lib.so $B --config C4 --spp 625
    run "C2 $lib"         RTW_B200_LIB=$L/librtw_$lib.so $B --config C2
    run "C3 $lib"         RTW_B200_LIB=$L/librtw_$lib.so $B --config C3
    run "C1 $lib"         RTW_B200_LIB=$L/librtw_$lib.so $B --config C1
    run "C3 wavefront $lib" RTW_B200_LIB=$L/librtw_$lib.so $B --config C3 --flags 2
  done
done
