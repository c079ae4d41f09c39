#!/bin/bash
# Round 2, 1-GPU call: the no-allocation filter entry point (alone in a fresh process, then after the whole suite),
# and the megakernel split (plain / light-biased instances) against the previous build.
TAG=${1:-r2o}; OUT=gpurun_out; mkdir -p $OUT
L=$PWD/raytracing-in-a-weekend_b200/csrc
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q > $OUT/${TAG}_pytest_noalloc_alone.log 2>&1; echo "noalloc alone rc=$?"; tail -n 3 $OUT/${TAG}_pytest_noalloc_alone.log
timeout 1500 python -m pytest tests -m gpu -x -q > $OUT/${TAG}_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -n 5 $OUT/${TAG}_pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $OUT/${TAG}_smoke.log 2>&1; echo "smoke rc=$?"; tail -n 1 $OUT/${TAG}_smoke.log
AB=$OUT/${TAG}_ab.txt; : > $AB
run() { printf "%-52s " "$1" | tee -a $AB; shift; env "$@" 2>&1 | python tools/bench_line.py | tee -a $AB; }
B="timeout 300 python bench.py --steps 5 --warmup 3 --cpu-seconds 1 --flags 8"
for rep in 1 2; do
  run "C1 megakernel, previous build"            RTW_B200_LIB=$L/librtw_head.so $B --config C1
  run "C1 megakernel, split instances"           $B --config C1
  run "C2 100spp megakernel, previous build"     RTW_B200_LIB=$L/librtw_head.so $B --config C2 --spp 100
  run "C2 100spp megakernel, split instances"    $B --config C2 --spp 100
  run "C5 50spp megakernel, previous build"      RTW_B200_LIB=$L/librtw_head.so $B --spp 50
  run "C5 50spp megakernel, split instances"     $B --spp 50
done
for lib in librtw_head.so librtw_b200.so; do echo "== $lib" | tee -a $AB; RTW_B200_LIB=$L/$lib timeout 300 python tools/light_timing.py 2>&1 | tee -a $AB; done
