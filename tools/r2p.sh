#!/bin/bash
# Round 2, 1-GPU call: megakernel as two instances (plain / light-biased) against the single-instance build.
TAG=${1:-r2p}; OUT=gpurun_out; mkdir -p $OUT
L=$PWD/raytracing-in-a-weekend_b200/csrc
AB=$OUT/${TAG}_ab.txt; : > $AB
run() { printf "%-52s " "$1" | tee -a $AB; shift; env "$@" 2>&1 | python tools/bench_line.py | tee -a $AB; }
B="timeout 300 python bench.py --steps 5 --warmup 3 --cpu-seconds 1 --flags 8"
for rep in 1 2; do
  run "C1 megakernel, one instance"              RTW_B200_LIB=$L/librtw_head.so $B --config C1
  run "C1 megakernel, split instances"           $B --config C1
  run "C2 100spp megakernel, one instance"       RTW_B200_LIB=$L/librtw_head.so $B --config C2 --spp 100
  run "C2 100spp megakernel, split instances"    $B --config C2 --spp 100
  run "C5 50spp megakernel, one instance"        RTW_B200_LIB=$L/librtw_head.so $B --spp 50
  run "C5 50spp megakernel, split instances"     $B --spp 50
done
for lib in librtw_head.so librtw_b200.so; do echo "== $lib" | tee -a $AB; RTW_B200_LIB=$L/$lib timeout 300 python tools/light_timing.py 2>&1 | tee -a $AB; done
timeout 600 python -m pytest tests/test_lbvh.py -m gpu -x -q 2>&1 | tail -n 3 | tee $OUT/${TAG}_pytest_lbvh.log
for c in C5 C2 C4 C3; do timeout 300 python tools/lbvh_probe.py $c 100 2>&1 | tee -a $OUT/${TAG}_lbvh_probe.txt; done
