#!/bin/bash
# Round 2, 1-GPU call: megakernel (two instances) against the wavefront pipeline on every config; device rebuild builders.
TAG=${1:-r2q}; OUT=gpurun_out; mkdir -p $OUT
AB=$OUT/${TAG}_mega_vs_wavefront.txt; : > $AB
run() { printf "%-40s " "$1" | tee -a $AB; shift; env "$@" 2>&1 | python tools/bench_line.py | tee -a $AB; }
B="timeout 300 python bench.py --steps 3 --warmup 3 --cpu-seconds 1"
for c in C1 C2 C3; do
  run "$c megakernel"  $B --config $c --flags 8
  run "$c wavefront"   $B --config $c --flags 2
done
run "C4 625spp megakernel"  $B --config C4 --spp 625 --flags 8
run "C4 625spp wavefront"   $B --config C4 --spp 625 --flags 2
run "C5 625spp megakernel"  $B --spp 625 --flags 8
run "C5 625spp wavefront"   $B --spp 625 --flags 2
for c in C5 C2; do timeout 300 python tools/lbvh_probe.py $c 100 2>&1 | tee -a $OUT/${TAG}_lbvh_probe.txt; done
