#!/bin/bash
# Round 2, 1-GPU call: streaming (evict-first) cache operators on the path records, in extend (1), shade (2), both (3).
TAG=${1:-r3a}; OUT=gpurun_out; mkdir -p $OUT
L=$PWD/raytracing-in-a-weekend_b200/csrc
AB=$OUT/${TAG}_ab_streaming_records.txt; : > $AB
run() { printf "%-34s " "$1" | tee -a $AB; shift; env "$@" 2>&1 | python tools/bench_line.py | tee -a $AB; }
B="timeout 300 python bench.py --steps 3 --warmup 3 --cpu-seconds 1"
for rep in 1 2; do
  for lib in b200 st1 st2 st3; do
    run "C5 625spp $lib"   RTW_B200_LIB=$L/librtw_$lib.so $B --spp 625
    run "C5 2500spp $lib"  RTW_B200_LIB=$L/librtw_$lib.so $B --spp 2500
    run "C4 625spp $lib"   RTW_B200_LIB=$L/librtw_$lib.so $B --config C4 --spp 625
  done
done
