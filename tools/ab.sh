#!/bin/bash
# A/B tuning helper (GPU box): runs bench.py with each library given (names under csrc/, without
# "librtw_" and ".so"), interleaved REPS times, and prints the key numbers. Usage:
#   tools/ab.sh "base b200" [spp] [reps] [extra bench args]
LIBS=${1:-"b200"}; SPP=${2:-200}; REPS=${3:-2}; shift 3
for r in $(seq $REPS); do
  for l in $LIBS; do
    printf "%-10s " "$l"
    RTW_B200_LIB=$PWD/raytracing-in-a-weekend_b200/csrc/librtw_$l.so timeout 300 python bench.py --spp $SPP --steps 2 --warmup 3 --cpu-seconds 1 "$@" 2>&1 | python tools/bench_line.py
  done
done
