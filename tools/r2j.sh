#!/bin/bash
# Round 2, 1-GPU call: fewer exchanges in the node step's sorting network (the ALU pipe is the busiest: 62 %).
TAG=${1:-r2j}; OUT=gpurun_out; mkdir -p $OUT
L=$PWD/raytracing-in-a-weekend_b200/csrc
AB=$OUT/${TAG}_ab.txt; : > $AB
run() { printf "%-46s " "$1" | tee -a $AB; shift; env "$@" 2>&1 | python tools/bench_line.py | tee -a $AB; }
B="timeout 300 python bench.py --steps 3 --warmup 3 --cpu-seconds 1"
for rep in 1 2; do
  run "C5 1250spp 5 exchanges (default)"   X=1 $B --spp 1250
  run "C5 1250spp 4 exchanges"             RTW_B200_LIB=$L/librtw_sort4.so $B --spp 1250
  run "C5 1250spp 3 exchanges"             RTW_B200_LIB=$L/librtw_sort3.so $B --spp 1250
  run "C2 5 exchanges (default)"           X=1 $B --config C2
  run "C2 4 exchanges"                     RTW_B200_LIB=$L/librtw_sort4.so $B --config C2
  run "C2 3 exchanges"                     RTW_B200_LIB=$L/librtw_sort3.so $B --config C2
  run "C4 625spp 5 exchanges (default)"    X=1 $B --config C4 --spp 625
  run "C4 625spp 4 exchanges"              RTW_B200_LIB=$L/librtw_sort4.so $B --config C4 --spp 625
  run "C4 625spp 3 exchanges"              RTW_B200_LIB=$L/librtw_sort3.so $B --config C4 --spp 625
done
S=$OUT/${TAG}_sched.txt; : > $S
for v in b200 sort4 sort3; do printf "%-6s " $v | tee -a $S; RTW_B200_LIB=$L/librtw_$v.so timeout 300 python bench.py --steps 2 --warmup 3 --cpu-seconds 1 --spp 1250 2>&1 | python tools/sched_line.py | tee -a $S; done
