#!/bin/bash
# Round 2, 1-GPU call: uint32_t instead of bool lane flags in wf_extend (the compiler packs bools into bytes and PRMTs them apart).
TAG=${1:-r2l}; OUT=gpurun_out; mkdir -p $OUT
L=$PWD/raytracing-in-a-weekend_b200/csrc
AB=$OUT/${TAG}_ab.txt; : > $AB
run() { printf "%-46s " "$1" | tee -a $AB; shift; env "$@" 2>&1 | python tools/bench_line.py | tee -a $AB; }
B="timeout 300 python bench.py --steps 3 --warmup 3 --cpu-seconds 1"
for rep in 1 2; do
  run "C5 1250spp bool flags (default)"      X=1 $B --spp 1250
  run "C5 1250spp uint32 flags"              RTW_B200_LIB=$L/librtw_u32flags.so $B --spp 1250
  run "C5 1250spp uint32 flags, 7 blocks"    RTW_B200_LIB=$L/librtw_u32minb7.so $B --spp 1250
  run "C2 bool flags (default)"              X=1 $B --config C2
  run "C2 uint32 flags"                      RTW_B200_LIB=$L/librtw_u32flags.so $B --config C2
  run "C2 uint32 flags, 7 blocks"            RTW_B200_LIB=$L/librtw_u32minb7.so $B --config C2
  run "C4 625spp bool flags (default)"       X=1 $B --config C4 --spp 625
  run "C4 625spp uint32 flags"               RTW_B200_LIB=$L/librtw_u32flags.so $B --config C4 --spp 625
  run "C4 625spp uint32 flags, 7 blocks"     RTW_B200_LIB=$L/librtw_u32minb7.so $B --config C4 --spp 625
done
