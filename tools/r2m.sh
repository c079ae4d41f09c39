#!/bin/bash
# Round 2, 1-GPU call: stack pointer as a shared-memory address, on top of uint32 flags and 7 (8) resident extend blocks.
TAG=${1:-r2m}; OUT=gpurun_out; mkdir -p $OUT
L=$PWD/raytracing-in-a-weekend_b200/csrc
RTW_B200_LIB=$L/librtw_u32stack7.so timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_lbvh.py tests/test_triangle.py -m gpu -q -x > $OUT/${TAG}_pytest_u32stack7.log 2>&1; echo "pytest(u32stack7) rc=$?"; tail -n 2 $OUT/${TAG}_pytest_u32stack7.log
AB=$OUT/${TAG}_ab.txt; : > $AB
run() { printf "%-52s " "$1" | tee -a $AB; shift; env "$@" 2>&1 | python tools/bench_line.py | tee -a $AB; }
B="timeout 300 python bench.py --steps 3 --warmup 3 --cpu-seconds 1"
for rep in 1 2; do
  run "C5 1250spp uint32 flags, 7 blocks"               RTW_B200_LIB=$L/librtw_u32minb7.so $B --spp 1250
  run "C5 1250spp + stack address (72 regs, 6 blocks)"  RTW_B200_LIB=$L/librtw_u32stack.so $B --spp 1250
  run "C5 1250spp + stack address, 7 blocks"            RTW_B200_LIB=$L/librtw_u32stack7.so $B --spp 1250
  run "C5 1250spp + stack address, 8 blocks"            RTW_B200_LIB=$L/librtw_u32stack8.so $B --spp 1250
  run "C2 uint32 flags, 7 blocks"                       RTW_B200_LIB=$L/librtw_u32minb7.so $B --config C2
  run "C2 + stack address, 7 blocks"                    RTW_B200_LIB=$L/librtw_u32stack7.so $B --config C2
  run "C2 + stack address, 8 blocks"                    RTW_B200_LIB=$L/librtw_u32stack8.so $B --config C2
  run "C4 625spp uint32 flags, 7 blocks"                RTW_B200_LIB=$L/librtw_u32minb7.so $B --config C4 --spp 625
  run "C4 625spp + stack address, 7 blocks"             RTW_B200_LIB=$L/librtw_u32stack7.so $B --config C4 --spp 625
  run "C4 625spp + stack address, 8 blocks"             RTW_B200_LIB=$L/librtw_u32stack8.so $B --config C4 --spp 625
done
