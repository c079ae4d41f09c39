#!/bin/bash
# Round 2, 1-GPU call: the LBVH on a 30 000-leaf level; resident blocks of wf_shade after its one-tail rewrite
# (8 / 9 / 10 / 12 blocks = 64 / 56 / 48 / 40 registers).
TAG=${1:-r2z}; OUT=gpurun_out; mkdir -p $OUT
L=$PWD/raytracing-in-a-weekend_b200/csrc
timeout 600 python -m pytest tests/test_lbvh.py -m gpu -q -k "wider" > $OUT/${TAG}_pytest_lbvh_wide.log 2>&1; echo "pytest rc=$?"; tail -n 15 $OUT/${TAG}_pytest_lbvh_wide.log
AB=$OUT/${TAG}_ab_shade_resident_blocks.txt; : > $AB
run() { printf "%-34s " "$1" | tee -a $AB; shift; env "$@" 2>&1 | python tools/bench_line.py | tee -a $AB; }
B="timeout 300 python bench.py --steps 3 --warmup 3 --cpu-seconds 1"
for rep in 1 2; do
  for lib in sh8 sh9 b200 sh12; do
    run "C5 625spp $lib"  RTW_B200_LIB=$L/librtw_$lib.so $B --spp 625
    run "C4 625spp $lib"  RTW_B200_LIB=$L/librtw_$lib.so $B --config C4 --spp 625
    run "C2 $lib"         RTW_B200_LIB=$L/librtw_$lib.so $B --config C2
  done
done
