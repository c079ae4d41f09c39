#!/bin/bash
# Round 2, 1-GPU call: the two Hit flavours (one tail in wf_shade, split in the megakernel) against the build before
# either (librtw_head.so); ncu captures of wf_shade / wf_extend of this build.
TAG=${1:-r2w}; OUT=gpurun_out; mkdir -p $OUT
L=$PWD/raytracing-in-a-weekend_b200/csrc
timeout 900 python -m pytest tests -m gpu -q > $OUT/${TAG}_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -n 4 $OUT/${TAG}_pytest_gpu.log
AB=$OUT/${TAG}_ab_hit_flavours.txt; : > $AB
run() { printf "%-34s " "$1" | tee -a $AB; shift; env "$@" 2>&1 | python tools/bench_line.py | tee -a $AB; }
B="timeout 300 python bench.py --steps 3 --warmup 3 --cpu-seconds 1"
for rep in 1 2; do
  for lib in head b200; do
    run "C5 625spp $lib"  RTW_B200_LIB=$L/librtw_$lib.so $B --spp 625
    run "C4 625spp $lib"  RTW_B200_LIB=$L/librtw_$lib.so $B --config C4 --spp 625
    run "C3 $lib"         RTW_B200_LIB=$L/librtw_$lib.so $B --config C3
    run "C1 $lib"         RTW_B200_LIB=$L/librtw_$lib.so $B --config C1
  done
done
timeout 600 ncu --set full --clock-control none --import-source on -k regex:wf_shade_kernel -s 150 -c 1 -o $OUT/${TAG}_wf_shade \
    python bench.py --spp 400 --steps 2 --warmup 3 --cpu-seconds 1 > $OUT/${TAG}_ncu_full_shade.log 2>&1; echo "ncu full shade rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:wf_extend_kernel -s 150 -c 1 -o $OUT/${TAG}_wf_extend \
    python bench.py --spp 400 --steps 2 --warmup 3 --cpu-seconds 1 > $OUT/${TAG}_ncu_full.log 2>&1; echo "ncu full rc=$?"
