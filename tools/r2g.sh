#!/bin/bash
# Round 2, sixth 1-GPU call: the warp-window ray ordering (camera rays first, then bounced rays by direction octant) A/B.
TAG=${1:-r2g}; OUT=gpurun_out; mkdir -p $OUT
L=$PWD/raytracing-in-a-weekend_b200/csrc
RTW_B200_LIB=$L/librtw_win256.so timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -q -x -k "path_parity or pipelines_agree or statistically" > $OUT/${TAG}_pytest_win256.log 2>&1; echo "pytest(win256) rc=$?"; tail -3 $OUT/${TAG}_pytest_win256.log
AB=$OUT/${TAG}_ab.txt; : > $AB
run() { printf "%-46s " "$1" | tee -a $AB; shift; env "$@" 2>&1 | python tools/bench_line.py | tee -a $AB; }
B="timeout 300 python bench.py --steps 3 --warmup 3 --cpu-seconds 1"
for rep in 1 2; do
  run "C5 1250spp slot order (default)"   X=1 $B --spp 1250
  run "C5 1250spp window 128"             RTW_B200_LIB=$L/librtw_win128.so $B --spp 1250
  run "C5 1250spp window 256"             RTW_B200_LIB=$L/librtw_win256.so $B --spp 1250
  run "C5 1250spp window 512"             RTW_B200_LIB=$L/librtw_win512.so $B --spp 1250
  run "C2 slot order (default)"           X=1 $B --config C2
  run "C2 window 256"                     RTW_B200_LIB=$L/librtw_win256.so $B --config C2
  run "C4 625spp slot order (default)"    X=1 $B --config C4 --spp 625
  run "C4 625spp window 256"              RTW_B200_LIB=$L/librtw_win256.so $B --config C4 --spp 625
  run "C4 625spp pool 3.5Mi"              RTW_WF_POOL=3670016 $B --config C4 --spp 625
  run "C4 625spp pool 1Mi"                RTW_WF_POOL=1048576 $B --config C4 --spp 625
done
S=$OUT/${TAG}_sched.txt; : > $S
printf "slot order  " | tee -a $S; timeout 300 python bench.py --steps 2 --warmup 3 --cpu-seconds 1 --spp 1250 2>&1 | python tools/sched_line.py | tee -a $S
printf "window 256  " | tee -a $S; RTW_B200_LIB=$L/librtw_win256.so timeout 300 python bench.py --steps 2 --warmup 3 --cpu-seconds 1 --spp 1250 2>&1 | python tools/sched_line.py | tee -a $S
timeout 600 python bench.py --steps 3 --warmup 3 --cpu-seconds 2 > $OUT/${TAG}_bench_C5_n1.json 2>/dev/null; printf "headline (pool 12Mi): "; python tools/bench_line.py < $OUT/${TAG}_bench_C5_n1.json
RTW_BUILD_TIMING=1 RTW_BENCH_E2E_TRACE=1 timeout 300 python bench.py --spp 200 --steps 2 --warmup 3 --cpu-seconds 1 > /dev/null 2> $OUT/${TAG}_build_timing.txt; tail -12 $OUT/${TAG}_build_timing.txt
