#!/bin/bash
# Round 2, fifth 1-GPU call: tuning sweep of the baked build (vote weights, fetch threshold, pool, resident blocks, node
# bursts, shade occupancy) and the coherent-ray ceiling (depth 1: camera rays only) from the instrumented build's counters.
TAG=${1:-r2f}; OUT=gpurun_out; mkdir -p $OUT
AB=$OUT/${TAG}_sweep.txt; : > $AB
run() { printf "%-46s " "$1" | tee -a $AB; shift; env "$@" 2>&1 | python tools/bench_line.py | tee -a $AB; }
L=$PWD/raytracing-in-a-weekend_b200/csrc
B="timeout 300 python bench.py --steps 3 --warmup 3 --cpu-seconds 1 --spp 1250"
for rep in 1 2; do
  run "base"                          X=1 $B
  run "RTW_WF_PRIM_W=4"               RTW_WF_PRIM_W=4 $B
  run "RTW_WF_PRIM_W=6"               RTW_WF_PRIM_W=6 $B
  run "RTW_WF_PRIM_W=12"              RTW_WF_PRIM_W=12 $B
  run "RTW_WF_PRIM_W=16"              RTW_WF_PRIM_W=16 $B
  run "RTW_WF_EVENT_W=4"              RTW_WF_EVENT_W=4 $B
  run "RTW_WF_EVENT_W=16"             RTW_WF_EVENT_W=16 $B
  run "RTW_WF_FETCH_MIN=8"            RTW_WF_FETCH_MIN=8 $B
  run "RTW_WF_FETCH_MIN=12"           RTW_WF_FETCH_MIN=12 $B
  run "RTW_WF_FETCH_MIN=20"           RTW_WF_FETCH_MIN=20 $B
  run "RTW_WF_FETCH_MIN=24"           RTW_WF_FETCH_MIN=24 $B
  run "RTW_WF_CHUNK=128"              RTW_WF_CHUNK=128 $B
  run "extend 7 resident blocks (72 regs)" RTW_B200_LIB=$L/librtw_minb7.so $B
  run "node burst 2"                  RTW_B200_LIB=$L/librtw_burst2.so $B
  run "node burst 5"                  RTW_B200_LIB=$L/librtw_burst5.so $B
  run "shade 12 resident blocks (40 regs)" RTW_B200_LIB=$L/librtw_shade12.so $B
  run "shade 8 resident blocks (64 regs)"  RTW_B200_LIB=$L/librtw_shade8.so $B
done
B2="timeout 300 python bench.py --steps 2 --warmup 3 --cpu-seconds 1"
run "10000spp pool 6Mi (default)"     X=1 $B2
run "10000spp pool 8Mi"               RTW_WF_POOL=8388608 $B2
run "10000spp pool 12Mi"              RTW_WF_POOL=12582912 $B2
run "10000spp pool 4Mi"               RTW_WF_POOL=4194304 $B2
# coherent-ray ceiling: what the scheduler achieves on camera rays alone (depth 1) against the full path mix (depth 50)
S=$OUT/${TAG}_sched_depth.txt; : > $S
for d in 1 2 50; do printf "depth %-3s " $d | tee -a $S; timeout 300 python bench.py --steps 2 --warmup 3 --cpu-seconds 1 --spp 1250 --depth $d 2>&1 | python tools/sched_line.py | tee -a $S; done
