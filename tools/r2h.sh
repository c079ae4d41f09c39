#!/bin/bash
# Round 2, seventh 1-GPU call: streaming key loads for the window ordering, pool sizes for short shares, the microbenchmarks
# behind the roofline denominators with an ncu capture of the L2 one.
TAG=${1:-r2h}; OUT=gpurun_out; mkdir -p $OUT
L=$PWD/raytracing-in-a-weekend_b200/csrc
AB=$OUT/${TAG}_ab.txt; : > $AB
run() { printf "%-46s " "$1" | tee -a $AB; shift; env "$@" 2>&1 | python tools/bench_line.py | tee -a $AB; }
B="timeout 300 python bench.py --steps 3 --warmup 3 --cpu-seconds 1"
for rep in 1 2; do
  run "C5 1250spp slot order (default)"       X=1 $B --spp 1250
  run "C5 1250spp window 256, streaming keys" RTW_B200_LIB=$L/librtw_win256cs.so $B --spp 1250
  run "C5 1250spp window 512, streaming keys" RTW_B200_LIB=$L/librtw_win512cs.so $B --spp 1250
  run "C5 1250spp pool 9Mi"                   RTW_WF_POOL=9437184 $B --spp 1250
  run "C5 1250spp pool 12Mi"                  RTW_WF_POOL=12582912 $B --spp 1250
  run "C4 625spp pool 1.76Mi (default)"       X=1 $B --config C4 --spp 625
  run "C4 625spp pool 3.5Mi"                  RTW_WF_POOL=3670016 $B --config C4 --spp 625
  run "C4 625spp pool 5Mi"                    RTW_WF_POOL=5242880 $B --config C4 --spp 625
  run "C4 625spp pool 7Mi"                    RTW_WF_POOL=7340032 $B --config C4 --spp 625
  run "C2 pool 3.2Mi (default)"               X=1 $B --config C2
  run "C2 pool 6Mi"                           RTW_WF_POOL=6291456 $B --config C2
done
timeout 300 python tools/microbench.py 5 > $OUT/${TAG}_microbench.json 2>&1; cat $OUT/${TAG}_microbench.json
timeout 600 ncu --set full --clock-control none -k regex:bench_read_kernel -s 1 -c 1 -o $OUT/${TAG}_l2_microbench python tools/microbench.py 1 > $OUT/${TAG}_ncu_microbench.log 2>&1; echo "ncu microbench rc=$?"
