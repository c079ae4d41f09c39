#!/bin/bash
# Round 2, 1-GPU call: host batch size and supply chunk at one GPU's share of eight (C5, 1 250 spp), env knobs only.
TAG=${1:-r3l}; OUT=gpurun_out; mkdir -p $OUT
AB=$OUT/${TAG}_ab_batch_supply.txt; : > $AB
run() { printf "%-34s " "$1" | tee -a $AB; shift; env "$@" 2>&1 | python tools/bench_line.py | tee -a $AB; }
B="timeout 300 python bench.py --steps 3 --warmup 3 --cpu-seconds 1 --spp 1250"
for rep in 1 2; do
  run "default"              $B
  run "RTW_WF_BATCH=4"       RTW_WF_BATCH=4 $B
  run "RTW_WF_BATCH=16"      RTW_WF_BATCH=16 $B
  run "SUPPLY_CHUNK=128"     RTW_WF_SUPPLY_CHUNK=128 $B
  run "SUPPLY_CHUNK=512"     RTW_WF_SUPPLY_CHUNK=512 $B
done
