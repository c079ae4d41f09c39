#!/bin/bash
# Env sweep helper (GPU box): runs bench.py once per quoted "VAR=val VAR2=val" group and prints the key numbers.
#   tools/sweep.sh SPP "RTW_WF_POOL=1048576" "RTW_WF_POOL=4194304 RTW_WF_SAMPLE_MAJOR=1" ...
SPP=$1; shift
for e in "$@"; do
  printf "%-50s " "$e"
  env $e timeout 300 python bench.py --spp $SPP --steps 2 --warmup 3 --cpu-seconds 1 2>&1 | python tools/bench_line.py
done
