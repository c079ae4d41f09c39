#!/bin/bash
# Round 2, 1-GPU call: ncu --set full captures of wf_extend / wf_shade of the final build (streaming cache operators on the records).
TAG=${1:-r3m}; OUT=gpurun_out; mkdir -p $OUT
timeout 200 ncu --set full --clock-control none --import-source on -k regex:wf_shade_kernel -s 150 -c 1 -o $OUT/${TAG}_wf_shade \
    python bench.py --spp 400 --steps 1 --warmup 1 --cpu-seconds 1 > $OUT/${TAG}_ncu_full_shade.log 2>&1; echo "ncu full shade rc=$?"
timeout 200 ncu --set full --clock-control none --import-source on -k regex:wf_extend_kernel -s 150 -c 1 -o $OUT/${TAG}_wf_extend \
    python bench.py --spp 400 --steps 1 --warmup 1 --cpu-seconds 1 > $OUT/${TAG}_ncu_full.log 2>&1; echo "ncu full rc=$?"
