#!/bin/bash
# Round 2, first GPU call: smoke, the whole `-m gpu` suite (recording the per-scene L1 figures), then A/Bs of this
# round's changes on C5 and C4. Everything into gpurun_out/ under the tag.  gpurun --timeout 2400 -- 'bash tools/r2a.sh r2a'
TAG=${1:-r2a}; OUT=gpurun_out; mkdir -p $OUT
nvidia-smi --query-gpu=name,clocks.max.sm,memory.total --format=csv,noheader > $OUT/${TAG}_gpu.txt; nproc >> $OUT/${TAG}_gpu.txt
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $OUT/${TAG}_smoke.log 2>&1; echo "smoke rc=$?"; tail -2 $OUT/${TAG}_smoke.log
rm -f $OUT/${TAG}_l1.jsonl
RTW_L1_RECORD=$PWD/$OUT/${TAG}_l1.jsonl timeout 1500 python -m pytest tests -m gpu -q > $OUT/${TAG}_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -15 $OUT/${TAG}_pytest_gpu.log
AB=$OUT/${TAG}_ab.txt; : > $AB
run() { printf "%-44s " "$1" | tee -a $AB; shift; env "$@" 2>&1 | python tools/bench_line.py | tee -a $AB; }
B="timeout 300 python bench.py --steps 2 --warmup 3 --cpu-seconds 1"
for rep in 1 2; do
  run "C5 1000spp base (baked, fastvote)"      X=1 $B --spp 1000
  run "C5 1000spp RTW_NO_BAKE=1"               RTW_NO_BAKE=1 $B --spp 1000
  run "C5 1000spp RTW_BVH_COLLAPSE=dp"         RTW_BVH_COLLAPSE=dp $B --spp 1000
  run "C5 1000spp no fastvote"                 RTW_B200_LIB=$PWD/raytracing-in-a-weekend_b200/csrc/librtw_novote.so $B --spp 1000
  run "C5 1000spp NO_BAKE + dp"                RTW_NO_BAKE=1 RTW_BVH_COLLAPSE=dp $B --spp 1000
done
for rep in 1 2; do
  run "C4 500spp base (baked)"                 X=1 $B --config C4 --spp 500
  run "C4 500spp RTW_NO_BAKE=1"                RTW_NO_BAKE=1 $B --config C4 --spp 500
  run "C4 500spp dp"                           RTW_BVH_COLLAPSE=dp $B --config C4 --spp 500
  run "C2 500spp base"                         X=1 $B --config C2
  run "C2 500spp dp"                           RTW_BVH_COLLAPSE=dp $B --config C2
done
RTW_BUILD_TIMING=1 RTW_BENCH_E2E_TRACE=1 timeout 300 python bench.py --spp 200 --steps 2 --warmup 3 --cpu-seconds 1 > $OUT/${TAG}_bench_trace.json 2> $OUT/${TAG}_build_timing.txt
tail -40 $OUT/${TAG}_build_timing.txt
timeout 600 python bench.py --steps 3 --warmup 3 > $OUT/${TAG}_bench_C5_n1.json 2> $OUT/${TAG}_bench_C5_n1.err; echo "bench rc=$?"; python tools/bench_line.py < $OUT/${TAG}_bench_C5_n1.json
