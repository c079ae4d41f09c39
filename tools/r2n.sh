#!/bin/bash
# Round 2, validation call of the final build: the whole `-m gpu` suite (recording the per-scene L1 figures, then asserting
# against them), smoke, the headline bench and the per-config lines, ncu launch list + full captures.
#   gpurun --timeout 2700 -- 'bash tools/r2n.sh r2n'
TAG=${1:-r2n}; OUT=gpurun_out; mkdir -p $OUT
nvidia-smi --query-gpu=name,clocks.max.sm,memory.total --format=csv,noheader > $OUT/${TAG}_gpu.txt; nproc >> $OUT/${TAG}_gpu.txt
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $OUT/${TAG}_smoke.log 2>&1; echo "smoke rc=$?"; tail -2 $OUT/${TAG}_smoke.log
rm -f $OUT/${TAG}_l1.jsonl
RTW_L1_RECORD=$PWD/$OUT/${TAG}_l1.jsonl timeout 1500 python -m pytest tests -m gpu -q > $OUT/${TAG}_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -12 $OUT/${TAG}_pytest_gpu.log
python tools/l1_collect.py $OUT/${TAG}_l1.jsonl tests/golden/l1_measured.json && cp tests/golden/l1_measured.json $OUT/${TAG}_l1_measured.json
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -q -k first_hit > $OUT/${TAG}_pytest_l1_verify.log 2>&1; echo "L1 verify rc=$?"; tail -3 $OUT/${TAG}_pytest_l1_verify.log
RTW_BUILD_TIMING=1 RTW_BENCH_E2E_TRACE=1 timeout 300 python bench.py --spp 200 --steps 2 --warmup 3 --cpu-seconds 1 > $OUT/${TAG}_bench_trace.json 2> $OUT/${TAG}_build_timing.txt
tail -24 $OUT/${TAG}_build_timing.txt
timeout 600 python bench.py --steps 3 --warmup 3 > $OUT/${TAG}_bench_C5_n1.json 2> $OUT/${TAG}_bench_C5_n1.err; echo "bench rc=$?"; python tools/bench_line.py < $OUT/${TAG}_bench_C5_n1.json
for c in C1 C2 C3 C4; do timeout 300 python bench.py --config $c --steps 3 --warmup 3 --cpu-seconds 3 > $OUT/${TAG}_bench_$c.json 2>/dev/null; printf "$c "; python tools/bench_line.py < $OUT/${TAG}_bench_$c.json; done
timeout 300 python bench.py --spp 400 --steps 2 --warmup 3 --cpu-seconds 1 > $OUT/${TAG}_bench_spp400.json 2>/dev/null &&
  timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 1400 --csv --log-file $OUT/${TAG}_launches_bench_spp400.csv \
    python bench.py --spp 400 --steps 1 --warmup 1 --cpu-seconds 1 > $OUT/${TAG}_ncu_launches.log 2>&1; echo "ncu launches rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:wf_extend_kernel -s 150 -c 1 -o $OUT/${TAG}_wf_extend \
    python bench.py --spp 400 --steps 2 --warmup 3 --cpu-seconds 1 > $OUT/${TAG}_ncu_full.log 2>&1; echo "ncu full rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:wf_shade_kernel -s 150 -c 1 -o $OUT/${TAG}_wf_shade \
    python bench.py --spp 400 --steps 2 --warmup 3 --cpu-seconds 1 > $OUT/${TAG}_ncu_full_shade.log 2>&1; echo "ncu full shade rc=$?"
ls -la $OUT | tail -30
