#!/bin/bash
# Round 2, 1-GPU call: the pipeline rule by scene (flat scenes -> megakernel) and the rotated / largest-child-first LBVH.
#   gpurun --timeout 1500 -- 'bash tools/r2r.sh r2r'
TAG=${1:-r2r}; OUT=gpurun_out; mkdir -p $OUT
nvidia-smi --query-gpu=name,clocks.max.sm,memory.total --format=csv,noheader > $OUT/${TAG}_gpu.txt; nproc >> $OUT/${TAG}_gpu.txt
timeout 900 python -m pytest tests -m gpu -q > $OUT/${TAG}_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -n 12 $OUT/${TAG}_pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $OUT/${TAG}_smoke.log 2>&1; echo "smoke rc=$?"; tail -n 2 $OUT/${TAG}_smoke.log
timeout 600 python tools/crossover.py C1,C2,C3,C4,C5 1,4,16,64,256 3 2>&1 | tee $OUT/${TAG}_crossover.txt
for c in C5 C2 C4; do timeout 300 python tools/lbvh_probe.py $c 200 2>&1 | tee -a $OUT/${TAG}_lbvh_vs_sah.txt; done
for c in C1 C3; do timeout 300 python bench.py --config $c --steps 5 --warmup 3 --cpu-seconds 3 > $OUT/${TAG}_bench_$c.json 2>/dev/null; printf "$c "; python tools/bench_line.py < $OUT/${TAG}_bench_$c.json; done
timeout 600 python bench.py --steps 3 --warmup 3 > $OUT/${TAG}_bench_C5_n1.json 2> $OUT/${TAG}_bench_C5_n1.err; echo "bench rc=$?"; python tools/bench_line.py < $OUT/${TAG}_bench_C5_n1.json
timeout 600 ncu --set full --clock-control none --import-source on -k regex:mega_kernel -s 3 -c 1 -o $OUT/${TAG}_mega_C3 \
    python bench.py --config C3 --spp 200 --steps 2 --warmup 3 --cpu-seconds 1 > $OUT/${TAG}_ncu_mega.log 2>&1; echo "ncu mega rc=$?"
ls -la $OUT | tail -n 15
