#!/bin/bash
# Round 2, multi-GPU call (gpurun --gpus 8): the three partitions north_star names, under torchrun (one rank per GPU, NCCL) and
# behind one process (rtw_render_multi), plus the multi-device tests over real devices.  bash tools/r2e_multi.sh r2e 8
TAG=${1:-r2e}; N=${2:-8}; OUT=gpurun_out; mkdir -p $OUT
nvidia-smi --query-gpu=index,name --format=csv,noheader > $OUT/${TAG}_gpus.txt; nvidia-smi topo -m >> $OUT/${TAG}_gpus.txt 2>&1
timeout 600 python -m pytest tests/test_multi_device.py -m gpu -q > $OUT/${TAG}_pytest_multi_device.log 2>&1; echo "pytest multi rc=$?"; tail -3 $OUT/${TAG}_pytest_multi_device.log
T() { local n=$1; shift; if [ "$n" = 1 ]; then timeout 600 python bench.py --gpus 1 "$@"; else timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29500 + n)) bench.py --gpus $n "$@"; fi; }
line() { python tools/bench_line.py; }
# --- A: sample ranges + NCCL reduce (C5 headline), torchrun
for n in 1 $N; do
  T $n --steps 3 --warmup 3 --cpu-seconds 2 > $OUT/${TAG}_C5_samples_n$n.json 2> $OUT/${TAG}_C5_samples_n$n.err; printf "C5 samples torchrun n=$n: "; line < $OUT/${TAG}_C5_samples_n$n.json
done
# --- B: tiles (C4 as BASELINE.json names it), torchrun, N = 1 and N
for n in 1 $N; do
  T $n --config C4 --partition tiles --band 4 --steps 3 --warmup 3 --cpu-seconds 2 > $OUT/${TAG}_C4_tiles_n$n.json 2> $OUT/${TAG}_C4_tiles_n$n.err; printf "C4 tiles torchrun n=$n: "; line < $OUT/${TAG}_C4_tiles_n$n.json
done
T $N --config C4 --steps 3 --warmup 3 --cpu-seconds 2 > $OUT/${TAG}_C4_samples_n$N.json 2>/dev/null; printf "C4 samples torchrun n=$N: "; line < $OUT/${TAG}_C4_samples_n$N.json
# --- C: frames, one animation frame per GPU per step (C5, frame = k)
T $N --partition frames --steps 2 --warmup 3 --cpu-seconds 2 > $OUT/${TAG}_C5_frames_n$N.json 2> $OUT/${TAG}_C5_frames_n$N.err; printf "C5 frames torchrun n=$N: "; line < $OUT/${TAG}_C5_frames_n$N.json
# --- one process, rtw_render_multi
timeout 600 python bench.py --single-process --gpus $N --steps 3 --warmup 3 > $OUT/${TAG}_C5_samples_single_process_n$N.json 2> $OUT/${TAG}_sp.err; printf "C5 samples single process n=$N: "; line < $OUT/${TAG}_C5_samples_single_process_n$N.json
timeout 600 python bench.py --single-process --gpus $N --config C4 --partition tiles --band 4 --steps 3 --warmup 3 > $OUT/${TAG}_C4_tiles_single_process_n$N.json 2>> $OUT/${TAG}_sp.err; printf "C4 tiles single process n=$N: "; line < $OUT/${TAG}_C4_tiles_single_process_n$N.json
ls $OUT | grep $TAG
