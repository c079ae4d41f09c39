#!/bin/bash
# Round 2, 1-GPU call: the default bench line of the final build (CPU baseline timed first); BVH node loads with the
# L1 evict_last priority (librtw_keep.so) against the build.
TAG=${1:-r3d}; OUT=gpurun_out; mkdir -p $OUT
L=$PWD/raytracing-in-a-weekend_b200/csrc
timeout 600 python bench.py > $OUT/${TAG}_bench_default.json 2> $OUT/${TAG}_bench_default.err; echo "bench rc=$?"; python tools/bench_line.py < $OUT/${TAG}_bench_default.json
python - <<'PY'
import json
d=[json.loads(l) for l in open('gpurun_out/r3d_bench_default.json') if l.startswith('{')][-1]
print('cpu_baseline', d['cpu_baseline'], 'steps', d['steps'], 'warmup', d['warmup'])
PY
AB=$OUT/${TAG}_ab_node_evict_last.txt; : > $AB
run() { printf "%-34s " "$1" | tee -a $AB; shift; env "$@" 2>&1 | python tools/bench_line.py | tee -a $AB; }
B="timeout 300 python bench.py --steps 3 --warmup 3 --cpu-seconds 1"
for rep in 1 2; do
  for lib in b200 keep; do
    run "C5 625spp $lib"   RTW_B200_LIB=$L/librtw_$lib.so $B --spp 625
    run "C5 2500spp $lib"  RTW_B200_LIB=$L/librtw_$lib.so $B --spp 2500
    run "C4 625spp $lib"   RTW_B200_LIB=$L/librtw_$lib.so $B --config C4 --spp 625
    run "C2 $lib"          RTW_B200_LIB=$L/librtw_$lib.so $B --config C2
  done
done
