#!/bin/bash
# Round 2, fourth GPU call (1 GPU): the whole gpu suite with the light-biased integrators and interleaved-row tiles, then the
# tile partition through one device twice (devices 0,0) and the headline line.
TAG=${1:-r2d}; OUT=gpurun_out; mkdir -p $OUT
timeout 1500 python -m pytest tests -m gpu -q > $OUT/${TAG}_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -8 $OUT/${TAG}_pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $OUT/${TAG}_smoke.log 2>&1; echo "smoke rc=$?"; tail -2 $OUT/${TAG}_smoke.log
AB=$OUT/${TAG}_ab.txt; : > $AB
run() { printf "%-52s " "$1" | tee -a $AB; shift; env "$@" 2>&1 | python tools/bench_line.py | tee -a $AB; }
B="timeout 300 python bench.py --steps 3 --warmup 3 --cpu-seconds 1"
run "C4 500spp samples (1 GPU)"                         X=1 $B --config C4 --spp 500
run "C4 500spp partition tiles (1 GPU, band 8)"         X=1 $B --config C4 --spp 500 --partition tiles
run "C4 500spp single-process tiles devices 0,0"        X=1 $B --config C4 --spp 500 --single-process --devices 0,0 --partition tiles
run "C4 500spp single-process tiles devices 0,0,0,0 band 4" X=1 $B --config C4 --spp 500 --single-process --devices 0,0,0,0 --partition tiles --band 4
run "C4 500spp single-process samples devices 0,0"      X=1 $B --config C4 --spp 500 --single-process --devices 0,0
python - <<'PY' 2>&1 | tee $OUT/${TAG}_light_biased_timing.txt
import sys, time; sys.path[:0] = ['.', 'tests']
import numpy as np, rtw_b200
from rtw_b200 import scenes
be = rtw_b200.get_default_backend()
for cast in (True, False):
    job = scenes.light_biased_test(cast)            # Rust2's own settings: 400x300, 100 spp, depth 9, weight 100
    img = job.render(be); st = be.scene(job.scene).stats()
    img = job.render(be); st = be.scene(job.scene).stats()
    print(f"{job.name}: {st.paths} paths {st.rays} rays  kernel {st.kernel_ms:.2f} ms  {st.paths / st.kernel_ms / 1e3:.1f} Mpaths/s {st.rays / st.kernel_ms / 1e3:.1f} Mrays/s  mean {np.nanmean(img):.4f}")
PY
