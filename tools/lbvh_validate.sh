#!/bin/bash
# Round 2, 1-GPU call: the LBVH with treelet restructuring (6 leaves, 3 passes): its gpu tests, throughput against the SAH tree.
TAG=${1:-r3i}; OUT=gpurun_out; mkdir -p $OUT
for c in C5 C2 C4; do timeout 300 python tools/lbvh_probe.py $c 200 2>&1 | tee -a $OUT/${TAG}_lbvh_vs_sah.txt; done
timeout 900 python -m pytest tests/test_lbvh.py tests/test_zz_no_alloc.py -m gpu -q > $OUT/${TAG}_pytest_lbvh.log 2>&1; echo "pytest rc=$?"; tail -n 3 $OUT/${TAG}_pytest_lbvh.log
