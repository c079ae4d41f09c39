#!/bin/bash
# Round 2, 1-GPU call: the LBVH with the dynamic-programming collapse: gpu tests of it, throughput against the SAH tree, whole suite.
TAG=${1:-r3g}; OUT=gpurun_out; mkdir -p $OUT
for c in C5 C2 C4; do timeout 300 python tools/lbvh_probe.py $c 200 2>&1 | tee -a $OUT/${TAG}_lbvh_vs_sah.txt; done
timeout 900 python -m pytest tests -m gpu -q > $OUT/${TAG}_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -n 4 $OUT/${TAG}_pytest_gpu.log
