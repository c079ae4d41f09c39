#!/bin/bash
# Round 2, 1-GPU call: re-record the per-scene first-hit figures of the fast build (tests/golden/l1_measured.json), verify, whole suite.
TAG=${1:-r3c}; OUT=gpurun_out; mkdir -p $OUT
rm -f $OUT/${TAG}_l1.jsonl
RTW_L1_RECORD=$PWD/$OUT/${TAG}_l1.jsonl timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -q -k first_hit > $OUT/${TAG}_pytest_l1_record.log 2>&1; echo "record rc=$?"; tail -n 2 $OUT/${TAG}_pytest_l1_record.log
python tools/l1_collect.py $OUT/${TAG}_l1.jsonl tests/golden/l1_measured.json && cp tests/golden/l1_measured.json $OUT/${TAG}_l1_measured.json
timeout 900 python -m pytest tests -m gpu -q > $OUT/${TAG}_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -n 4 $OUT/${TAG}_pytest_gpu.log
