"""Prints the device micro-benchmarks behind the roofline's FP32 and L2 denominators (rtw_microbench) next to the
driver-measured HBM peak of MEASURED_PEAKS.json, which anchors the self-measured figures: the same read kernel over a
2 GiB buffer (RTW_BENCH_HBM) must land on the driver's HBM number.   python tools/microbench.py [reps]"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT]
import rtw_b200  # noqa: E402
from rtw_b200 import _abi as A  # noqa: E402

be = rtw_b200.get_default_backend()
reps = int(sys.argv[1]) if len(sys.argv) > 1 else 5
out = {"fp32_tflops": be.microbench(A.RTW_BENCH_FP32, reps), "l2_gbs": be.microbench(A.RTW_BENCH_L2, reps),
       "hbm_gbs": be.microbench(A.RTW_BENCH_HBM, reps)}
p = os.path.join(ROOT, "MEASURED_PEAKS.json")
if os.path.exists(p):
    out["measured_peaks_hbm_gbs"] = json.load(open(p)).get("hbm_gbs")
    out["hbm_read_over_driver_copy"] = out["hbm_gbs"] / out["measured_peaks_hbm_gbs"]
out["note"] = ("fp32: 8 independent FFMA chains per thread on every SM; l2: 128-bit read-only loads over a 32 MiB buffer, block->slice "
               "mapping rotated every pass so no SM re-reads what its L1 holds; hbm: the same loads over 2 GiB (read-only: a read "
               "stream reaches more than the driver's copy figure, which moves read + write bytes)")
print(json.dumps(out))
