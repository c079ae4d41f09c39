"""Times the bilateral filter (Rust2/src/postprocessing.rs:70-135) on the device next to the CPU oracle.
Usage: python tests/bench_post.py [--width 800 --height 600 --size 10 --reps 5]
The CPU leg runs the oracle restatement (one thread, like the reference's from_fn loop) on the same image."""
import argparse, json, os, sys, time
import numpy as np
sys.path.insert(0, os.path.join(os.path.dirname(__file__), ".."))
sys.path.insert(0, os.path.dirname(__file__))
import rtw_b200
from rtw_b200.backend import CudaBackend


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--width", type=int, default=800); ap.add_argument("--height", type=int, default=600)
    ap.add_argument("--size", type=int, default=10); ap.add_argument("--kind", type=int, default=0)
    ap.add_argument("--reps", type=int, default=5); ap.add_argument("--no-cpu", action="store_true")
    a = ap.parse_args()
    rng = np.random.default_rng(1)
    yy, xx = np.mgrid[0:a.height, 0:a.width]
    base = np.stack([(xx * 255 // a.width), (yy * 255 // a.height), ((xx // 64 + yy // 64) % 2) * 180 + 30], -1).astype(np.float64)
    img = np.clip(base + rng.normal(0, 10, base.shape), 0, 255).astype(np.uint8)
    be = CudaBackend(0)
    info = {}
    be.bilateral_filter(img, a.size, a.kind, info)                    # warm-up
    ms, wall = [], []
    for _ in range(a.reps):
        t0 = time.perf_counter(); out = be.bilateral_filter(img, a.size, a.kind, info); wall.append((time.perf_counter() - t0) * 1e3)
        ms.append(info["kernel_ms"])
    taps = 0
    s = a.size
    for x in (range(a.width)):
        nx = min(a.width - x - 1, s) + min(x, s)
        taps += nx
    # separable count of the square window; the diamond is reported as the square's bound
    ty = sum(min(a.height - y - 1, s) + min(y, s) for y in range(a.height))
    n_taps = taps * ty * 3
    line = {"op": "bilateral_filter", "width": a.width, "height": a.height, "size": a.size, "kind": a.kind,
            "kernel_ms": float(np.median(ms)), "host_call_ms": float(np.median(wall)), "taps": int(n_taps),
            "gtaps_per_s": n_taps / (np.median(ms) * 1e-3) / 1e9, "params": {k: float(v) for k, v in info.items() if k != "kernel_ms"}}
    if not a.no_cpu:
        from oracle_backend import OracleBackend
        ob = OracleBackend(fast=True)
        t0 = time.perf_counter(); want = ob.bilateral_filter(img, a.size, a.kind); cpu = time.perf_counter() - t0
        line["cpu_oracle_ms"] = cpu * 1e3
        line["identical"] = bool(np.array_equal(out, want))
    print(json.dumps(line))


if __name__ == "__main__":
    main()
