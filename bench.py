#!/usr/bin/env python
"""bench.py — the headline measurement: Mpaths/s (+ Mrays/s) on the Next-Week final scene (C5),
800x800, 10000 spp, depth 50, on N B200s of one node, next to the host-CPU renderer.

    python bench.py --gpus N --steps K --warmup W            (N>1: launched by torchrun, one rank per GPU)
    python bench.py --impl reference --gpus N --steps K --warmup W   (CPU arm: the oracle port, all host threads)

A "step" is one whole frame of the workload: every rank renders its sample range of every pixel
into a per-GPU f32 sum buffer (wavefront pipeline: wf_extend + wf_shade kernels over a pool of path
slots in HBM), one NCCL reduce(sum) to rank 0, resolve (/spp, powf(1/gamma)) there. Total work per step
is fixed as N grows => "scaling": "strong". --partition tiles (row bands dealt round-robin, gathered,
no reduction; C4 in BASELINE.json) and --partition frames (one animation frame per GPU per step, no
exchange; "scaling": "weak") are the two other partitions north_star names. --single-process runs the
sample / tile split over --gpus N devices inside ONE process through rtw_render_multi (no torchrun).
  value  = paths of K frames / device time (CUDA events on the launching stream, max over ranks),
           scene resident in HBM, image left on the device.
  e2e    = the same frame through the C ABI with HOST buffers: rtw_scene_create (flatten, SAH BVH,
           H2D of the scene arrays and textures) + render + D2H of the image + destroy, every step.
One JSON line on stdout (rank 0).
"""
from __future__ import annotations

import argparse
import ctypes as C
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
for _p in (ROOT, os.path.join(ROOT, "tests")):
    if _p not in sys.path:
        sys.path.insert(0, _p)

import numpy as np  # noqa: E402

METRIC = "Mpaths/s, Next-Week final scene 800x800 (C5), 10000 spp, depth 50"
L2_NOTE = "GPU arm: 256 MiB memset between steps, inside the timed region (L2 is 126 MB); CPU arm: not applicable"


def config_of(job) -> dict:
    """The `config` object: the same on both arms (what differs per arm is under "run")."""
    return {"workload": workload_name(job), "l2": L2_NOTE}

UNIT = "Mpaths/s"

# Algorithmic work per operation (DESIGN.md "Roofline"): flop and bytes of scene data.
COST = {"node": (112.0, 64.0),       # four child-box slab tests (28 flop each, objects.rs:125-140), one 64-B half-float node
        "sphere": (36.0, 16.0),      # sphere.rs:100-123, (origin, radius)
        "quad": (49.0, 64.0),        # quad.rs:38-60, 64-B record
        "instance": (60.0, 160.0),   # ray -> local and hit -> world (two 3x3 Euler maps + T)
        "hit": (190.0, 64.0)}        # Hit evaluation + uber-material scatter (+70 on accept, +120 shade)


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--config", default="C5", help="C1..C5 (C5 is the headline; others for exploration)")
    ap.add_argument("--spp", type=int, default=0, help="override samples per pixel (0 = the config's own)")
    ap.add_argument("--depth", type=int, default=0, help="override the path depth (0 = the config's own; exploration only)")
    ap.add_argument("--e2e-steps", type=int, default=0, help="timed e2e steps (0 = same as --steps)")
    ap.add_argument("--cpu-seconds", type=float, default=15.0, help="target CPU time of the cpu_baseline sample")
    ap.add_argument("--flags", type=int, default=0)
    ap.add_argument("--partition", default="samples", choices=["samples", "tiles", "frames"],
                    help="how the frame is split over the GPUs (SURVEY 8e A / B / C)")
    ap.add_argument("--band", type=int, default=8, help="rows per band of --partition tiles")
    ap.add_argument("--single-process", action="store_true",
                    help="one process drives --gpus N devices through rtw_render_multi (samples / tiles)")
    ap.add_argument("--devices", default="", help="--single-process: explicit device list, e.g. 0,0 (two shares side by side on device 0)")
    ap.add_argument("--synthetic-textures", action="store_true",
                    help="generated stand-ins instead of the reference's earthmap.jpg / default.png (named in `data`)")
    return ap.parse_args()


# ----------------------------------------------------------------------------- clocks
class ClockSampler:
    """nvidia-smi clocks + throttle reasons sampled during the timed region (B200_PROFILING.md)."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index, self.proc, self.lines = index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "200"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except OSError:
            self.proc = None

    def _read(self):
        for ln in self.proc.stdout:
            self.lines.append(ln.strip())

    def stop(self) -> dict:
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0])); mx.append(float(f[1]))
            except ValueError:
                continue
            for n, v in zip(names, f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "samples": len(sm), "reasons": sorted(reasons)}


# ----------------------------------------------------------------------------- workload
def make_job(args):
    from rtw_b200 import scenes
    if args.synthetic_textures:
        scenes.SYNTHETIC_TEXTURES = True
    kw = {}
    if args.spp:
        kw["samples"] = args.spp
    if args.depth:
        kw["depth"] = args.depth
    job = scenes.CONFIGS[args.config](**kw)
    job.viewport.flags = args.flags
    return job


def data_label() -> str:
    from rtw_b200 import scenes
    return scenes.texture_data_label()


def workload_name(job) -> str:
    vp = job.viewport
    return f"{job.name}: {job.note or job.name} ({vp.width}x{vp.height}, {vp.samples} spp, depth {vp.depth})"


# ----------------------------------------------------------------------------- CPU arm
def cpu_sample(job, target_s: float, fast: bool = True):
    """Time the CPU oracle (the port of the reference's renderer; row-parallel over every host thread
    like async_render's one-task-per-row, viewport.rs:236-240) on a bounded sample of the workload:
    the full frame at n spp, n chosen so the sample costs about `target_s` seconds."""
    from oracle_backend import OracleBackend
    from rtw_b200 import _abi as A
    orc = OracleBackend(threads=0, fast=fast)
    vp = job.viewport
    h = orc.scene(job.scene)
    cam = vp.camera(orc)
    o = vp.opts(job.ray_color, job.mode)
    o.output = A.RTW_OUT_SUM
    o.sample_begin, o.sample_count = 0, 1
    h.render(cam, o)
    t1 = h.stats().seconds
    paths, rays, secs, spp = h.stats().paths, h.stats().rays, t1, 1
    extra = int(min(vp.samples - 1, max(0, round(target_s / max(t1, 1e-6)) - 1)))
    if extra > 0:
        o.sample_begin, o.sample_count = 1, extra
        h.render(cam, o)
        st = h.stats()
        paths += st.paths; rays += st.rays; secs += st.seconds; spp += extra
    threads = h.stats().threads
    return {"paths": paths, "rays": rays, "seconds": secs, "spp": spp, "threads": int(threads),
            "sample": f"{vp.width}x{vp.height} full frame x {spp} of {vp.samples} spp, depth {vp.depth}"}


def cpu_baseline_obj(s) -> dict:
    return {"value": s["paths"] / s["seconds"] / 1e6, "unit": UNIT, "cores": s["threads"], "kind": "port",
            "sample": s["sample"], "mrays_per_s": s["rays"] / s["seconds"] / 1e6, "seconds": s["seconds"]}


def run_reference(args):
    """--impl reference: the reference's CPU renderer for this path. The Rust crate cannot be built
    here (no cargo/rustc on the image), so this is the oracle port (oracle/, -O3 -march=native),
    with every host thread, each step a bounded sample of the same workload."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    job = make_job(args)
    per_step = max(2.0, min(args.cpu_seconds, 120.0 / max(1, args.steps + args.warmup)))
    for _ in range(args.warmup):
        cpu_sample(job, per_step)
    tot_paths = tot_rays = 0
    tot_s = 0.0
    last = None
    for _ in range(args.steps):
        last = cpu_sample(job, per_step)
        tot_paths += last["paths"]; tot_rays += last["rays"]; tot_s += last["seconds"]
    val = tot_paths / tot_s / 1e6
    vp = job.viewport
    line = {"impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": tot_s / args.steps * 1e3, "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": data_label(),
            "config": config_of(job),
            "run": {"step": last["sample"], "host_threads": last["threads"]},
            "mrays_per_s": tot_rays / tot_s / 1e6,
            "cpu_baseline": {"value": val, "unit": UNIT, "cores": last["threads"], "kind": "port", "sample": last["sample"]},
            "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "full_frame_seconds_extrapolated": vp.width * vp.height * vp.samples / (val * 1e6)}
    print(json.dumps(line), flush=True)


# ----------------------------------------------------------------------------- GPU arm
def wavefront_probe(per_step) -> bool:
    return bool(per_step) and per_step[0][5] > 0


def run_b200(args):
    import torch
    import torch.distributed as dist
    import rtw_b200
    from rtw_b200 import _abi as A
    from rtw_b200.backend import CudaBackend
    from rtw_b200.sharding import band_rows_for_rank, sample_range

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the product has no CPU path (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    dev = torch.device("cuda", local)

    def barrier():
        if world > 1:
            dist.barrier()

    job = make_job(args)
    vp = job.viewport
    be = CudaBackend(device=local)
    rtw_b200.set_default_backend(be)
    h = be.scene(job.scene)
    cam = vp.camera(be)
    W, H, spp = cam.width, cam.height, vp.samples
    begin, count = sample_range(spp, world, rank)
    part = args.partition
    o = vp.opts(job.ray_color, job.mode)
    o.output = A.RTW_OUT_SUM
    o.sample_begin, o.sample_count = begin, count
    buf = torch.empty((H, W, 3), dtype=torch.float32, device=dev)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)      # > 126 MB L2
    stream = torch.cuda.current_stream().cuda_stream
    # tiles: this rank's bands as one render of interleaved rows into a compact buffer, then put in place in the frame
    my_rows = band_rows_for_rank(H, world, rank, args.band) if part == "tiles" else np.zeros(0, dtype=np.int64)
    row_idx = torch.from_numpy(my_rows.astype(np.int64)).to(dev)
    cbuf = torch.empty((max(1, my_rows.size), W, 3), dtype=torch.float32, device=dev)
    frame_no = [0]

    def tile_opts():
        ob = vp.opts(job.ray_color, job.mode)
        ob.row_begin, ob.row_count = rank * args.band, int(my_rows.size)
        ob.row_band, ob.row_stride = args.band, world * args.band
        return ob

    def collect_stats(collect):
        st = h.stats()                                  # syncs on the last render's events
        collect.append((st.kernel_ms, st.rays, st.paths, st.nan_paths, st.gpu_launches, st.iterations, st.extend_ms, st.shade_ms))

    def step(collect=None):
        flush.zero_()                                   # L2 flush between steps
        if part == "samples":
            if count > 0:
                h.render_device(cam, o, buf.data_ptr(), stream)
            else:
                buf.zero_()
            if world > 1:
                dist.reduce(buf, dst=0, op=dist.ReduceOp.SUM)
            if rank == 0:
                be.resolve_device(buf.data_ptr(), H * W, spp, vp.gamma, stream)
            if collect is not None and count > 0:
                collect_stats(collect)
        elif part == "tiles":
            # row bands dealt round-robin: every rank renders and RESOLVES its own bands in place in a
            # zeroed frame; the frames are disjoint, so the gather to rank 0 is a reduce(sum) of them with zeros
            # (x + 0 = x: the gather changes no value). No sample reduction across GPUs anywhere.
            if world > 1:
                buf.zero_()
            if my_rows.size:
                h.render_device(cam, tile_opts(), cbuf.data_ptr(), stream)
                buf[row_idx] = cbuf[:my_rows.size]
            if world > 1:
                dist.reduce(buf, dst=0, op=dist.ReduceOp.SUM)
            if collect is not None and my_rows.size:
                collect_stats(collect)
        else:
            # frames (render_multi viewport.rs:249-269): frame k of the animation on rank k mod n, every
            # rank renders ONE whole frame per step; nothing is exchanged (rank 0's frame is the one that is read back)
            of = vp.opts(job.ray_color, job.mode)
            of.frame = vp.frame + frame_no[0] * world + rank
            frame_no[0] += 1
            h.render_device(cam, of, buf.data_ptr(), stream)
            if collect is not None:
                collect_stats(collect)

    # The CPU baseline leg runs BEFORE the GPU legs: measured right after a minute of the GPU at 700 W the same 16 host
    # threads ran the same sample at half their rate on some boxes (0.65 against 1.30 Mpaths/s; `--impl reference`, a
    # process of its own, read 1.33 on the same box: profiles/r2x_bench_C5_n1.json, r3b_*; tools/cpu_probe.py rules out
    # torch, the CUDA context and the library). Nothing of the GPU arm runs while it is timed.
    cpu_pre = cpu_baseline_obj(cpu_sample(job, args.cpu_seconds)) if (world == 1 and rank == 0) else None
    for _ in range(args.warmup):
        step()
    torch.cuda.synchronize()
    sampler = ClockSampler(local) if rank == 0 else None
    if sampler:
        sampler.start()
    per_step = []
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier(); torch.cuda.synchronize()
    ev0.record()
    for _ in range(args.steps):
        step(per_step)
    ev1.record()
    torch.cuda.synchronize(); barrier()
    clocks = sampler.stop() if sampler else None
    ms = torch.tensor([ev0.elapsed_time(ev1)], dtype=torch.float64, device=dev)
    tot = torch.tensor([sum(p[1] for p in per_step), sum(p[2] for p in per_step), sum(p[3] for p in per_step)],
                       dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        dist.all_reduce(tot, op=dist.ReduceOp.SUM)
    elapsed_ms = float(ms.item())
    rays_total, paths_total, nan_total = (float(x) for x in tot.tolist())
    image_mean = None
    if rank == 0:
        img = buf.cpu().numpy()
        fin = np.isfinite(img)
        image_mean = float(img[fin].mean())

    # ---- e2e: the C-ABI call a host program makes, host buffers, every step from scratch
    fs = job.scene.flatten()
    host_out = torch.empty((H, W, 3), dtype=torch.float32).pin_memory()
    lib = be.lib
    e2e_steps = args.e2e_steps or args.steps

    e2e_trace = []

    def e2e_step():
        hd = C.c_void_p()
        t_a = time.perf_counter()
        A.check(lib.rtw_scene_create(C.byref(fs.desc), local, C.byref(hd)))
        e2e_trace.append(("create_ms", (time.perf_counter() - t_a) * 1e3))
        try:
            if part == "frames":
                of = vp.opts(job.ray_color, job.mode)
                of.frame = vp.frame + rank
                A.check(lib.rtw_render(hd, C.byref(cam), C.byref(of), C.c_void_p(host_out.data_ptr())))
            elif part == "tiles":
                if world > 1:
                    buf.zero_()
                if my_rows.size:
                    A.check(lib.rtw_render_device(hd, C.byref(cam), C.byref(tile_opts()), C.c_void_p(cbuf.data_ptr()), C.c_void_p(stream)))
                    buf[row_idx] = cbuf[:my_rows.size]
                if world > 1:
                    dist.reduce(buf, dst=0, op=dist.ReduceOp.SUM)
                if rank == 0:
                    host_out.copy_(buf, non_blocking=True)
                torch.cuda.synchronize()
            elif world == 1:
                o2 = vp.opts(job.ray_color, job.mode)
                A.check(lib.rtw_render(hd, C.byref(cam), C.byref(o2), C.c_void_p(host_out.data_ptr())))
            else:
                if count > 0:
                    A.check(lib.rtw_render_device(hd, C.byref(cam), C.byref(o), C.c_void_p(buf.data_ptr()), C.c_void_p(stream)))
                else:
                    buf.zero_()
                dist.reduce(buf, dst=0, op=dist.ReduceOp.SUM)
                if rank == 0:
                    be.resolve_device(buf.data_ptr(), H * W, spp, vp.gamma, stream)
                    host_out.copy_(buf, non_blocking=True)
                torch.cuda.synchronize()
            e2e_trace.append(("render_ms", (time.perf_counter() - t_a) * 1e3))
            inf = A.rtw_scene_info()
            A.check(lib.rtw_scene_get_info(hd, C.byref(inf)))
            return int(inf.h2d_bytes), int(inf.device_bytes)
        finally:
            t_b = time.perf_counter()
            lib.rtw_scene_destroy(hd)
            e2e_trace.append(("destroy_ms", (time.perf_counter() - t_b) * 1e3))

    e2e_step()
    barrier(); torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        h2d, scene_device_bytes = e2e_step()
    torch.cuda.synchronize(); barrier()
    e2e_s = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(e2e_s, op=dist.ReduceOp.MAX)
    e2e_s = float(e2e_s.item())
    if os.environ.get("RTW_BENCH_E2E_TRACE") and rank == 0:
        print("e2e phases (cumulative ms within each step): " + " ".join(f"{k}={v:.1f}" for k, v in e2e_trace), file=sys.stderr)

    # ---- rank 0: algorithmic work per ray (instrumented kernel), roofline denominators, CPU baseline
    if rank == 0:
        oc = vp.opts(job.ray_color, job.mode)
        oc.output = A.RTW_OUT_SUM
        oc.flags |= A.RTW_FLAG_COUNT_WORK
        oc.sample_begin, oc.sample_count = 0, min(256, spp)
        h.render_device(cam, oc, buf.data_ptr(), stream)
        wc = h.stats()
        per_ray = {"node": wc.nodes / wc.rays, "sphere": wc.sphere_tests / wc.rays, "quad": wc.quad_tests / wc.rays,
                   "instance": wc.instance_tests / wc.rays, "hit": wc.hits / wc.rays}
        sched = h.sched_stats() if wavefront_probe(per_step) else None
        fp32_peak = be.microbench(A.RTW_BENCH_FP32)
        l2_peak = be.microbench(A.RTW_BENCH_L2)
        peaks_file = os.path.join(ROOT, "MEASURED_PEAKS.json")
        hbm_peak, hbm_src = 6650.0, "fallback"
        if os.path.exists(peaks_file):
            try:
                hbm_peak, hbm_src = float(json.load(open(peaks_file))["hbm_gbs"]), "measured"
            except Exception:
                pass
        # dominant kernel: average launch duration over the timed region, rank 0, CUDA events recorded by
        # the library on the launching stream (wavefront: every 8th extend/shade launch is bracketed)
        wavefront = bool(per_step) and per_step[0][5] > 0
        rays_step = statistics.mean(p[1] for p in per_step) if per_step else 0.0
        if wavefront:
            kname = "wf_extend_kernel"
            iters = statistics.mean(p[5] for p in per_step)
            k_ms = statistics.mean(p[6] for p in per_step)
            shade_ms = statistics.mean(p[7] for p in per_step)
            rays_launch = rays_step / iters
            work_keys = ("node", "sphere", "quad", "instance")     # the Hit/scatter work runs in wf_shade_kernel
        else:
            kname, iters, shade_ms = "mega_kernel", 0, 0.0
            k_ms = statistics.mean(p[0] for p in per_step) if per_step else float("nan")
            rays_launch = rays_step
            work_keys = tuple(COST)
        flop_per_ray = sum(per_ray[k] * COST[k][0] for k in work_keys)
        bytes_per_ray = sum(per_ray[k] * COST[k][1] for k in work_keys)
        tflops = flop_per_ray * rays_launch / (k_ms * 1e-3) / 1e12
        l2_gbs = bytes_per_ray * rays_launch / (k_ms * 1e-3) / 1e9
        # path records streamed per ray: extend 40 R + 16 W, shade 72 R + 48 W, + 8 W per new path
        state_bytes = (176.0 + 8.0 * paths_total / max(1.0, rays_total)) if wavefront else 0.0
        step_ms = elapsed_ms / args.steps
        hbm_gbs = (state_bytes * rays_step + H * W * 3 * 4 * 2.0) / (step_ms * 1e-3) / 1e9
        # the kernel's own HBM stream (wavefront: 40 B of ray record read + 16 B of hit record written per ray)
        k_hbm_bytes = 56.0 if wavefront else 0.0
        k_hbm_gbs = k_hbm_bytes * rays_launch / (k_ms * 1e-3) / 1e9
        cands = [("fp32", tflops, fp32_peak, "TFLOP/s"), ("l2", l2_gbs, l2_peak, "GB/s"), ("hbm", k_hbm_gbs, hbm_peak, "GB/s")]
        bname, bach, bpeak, bunit = max(cands, key=lambda c: c[1] / c[2])
        traffic, traffic_src, measured = None, None, None
        try:      # DRAM bytes of the dominant kernel per launch, from the committed ncu --set full capture
            tj = json.load(open(os.path.join(ROOT, "profiles", "ncu_traffic.json")))[kname]
            # the capture ran at slots_per_launch slots; this run's launches hold rays_launch rays
            traffic = float(tj["dram_bytes_per_launch"]) * rays_launch / float(tj.get("slots_per_launch", rays_launch))
            traffic_src = tj["source"]
            # what the memory system actually moved for this kernel in that capture, per launch, scaled the same way:
            # the algorithmic scene bytes (`per_ray.scene_bytes`) are mostly served by L1 - L2 sees the misses only
            scale = rays_launch / float(tj.get("slots_per_launch", rays_launch))
            measured = {"l2_bytes_per_launch": float(tj["l2_bytes_per_launch"]) * scale, "l1_hit_rate_pct": tj["l1_hit_rate_pct"],
                        "l2_hit_rate_pct": tj["l2_hit_rate_pct"], "l2_bytes_per_ray": float(tj["l2_bytes_per_launch"]) / float(tj["slots_per_launch"]),
                        "note": "lts__t_sectors.sum x 32 B and l1tex__t_sector_hit_rate.pct of the committed ncu --set full capture"}
        except Exception:
            pass
        roofline = {"kernel": kname, "bound": bname, "achieved": bach, "peak": bpeak, "unit": bunit,
                    "frac": bach / bpeak, "traffic": traffic, "traffic_source": traffic_src,
                    "traffic_note": "ncu DRAM bytes per slot (captured at a pool of 2 000 128 slots) x this run's rays per launch",
                    "algorithmic_hbm_bytes_per_launch": k_hbm_bytes * rays_launch, "measured_memory": measured,
                    "peak_source": "rtw_microbench on this GPU for fp32 and L2 (MEASURED_PEAKS.json has neither); MEASURED_PEAKS.json hbm_gbs for HBM" if hbm_src == "measured" else "rtw_microbench on this GPU for fp32 and L2; B200_PROFILING.md fallback for HBM",
                    "kernel_ms_per_launch": k_ms, "rays_per_launch": rays_launch,
                    "launches_per_step": iters if wavefront else 1,
                    "share_of_step": (k_ms * iters / step_ms) if wavefront else (k_ms / step_ms),
                    "shade_ms_per_launch": shade_ms,
                    "per_ray": {**per_ray, "flop": flop_per_ray, "scene_bytes": bytes_per_ray},
                    "extend_warp_scheduling": sched,
                    "all": {"fp32": {"achieved": tflops, "peak": fp32_peak, "unit": "TFLOP/s", "frac": tflops / fp32_peak},
                            "l2": {"achieved": l2_gbs, "peak": l2_peak, "unit": "GB/s", "frac": l2_gbs / l2_peak},
                            "hbm_kernel": {"achieved": k_hbm_gbs, "peak": hbm_peak, "unit": "GB/s", "frac": k_hbm_gbs / hbm_peak},
                            "hbm": {"achieved": hbm_gbs, "peak": hbm_peak, "unit": "GB/s", "frac": hbm_gbs / hbm_peak,
                                    "peak_source": hbm_src,
                                    "note": "whole step: path records streamed by extend+shade (176 B/ray + 8 B per new path) + framebuffer REDs; the scene (<10 MB) is L1/L2-resident"}}}
        cpu = cpu_pre
        paths_frame = W * H * spp * (world if part == "frames" else 1)     # paths per step
        val = paths_frame * args.steps / (elapsed_ms * 1e-3) / 1e6
        part_text = {"samples": f"sample range x{world} + NCCL reduce(sum) to rank 0",
                     "tiles": f"row bands of {args.band} rows dealt round-robin over {world} GPU(s), one interleaved-row render per GPU, resolved in place, gathered to rank 0 (no reduction)",
                     "frames": f"one animation frame per GPU per step (frame k on rank k mod {world}), no exchange"}[part]
        line = {"metric": METRIC, "value": val, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
                "ms_per_step": elapsed_ms / args.steps, "higher_is_better": True, "scaling": "weak" if part == "frames" else "strong", "vs_baseline": None,
                "dtype": "f32", "data": data_label(),
                "config": config_of(job),
                "run": {"partition": part_text,
                           "pipeline": "wavefront (extend + shade kernels over path queues in HBM)" if wavefront else "megakernel",
                           "fp": "strict (-fmad=false)" if args.flags & 1 else "fast (FFMA contraction)"},
                "mrays_per_s": rays_total / (elapsed_ms * 1e-3) / 1e6, "rays_per_path": rays_total / max(1.0, paths_total),
                "nan_paths": nan_total, "image_mean": image_mean,
                "e2e": {"value": paths_frame * e2e_steps / e2e_s / 1e6, "unit": UNIT, "h2d_bytes_per_step": h2d * world,
                        "d2h_bytes_per_step": H * W * 3 * 4 * (world if part == "frames" else 1), "scene_device_bytes": scene_device_bytes, "ms_per_step": e2e_s / e2e_steps * 1e3, "steps": e2e_steps,
                        "what": "rtw_scene_create + rtw_render(host out) + rtw_scene_destroy per step"},
                "gpu_launches": int(sum(p[4] for p in per_step)) + args.steps, "clocks": clocks, "roofline": roofline}
        if cpu is not None:
            line["cpu_baseline"] = cpu
        print(json.dumps(line), flush=True)
    barrier()
    if world > 1:
        dist.destroy_process_group()


def run_single_process(args):
    """--single-process: ONE process drives --gpus N devices through the C ABI's multi-device entry points
    (rtw_scene_multi_create / rtw_render_multi: a host thread and a stream per device, shares combined on device 0
    by one peer-memory kernel) - what the reference's single-process caller (main.rs:76-80,412-414) would bind.
    value: paths / (max over devices of the share's device time + the combine kernel), CUDA events inside the
    library; e2e: wall clock of create + render into a pinned host buffer + destroy, every step."""
    import torch
    import rtw_b200
    from rtw_b200 import _abi as A
    from rtw_b200.backend import CudaBackend, CudaMultiSceneHandle
    if args.partition == "frames":
        raise SystemExit("--single-process: samples or tiles (frames: rtw_render_frames_multi is exercised by the tests)")
    be = CudaBackend(device=0)
    n = args.gpus
    if args.devices:
        devices = [int(x) for x in args.devices.split(",")]
        n = len(set(devices))
    else:
        devices = list(range(n))
    if be.device_count() <= max(devices):
        raise SystemExit(f"--single-process: device {max(devices)} asked for, only {be.device_count()} device(s) visible")
    job = make_job(args)
    vp = job.viewport
    cam = vp.camera(be)
    W, H, spp = cam.width, cam.height, vp.samples
    o = vp.opts(job.ray_color, job.mode)
    split = A.RTW_SPLIT_SAMPLES if args.partition == "samples" else A.RTW_SPLIT_TILES
    host_out = torch.empty((H, W, 3), dtype=torch.float32).pin_memory().numpy()
    ms = be.scene_multi(job.scene, devices)
    for _ in range(args.warmup):
        ms.render(cam, o, split, args.band, out=host_out)
    sampler = ClockSampler(0)
    sampler.start()
    dev_ms = 0.0
    rays = paths = launches = 0
    t0 = time.perf_counter()
    for _ in range(args.steps):
        ms.render(cam, o, split, args.band, out=host_out)
        tot, per, comb = ms.stats()
        dev_ms += tot.total_ms + comb
        rays += tot.rays; paths += tot.paths; launches += tot.gpu_launches
    wall_s = time.perf_counter() - t0
    clocks = sampler.stop()
    ms.close()
    del job.scene._backend_cache[("cuda-multi", tuple(devices))]
    fs = job.scene.flatten()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        m2 = CudaMultiSceneHandle(be.lib, fs, devices)
        m2.render(cam, o, split, args.band, out=host_out)
        m2.close()
    e2e_s = time.perf_counter() - t0
    paths_frame = W * H * spp
    fin = np.isfinite(host_out)
    line = {"metric": METRIC, "value": paths_frame * args.steps / (dev_ms * 1e-3) / 1e6, "unit": UNIT, "n_gpus": n, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": dev_ms / args.steps, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "f32", "data": data_label(),
            "config": config_of(job),
            "run": {"launch": "single process, rtw_render_multi (one host thread + stream per share)", "devices": devices,
                       "partition": ("sample ranges, combined on device 0 by one peer-memory kernel" if split == A.RTW_SPLIT_SAMPLES
                                     else f"row bands of {args.band or 8} rows dealt round-robin, copied straight to the host image"),
                       "timing": "value: max over devices of the share's device time (CUDA events) + combine kernel; wall_ms_per_step: host clock around the call"},
            "wall_ms_per_step": wall_s / args.steps * 1e3, "mrays_per_s": rays / (dev_ms * 1e-3) / 1e6,
            "rays_per_path": rays / max(1, paths), "image_mean": float(host_out[fin].mean()),
            "e2e": {"value": paths_frame * args.steps / e2e_s / 1e6, "unit": UNIT, "ms_per_step": e2e_s / args.steps * 1e3,
                    "d2h_bytes_per_step": H * W * 3 * 4, "what": "rtw_scene_multi_create + rtw_render_multi(host out) + rtw_scene_multi_destroy per step"},
            "gpu_launches": int(launches), "clocks": clocks}
    print(json.dumps(line), flush=True)


if __name__ == "__main__":
    a = parse_args()
    if a.impl == "reference":
        run_reference(a)
    elif a.single_process:
        run_single_process(a)
    else:
        run_b200(a)
