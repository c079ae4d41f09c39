"""Splits one end-to-end frame (rtw_scene_create + rtw_render to a host buffer + rtw_scene_destroy) into its phases.
Usage: python tools/e2e_phases.py [--spp 2000] [--config C5]"""
import argparse, ctypes as C, json, os, sys, time
import numpy as np
sys.path.insert(0, os.path.join(os.path.dirname(__file__), ".."))
import torch
import rtw_b200
from rtw_b200 import _abi as A, scenes
from rtw_b200.backend import CudaBackend

ap = argparse.ArgumentParser()
ap.add_argument("--spp", type=int, default=2000); ap.add_argument("--config", default="C5"); ap.add_argument("--reps", type=int, default=3)
a = ap.parse_args()
job = getattr(scenes, "config_" + a.config.lower())(samples=a.spp)
be = CudaBackend(0); lib = be.lib
vp = job.viewport; cam = vp.camera(be)
fs = job.scene.flatten()
H, W = cam.height, cam.width
host_out = torch.empty((H, W, 3), dtype=torch.float32).pin_memory()
pageable = np.empty((H, W, 3), dtype=np.float32)
for rep in range(a.reps):
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    hd = C.c_void_p(); A.check(lib.rtw_scene_create(C.byref(fs.desc), 0, C.byref(hd)))
    t1 = time.perf_counter()
    o = vp.opts(job.ray_color, job.mode)
    A.check(lib.rtw_render(hd, C.byref(cam), C.byref(o), C.c_void_p(host_out.data_ptr())))
    t2 = time.perf_counter()
    st = A.rtw_stats(); A.check(lib.rtw_get_stats(hd, C.byref(st)))
    A.check(lib.rtw_render(hd, C.byref(cam), C.byref(o), C.c_void_p(host_out.data_ptr())))
    t3 = time.perf_counter()
    st2 = A.rtw_stats(); A.check(lib.rtw_get_stats(hd, C.byref(st2)))
    lib.rtw_scene_destroy(hd)
    t4 = time.perf_counter()
    print(json.dumps({"create_ms": (t1 - t0) * 1e3, "render1_ms": (t2 - t1) * 1e3, "render1_device_ms": st.total_ms,
                      "render2_ms": (t3 - t2) * 1e3, "render2_device_ms": st2.total_ms, "destroy_ms": (t4 - t3) * 1e3}))
