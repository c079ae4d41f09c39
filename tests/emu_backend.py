"""ctypes loader for tests/emu/libemu.so: the kernels' own headers (rtw_trace.h, rtw_shade.h) and the
host scene builder compiled for the CPU. TEST INFRASTRUCTURE ONLY — lets the CPU suite check the
flattening / SAH BVH / traversal / shading logic against the oracle without a GPU. The package never
imports this and librtw_b200.so does not contain it."""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

import rtw_b200  # noqa: F401
from rtw_b200 import _abi as A

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
EMU_DIR = os.path.join(ROOT, "tests", "emu")
CSRC = os.path.join(ROOT, "raytracing-in-a-weekend_b200", "csrc")


def build_emu() -> str:
    path = os.path.join(EMU_DIR, "libemu.so")
    srcs = [os.path.join(EMU_DIR, "emu.cpp"), os.path.join(ROOT, "include", "rtw.h")]
    srcs += [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".h", ".cpp"))]
    if not os.path.exists(path) or any(os.path.getmtime(s) > os.path.getmtime(path) for s in srcs):
        subprocess.check_call(["g++", "-std=c++17", "-O2", "-pthread", "-ffp-contract=off", "-fno-fast-math", "-fPIC", "-shared", "-DRTW_COUNT_WORK=1",
                               "-I/usr/local/cuda/include", "-o", path, os.path.join(EMU_DIR, "emu.cpp"),
                               os.path.join(CSRC, "scene_build.cpp")])
    return path


_lib = None


def load_emu():
    global _lib
    if _lib is None:
        _lib = C.CDLL(build_emu())
        _lib.emu_render.restype = C.c_int
        _lib.emu_render.argtypes = [C.POINTER(A.rtw_scene_desc), C.POINTER(A.rtw_camera), C.POINTER(A.rtw_render_opts),
                                    C.c_void_p, C.POINTER(C.c_uint64), C.POINTER(A.rtw_scene_info)]
        _lib.emu_render_ex.restype = C.c_int
        _lib.emu_render_ex.argtypes = _lib.emu_render.argtypes + [C.c_int]
    return _lib


class EmuSceneHandle:
    def __init__(self, lib, fs, lbvh: bool = False, baked: bool = False, inside: bool = False, split_hit: bool = False):
        self._lib, self._fs = lib, fs
        self._lbvh = (1 if lbvh else 0) | (2 if baked else 0) | (4 if inside else 0) | (8 if split_hit else 0)
        self.last = (C.c_uint64 * 8)()
        self.last_info = A.rtw_scene_info()

    def close(self):
        pass

    def render(self, cam, o) -> np.ndarray:
        rows = o.row_count if o.row_count else cam.height
        out = np.empty((rows, cam.width, 3), dtype=np.float32)
        rc = self._lib.emu_render_ex(C.byref(self._fs.desc), C.byref(cam), C.byref(o), out.ctypes.data_as(C.c_void_p),
                                     self.last, C.byref(self.last_info), self._lbvh)
        if rc != 0:
            raise ValueError(f"emu_render failed: {rc}")
        return out

    def stats(self):
        keys = ("paths", "rays", "nan_paths", "nodes", "sphere_tests", "quad_tests", "instance_tests", "hits")
        return {k: self.last[i] for i, k in enumerate(keys)}

    def info(self):
        return self.last_info


class EmuBackend:
    name = "emu"

    def __init__(self, lbvh: bool = False, baked: bool = False, inside: bool = False, split_hit: bool = False):
        """lbvh=True: every level is rebuilt with the LBVH steps of rtw_lbvh.h (run in host loops)
        before rendering, as rtw_scene_rebuild_lbvh does on the device. baked=True: walk the top level in which
        rigidly placed surface instances are replaced by world-space copies of their primitives (rtw_host.h)."""
        self.lib = load_emu()
        self.lbvh, self.baked, self.inside = lbvh, baked, inside   # inside: node-count probes only (tools/lbvh_quality.py)
        self.split_hit = split_hit     # the Hit evaluation as the megakernel compiles it (rtw_shade.h eval_hit_split)

    def camera_init(self, width, aspect_ratio, vfov, origin, direction, vup, lens_radius):
        from oracle_backend import OracleBackend
        return OracleBackend().camera_init(width, aspect_ratio, vfov, origin, direction, vup, lens_radius)

    def scene(self, scene) -> EmuSceneHandle:
        key = ("emu", self.lbvh, self.baked, self.inside, self.split_hit)
        h = scene._backend_cache.get(key)
        if h is None:
            h = EmuSceneHandle(self.lib, scene.flatten(), self.lbvh, self.baked, self.inside, self.split_hit)
            scene._backend_cache[key] = h
        return h
