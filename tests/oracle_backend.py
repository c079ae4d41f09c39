"""ctypes loader for the CPU oracle (oracle/liboracle.so) exposing the same backend interface as
rtw_b200.backend.CudaBackend, so a test renders one Scene/Viewport through both and compares.

TEST INFRASTRUCTURE ONLY: nothing under raytracing-in-a-weekend_b200/ imports this.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

import rtw_b200  # noqa: F401  (registers the package)
from rtw_b200 import _abi as A

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")


class oracle_stats(C.Structure):
    _fields_ = [("paths", C.c_uint64), ("rays", C.c_uint64), ("nan_paths", C.c_uint64),
                ("slab_tests", C.c_uint64), ("sphere_tests", C.c_uint64), ("quad_tests", C.c_uint64),
                ("instance_tests", C.c_uint64), ("seconds", C.c_double), ("threads", C.c_uint32)]


class oracle_hit(C.Structure):
    _fields_ = [("hit", C.c_int32), ("t", C.c_float), ("normal", C.c_float * 3), ("point", C.c_float * 3),
                ("col_mod", C.c_float * 3), ("metallicness", C.c_float), ("opacity", C.c_float),
                ("ir", C.c_float), ("emitted", C.c_float * 3)]


def build_oracle(fast: bool = False) -> str:
    name = "liboracle_fast.so" if fast else "liboracle.so"
    path = os.path.join(ORACLE_DIR, name)
    srcs = [os.path.join(ORACLE_DIR, f) for f in os.listdir(ORACLE_DIR) if f.endswith((".hpp", ".cpp"))]
    srcs.append(os.path.join(ROOT, "include", "rtw.h"))
    if not os.path.exists(path) or any(os.path.getmtime(s) > os.path.getmtime(path) for s in srcs):
        subprocess.check_call(["make", "-s", "-C", ORACLE_DIR, name])
    return path


_libs = {}


def load_oracle(fast: bool = False):
    if fast in _libs:
        return _libs[fast]
    lib = C.CDLL(build_oracle(fast))
    P = C.POINTER
    f32p = P(C.c_float)
    lib.oracle_scene_create.restype = C.c_int
    lib.oracle_scene_create.argtypes = [P(A.rtw_scene_desc), C.c_uint32, P(C.c_void_p)]
    lib.oracle_scene_destroy.restype = None
    lib.oracle_scene_destroy.argtypes = [C.c_void_p]
    lib.oracle_render.restype = C.c_int
    lib.oracle_render.argtypes = [C.c_void_p, P(A.rtw_camera), P(A.rtw_render_opts), C.c_void_p, C.c_uint32,
                                  P(oracle_stats)]
    lib.oracle_render_ex.restype = C.c_int
    lib.oracle_render_ex.argtypes = [C.c_void_p, P(A.rtw_camera), P(A.rtw_render_opts), C.c_void_p, C.c_uint32,
                                     P(oracle_stats), C.c_void_p]
    lib.oracle_camera_init.restype = C.c_int
    lib.oracle_camera_init.argtypes = [P(A.rtw_camera), C.c_uint32, C.c_float, C.c_float, f32p, f32p, f32p,
                                       C.c_float]
    lib.oracle_vec3_rotated.restype = None
    lib.oracle_vec3_rotated.argtypes = [f32p, f32p, f32p]
    lib.oracle_philox4x32_10.restype = None
    lib.oracle_philox4x32_10.argtypes = [P(C.c_uint32), P(C.c_uint32), P(C.c_uint32)]
    lib.oracle_perlin_noise.restype = C.c_float
    lib.oracle_perlin_noise.argtypes = [P(A.rtw_noise), f32p]
    lib.oracle_reflectance.restype = C.c_float
    lib.oracle_reflectance.argtypes = [C.c_float, C.c_float]
    lib.oracle_refract.restype = None
    lib.oracle_refract.argtypes = [f32p, f32p, C.c_float, f32p]
    lib.oracle_interval_intersect.restype = C.c_int
    lib.oracle_interval_intersect.argtypes = [C.c_float, C.c_float, C.c_float, C.c_float, f32p]
    lib.oracle_quantize.restype = None
    lib.oracle_quantize.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64]
    lib.oracle_bilateral_filter.restype = None
    lib.oracle_bilateral_filter.argtypes = [C.c_void_p, C.c_uint32, C.c_uint32, C.c_uint32, C.c_int, C.c_int, C.c_void_p, f32p]
    lib.oracle_collide.restype = C.c_int
    lib.oracle_collide.argtypes = [C.c_void_p, f32p, f32p, C.c_float, C.c_float, C.c_float, C.c_uint64,
                                   C.c_uint32, C.c_uint32, C.c_uint32, P(oracle_hit)]
    _libs[fast] = lib
    return lib


def _f3(v):
    return (C.c_float * 3)(*[float(x) for x in v])


class OracleSceneHandle:
    def __init__(self, lib, fs, build_seed: int, threads: int):
        self._lib, self._fs, self.threads = lib, fs, threads
        self._h = C.c_void_p()
        rc = lib.oracle_scene_create(C.byref(fs.desc), build_seed, C.byref(self._h))
        if rc != 0:
            raise ValueError(f"oracle_scene_create failed: {rc}")
        self.last = oracle_stats()

    def close(self):
        if self._h:
            self._lib.oracle_scene_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def render(self, cam, o) -> np.ndarray:
        rows = o.row_count if o.row_count else cam.height
        out = np.empty((rows, cam.width, 3), dtype=np.float32)
        # per pixel: was some closest-hit decision taken between equal t (unpinned by the reference)?
        self.last_tie = np.zeros((rows, cam.width), dtype=np.uint8)
        rc = self._lib.oracle_render_ex(self._h, C.byref(cam), C.byref(o), out.ctypes.data_as(C.c_void_p),
                                        self.threads, C.byref(self.last), self.last_tie.ctypes.data_as(C.c_void_p))
        if rc != 0:
            raise ValueError(f"oracle_render failed: {rc}")
        return out

    def render_frames(self, cam, o, first_frame: int, n_frames: int) -> np.ndarray:
        # render_multi (viewport.rs:249-269): sequential frames, only `frame` changes
        outs = []
        o2 = A.rtw_render_opts.from_buffer_copy(o)
        for k in range(first_frame, first_frame + n_frames):
            o2.frame = k
            outs.append(self.render(cam, o2))
        return np.stack(outs)

    def stats(self) -> oracle_stats:
        return self.last

    def collide(self, origin, direction, time=0.0, mint=0.001, maxt=1e5, seed=1, pixel=0, sample=0, bounce=0):
        h = oracle_hit()
        self._lib.oracle_collide(self._h, _f3(origin), _f3(direction), time, mint, maxt, seed, pixel, sample,
                                 bounce, C.byref(h))
        return h


class OracleBackend:
    name = "oracle"

    def __init__(self, threads: int = 0, build_seed: int = 12345, fast: bool = False):
        self.lib = load_oracle(fast)
        self.threads, self.build_seed, self.fast = threads, build_seed, fast

    def camera_init(self, width, aspect_ratio, vfov, origin, direction, vup, lens_radius) -> A.rtw_camera:
        cam = A.rtw_camera()
        self.lib.oracle_camera_init(C.byref(cam), width, aspect_ratio, vfov, _f3(origin), _f3(direction),
                                    _f3(vup), lens_radius)
        return cam

    def quantize(self, img: np.ndarray) -> np.ndarray:
        """write_img_f32's conversion (write_img.rs:12-14), CPU restatement."""
        src = np.ascontiguousarray(img, dtype=np.float32)
        out = np.empty(src.shape, dtype=np.uint8)
        self.lib.oracle_quantize(src.ctypes.data_as(C.c_void_p), out.ctypes.data_as(C.c_void_p), src.size)
        return out

    def bilateral_filter(self, img_u8: np.ndarray, size: int, kind: int = 0, info: dict | None = None,
                         exp_mode: int = 1) -> np.ndarray:
        """Rust2/src/postprocessing.rs:70-135, CPU restatement. exp_mode 0: libm expf (what f32::exp calls);
        1: double exp rounded to f32 (what the CUDA path computes)."""
        src = np.ascontiguousarray(img_u8, dtype=np.uint8)
        out = np.empty_like(src)
        par = (C.c_float * 4)()
        self.lib.oracle_bilateral_filter(src.ctypes.data_as(C.c_void_p), src.shape[1], src.shape[0], int(size), int(kind),
                                         int(exp_mode), out.ctypes.data_as(C.c_void_p), par)
        if info is not None:
            info.update(spatial=par[0], avg_gradient=par[1], inv_range=par[2], inv_spatial=par[3])
        return out

    def scene(self, scene) -> OracleSceneHandle:
        key = ("oracle", self.build_seed, self.fast)
        h = scene._backend_cache.get(key)
        if h is None:
            h = OracleSceneHandle(self.lib, scene.flatten(), self.build_seed, self.threads)
            scene._backend_cache[key] = h
        h.threads = self.threads
        return h
