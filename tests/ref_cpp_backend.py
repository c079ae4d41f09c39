"""ctypes loader for oracle/_ref/libref_cpp.so — the REFERENCE'S OWN C++ renderer (Book-1 spheres),
built from /root/reference/C++ by oracle/Makefile (`make ref`). TEST INFRASTRUCTURE ONLY. Present
only where the reference tree was available at build time; the committed fixtures under
tests/golden/ carry its outputs everywhere else."""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "oracle", "_ref", "libref_cpp.so")

MODE_NORMAL, MODE_T, MODE_GLASS_TEST, MODE_MIRROR_SKY = 0, 1, 2, 3


def available() -> bool:
    return os.path.exists(LIB)


_lib = None


def load():
    global _lib
    if _lib is None:
        _lib = C.CDLL(LIB)
        f32p = C.POINTER(C.c_float)
        _lib.ref_cpp_render_no_rand.restype = C.c_int
        _lib.ref_cpp_render_no_rand.argtypes = [f32p, C.c_int, C.c_uint, C.c_float, C.c_float, f32p, f32p, f32p, C.c_int,
                                                C.c_float, C.c_float, C.c_uint, C.c_void_p, C.POINTER(C.c_uint)]
    return _lib


def render_no_rand(spheres, width, aspect, vfov, origin, vup, direction, mode, mint, maxt, max_depth=10) -> np.ndarray:
    """spheres: rows of (ox, oy, oz, r, cr, cg, cb, metallicness, opacity, ir)."""
    lib = load()
    sp = np.ascontiguousarray(spheres, dtype=np.float32).reshape(-1, 10)
    height = int(np.float32(width) / np.float32(aspect))
    out = np.zeros((height, width, 3), dtype=np.float32)
    h = C.c_uint()
    f3 = lambda v: (C.c_float * 3)(*[float(x) for x in v])
    rc = lib.ref_cpp_render_no_rand(sp.ctypes.data_as(C.POINTER(C.c_float)), sp.shape[0], width, aspect, vfov, f3(origin),
                                    f3(vup), f3(direction), mode, mint, maxt, max_depth, out.ctypes.data_as(C.c_void_p),
                                    C.byref(h))
    if rc != 0:
        raise RuntimeError("reference C++ renderer threw")
    assert h.value == height
    return out
