"""ctypes mirror of include/rtw.h and loader for the CUDA library (librtw_b200.so).

There is no CPU fallback: if the library has not been built (`python -c "import __graft_entry__ as g;
g.build()"`) importing the render entry points raises, and on a box without a CUDA device every
render call returns RTW_E_CUDA (surfaced as RtwError).
"""
from __future__ import annotations

import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("RTW_B200_LIB") or os.path.join(HERE, "csrc", "librtw_b200.so")   # env override: tuning builds

RTW_ABI_VERSION = 3

# error codes
RTW_OK, RTW_E_INVALID, RTW_E_CUDA, RTW_E_NOMEM, RTW_E_UNSUPPORTED = 0, -1, -2, -3, -4
RTW_NOALLOC_MAX_INTERIOR_PIXELS, RTW_NOALLOC_MAX_WINDOW = (1 << 20) - 2048, 16       # rtw_bilateral_filter_device_noalloc
# density kinds (instance.rs:21-26)
RTW_DENSITY_SURFACE, RTW_DENSITY_CONST = 0, 1
RTW_SHAPE_PARALLELOGRAM, RTW_SHAPE_TRIANGLE = 0, 1
# integrators
RTW_INT_GRADIENT, RTW_INT_D, RTW_INT_BG_COLOR, RTW_INT_NORMAL, RTW_INT_ALBEDO, RTW_INT_GLASS_TEST = range(6)
RTW_INT_LIGHT_CAST, RTW_INT_LIGHT_BIASED = 6, 7
RTW_LIGHT_SPHERE, RTW_LIGHT_QUAD = 0, 1
# modes
RTW_MODE_RANDOM, RTW_MODE_STRATIFIED, RTW_MODE_NO_RAND = range(3)
# outputs
RTW_OUT_RESOLVED, RTW_OUT_SUM = 0, 1
# flags
RTW_FLAG_STRICT_FP, RTW_FLAG_WAVEFRONT, RTW_FLAG_COUNT_WORK, RTW_FLAG_MEGAKERNEL = 1, 2, 4, 8
RTW_BENCH_FP32, RTW_BENCH_L2, RTW_BENCH_HBM = 0, 1, 2
RTW_SPLIT_SAMPLES, RTW_SPLIT_TILES, RTW_SPLIT_FRAMES = 0, 1, 2

f32, u32, i32, u64 = C.c_float, C.c_uint32, C.c_int32, C.c_uint64
f3 = f32 * 3


class rtw_material(C.Structure):
    _fields_ = [("metallicness", f32), ("opacity", f32), ("ir", f32), ("emitted", f3)]


class rtw_noise(C.Structure):
    _fields_ = [("ranvec", (f32 * 3) * 256), ("perm_x", u32 * 256), ("perm_y", u32 * 256), ("perm_z", u32 * 256)]


class rtw_texture(C.Structure):
    _fields_ = [("width", u32), ("height", u32), ("rgb", C.POINTER(f32)), ("noise_id", i32), ("noise_scale", f32)]


class rtw_sphere(C.Structure):
    _fields_ = [("origin", f3), ("radius", f32), ("col_mod", f3), ("velocity", f3), ("mat_id", u32), ("tex_id", u32)]


class rtw_quad(C.Structure):
    _fields_ = [("origin", f3), ("u", f3), ("v", f3), ("velocity", f3), ("mat_id", u32), ("tex_id", u32), ("shape", u32)]


class rtw_instance(C.Structure):
    _fields_ = [("first_sphere", u32), ("n_spheres", u32), ("first_quad", u32), ("n_quads", u32),
                ("translation", f3), ("rotation", f3), ("density_kind", u32), ("density", f32)]


class rtw_light(C.Structure):
    _fields_ = [("kind", u32), ("index", u32)]


class rtw_scene_desc(C.Structure):
    _fields_ = [("abi_version", u32),
                ("n_spheres", u32), ("n_top_spheres", u32),
                ("n_quads", u32), ("n_top_quads", u32),
                ("n_instances", u32),
                ("n_materials", u32), ("n_textures", u32), ("n_noises", u32),
                ("spheres", C.POINTER(rtw_sphere)), ("quads", C.POINTER(rtw_quad)),
                ("instances", C.POINTER(rtw_instance)), ("materials", C.POINTER(rtw_material)),
                ("textures", C.POINTER(rtw_texture)), ("noises", C.POINTER(rtw_noise)),
                ("background", f3), ("n_lights", u32), ("lights", C.POINTER(rtw_light))]


class rtw_camera(C.Structure):
    _fields_ = [("width", u32), ("height", u32), ("origin", f3), ("pixel00", f3), ("delta_u", f3),
                ("delta_v", f3), ("u", f3), ("v", f3), ("lens_radius", f32)]


class rtw_render_opts(C.Structure):
    _fields_ = [("samples", u32), ("depth", u32), ("gamma", f32), ("integrator", u32), ("mode", u32),
                ("shutter_speed", f32), ("fps", f32), ("frame", u32), ("seed", u64),
                ("mint", f32), ("maxt", f32),
                ("sample_begin", u32), ("sample_count", u32), ("row_begin", u32), ("row_count", u32),
                ("output", u32), ("flags", u32), ("light_weight", f32), ("row_band", u32), ("row_stride", u32)]


class rtw_stats(C.Structure):
    _fields_ = [("paths", u64), ("rays", u64), ("nan_paths", u64), ("gpu_launches", u64),
                ("kernel_ms", f32), ("total_ms", f32),
                ("nodes", u64), ("sphere_tests", u64), ("quad_tests", u64), ("instance_tests", u64), ("hits", u64),
                ("iterations", u64), ("extend_ms", f32), ("shade_ms", f32)]


class rtw_scene_info(C.Structure):
    _fields_ = [("n_tlas_nodes", u32), ("n_blas_nodes", u32), ("n_tlas_leaves", u32),
                ("n_flattened_instances", u32), ("bvh_max_depth", u32), ("n_baked_instances", u32), ("device_bytes", u64),
                ("h2d_bytes", u64)]


# every symbol include/rtw.h declares: name -> (restype, argtypes)
P = C.POINTER
SYMBOLS = {
    "rtw_abi_version": (C.c_int, []),
    "rtw_last_error": (C.c_char_p, []),
    "rtw_device_count": (C.c_int, []),
    "rtw_camera_init": (C.c_int, [P(rtw_camera), u32, f32, f32, P(f32), P(f32), P(f32), f32]),
    "rtw_camera_init_res": (C.c_int, [P(rtw_camera), u32, u32, f32, P(f32), P(f32), P(f32), f32]),
    "rtw_render_opts_default": (None, [P(rtw_render_opts)]),
    "rtw_scene_create": (C.c_int, [P(rtw_scene_desc), C.c_int, P(C.c_void_p)]),
    "rtw_scene_destroy": (None, [C.c_void_p]),
    "rtw_trim_memory": (C.c_int, [C.c_int]),
    "rtw_scene_rebuild_lbvh": (C.c_int, [C.c_void_p, P(f32)]),
    "rtw_scene_use_sah": (C.c_int, [C.c_void_p]),
    "rtw_render": (C.c_int, [C.c_void_p, P(rtw_camera), P(rtw_render_opts), C.c_void_p]),
    "rtw_render_device": (C.c_int, [C.c_void_p, P(rtw_camera), P(rtw_render_opts), C.c_void_p, C.c_void_p]),
    "rtw_render_frames": (C.c_int, [C.c_void_p, P(rtw_camera), P(rtw_render_opts), u32, u32, C.c_void_p]),
    "rtw_resolve_device": (C.c_int, [C.c_void_p, u64, u32, f32, C.c_int, C.c_void_p]),
    "rtw_quantize_device": (C.c_int, [C.c_void_p, C.c_void_p, u64, C.c_int, C.c_void_p]),
    "rtw_quantize": (C.c_int, [C.c_void_p, C.c_void_p, u64, C.c_int]),
    "rtw_render_rgb8": (C.c_int, [C.c_void_p, P(rtw_camera), P(rtw_render_opts), C.c_void_p]),
    "rtw_bilateral_filter_device": (C.c_int, [C.c_void_p, u32, u32, u32, C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p]),
    "rtw_bilateral_filter_device_noalloc": (C.c_int, [C.c_void_p, u32, u32, u32, C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p]),
    "rtw_bilateral_filter": (C.c_int, [C.c_void_p, u32, u32, u32, C.c_int, C.c_void_p, P(f32), P(f32), C.c_int]),
    "rtw_scene_multi_create": (C.c_int, [P(rtw_scene_desc), P(C.c_int), C.c_int, P(C.c_void_p)]),
    "rtw_scene_multi_destroy": (None, [C.c_void_p]),
    "rtw_render_multi": (C.c_int, [C.c_void_p, P(rtw_camera), P(rtw_render_opts), u32, u32, C.c_void_p]),
    "rtw_render_frames_multi": (C.c_int, [C.c_void_p, P(rtw_camera), P(rtw_render_opts), u32, u32, C.c_void_p]),
    "rtw_multi_get_stats": (C.c_int, [C.c_void_p, P(rtw_stats), P(rtw_stats), P(f32)]),
    "rtw_scene_multi_device_count": (C.c_int, [C.c_void_p]),
    "rtw_get_stats": (C.c_int, [C.c_void_p, P(rtw_stats)]),
    "rtw_get_sched_stats": (C.c_int, [C.c_void_p, P(u64)]),
    "rtw_microbench": (C.c_int, [C.c_int, C.c_int, C.c_int, P(C.c_double)]),
    "rtw_scene_get_info": (C.c_int, [C.c_void_p, P(rtw_scene_info)]),
}


class RtwError(RuntimeError):
    def __init__(self, code: int, msg: str):
        super().__init__(f"rtw error {code}: {msg}")
        self.code = code


_lib = None


def load_library(path: str | None = None):
    """dlopen librtw_b200.so and bind every declared symbol. Raises if it is missing — the product
    has no other implementation to fall back to."""
    global _lib
    if _lib is not None and path is None:
        return _lib
    p = path or LIB_PATH
    if not os.path.exists(p):
        raise ImportError(f"{p} not built: run `python -c 'import __graft_entry__ as g; g.build()'` "
                          "(nvcc, sm_100a). There is no CPU fallback.")
    lib = C.CDLL(p)
    for name, (res, args) in SYMBOLS.items():
        fn = getattr(lib, name)  # AttributeError if the .so does not export it
        fn.restype = res
        fn.argtypes = args
    if lib.rtw_abi_version() != RTW_ABI_VERSION:
        raise ImportError("librtw_b200.so ABI version mismatch")
    if path is None:
        _lib = lib
    return lib


def check(code: int):
    if code != RTW_OK:
        raise RtwError(code, load_library().rtw_last_error().decode("utf-8", "replace"))
