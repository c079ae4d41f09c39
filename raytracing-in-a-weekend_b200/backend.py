"""CudaBackend: the product's only renderer — a thin ctypes layer over librtw_b200.so (include/rtw.h).

`torch` is optional here and only used for device buffers / streams when the caller wants the
image to stay on the GPU (multi-GPU reduce); plain renders go host buffer -> rtw_render.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _abi as A


def _f3(v):
    return (C.c_float * 3)(*[float(x) for x in v])


class CudaSceneHandle:
    def __init__(self, lib, fs, device: int):
        self._lib = lib
        self._fs = fs
        self._h = C.c_void_p()
        A.check(lib.rtw_scene_create(C.byref(fs.desc), device, C.byref(self._h)))
        self.device = device

    def close(self):
        if self._h:
            self._lib.rtw_scene_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    @staticmethod
    def out_shape(cam, o):
        rows = o.row_count if o.row_count else cam.height
        return (rows, cam.width, 3)

    def render(self, cam, o) -> np.ndarray:
        out = np.empty(self.out_shape(cam, o), dtype=np.float32)
        A.check(self._lib.rtw_render(self._h, C.byref(cam), C.byref(o), out.ctypes.data_as(C.c_void_p)))
        return out

    def render_into(self, cam, o, host_out: np.ndarray) -> None:
        """Render into a caller-owned host buffer (e.g. pinned memory)."""
        A.check(self._lib.rtw_render(self._h, C.byref(cam), C.byref(o), C.c_void_p(host_out.ctypes.data)))

    def render_rgb8(self, cam, o) -> np.ndarray:
        """Render + write_img_f32's quantiser fused on the device (write_img.rs:6-19): H x W x 3 uint8."""
        rows = o.row_count if o.row_count else cam.height
        out = np.empty((rows, cam.width, 3), dtype=np.uint8)
        A.check(self._lib.rtw_render_rgb8(self._h, C.byref(cam), C.byref(o), C.c_void_p(out.ctypes.data)))
        return out

    def render_device(self, cam, o, d_ptr: int, stream: int = 0) -> None:
        """Enqueue a render into device memory `d_ptr` on CUDA stream `stream` (no host sync)."""
        A.check(self._lib.rtw_render_device(self._h, C.byref(cam), C.byref(o), C.c_void_p(d_ptr),
                                            C.c_void_p(stream)))

    def render_frames(self, cam, o, first_frame: int, n_frames: int) -> np.ndarray:
        rows = o.row_count if o.row_count else cam.height
        out = np.empty((n_frames, rows, cam.width, 3), dtype=np.float32)
        A.check(self._lib.rtw_render_frames(self._h, C.byref(cam), C.byref(o), first_frame, n_frames,
                                            out.ctypes.data_as(C.c_void_p)))
        return out

    def rebuild_lbvh(self) -> float:
        """Rebuild every BVH level on the device (LBVH) and walk it from now on; returns the build's device ms."""
        ms = C.c_float(0.0)
        A.check(self._lib.rtw_scene_rebuild_lbvh(self._h, C.byref(ms)))
        return float(ms.value)

    def use_sah(self) -> None:
        """Back to the host SAH tree built at scene creation."""
        A.check(self._lib.rtw_scene_use_sah(self._h))

    def stats(self) -> A.rtw_stats:
        st = A.rtw_stats()
        A.check(self._lib.rtw_get_stats(self._h, C.byref(st)))
        return st

    def sched_stats(self) -> dict:
        """Warp-scheduling counters of the last COUNT_WORK | WAVEFRONT render (tuning aid)."""
        out = (C.c_uint64 * 14)()
        A.check(self._lib.rtw_get_sched_stats(self._h, out))
        kinds = ("node", "prim", "event", "fetch")
        d = {"iterations": int(out[0])}
        for k, name in enumerate(kinds):
            steps, lanes = int(out[1 + k]), int(out[5 + k])
            d[name] = {"steps": steps, "lanes": lanes, "lanes_per_step": lanes / steps if steps else 0.0,
                       "cycles": int(out[9 + k]), "cycles_per_step": int(out[9 + k]) / steps if steps else 0.0}
        d["loop_cycles"] = int(out[13])
        return d

    def info(self) -> A.rtw_scene_info:
        inf = A.rtw_scene_info()
        A.check(self._lib.rtw_scene_get_info(self._h, C.byref(inf)))
        return inf


class CudaMultiSceneHandle:
    """One scene replicated on a list of CUDA devices of THIS process (rtw_scene_multi, include/rtw.h): what the
    reference's single-process caller (main.rs:76-80,412-414) gets instead of one tokio runtime. A device may be
    listed more than once (its shares then run side by side on separate streams)."""

    def __init__(self, lib, fs, devices):
        self._lib, self._fs = lib, fs
        self.devices = [int(d) for d in devices]
        self._h = C.c_void_p()
        arr = (C.c_int * len(self.devices))(*self.devices)
        A.check(lib.rtw_scene_multi_create(C.byref(fs.desc), arr, len(self.devices), C.byref(self._h)))

    def close(self):
        if self._h:
            self._lib.rtw_scene_multi_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def render(self, cam, o, split: int = A.RTW_SPLIT_SAMPLES, band_rows: int = 0, out: np.ndarray | None = None) -> np.ndarray:
        """async_render / Viewport::render over every device: sample ranges (one peer-memory combine on devices[0])
        or round-robin row bands (no exchange)."""
        if out is None:
            out = np.empty(CudaSceneHandle.out_shape(cam, o), dtype=np.float32)
        A.check(self._lib.rtw_render_multi(self._h, C.byref(cam), C.byref(o), split, band_rows, C.c_void_p(out.ctypes.data)))
        return out

    def render_frames(self, cam, o, first_frame: int, n_frames: int) -> np.ndarray:
        """render_multi (viewport.rs:249-269), frame k on devices[k mod n]."""
        rows = o.row_count if o.row_count else cam.height
        out = np.empty((n_frames, rows, cam.width, 3), dtype=np.float32)
        A.check(self._lib.rtw_render_frames_multi(self._h, C.byref(cam), C.byref(o), first_frame, n_frames,
                                                  C.c_void_p(out.ctypes.data)))
        return out

    def stats(self):
        """(totals, per-share list, combine kernel ms) of the last multi render."""
        total = A.rtw_stats()
        per = (A.rtw_stats * len(self.devices))()
        ms = C.c_float(0.0)
        A.check(self._lib.rtw_multi_get_stats(self._h, C.byref(total), per, C.byref(ms)))
        return total, list(per), float(ms.value)


class CudaBackend:
    name = "cuda"

    def __init__(self, device: int = -1):
        self.lib = A.load_library()
        self.device = device

    def device_count(self) -> int:
        return int(self.lib.rtw_device_count())

    def camera_init(self, width, aspect_ratio, vfov, origin, direction, vup, lens_radius) -> A.rtw_camera:
        cam = A.rtw_camera()
        A.check(self.lib.rtw_camera_init(C.byref(cam), width, aspect_ratio, vfov, _f3(origin), _f3(direction),
                                         _f3(vup), lens_radius))
        return cam

    def scene(self, scene) -> CudaSceneHandle:
        key = ("cuda", self.device)
        h = scene._backend_cache.get(key)
        if h is None:
            h = CudaSceneHandle(self.lib, scene.flatten(), self.device)
            scene._backend_cache[key] = h
        return h

    def scene_multi(self, scene, devices) -> CudaMultiSceneHandle:
        key = ("cuda-multi", tuple(int(d) for d in devices))
        h = scene._backend_cache.get(key)
        if h is None:
            h = CudaMultiSceneHandle(self.lib, scene.flatten(), devices)
            scene._backend_cache[key] = h
        return h

    def trim_memory(self) -> None:
        """Give the device / pinned blocks cached from destroyed scenes back to the driver."""
        A.check(self.lib.rtw_trim_memory(self.device))

    def microbench(self, which: int, reps: int = 3) -> float:
        """FP32 TFLOP/s / L2 GB/s / HBM GB/s of this device (roofline denominators)."""
        v = C.c_double()
        A.check(self.lib.rtw_microbench(self.device, which, reps, C.byref(v)))
        return float(v.value)

    def resolve_device(self, d_ptr: int, n_pixels: int, total_samples: int, gamma: float, stream: int = 0):
        A.check(self.lib.rtw_resolve_device(C.c_void_p(d_ptr), n_pixels, total_samples, gamma, self.device,
                                            C.c_void_p(stream)))

    def quantize(self, img: np.ndarray) -> np.ndarray:
        """write_img_f32's f32 -> u8 conversion (write_img.rs:12-14) of a host image, on the device."""
        src = np.ascontiguousarray(img, dtype=np.float32)
        out = np.empty(src.shape, dtype=np.uint8)
        A.check(self.lib.rtw_quantize(C.c_void_p(src.ctypes.data), C.c_void_p(out.ctypes.data), src.size, self.device))
        return out

    def bilateral_filter(self, img_u8: np.ndarray, size: int, kind: int = 0, info: dict | None = None) -> np.ndarray:
        """Rust2/src/postprocessing.rs:70-135 on the device: H x W x 3 uint8 in and out. kind 0 = Square, 1 = Edges.
        `info` (optional dict) receives spatial / avg_gradient / inv_range / inv_spatial / kernel_ms."""
        src = np.ascontiguousarray(img_u8, dtype=np.uint8)
        if src.ndim != 3 or src.shape[2] != 3:
            raise ValueError("expected an H x W x 3 uint8 image")
        out = np.empty_like(src)
        par = (C.c_float * 4)()
        ms = C.c_float()
        A.check(self.lib.rtw_bilateral_filter(C.c_void_p(src.ctypes.data), src.shape[1], src.shape[0], int(size), int(kind),
                                              C.c_void_p(out.ctypes.data), par, C.byref(ms), self.device))
        if info is not None:
            info.update(spatial=par[0], avg_gradient=par[1], inv_range=par[2], inv_spatial=par[3], kernel_ms=ms.value)
        return out

    def bilateral_filter_device(self, d_in: int, width: int, height: int, size: int, kind: int, d_out: int,
                                d_params: int = 0, stream: int = 0):
        A.check(self.lib.rtw_bilateral_filter_device(C.c_void_p(d_in), width, height, int(size), int(kind), C.c_void_p(d_out),
                                                     C.c_void_p(d_params), self.device, C.c_void_p(stream)))

    def quantize_device(self, d_rgb: int, d_rgb8: int, n_values: int, stream: int = 0):
        A.check(self.lib.rtw_quantize_device(C.c_void_p(d_rgb), C.c_void_p(d_rgb8), n_values, self.device,
                                             C.c_void_p(stream)))
