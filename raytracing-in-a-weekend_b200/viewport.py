"""Mirror of the reference's camera + render entry points (Rust/src/viewport.rs).

    Viewport::new / new_from_res   viewport.rs:308-428
    Viewport::render               viewport.rs:430-478   stratified, time = 0
    Viewport::render_no_rand       viewport.rs:479-516   one centre ray per pixel
    async_render                   viewport.rs:215-248   random jitter, lens, motion blur
    render_multi                   viewport.rs:249-269   frames start_frame..start_frame+n

The reference passes the integrator as a closure `Fn(Ray, &Scene, usize) -> Rgb<f32>`; a closure
cannot cross the ABI, so the shipped closures are named constants here:
ray_color_gradient, ray_color_bg_color, ray_color_d, plus the test closures ray_color_normal
(viewport.rs:524-554), ray_color_glass_test (glass_tests.rs:8-55) and ray_color_albedo.

`Img` is a numpy float32 array [height, width, 3], rows top to bottom, gamma already applied —
the reference's Vec<Vec<Rgb<f32>>> (viewport.rs:46).
"""
from __future__ import annotations

import copy
import ctypes as C
from typing import List, Optional

import numpy as np

from . import _abi as A
from .scene import Scene, Vec3, _v


class Integrator(int):
    """An RTW_INT_* code that prints like the reference function it stands for."""
    _names = {A.RTW_INT_GRADIENT: "ray_color_gradient", A.RTW_INT_D: "ray_color_d",
              A.RTW_INT_BG_COLOR: "ray_color_bg_color", A.RTW_INT_NORMAL: "ray_color_normal",
              A.RTW_INT_ALBEDO: "ray_color_albedo", A.RTW_INT_GLASS_TEST: "ray_color_glass_test",
              A.RTW_INT_LIGHT_CAST: "light_biased_ray_cast", A.RTW_INT_LIGHT_BIASED: "light_biased_ray_color"}

    def __repr__(self):
        return self._names.get(int(self), f"Integrator({int(self)})")


ray_color_gradient = Integrator(A.RTW_INT_GRADIENT)     # viewport/ray_color.rs:12-41
ray_color_d = Integrator(A.RTW_INT_D)                   # main.rs:17-46
ray_color_bg_color = Integrator(A.RTW_INT_BG_COLOR)     # viewport/ray_color.rs:43-92
ray_color_normal = Integrator(A.RTW_INT_NORMAL)         # viewport.rs:524-554
ray_color_albedo = Integrator(A.RTW_INT_ALBEDO)
ray_color_glass_test = Integrator(A.RTW_INT_GLASS_TEST)  # viewport/glass_tests.rs:8-55
light_biased_ray_cast = Integrator(A.RTW_INT_LIGHT_CAST)      # Rust2/src/viewport/ray_color.rs:55-109 (lights = Scene.lights)
light_biased_ray_color = Integrator(A.RTW_INT_LIGHT_BIASED)   # Rust2/src/viewport/ray_color.rs:111-164 (weight = Viewport.light_weight)

_default_backend = None


def set_default_backend(backend) -> None:
    global _default_backend
    _default_backend = backend


def get_default_backend():
    """The CUDA backend. Raises (ImportError / RtwError) if the library or a device is missing."""
    global _default_backend
    if _default_backend is None:
        from .backend import CudaBackend
        _default_backend = CudaBackend()
    return _default_backend


class Viewport:
    """viewport.rs:48-77. Public fields keep the reference's names; the private camera basis is
    derived by the backend's `camera_init` (a restatement of Viewport::new, viewport.rs:342-400)."""

    def __init__(self, width: int, aspect_ratio: float, samples: int, depth: int, gamma: float,
                 vfov: Optional[float] = None, origin=None, direction=None, vup=None,
                 msg: Optional[str] = None, lens_radius: Optional[float] = None):
        self.samples = int(samples)
        self.aspect_ratio = float(np.float32(aspect_ratio))
        self.width = int(width)
        # viewport.rs:346: height = (width as f32 / aspect_ratio) as u64
        self.height = int(np.float32(self.width) / np.float32(aspect_ratio))
        self.vfov = 90.0 if vfov is None else float(vfov)
        self.origin = Vec3(0.0, 0.0, 0.0) if origin is None else _v(origin)
        self.direction = Vec3(0.0, 0.0, -1.0) if direction is None else _v(direction)
        self.vup = Vec3(0.0, 1.0, 0.0) if vup is None else _v(vup)
        self.lens_radius = 0.0 if lens_radius is None else float(lens_radius)
        self.depth = int(depth)
        self.gamma = float(gamma)
        self.msg = msg or ""
        self.shutter_speed = 0.0     # viewport.rs:394-398 defaults
        self.fps = 30.0
        self.frame = 0
        self.number_of_frames = 1
        self.start_frame = 0
        self.seed = 1                # the reference is unseeded; SURVEY §8d "Render RNG"
        self.flags = 0               # RTW_FLAG_* (STRICT_FP for parity runs, WAVEFRONT)
        self.light_weight = 1.0      # `biased_weight` of light_biased_ray_color (Rust2 ray_color.rs:116)

    @staticmethod
    def new(width, aspect_ratio, samples, depth, gamma, vfov=None, origin=None, direction=None,
            vup=None, msg=None, lens_radius=None) -> "Viewport":
        return Viewport(width, aspect_ratio, samples, depth, gamma, vfov, origin, direction, vup, msg,
                        lens_radius)

    @staticmethod
    def new_from_res(width, height, samples, depth, gamma, vfov=None, origin=None, direction=None,
                     vup=None, msg=None, lens_radius=None) -> "Viewport":
        # viewport.rs:402-428: aspect = width as f32 / height as f32
        return Viewport(width, float(np.float32(width) / np.float32(height)), samples, depth, gamma, vfov,
                        origin, direction, vup, msg, lens_radius)

    # ---- ABI structs
    def camera(self, backend=None) -> A.rtw_camera:
        be = backend or get_default_backend()
        return be.camera_init(self.width, self.aspect_ratio, self.vfov, self.origin, self.direction,
                              self.vup, self.lens_radius)

    def opts(self, ray_color: int, mode: int) -> A.rtw_render_opts:
        o = A.rtw_render_opts()
        o.samples, o.depth, o.gamma = self.samples, self.depth, self.gamma
        o.integrator, o.mode = int(ray_color), mode
        o.shutter_speed, o.fps, o.frame = self.shutter_speed, self.fps, self.frame
        o.seed = self.seed
        o.mint, o.maxt = 0.0, 0.0
        o.sample_begin = o.sample_count = o.row_begin = o.row_count = 0
        o.output, o.flags = A.RTW_OUT_RESOLVED, int(self.flags)
        o.light_weight = float(self.light_weight)
        o.row_band = o.row_stride = 0
        return o

    # ---- the reference's render entry points
    def _run(self, ray_color, scene: Scene, mode: int, backend=None) -> np.ndarray:
        be = backend or get_default_backend()
        return be.scene(scene).render(self.camera(be), self.opts(ray_color, mode))

    def render(self, ray_color, scene: Scene, backend=None) -> np.ndarray:
        """Viewport::render (viewport.rs:430-478): ceil(sqrt(samples))^2 stratified samples."""
        return self._run(ray_color, scene, A.RTW_MODE_STRATIFIED, backend)

    def render_rgb8(self, ray_color, scene: Scene, mode: int = A.RTW_MODE_RANDOM, backend=None) -> np.ndarray:
        """A render followed by write_img_f32's quantiser (write_img.rs:6-19), fused on the device:
        the H x W x 3 uint8 image that `img.save` would encode."""
        be = backend or get_default_backend()
        return be.scene(scene).render_rgb8(self.camera(be), self.opts(ray_color, mode))

    def render_no_rand(self, ray_color, scene: Scene, backend=None) -> np.ndarray:
        """Viewport::render_no_rand (viewport.rs:479-516)."""
        return self._run(ray_color, scene, A.RTW_MODE_NO_RAND, backend)


def async_render(viewport: Viewport, ray_color, scene: Scene, backend=None) -> np.ndarray:
    """async_render (viewport.rs:215-248). Blocking here: the tokio row tasks become the GPU grid."""
    return viewport._run(ray_color, scene, A.RTW_MODE_RANDOM, backend)


def render_multi(viewport: Viewport, ray_color, scene: Scene, backend=None) -> List[np.ndarray]:
    """render_multi (viewport.rs:249-269): one Img per frame in start_frame..start_frame+number_of_frames."""
    be = backend or get_default_backend()
    h = be.scene(scene)
    o = viewport.opts(ray_color, A.RTW_MODE_RANDOM)
    frames = h.render_frames(viewport.camera(be), o, viewport.start_frame, viewport.number_of_frames)
    return [frames[k] for k in range(frames.shape[0])]
