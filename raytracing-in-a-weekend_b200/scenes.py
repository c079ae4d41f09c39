"""Scenes authored against the mirrored reference API.

Two groups:
  * the reference's own test scenes — every render test of Rust/src is re-created here
    with its exact parameters, cited per function;
  * the five BASELINE.json configs C1..C5, which do not exist as code in the reference and are
    authored against its API exactly as SURVEY §8(d) specifies.

Each builder returns a `Job`: scene, viewport, integrator and which reference entry point to use.
Scene-building randomness is a Philox stream (key 2024) so every rank and every backend sees the
same scene. Image textures are the reference's own assets (Rust/assets/earthmap.jpg 1024x512,
default.png 128x64, squares.png), decoded once in the build container by tests/golden/make_textures.py
and committed as lossless PNGs under tests/golden/assets/ (RTW_ASSET_DIR overrides); they enter the
scene through ImageTexture.from_path, as in texture_test.rs:63,114. Nothing under /root/reference is
read at run time. `SYNTHETIC_TEXTURES = True` (bench.py --synthetic-textures) swaps in generated
stand-ins of the same sizes.
"""
from __future__ import annotations

import math
import os
from dataclasses import dataclass
from typing import Callable, Dict

import numpy as np

from . import _abi as A
from .philox import PhiloxStream
from .scene import (EMPTY_M, FUZZY3_M, GLASS_M, GLASSR_M, METALLIC_M, SCATTER_M, ImageTexture, Instance,
                    Material, PerlinNoise, Quad, Scene, Sphere, Triangle, Vec3, const_density)
from .viewport import (Viewport, light_biased_ray_cast, light_biased_ray_color, ray_color_bg_color, ray_color_d,
                       ray_color_glass_test, ray_color_gradient, ray_color_normal)

F = np.float32
PI = float(F(math.pi))
Z = (0.0, 0.0, 0.0)


@dataclass
class Job:
    name: str
    scene: Scene
    viewport: Viewport
    ray_color: int
    mode: int            # RTW_MODE_*: which reference loop the test/config used
    note: str = ""

    def render(self, backend=None) -> np.ndarray:
        return self.viewport._run(self.ray_color, self.scene, self.mode, backend)


# --------------------------------------------------------------------------- textures
ASSET_DIR = os.environ.get("RTW_ASSET_DIR") or os.path.join(
    os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden", "assets")
SYNTHETIC_TEXTURES = False     # True: generated stand-ins instead of the reference's assets
_asset_cache = {}


def asset_path(name: str) -> str:
    """Path of the committed lossless copy of Rust/assets/<name>.{jpg,png} (tests/golden/make_textures.py)."""
    return os.path.join(ASSET_DIR, name + ".png")


def _asset_texture(name: str) -> ImageTexture:
    """ImageTexture::from_path("assets/...") (texture.rs:215-226), decoded once per process."""
    if name not in _asset_cache:
        path = asset_path(name)
        if not os.path.exists(path):
            raise FileNotFoundError(f"{path}: run tests/golden/make_textures.py where /root/reference exists, "
                                    "or set scenes.SYNTHETIC_TEXTURES = True")
        _asset_cache[name] = ImageTexture.from_path(path)
    return _asset_cache[name]


def texture_data_label() -> str:
    return ("synthetic scene + synthetic textures" if SYNTHETIC_TEXTURES else
            "synthetic scene (authored per SURVEY 8d), reference texture assets (earthmap.jpg, default.png)")


def checker_texture(width=512, height=256, cells_x=32, cells_y=16, c0=(0.2, 0.3, 0.1), c1=(0.9, 0.9, 0.9)) -> ImageTexture:
    """A checker is just an image in the reference; C3's generated 512x256 one."""
    x = (np.arange(width) * cells_x // width)[None, :]
    y = (np.arange(height) * cells_y // height)[:, None]
    m = ((x + y) & 1).astype(bool)
    img = np.where(m[..., None], np.array(c1, dtype=F), np.array(c0, dtype=F)).astype(F)
    return ImageTexture.new(img.reshape(-1, 3), width, height)


def default_png_texture(synthetic=None) -> ImageTexture:
    """Rust/assets/default.png (128x64 magenta/black checker); synthetic: a generated one of that size."""
    if not (SYNTHETIC_TEXTURES if synthetic is None else synthetic):
        return _asset_texture("default")
    return checker_texture(128, 64, 16, 8, (0.0, 0.0, 0.0), (1.0, 0.0, 1.0))


def squares_png_texture() -> ImageTexture:
    """Rust/assets/squares.png (128x64 colour grid)."""
    return _asset_texture("squares")


_earth_cache = {}


def earth_texture(width=1024, height=512, synthetic=None) -> ImageTexture:
    """Rust/assets/earthmap.jpg through from_path (texture_test.rs:63,114). synthetic: a stand-in of the
    same size (1024x512 RGB, u8/255 texels): a few octaves of lattice value-noise thresholded into
    sea / land / ice."""
    if not (SYNTHETIC_TEXTURES if synthetic is None else synthetic):
        return _asset_texture("earthmap")
    key = (width, height)
    if key in _earth_cache:
        return _earth_cache[key]
    rs = PhiloxStream(0xEA27)
    acc = np.zeros((height, width), dtype=np.float64)
    amp = 1.0
    for octave in range(5):
        n = 4 << octave
        lat = rs.uniform_f32((n + 1) * (2 * n + 1)).reshape(n + 1, 2 * n + 1).astype(np.float64)
        lat[:, -1] = lat[:, 0]                      # wrap in longitude
        ys = np.linspace(0, n, height, endpoint=False)
        xs = np.linspace(0, 2 * n, width, endpoint=False)
        y0, x0 = ys.astype(int), xs.astype(int)
        fy, fx = (ys - y0)[:, None], (xs - x0)[None, :]
        fy, fx = fy * fy * (3 - 2 * fy), fx * fx * (3 - 2 * fx)
        a = lat[y0][:, x0]; b = lat[y0][:, x0 + 1]; c = lat[y0 + 1][:, x0]; d = lat[y0 + 1][:, x0 + 1]
        acc += amp * ((a * (1 - fx) + b * fx) * (1 - fy) + (c * (1 - fx) + d * fx) * fy)
        amp *= 0.5
    acc /= acc.max()
    lat_abs = np.abs(np.linspace(-1, 1, height))[:, None]
    img = np.zeros((height, width, 3), dtype=np.float64)
    sea = acc < 0.52
    img[sea] = np.stack([0.02 + 0 * acc, 0.10 + 0.2 * acc, 0.30 + 0.5 * acc], -1)[sea]
    land = ~sea
    img[land] = np.stack([0.15 + 0.5 * (acc - 0.5), 0.35 + 0.2 * acc, 0.08 + 0.1 * acc], -1)[land]
    ice = (lat_abs + 0.15 * acc) > 0.93
    img[np.broadcast_to(ice, acc.shape)] = (0.92, 0.94, 0.96)
    u8 = np.clip(np.round(img * 255.0), 0, 255).astype(np.uint8)
    tex = ImageTexture.new((u8.astype(F) / F(255.0)).reshape(-1, 3), width, height)
    _earth_cache[key] = tex
    return tex


def _solid(c) -> ImageTexture:
    return ImageTexture.from_color(c)


def _albedo_sphere(origin, r, albedo, mat) -> Sphere:
    """Exact-albedo sphere: new_with_texture(.., None, ..) avoids the c*c of Sphere::new."""
    return Sphere.new_with_texture(origin, r, None, mat, _solid(albedo))


# --------------------------------------------------------------------------- reference test scenes
def viewport_object_test(width=800, height=600, samples=100) -> Job:
    """viewport.rs:556-619 `test_viewport_object`: four default spheres, normal shading."""
    spheres = [Sphere.new(Vec3(-0.50, 0.0, -1.0), 0.5), Sphere.new(Vec3(0.50, 0.0, -1.0), 0.5),
               Sphere.new(Vec3(0.0, 0.0, -2.0), 1.0), Sphere.new(Vec3(0.0, 0.0, 0.0), 1.0)]
    vp = Viewport.new_from_res(width, height, samples, 10, 2.0, msg="Viewport object test")
    return Job("viewport_object", Scene.new_sphere(spheres), vp, ray_color_normal, A.RTW_MODE_STRATIFIED)


def camera_test_scene() -> Scene:
    """viewport/camera_tests.rs:56-89."""
    return Scene.new_sphere([
        Sphere.new(Vec3(-0.5, 0.0, -1.0), 0.5, Vec3(0.6, 0.6, 0.6), SCATTER_M),
        Sphere.new(Vec3(0.5, 0.0, -1.0), 0.5, Vec3(1.0, 1.0, 1.0), SCATTER_M),
        Sphere.new(Vec3(0.0, 0.0, -2.0), 1.0, Vec3(0.5, 1.0, 0.0), METALLIC_M)])


def camera_test(which="default", width=400, height=300, samples=100) -> Job:
    """viewport/camera_tests.rs:91-178: default / fov_120 / upside_down / depth_of_field."""
    kw = {"default": {}, "fov_120": {"vfov": 120.0}, "upside_down": {"vup": Vec3(0.0, -1.0, 0.0)},
          "depth_of_field": {"vup": Vec3(0.0, 1.0, 0.0), "lens_radius": 0.015}}[which]
    vp = Viewport.new_from_res(width, height, samples, 10, 2.0, **kw)
    return Job(f"camera_{which}", camera_test_scene(), vp, ray_color_d, A.RTW_MODE_STRATIFIED)


def diffuse_test(width=800, height=600, samples=100) -> Job:
    """viewport/material_tests.rs:50-95."""
    sc = Scene.new_sphere([Sphere.new(Vec3(-0.5, 0.0, -1.0), 0.5, Vec3(0.6, 0.6, 0.6), SCATTER_M),
                           Sphere.new(Vec3(0.5, 0.0, -1.0), 0.5, Vec3(1.0, 1.0, 1.0), SCATTER_M),
                           Sphere.new(Vec3(0.0, 0.0, -2.0), 1.0, Vec3(0.5, 1.0, 0.0), SCATTER_M)])
    return Job("diffuse", sc, Viewport.new_from_res(width, height, samples, 10, 2.0), ray_color_d, A.RTW_MODE_STRATIFIED)


def metal_test(width=400, height=225, samples=100) -> Job:
    """viewport/material_tests.rs:97-150."""
    sc = Scene.new_sphere([Sphere.new(Vec3(-1.0, 0.0, -1.0), 0.5, Vec3(0.8, 0.8, 0.8), FUZZY3_M),
                           Sphere.new(Vec3(1.0, 0.0, -1.0), 0.5, Vec3(0.8, 0.6, 0.2), METALLIC_M),
                           Sphere.new(Vec3(0.0, 0.0, -1.0), 0.5, Vec3(0.7, 0.30, 0.30), SCATTER_M),
                           Sphere.new(Vec3(0.0, -100.5, -1.0), 100.0, Vec3(0.8, 0.8, 0.0), SCATTER_M)])
    return Job("metal", sc, Viewport.new_from_res(width, height, samples, 10, 2.0), ray_color_d, A.RTW_MODE_STRATIFIED)


def hollow_glass_test(width=300, height=200, samples=100) -> Job:
    """viewport/material_tests.rs:196-269: glass r 0.5 + inverted-ir r 0.35 bubble."""
    sc = Scene.new_sphere([Sphere.new(Vec3(0.0, 0.0, -1.0), 0.5, Vec3(1.0, 1.0, 1.0), GLASS_M),
                           Sphere.new(Vec3(0.0, 0.0, -1.0), 0.35, Vec3(1.0, 1.0, 1.0), GLASSR_M),
                           Sphere.new(Vec3(0.0, -100.5, -1.0), 100.0, Vec3(0.8, 0.5, 1.0), EMPTY_M)])
    return Job("hollow_glass", sc, Viewport.new_from_res(width, height, samples, 10, 2.0), ray_color_d, A.RTW_MODE_STRATIFIED)


def glass_test_10x10(control=True) -> Job:
    """viewport/glass_tests.rs:57-144: the deterministic 10x10 render_no_rand pair that the
    reference's (ignored) `runner` compares pixel-for-pixel with the C++ tree (glass_tests.rs:146-193,
    C++/src/tests.cpp:275-294)."""
    if control:
        sp = [Sphere.new(Vec3(0.0, 0.0, -1.0), 0.5, Vec3(1.0, 1.0, 1.0), METALLIC_M),
              Sphere.new(Vec3(0.0, -100.5, -1.0), 100.0, Vec3(0.8, 0.5, 1.0), EMPTY_M)]
    else:
        sp = [Sphere.new(Vec3(0.0, 0.0, -1.0), 0.5, Vec3(1.0, 1.0, 1.0), GLASS_M),
              Sphere.new(Vec3(0.0, -100.5, -1.0), 100.0, Vec3(1.0, 1.0, 1.0), EMPTY_M)]
    vp = Viewport.new_from_res(10, 10, 100, 10, 2.0)
    return Job("glass10_control" if control else "glass10", Scene.new_sphere(sp), vp, ray_color_glass_test,
               A.RTW_MODE_NO_RAND)


def noise_test(width=400, height=225, samples=100) -> Job:
    """viewport/texture_test.rs:131-196 (the closure there uses scene.aabb with mint 1e-4 and nudges
    the origin; ray_color_d is the shipped equivalent)."""
    sc = Scene.new_sphere([
        Sphere.new(Vec3(-1.0, 0.0, -1.0), 0.5, Vec3(0.8, 0.8, 0.8), METALLIC_M),
        Sphere.new(Vec3(1.0, 0.0, -1.0), 0.5, Vec3(0.8, 0.6, 0.2), METALLIC_M),
        Sphere.new_with_texture(Vec3(0.0, -100.5, -1.0), 100.0, Vec3(1.0, 1.0, 1.0), SCATTER_M,
                                ImageTexture.from_color_noise(Vec3(1.0, 1.0, 1.0), 0.01, PerlinNoise.new(11))),
        Sphere.new(Vec3(0.0, 0.0, -1.0), 0.5, Vec3(0.8, 0.8, 0.0), SCATTER_M)])
    return Job("noise", sc, Viewport.new_from_res(width, height, samples, 10, 2.0), ray_color_d, A.RTW_MODE_STRATIFIED)


def texture_test(width=2000, height=2000, samples=100) -> Job:
    """viewport/texture_test.rs:45-86 `default_test`: a radius-0.05 mirror sphere carrying
    assets/earthmap.jpg, seen through a 7 degree lens from the origin, async_render. (The closure there is
    the file's own ray_color_d with mint 1e-4 and an origin nudge; ray_color_d is the shipped equivalent.)"""
    sc = Scene.new_sphere([Sphere.new_with_texture(Vec3(-1.0, 0.0, -0.0), 0.05, Vec3(1.0, 1.0, 1.0), METALLIC_M,
                                                   earth_texture())])
    vp = Viewport.new_from_res(width, height, samples, 10, 2.0, 7.0, None, Vec3(-1.0, 0.0, -0.0), None, "texture test")
    return Job("texture", sc, vp, ray_color_d, A.RTW_MODE_RANDOM)


def reflection_test(width=400, height=300, samples=100) -> Job:
    """viewport/texture_test.rs:88-129: mirror ball reflecting a radius-1000 earth-mapped sphere."""
    sc = Scene.new_sphere([
        Sphere.new(Vec3(0.520, 0.0, -1.0), 0.45, Vec3(0.95, 0.95, 0.95), METALLIC_M),
        Sphere.new_with_texture(Vec3(-1001.0, 0.0, 0.0), 1000.0, None, METALLIC_M, earth_texture())])
    vp = Viewport.new_from_res(width, height, samples, 10, 2.0, 90.0, Vec3(0.0, 0.0, 0.0))
    return Job("reflection", sc, vp, ray_color_d, A.RTW_MODE_STRATIFIED)


def _room_quads(light_quad: Quad, extra=()):
    """The four walls shared by instance.rs:313-600 / main.rs:107-300 scenes."""
    return [
        Quad.new(Vec3(-2.0, -2.0, 5.0), Vec3(0.0, 0.0, -4.0), Vec3(0.0, 4.0, 0.0), METALLIC_M, Z, _solid((0.85, 0.85, 0.85))),
        Quad.new(Vec3(2.0, -2.0, 1.0), Vec3(0.0, 0.0, 4.0), Vec3(0.0, 4.0, 0.0), SCATTER_M, Z, _solid((0.2, 0.2, 1.0))),
        Quad.new(Vec3(-2.0, 2.0, 1.0), Vec3(4.0, 0.0, 0.0), Vec3(0.0, 0.0, 4.0), SCATTER_M, Z, _solid((1.0, 0.5, 0.0))),
        Quad.new(Vec3(-2.0, -2.0, 5.0), Vec3(4.0, 0.0, 0.0), Vec3(0.0, 0.0, -4.0), SCATTER_M, Z, _solid((0.2, 0.8, 0.8))),
        light_quad, *extra]


def _room_viewport(samples, depth=20, size=400):
    return Viewport.new_from_res(size, size, samples, depth, 2.0, 90.0, Vec3(0.0, -0.0, 7.0), Vec3(0.0, 0.0, -1.0))


def instance_box_test(rotated=True, samples=100, size=400) -> Job:
    """objects/instance.rs:322-520 `box_test`: translated (and pi/4 y-rotated) unit box in the room."""
    light = Quad.new(Vec3(-0.5, -0.5, 1.0), Vec3(1.0, 0.0, 0.0), Vec3(0.0, 1.0, 0.0),
                     Material.new_emmiting(0.0, 0.0, 1.0, Vec3(4.0, 4.0, 4.0)), Z, _solid((0.2, 1.0, 0.2)))
    back = Quad.new(Vec3(-2.0, -2.0, 0.999), Vec3(4.0, 0.0, 0.0), Vec3(0.0, 4.0, 0.0),
                    Material.new_emmiting(0.0, 0.0, 1.0, Z), Z, _solid((0.2, 1.0, 0.2)))
    inst = Instance.new_box(Vec3(-0.5, -0.5, -0.5), Vec3(0.5, 0.5, 0.5), _solid((0.0, 0.5, 0.0)), SCATTER_M)
    inst.translate(Vec3(-0.0, -1.5, 2.0))
    if rotated:
        inst.rotate(Vec3(0.0, PI / 4.0, 0.0))
    sc = Scene.new([], _room_quads(light, [back]), [inst])
    return Job("instance_box" + ("_rot" if rotated else ""), sc, _room_viewport(samples, 20, size), ray_color_bg_color,
               A.RTW_MODE_RANDOM)


def volume_test(samples=100, size=400) -> Job:
    """objects/instance.rs:602-880 `volume_test`: const-density 2.0 smoke box rotated pi/4."""
    light = Quad.new(Vec3(-2.0, -2.0, 1.0), Vec3(4.0, 0.0, 0.0), Vec3(0.0, 4.0, 0.0),
                     Material.new_emmiting(0.0, 0.0, 1.0, Vec3(4.0, 4.0, 4.0)), Z, _solid((0.2, 1.0, 0.2)))
    smoke = Instance.new_box(Vec3(-1.5, -0.5, -0.5), Vec3(1.5, 0.5, 0.5), _solid((0.2, 0.2, 0.2)), SCATTER_M)
    smoke.translate(Vec3(-0.0, -1.5, 2.0))
    smoke.rotate(Vec3(0.0, PI / 4.0, 0.0))
    smoke.dist_fn = const_density
    smoke.density = 2.0
    sc = Scene.new([], _room_quads(light), [smoke])
    return Job("volume", sc, _room_viewport(samples, 20, size), ray_color_bg_color, A.RTW_MODE_RANDOM)


def presentation_image(samples=2500, size=400) -> Job:
    """main.rs:89-419 `presentation_image()` — what `cargo run` renders into Rust/Presentation.png."""
    spheres = [Sphere.new(Vec3(-1.6, -1.6, 3.0), 0.4, Vec3(1.0, 0.0, 0.0), SCATTER_M)]
    light = Quad.new(Vec3(-1.5, -1.5, 1.001), Vec3(3.0, 0.0, 0.0), Vec3(0.0, 3.0, 0.0),
                     Material.new_emmiting(0.0, 0.0, 1.0, Vec3(4.0, 4.0, 4.0)), Z, _solid((1.0, 1.0, 1.0)))
    green = Quad.new(Vec3(-2.0, -2.0, 1.0), Vec3(4.0, 0.0, 0.0), Vec3(0.0, 4.0, 0.0),
                     Material.new_emmiting(0.0, 0.0, 1.0, Z), Z, _solid((0.2, 1.0, 0.2)))
    # main.rs:107-300 lists Red(metal), Light, Blue, Orange, Teal, Green
    walls = _room_quads(light)
    quads = [walls[0], light, walls[1], walls[2], walls[3], green]
    smoke = Instance.new_box(Vec3(-1.0, -0.5, -0.5), Vec3(1.0, 0.5, 0.5), _solid((0.2, 0.2, 0.2)), SCATTER_M)
    glass_color = _solid((0.8, 0.8, 0.8))
    glass = Instance.new_quads([
        Quad.new(Vec3(0.0, 0.0, 0.0), Vec3(0.0, 4.0, 0.0), Vec3(2.0, 0.0, 0.0), Material(1.0, 0.8, 1.50), Z, glass_color),
        Quad.new(Vec3(0.0, 0.0, -0.001), Vec3(0.0, 4.0, 0.0), Vec3(2.0, 0.0, 0.0),
                 Material(1.0, 0.8, float(F(2.0) / F(3.0))), Z, glass_color)])
    smoke.translate(Vec3(-0.0, -1.5, 2.0))
    smoke.rotate(Vec3(0.0, PI / 4.0, 0.0))
    smoke.dist_fn = const_density
    smoke.density = 2.0
    glass.translate(Vec3(float(F(2.0) - np.sqrt(F(3.0))), -2.0, 3.0))
    glass.rotate(Vec3(0.0, -PI / 6.0, 0.0))
    sc = Scene.new(spheres, quads, [smoke, glass])
    return Job("presentation", sc, _room_viewport(samples, 20, size), ray_color_bg_color, A.RTW_MODE_RANDOM)


def first_frame(samples=100, size=400) -> Job:
    """main.rs:427-521 `_scene` + viewport (fps 15, shutter 0) -> Rust/First frame.png, ray_color_d."""
    spheres = [
        Sphere.new(Vec3(-0.8, 0.0, -1.0), 0.4, Vec3(1.0, 0.6, 0.6), FUZZY3_M),
        Sphere.new(Vec3(0.6, 0.0, -1.2), 0.3, Vec3(0.5, 0.9, 0.9), METALLIC_M),
        Sphere.new_moving(Vec3(2.4, 0.0, -0.8), 1.4, Vec3(0.9, 0.9, 0.9), METALLIC_M, Vec3(0.0, 60.0, 0.0)),
        Sphere.new(Vec3(0.0, 0.0, -0.7), 0.3, Vec3(1.0, 1.0, 1.0), GLASS_M),
        Sphere.new(Vec3(0.0, 0.0, -0.7), 0.2, Vec3(1.0, 1.0, 1.0), GLASSR_M),
        Sphere.new(Vec3(0.0, 0.0, -2.0), 1.0, Vec3(0.5, 1.0, 0.0), SCATTER_M),
        Sphere.new(Vec3(0.0, -1000.9, -5.0), 1000.0, Vec3(0.8, 0.5, 1.0), SCATTER_M)]
    vp = Viewport.new_from_res(size, size, samples, 100, 2.0)
    vp.fps, vp.shutter_speed = 15.0, 0.0
    return Job("first_frame", Scene.new_sphere(spheres), vp, ray_color_d, A.RTW_MODE_RANDOM)


def ltr_scene(samples=100, width=1280, height=720, frames=20) -> Job:
    """main.rs:524-553 `ltr_scene`: the animation behind Rust/video/img000NN.png (fps 15). The exact
    viewport of the shipped frames is not in the tree."""
    spheres = [Sphere.new_moving(Vec3(-1.0, 0.0, -1.0), 0.4, Vec3(0.9, 0.9, 0.9), METALLIC_M, Vec3(1.0, 0.0, 0.0)),
               Sphere.new_moving(Vec3(0.0, 1.0, -0.9), 0.5, Vec3(1.0, 0.0, 1.0), SCATTER_M, Vec3(0.0, 0.0, 0.0))]
    vp = Viewport.new_from_res(width, height, samples, 100, 2.0)
    vp.fps, vp.shutter_speed, vp.number_of_frames = 15.0, 0.0, frames
    return Job("ltr", Scene.new_sphere(spheres), vp, ray_color_d, A.RTW_MODE_RANDOM)


# --------------------------------------------------------------------------- BASELINE configs
def config_c1(width=400, height=225, samples=100, depth=50) -> Job:
    """C1: Book-1 three spheres + ground, glass with inverted-ir bubble."""
    sc = Scene.new_sphere([
        _albedo_sphere(Vec3(0.0, -100.5, -1.0), 100.0, (0.8, 0.8, 0.0), SCATTER_M),
        _albedo_sphere(Vec3(0.0, 0.0, -1.2), 0.5, (0.1, 0.2, 0.5), SCATTER_M),
        _albedo_sphere(Vec3(-1.0, 0.0, -1.0), 0.5, (1.0, 1.0, 1.0), GLASS_M),
        _albedo_sphere(Vec3(-1.0, 0.0, -1.0), 0.4, (1.0, 1.0, 1.0), GLASSR_M),
        _albedo_sphere(Vec3(1.0, 0.0, -1.0), 0.5, (0.8, 0.6, 0.2), Material.new_m(1.0))])
    vp = Viewport.new_from_res(width, height, samples, depth, 2.0, 90.0, Vec3(0.0, 0.0, 0.0), Vec3(0.0, 0.0, -1.0),
                               Vec3(0.0, 1.0, 0.0), None, 0.0)
    return Job("C1", sc, vp, ray_color_gradient, A.RTW_MODE_RANDOM, "Book-1 three spheres 400x225 100spp d50")


def _unit(v):
    v = np.asarray(v, dtype=F)
    return (v / np.sqrt((v * v).sum(dtype=F), dtype=F)).astype(F)


def config_c2(width=1200, height=675, samples=500, depth=50) -> Job:
    """C2: Book-1 cover, ~488 spheres, lens radius 0.0524."""
    rs = PhiloxStream(2024)
    spheres = [_albedo_sphere(Vec3(0.0, -1000.0, 0.0), 1000.0, (0.5, 0.5, 0.5), SCATTER_M)]
    for a in range(-11, 11):
        for b in range(-11, 11):
            choose = rs.random()
            c = Vec3(a + 0.9 * rs.random(), 0.2, b + 0.9 * rs.random())
            d = c - Vec3(4.0, 0.2, 0.0)
            if float(np.sqrt((d * d).sum())) <= 0.9:
                continue
            if choose < 0.8:
                alb = rs.uniform_f32(3) * rs.uniform_f32(3)
                spheres.append(_albedo_sphere(c, 0.2, alb, SCATTER_M))
            elif choose < 0.95:
                alb = [rs.uniform(0.5, 1.0) for _ in range(3)]
                spheres.append(_albedo_sphere(c, 0.2, alb, Material.new_m(1.0 - rs.uniform(0.0, 0.5))))
            else:
                spheres.append(_albedo_sphere(c, 0.2, (1.0, 1.0, 1.0), GLASS_M))
    spheres += [_albedo_sphere(Vec3(0.0, 1.0, 0.0), 1.0, (1.0, 1.0, 1.0), GLASS_M),
                _albedo_sphere(Vec3(-4.0, 1.0, 0.0), 1.0, (0.4, 0.2, 0.1), SCATTER_M),
                _albedo_sphere(Vec3(4.0, 1.0, 0.0), 1.0, (0.7, 0.6, 0.5), Material.new_m(1.0))]
    origin = Vec3(13.0, 2.0, 3.0)
    vp = Viewport.new_from_res(width, height, samples, depth, 2.0, 20.0, origin, _unit(-origin), Vec3(0.0, 1.0, 0.0),
                               None, 0.0524)
    return Job("C2", Scene.new_sphere(spheres), vp, ray_color_gradient, A.RTW_MODE_RANDOM,
               "Book-1 cover random spheres 1200x675 500spp d50")


def config_c3(width=800, height=450, samples=1000, depth=50) -> Job:
    """C3: checker ground + Perlin sphere + earth sphere + checker sphere."""
    sc = Scene.new_sphere([
        Sphere.new_with_texture(Vec3(0.0, -1000.0, 0.0), 1000.0, None, SCATTER_M, checker_texture()),
        Sphere.new_with_texture(Vec3(0.0, 2.0, 0.0), 2.0, None, SCATTER_M,
                                ImageTexture.from_color_noise(Vec3(1.0, 1.0, 1.0), 0.25, PerlinNoise.new(7))),
        Sphere.new_with_texture(Vec3(-4.0, 2.0, 2.0), 2.0, None, SCATTER_M, earth_texture()),
        Sphere.new_with_texture(Vec3(4.0, 2.0, -3.0), 2.0, None, SCATTER_M, default_png_texture())])
    origin = Vec3(13.0, 2.0, 3.0)
    vp = Viewport.new_from_res(width, height, samples, depth, 2.0, 20.0, origin, _unit(Vec3(0.0, 1.0, 0.0) - origin),
                               Vec3(0.0, 1.0, 0.0), None, 0.0)
    return Job("C3", sc, vp, ray_color_gradient, A.RTW_MODE_RANDOM, "checker+Perlin+earth 800x450 1000spp d50")


def config_c4(width=600, height=600, samples=5000, depth=50, smoke=False) -> Job:
    """C4: Book-2 Cornell box, two rotated/translated box instances (+ smoke variant)."""
    white, red, green = _solid((0.73, 0.73, 0.73)), _solid((0.65, 0.05, 0.05)), _solid((0.12, 0.45, 0.15))
    # The reference never flips the geometric normal u x v (quad.rs:64) and scatters about
    # it on either side, so a room only lights up when every wall's u x v faces inward: where the
    # book's (u, v) order points outward (green wall, floor, back wall) u and v are swapped.
    quads = [
        Quad.new(Vec3(555.0, 0.0, 0.0), Vec3(0.0, 0.0, 555.0), Vec3(0.0, 555.0, 0.0), SCATTER_M, Z, green),
        Quad.new(Vec3(0.0, 0.0, 0.0), Vec3(0.0, 555.0, 0.0), Vec3(0.0, 0.0, 555.0), SCATTER_M, Z, red),
        Quad.new(Vec3(343.0, 554.0, 332.0), Vec3(-130.0, 0.0, 0.0), Vec3(0.0, 0.0, -105.0),
                 Material.new_emmiting(0.0, 0.0, 1.0, Vec3(15.0, 15.0, 15.0)), Z, _solid((1.0, 1.0, 1.0))),
        Quad.new(Vec3(0.0, 0.0, 0.0), Vec3(0.0, 0.0, 555.0), Vec3(555.0, 0.0, 0.0), SCATTER_M, Z, white),
        Quad.new(Vec3(555.0, 555.0, 555.0), Vec3(-555.0, 0.0, 0.0), Vec3(0.0, 0.0, -555.0), SCATTER_M, Z, white),
        Quad.new(Vec3(0.0, 0.0, 555.0), Vec3(0.0, 555.0, 0.0), Vec3(555.0, 0.0, 0.0), SCATTER_M, Z, white)]
    b1 = Instance.new_box(Vec3(0.0, 0.0, 0.0), Vec3(165.0, 330.0, 165.0), _solid((0.0, 0.0, 0.0)) if smoke else white, SCATTER_M)
    b1.rotate(Vec3(0.0, float(F(15.0) * F(PI) / F(180.0)), 0.0))
    b1.translate(Vec3(265.0, 0.0, 295.0))
    b2 = Instance.new_box(Vec3(0.0, 0.0, 0.0), Vec3(165.0, 165.0, 165.0), _solid((1.0, 1.0, 1.0)) if smoke else white, SCATTER_M)
    b2.rotate(Vec3(0.0, float(F(-18.0) * F(PI) / F(180.0)), 0.0))
    b2.translate(Vec3(130.0, 0.0, 65.0))
    if smoke:
        for b in (b1, b2):
            b.dist_fn, b.density = const_density, 0.01
    sc = Scene.new([], quads, [b1, b2])
    vp = Viewport.new_from_res(width, height, samples, depth, 2.0, 40.0, Vec3(278.0, 278.0, -800.0), Vec3(0.0, 0.0, 1.0),
                               Vec3(0.0, 1.0, 0.0), None, 0.0)
    return Job("C4" + ("_smoke" if smoke else ""), sc, vp, ray_color_bg_color, A.RTW_MODE_RANDOM,
               "Cornell box 600x600 5000spp d50")


def config_c5(width=800, height=800, samples=10000, depth=50, frame=0) -> Job:
    """C5: Book-2 final scene through the reference API — 400 box instances, moving
    sphere, two volumes, 1000-sphere instance, earth + Perlin textures. The headline config."""
    rs = PhiloxStream(2024)
    ground = _solid((0.48, 0.83, 0.53))
    instances = []
    for i in range(20):
        for j in range(20):
            w = 100.0
            x0, z0, y0 = -1000.0 + i * w, -1000.0 + j * w, 0.0
            x1, y1, z1 = x0 + w, rs.uniform(1.0, 101.0), z0 + w
            instances.append(Instance.new_box(Vec3(x0, y0, z0), Vec3(x1, y1, z1), ground, SCATTER_M))
    quads = [Quad.new(Vec3(123.0, 554.0, 147.0), Vec3(300.0, 0.0, 0.0), Vec3(0.0, 0.0, 265.0),
                      Material.new_emmiting(0.0, 0.0, 1.0, Vec3(7.0, 7.0, 7.0)), Z, _solid((1.0, 1.0, 1.0)))]
    spheres = [
        Sphere.new_moving_with_texture(Vec3(400.0, 400.0, 200.0), 50.0, None, SCATTER_M, Vec3(30.0, 0.0, 0.0),
                                       _solid((0.7, 0.3, 0.1))),
        _albedo_sphere(Vec3(260.0, 150.0, 45.0), 50.0, (1.0, 1.0, 1.0), GLASS_M),
        _albedo_sphere(Vec3(0.0, 150.0, 145.0), 50.0, (0.8, 0.8, 0.9), Material.new_m(0.0)),
        _albedo_sphere(Vec3(360.0, 150.0, 145.0), 70.0, (1.0, 1.0, 1.0), GLASS_M),
        Sphere.new_with_texture(Vec3(400.0, 200.0, 400.0), 100.0, None, SCATTER_M, earth_texture()),
        Sphere.new_with_texture(Vec3(220.0, 280.0, 300.0), 80.0, None, SCATTER_M,
                                ImageTexture.from_color_noise(Vec3(1.0, 1.0, 1.0), 5.0, PerlinNoise.new(7)))]
    blue_fog = Instance.new_sphere([_albedo_sphere(Vec3(360.0, 150.0, 145.0), 70.0, (0.2, 0.4, 0.9), SCATTER_M)])
    blue_fog.dist_fn, blue_fog.density = const_density, 0.2
    fog = Instance.new_sphere([_albedo_sphere(Vec3(0.0, 0.0, 0.0), 5000.0, (1.0, 1.0, 1.0), SCATTER_M)])
    fog.dist_fn, fog.density = const_density, 0.0001
    white = _solid((0.73, 0.73, 0.73))
    cluster = Instance.new_sphere([
        Sphere.new_with_texture(Vec3(rs.uniform(0.0, 165.0), rs.uniform(0.0, 165.0), rs.uniform(0.0, 165.0)), 10.0,
                                None, SCATTER_M, white) for _ in range(1000)])
    cluster.rotate(Vec3(0.0, float(F(15.0) * F(PI) / F(180.0)), 0.0))
    cluster.translate(Vec3(-100.0, 270.0, 395.0))
    instances += [blue_fog, fog, cluster]
    sc = Scene.new(spheres, quads, instances)
    vp = Viewport.new_from_res(width, height, samples, depth, 2.0, 40.0, Vec3(478.0, 278.0, -600.0),
                               _unit(Vec3(-200.0, 0.0, 600.0)), Vec3(0.0, 1.0, 0.0), None, 0.0)
    vp.shutter_speed, vp.fps, vp.frame = 1.0, 30.0, frame
    return Job("C5", sc, vp, ray_color_bg_color, A.RTW_MODE_RANDOM, "Next-Week final scene 800x800 10000spp d50")


def _five_quad_room(green_mat=SCATTER_M):
    """The open room of objects/quad.rs:166-292 and viewport/ray_color.rs:115-213: red left, green back, blue
    right, orange top, teal bottom."""
    q = lambda o, u, v, m, c: Quad.new(Vec3(*o), Vec3(*u), Vec3(*v), m, Z, _solid(c))
    return [q((-3.0, -2.0, 5.0), (0.0, 0.0, -4.0), (0.0, 4.0, 0.0), SCATTER_M, (1.0, 0.2, 0.2)),
            q((-2.0, -2.0, 0.0), (4.0, 0.0, 0.0), (0.0, 4.0, 0.0), green_mat, (0.2, 1.0, 0.2)),
            q((3.0, -2.0, 1.0), (0.0, 0.0, 4.0), (0.0, 4.0, 0.0), SCATTER_M, (0.2, 0.2, 1.0)),
            q((-2.0, 3.0, 1.0), (4.0, 0.0, 0.0), (0.0, 0.0, 4.0), SCATTER_M, (1.0, 0.5, 0.0)),
            q((-2.0, -3.0, 5.0), (4.0, 0.0, 0.0), (0.0, 0.0, -4.0), SCATTER_M, (0.2, 0.8, 0.8))]


def quad_test(samples=100, size=400) -> Job:
    """objects/quad.rs:153-321 `quad_test`: the five-quad room under the sky gradient, Viewport::render (stratified)."""
    vp = Viewport.new_from_res(size, size, samples, 10, 2.0, 80.0, Vec3(0.0, 0.0, 9.0), Vec3(0.0, 0.0, -1.0), None, "Quad test")
    return Job("quad_room", Scene.new_quad(_five_quad_room()), vp, ray_color_d, A.RTW_MODE_STRATIFIED)


def simple_quad_test(samples=100, size=200) -> Job:
    """objects/quad.rs:323-390 `simple_test`: one green quad two units in front of the camera."""
    quads = [Quad.new(Vec3(-1.0, -0.5, -2.0), Vec3(2.0, 0.0, 0.0), Vec3(0.0, 1.0, 0.0), SCATTER_M, Z, _solid((0.20, 1.0, 0.20)))]
    vp = Viewport.new_from_res(size, size, samples, 10, 2.0, 80.0, Vec3(0.0, 0.0, 0.0), Vec3(0.0, 0.0, -1.0), None, "Quad test")
    return Job("simple_quad", Scene.new([], quads, []), vp, ray_color_d, A.RTW_MODE_STRATIFIED)


def light_test(samples=100, size=400) -> Job:
    """viewport/ray_color.rs:107-240 `light_test`: the same room with an emissive (4, 2, 4) back wall on a black
    background, ray_color_bg_color."""
    room = _five_quad_room(Material.new_emmiting(0.0, 0.0, 1.0, Vec3(4.0, 2.0, 4.0)))
    vp = Viewport.new_from_res(size, size, samples, 10, 2.0, 80.0, Vec3(0.0, 0.0, 9.0), Vec3(0.0, 0.0, -1.0), None, "Quad test")
    return Job("light", Scene.new_quad(room), vp, ray_color_bg_color, A.RTW_MODE_STRATIFIED)


def light_box_test(samples=100, size=400) -> Job:
    """viewport/ray_color.rs:242-518 `box_test`: an emissive mirror sphere (2, 0.5, 2) inside a five-quad box whose
    back wall carries a zero emission, async_render."""
    q = lambda o, u, v, m, c: Quad.new(Vec3(*o), Vec3(*u), Vec3(*v), m, Z, _solid(c))
    quads = [q((-2.0, -2.0, 5.0), (0.0, 0.0, -4.0), (0.0, 4.0, 0.0), SCATTER_M, (1.0, 0.2, 0.2)),
             q((2.0, -2.0, 1.0), (0.0, 0.0, 4.0), (0.0, 4.0, 0.0), SCATTER_M, (0.2, 0.2, 1.0)),
             q((-2.0, 2.0, 1.0), (4.0, 0.0, 0.0), (0.0, 0.0, 4.0), SCATTER_M, (1.0, 0.5, 0.0)),
             q((-2.0, -2.0, 5.0), (4.0, 0.0, 0.0), (0.0, 0.0, -4.0), SCATTER_M, (0.2, 0.8, 0.8)),
             q((-2.0, -2.0, 1.0), (4.0, 0.0, 0.0), (0.0, 4.0, 0.0), Material.new_emmiting(0.0, 0.0, 1.0, Vec3(0.0, 0.0, 0.0)), (0.2, 1.0, 0.2))]
    lamp = Sphere.new(Vec3(0.0, 0.0, 2.0), 0.4, Vec3(1.0, 1.0, 1.0), Material.new_emmiting(1.0, 0.0, 0.0, Vec3(2.0, 0.5, 2.0)))
    vp = Viewport.new_from_res(size, size, samples, 10, 2.0, 90.0, Vec3(0.0, 0.0, 4.0), Vec3(0.0, 0.0, -1.0), None, "Quad test")
    return Job("light_box", Scene.new([lamp], quads, []), vp, ray_color_bg_color, A.RTW_MODE_RANDOM)


def triangle_test(width=400, height=300, samples=25, depth=2) -> Job:
    """Rust2/src/objects/triangle.rs:149-199 `triangle_test`, in the Rust/src object model: one grey
    lambertian triangle three units in front of the camera, background 0.6 white. Rust2's axes:
    LEFT = +x, RIGHT = -x, FORWARD = +z (Rust2/src/vec3.rs:126-155). Rendered with ray_color_bg_color
    (Rust2's ray_color returns the background where this one returns black at depth 0)."""
    tri = Triangle.new(Vec3(1.0, -1.0, 3.0), Vec3(-1.0, 2.0, 0.0), Vec3(-2.0, 1.0, 0.0), SCATTER_M, _solid((0.5, 0.5, 0.5)))
    sc = Scene.new([], [], [Instance.new_quads([tri])])
    sc.background_color = Vec3(0.6, 0.6, 0.6)
    vp = Viewport.new_from_res(width, height, samples, depth, 1.0, 50.0, Vec3(0.0, 0.0, 0.0), Vec3(0.0, 0.0, 1.0))
    return Job("triangle", sc, vp, ray_color_bg_color, A.RTW_MODE_RANDOM)


def triangle_mesh_test(samples=100, size=400) -> Job:
    """Triangles everywhere a quad can stand (Rust2/src/objects/triangle.rs): at top level
    (a textured one and the two halves of a split parallelogram), as a rotated four-face mesh instance,
    as a one-triangle instance, and as the boundary of a constant-density volume, inside the room of
    instance.rs:313-600."""
    light = Quad.new(Vec3(-0.5, 1.99, 2.5), Vec3(1.0, 0.0, 0.0), Vec3(0.0, 0.0, 1.0),
                     Material.new_emmiting(0.0, 0.0, 1.0, Vec3(6.0, 6.0, 6.0)), Z, _solid((1.0, 1.0, 1.0)))
    back = Quad.new(Vec3(-2.0, -2.0, 0.999), Vec3(4.0, 0.0, 0.0), Vec3(0.0, 4.0, 0.0), SCATTER_M, Z, _solid((0.7, 0.7, 0.7)))
    tex_tri = Triangle.new(Vec3(-1.9, -0.3, 1.2), Vec3(1.7, 0.1, 0.3), Vec3(0.2, 1.9, 0.1), SCATTER_M,
                           checker_texture(64, 32, 8, 4))
    half_a = Triangle.new(Vec3(0.3, 0.2, 1.4), Vec3(1.3, 0.1, 0.0), Vec3(0.1, 1.2, 0.2), SCATTER_M, _solid((0.9, 0.2, 0.2)))
    # the other half of the same parallelogram: from the opposite corner, edges reversed
    half_b = Triangle.new(Vec3(0.3 + 1.3 + 0.1, 0.2 + 0.1 + 1.2, 1.4 + 0.2), Vec3(-1.3, -0.1, 0.0), Vec3(-0.1, -1.2, -0.2),
                          METALLIC_M, _solid((0.2, 0.9, 0.2)))

    def tetra(scale, tex, mat):
        a, b, c, d = (np.array(p, F) * F(scale) for p in ((0.0, 0.0, 0.0), (1.0, 0.0, 0.1), (0.4, 0.0, 0.9), (0.5, 1.1, 0.4)))
        face = lambda p, q, r: Triangle.new(p, q - p, r - p, mat, tex)
        return [face(a, c, b), face(a, b, d), face(b, c, d), face(c, a, d)]

    mesh = Instance.new_quads(tetra(1.2, _solid((0.3, 0.4, 0.9)), SCATTER_M))
    mesh.rotate(Vec3(0.0, PI / 5.0, 0.0))
    mesh.translate(Vec3(-1.4, -1.9, 2.2))
    single = Instance.new_quads([Triangle.new(Vec3(0.0, 0.0, 0.0), Vec3(0.9, 0.1, 0.0), Vec3(0.2, 0.0, 0.8), GLASS_M,
                                              _solid((1.0, 1.0, 1.0)))])
    single.translate(Vec3(0.6, -1.2, 2.6))
    smoke = Instance.new_quads(tetra(1.5, _solid((0.8, 0.8, 0.8)), SCATTER_M))
    smoke.translate(Vec3(0.2, -1.95, 1.3))
    smoke.dist_fn = const_density
    smoke.density = 1.5
    ball = _albedo_sphere(Vec3(1.3, -1.5, 3.6), 0.4, (0.8, 0.6, 0.2), METALLIC_M)
    sc = Scene.new([ball], _room_quads(light, [back, tex_tri, half_a, half_b]), [mesh, single, smoke])
    return Job("triangle_mesh", sc, _room_viewport(samples, 20, size), ray_color_bg_color, A.RTW_MODE_RANDOM)


def light_biased_test(cast=False, width=400, height=300, samples=100, depth=9, weight=100.0) -> Job:
    """Rust2/src/tests.rs:409-691 `make_scene` + `light_biased_ray_color_test`: a room of five quads (red, green with
    a faint emission, blue, an orange MIRROR ceiling, teal floor), a small emissive quad (2, 2, 2) and an emissive
    sphere (4, 2, 4) at its back; both are `lights`. Camera at the origin looking along +z (Vec3::FORWARD), vfov 90;
    SAMPLES 100, DEPTH 9 (the ray-cast control uses 10, ignored), BIASED_WEIGHT 100, background black.
    ConstColorTexture::new(multiplied, emitted) becomes col_mod via the texture and Material::new_emmiting."""
    def quad(o, u, v, colour, emitted=(0.0, 0.0, 0.0), m=0.0):
        return Quad.new(Vec3(*o), Vec3(*u), Vec3(*v), Material.new_emmiting(m, 0.0, 1.0, Vec3(*emitted)), Z, _solid(colour))
    light_quad = quad((-0.10, -0.10, 4.5), (0.2, 0.0, 0.0), (0.0, 0.2, 0.0), (1.0, 1.0, 1.0), (2.0, 2.0, 2.0))
    light_sphere = Sphere.new_with_texture(Vec3(-0.40, 0.0, 4.5), 0.2, None, Material.new_emmiting(0.0, 0.0, 1.0, Vec3(4.0, 2.0, 4.0)),
                                           _solid((1.0, 1.0, 1.0)))
    quads = [quad((-2.0, -2.0, 5.0), (0.0, 0.0, -4.0), (0.0, 4.0, 0.0), (1.0, 0.2, 0.2)),             # red
             light_quad,
             quad((-2.0, -2.0, 6.0), (4.0, 0.0, 0.0), (0.0, 4.0, 0.0), (0.2, 1.0, 0.2), (0.0, 0.0, 0.1)),   # green
             quad((2.0, -2.0, 1.0), (0.0, 0.0, 4.0), (0.0, 4.0, 0.0), (0.2, 0.2, 1.0)),                 # blue
             quad((-2.0, 2.0, 1.0), (4.0, 0.0, 0.0), (0.0, 0.0, 4.0), (1.0, 0.5, 0.0), m=1.0),          # orange, MIRROR
             quad((-2.0, -2.0, 5.0), (4.0, 0.0, 0.0), (0.0, 0.0, -4.0), (0.2, 0.8, 0.8))]               # teal
    sc = Scene.new([light_sphere], quads, [])
    sc.lights = [light_quad, light_sphere]
    vp = Viewport.new_from_res(width, height, samples, depth, 2.0, 90.0, Vec3(0.0, 0.0, 0.0), Vec3(0.0, 0.0, 1.0), Vec3(0.0, 1.0, 0.0))
    vp.light_weight = weight
    return Job("light_cast" if cast else "light_biased", sc, vp, light_biased_ray_cast if cast else light_biased_ray_color,
               A.RTW_MODE_RANDOM)


CONFIGS: Dict[str, Callable[..., Job]] = {"C1": config_c1, "C2": config_c2, "C3": config_c3, "C4": config_c4,
                                          "C5": config_c5}

REFERENCE_TESTS: Dict[str, Callable[..., Job]] = {
    "viewport_object": viewport_object_test, "camera_default": lambda **k: camera_test("default", **k),
    "camera_fov_120": lambda **k: camera_test("fov_120", **k),
    "camera_upside_down": lambda **k: camera_test("upside_down", **k),
    "camera_depth_of_field": lambda **k: camera_test("depth_of_field", **k),
    "diffuse": diffuse_test, "metal": metal_test, "hollow_glass": hollow_glass_test, "noise": noise_test,
    "reflection": reflection_test, "texture": texture_test, "ltr": ltr_scene, "instance_box": lambda **k: instance_box_test(False, **k),
    "instance_box_rot": lambda **k: instance_box_test(True, **k), "volume": volume_test,
    "presentation": presentation_image, "first_frame": first_frame,
    "triangle": triangle_test, "triangle_mesh": triangle_mesh_test,
    "quad_room": quad_test, "simple_quad": simple_quad_test, "light": light_test, "light_box": light_box_test,
}
