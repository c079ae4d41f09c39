"""The light-biased estimators of the Rust2 tree (Rust2/src/viewport/ray_color.rs:55-164
`light_biased_ray_cast`, `light_biased_ray_color`; `material_pdf` Rust2/src/objects/material.rs:17,44-63,93-101,199-235)
over the Rust/src scene model: RTW_INT_LIGHT_CAST / RTW_INT_LIGHT_BIASED, the light list of rtw_scene_desc, the weight
in rtw_render_opts.

Pins: the reference asserts nothing about these functions (its tests write PNGs: Rust2/src/tests.rs:649-691), and they
are re-expressed here over another object model (uber-material instead of `dyn Material`, emission in the material
instead of the texture), so parity is UNPINNED by the reference; what is held:
  * a closed form (one diffuse floor, one spherical light, float64) against the oracle and the kernels' headers;
  * oracle == emulated headers == CUDA strict build, sample for sample, on Rust2's own test scene (tests.rs:409-691);
  * CUDA fast build == strict build statistically;
  * the structure of the estimator: mirrors and glass get no direct light, `depth` exhausted -> background, weight 0
    is the plain recursion.
"""
import numpy as np
import pytest

import rtw_b200
from rtw_b200 import _abi as A, scenes
from rtw_b200.scene import Material, METALLIC_M, SCATTER_M, ImageTexture, Quad, Scene, Sphere, Vec3
from rtw_b200.viewport import Viewport, light_biased_ray_cast, light_biased_ray_color


@pytest.fixture(scope="module")
def emu():
    from emu_backend import EmuBackend
    return EmuBackend()


def _floor_and_lamp(albedo=(0.6, 0.5, 0.4), E=(3.0, 2.0, 1.0), floor_mat=SCATTER_M):
    floor = Quad.new(Vec3(-5.0, 0.0, -5.0), Vec3(0.0, 0.0, 10.0), Vec3(10.0, 0.0, 0.0), floor_mat, (0, 0, 0), ImageTexture.from_color(albedo))
    lamp = Sphere.new_with_texture(Vec3(0.0, 2.0, 0.0), 0.25, None, Material.new_emmiting(0.0, 0.0, 1.0, Vec3(*E)),
                                   ImageTexture.from_color((1.0, 1.0, 1.0)))
    sc = Scene.new([lamp], [floor], [])
    sc.lights = [lamp]
    sc.background_color = Vec3(0.1, 0.2, 0.3)
    return sc


def test_ray_cast_closed_form(oracle, emu):
    """Camera high above a diffuse floor, looking straight down (past the lamp's silhouette for most pixels): a floor
    pixel at p sees radiance albedo * E * (cos / pi) / t^2 with cos = n . unit(lamp - p) and t the distance to the
    lamp's surface along that direction; the lamp itself shows its emission; a miss the background."""
    albedo, E = np.array([0.6, 0.5, 0.4]), np.array([3.0, 2.0, 1.0])
    sc = _floor_and_lamp(tuple(albedo), tuple(E))
    vp = Viewport.new_from_res(41, 31, 1, 5, 1.0, 60.0, Vec3(0.0, 6.0, 0.0), Vec3(0.0, -1.0, 0.0), Vec3(0.0, 0.0, 1.0))
    cam = vp.camera(oracle)
    got_o = vp._run(light_biased_ray_cast, sc, A.RTW_MODE_NO_RAND, oracle)
    got_e = vp._run(light_biased_ray_cast, sc, A.RTW_MODE_NO_RAND, emu)
    assert np.allclose(got_e, got_o, rtol=2e-6, atol=1e-7)
    o = np.array(list(cam.origin), dtype=np.float64)
    c, r = np.array([0.0, 2.0, 0.0]), 0.25
    n_floor = n_lamp = 0
    for j in range(cam.height):
        for i in range(cam.width):
            d = np.array(list(cam.pixel00), np.float64) + np.array(list(cam.delta_u), np.float64) * i + np.array(list(cam.delta_v), np.float64) * j
            oc = o - c
            a, b, cc = d @ d, oc @ d, oc @ oc - r * r
            if b * b - a * cc >= 0:                   # the lamp is in front of the floor for these pixels
                assert np.allclose(got_o[j, i], E, rtol=1e-5), (i, j)
                n_lamp += 1
                continue
            t = -o[1] / d[1]
            p = o + d * t
            if abs(p[0]) > 5 or abs(p[2]) > 5:
                assert np.allclose(got_o[j, i], [0.1, 0.2, 0.3], rtol=1e-6)
                continue
            to = (c - p) / np.linalg.norm(c - p)
            tl = np.linalg.norm(c - p) - r
            want = albedo * E * (to[1] / np.pi) / (tl * tl)
            assert np.allclose(got_o[j, i], want, rtol=3e-5), (i, j, got_o[j, i], want)
            n_floor += 1
    assert n_floor > 800 and n_lamp > 3


def test_mirrors_and_glass_get_no_direct_light(oracle, emu):
    """material_pdf of Mirror / MirrorGlass is 0 for any direction but their own (Rust2 material.rs:93-101, 199-235):
    light_biased_ray_cast leaves such surfaces at their own emission, i.e. black."""
    for mat in (METALLIC_M, Material.new(0.0, 1.0, 1.5)):
        sc = _floor_and_lamp(floor_mat=mat)
        vp = Viewport.new_from_res(16, 12, 1, 5, 1.0, 40.0, Vec3(3.0, 4.0, 0.0), Vec3(0.0, -1.0, 0.0), Vec3(0.0, 0.0, 1.0))
        for be in (oracle, emu):
            img = vp._run(light_biased_ray_cast, sc, A.RTW_MODE_NO_RAND, be)
            assert np.array_equal(img, np.zeros_like(img))


def test_depth_exhausted_returns_the_background_and_weight_zero_is_the_plain_recursion(oracle, emu):
    """`if depth <= 0 { return vp.bg_color }` (ray_color.rs:118-120), unlike Rust/src's integrators, which return black;
    with biased_weight 0 the light terms vanish: L = emitted + col_mod * L(reflected)."""
    sc = _floor_and_lamp()
    vp = Viewport.new_from_res(25, 19, 4, 1, 1.0, 60.0, Vec3(0.0, 6.0, 0.0), Vec3(0.0, -1.0, 0.0), Vec3(0.0, 0.0, 1.0))
    vp.light_weight = 0.0
    for be in (oracle, emu):
        img = vp._run(light_biased_ray_color, sc, A.RTW_MODE_RANDOM, be)
        # depth 1: one hit, then the recursion is at depth 0 -> background: floor pixels = albedo * bg, lamp = E + 1 * bg
        assert np.allclose(img[2, 2], np.array([0.6, 0.5, 0.4]) * np.array([0.1, 0.2, 0.3]), rtol=1e-5)
        assert np.allclose(img[9, 12], np.array([3.0, 2.0, 1.0]) + np.array([0.1, 0.2, 0.3]), rtol=1e-5)


def _small(job, w, h, spp):
    vp = job.viewport
    nv = type(vp).new_from_res(w, h, spp, vp.depth, 1.0, vp.vfov, vp.origin, vp.direction, vp.vup, None, vp.lens_radius)
    nv.light_weight = vp.light_weight
    job.viewport = nv
    return job


@pytest.mark.parametrize("cast", [True, False], ids=["ray_cast", "ray_color"])
def test_headers_match_the_oracle_on_the_reference_scene(cast, oracle, emu):
    job = _small(scenes.light_biased_test(cast), 64, 48, 3)
    vp = job.viewport
    ref = vp._run(job.ray_color, job.scene, job.mode, oracle)
    got = vp._run(job.ray_color, job.scene, job.mode, emu)
    so, se = oracle.scene(job.scene).stats(), emu.scene(job.scene).stats()
    assert (se["paths"], se["rays"]) == (so.paths, so.rays)               # shadow rays are rays on both sides
    assert so.rays > 2 * so.paths                                         # two lights: at least two shadow rays per hit
    scale = float(np.abs(ref).max())
    assert np.all(np.abs(got - ref) <= 2e-6 * scale + 2e-5 * np.abs(ref))
    assert ref.mean() > 0.01


@pytest.mark.gpu
@pytest.mark.parametrize("cast", [True, False], ids=["ray_cast", "ray_color"])
def test_cuda_matches_the_oracle_sample_for_sample(cast, oracle, cuda):
    job = _small(scenes.light_biased_test(cast), 120, 90, 4)
    vp = job.viewport
    vp.flags = A.RTW_FLAG_STRICT_FP
    ref = vp._run(job.ray_color, job.scene, job.mode, oracle)
    got = vp._run(job.ray_color, job.scene, job.mode, cuda)
    so, sg = oracle.scene(job.scene).stats(), cuda.scene(job.scene).stats()
    assert sg.paths == so.paths and abs(int(sg.rays) - int(so.rays)) <= 2e-3 * so.rays + 2
    assert sg.iterations == 0                                             # the megakernel
    scale = float(np.abs(ref).max())
    close = np.abs(got - ref) <= 1e-4 * scale + 1e-4 * np.abs(ref)
    assert (~close).mean() < 5e-3, f"{(~close).sum()} of {close.size} values differ"


@pytest.mark.gpu
def test_cuda_fast_build_and_pipeline_choice(cuda):
    job = scenes.light_biased_test(False, 320, 240, 64)                   # 4.9 M paths (the light integrators take the megakernel at any size)
    vp = job.viewport
    vp.gamma = 1.0
    vp.flags = A.RTW_FLAG_STRICT_FP
    strict = vp._run(job.ray_color, job.scene, job.mode, cuda)
    vp.flags = 0
    fast = vp._run(job.ray_color, job.scene, job.mode, cuda)
    st = cuda.scene(job.scene).stats()
    assert st.iterations == 0 and st.paths == 320 * 240 * 64              # light integrators take the megakernel
    assert abs(float(fast.mean()) / float(strict.mean()) - 1.0) < 5e-3
    vp.flags = A.RTW_FLAG_WAVEFRONT
    with pytest.raises(rtw_b200.RtwError) as ei:
        vp._run(job.ray_color, job.scene, job.mode, cuda)
    assert ei.value.code == A.RTW_E_UNSUPPORTED
    # the direct-light terms cut the variance: at equal samples the image is far smoother than the plain recursion's
    vp.flags, vp.light_weight = 0, 0.0
    plain = vp._run(job.ray_color, job.scene, job.mode, cuda)
    box = lambda a: a.reshape(30, 8, 40, 8, 3).mean(axis=(1, 3))
    noise = lambda a: float(np.abs(a - np.repeat(np.repeat(box(a), 8, 0), 8, 1)).mean())
    assert noise(fast) < 0.9 * noise(plain)                               # 0.75 measured (profiles/r2d_pytest_gpu.log)
