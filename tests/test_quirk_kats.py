"""Known-answer tests for the reference behaviours of SURVEY 8a-notes that are not pinned elsewhere, each
computed from first principles (closed forms, float64 numpy) and held against BOTH the oracle and the
kernels' own headers compiled for the host (tests/emu) — and, in the `-m gpu` half, against the CUDA path.

  N1/N4  camera basis from an unnormalised view direction           viewport.rs:342-360
  N2     jitter footprint [centre, centre + 1 px)                    viewport.rs:294-296
  N7     L = emitted + col_mod * L(next); depth exhausted adds 0     viewport/ray_color.rs:43-92
  N9     opacity > 0 selects the dielectric branch, its value unused materials.rs:106
  N10    refract uses sqrt|1 - |r_perp|^2|, cos clamped from above   materials.rs:89-97,118-121
  N18    rows top to bottom, gamma without clamp                     viewport.rs:46,207-213,236-244
  N19    random::<f32>() = 24 bits in [0, 1)                         rand 0.8.5 Standard
"""
import ctypes as C

import numpy as np
import pytest

import rtw_b200
from rtw_b200 import _abi as A
from rtw_b200.philox import u32_to_f32
from rtw_b200.scene import GLASS_M, METALLIC_M, SCATTER_M, ImageTexture, Material, Quad, Scene, Sphere, Vec3
from rtw_b200.viewport import Viewport

F = np.float32


@pytest.fixture(scope="module")
def emu():
    from emu_backend import EmuBackend
    return EmuBackend()


@pytest.fixture(params=["oracle", "emu"])
def backend(request, oracle, emu):
    return oracle if request.param == "oracle" else emu


# ------------------------------------------------------------------------------------------ scenes
def lamp_scene(mat, albedo=0.5, emitted=1.0, bg=0.2):
    """The camera inside a closed emissive sphere: every path segment from inside hits it."""
    m = Material.new_emmiting(mat.metallicness, mat.opacity, mat.ir, Vec3(emitted, emitted, emitted))
    sc = Scene.new_sphere([Sphere.new_with_texture(Vec3(0.0, 0.0, 0.0), 10.0, None, m, ImageTexture.from_color((albedo,) * 3))])
    sc.background_color = Vec3(bg, bg, bg)
    return sc


def edge_scene(x_edge):
    """A white half-plane x < x_edge at z = -1 in front of a black background."""
    q = Quad.new(Vec3(x_edge - 50.0, -50.0, -1.0), Vec3(50.0, 0.0, 0.0), Vec3(0.0, 100.0, 0.0), SCATTER_M, (0.0, 0.0, 0.0),
                 ImageTexture.from_color((1.0, 1.0, 1.0)))
    return Scene.new_quad([q])


def check_estimator(be):
    # mirror walls (m = 1): the reflected ray stays inside, so all `depth` segments hit:
    # L = E (1 + a + ... + a^(depth-1)); the depth+1-th call returns 0 (ray_color.rs:44-46)
    vp = Viewport.new_from_res(8, 6, 3, 5, 1.0, 90.0, Vec3(1.0, 2.0, -3.0), Vec3(0.0, 0.0, -1.0))
    img = vp._run(A.RTW_INT_BG_COLOR, lamp_scene(METALLIC_M), A.RTW_MODE_RANDOM, be)
    assert np.allclose(img, 1.0 + 0.5 + 0.25 + 0.125 + 0.0625, rtol=1e-6)
    # diffuse walls scatter about the OUTWARD normal (N6), so the second segment leaves: L = E + a * background
    # (a scattered ray that leaves almost tangentially from a hit point rounded to just inside the sphere can meet
    # the wall once more beyond mint — about one sample in 400 — and then adds a second E: never less than the
    # closed form, and equal to it for nearly every pixel)
    img = vp._run(A.RTW_INT_BG_COLOR, lamp_scene(SCATTER_M), A.RTW_MODE_RANDOM, be)
    want = 1.0 + 0.5 * 0.2
    assert img.min() >= want * (1.0 - 1e-6) and np.isclose(img, want, rtol=1e-6).mean() >= 0.85
    # N18: gamma is applied without a clamp: sqrt(1.9375) > 1 is stored as such
    vp.gamma = 2.0
    img = vp._run(A.RTW_INT_BG_COLOR, lamp_scene(METALLIC_M), A.RTW_MODE_RANDOM, be)
    assert np.allclose(img, np.sqrt(1.9375), rtol=1e-6) and img.min() > 1.0


def check_rows_top_to_bottom(be):
    # ray_color_gradient's sky: t = 0.5 (unit(d).y + 1), red = (1 - t) + 0.5 t: smaller towards the zenith
    vp = Viewport.new_from_res(4, 16, 1, 3, 1.0, 90.0)
    behind = Scene.new_sphere([Sphere.new(Vec3(0.0, 0.0, 100.0), 1.0)])          # nothing in view
    img = vp._run(A.RTW_INT_GRADIENT, behind, A.RTW_MODE_NO_RAND, be)
    assert img.shape == (16, 4, 3) and np.all(np.diff(img[:, 0, 0]) > 0) and np.all(img[..., 2] == 1.0)
    cam = vp.camera(be)
    d = np.array(cam.pixel00[:], np.float64)                       # NoRand: pixel (0, 0) looks along pixel00 (viewport.rs:479-516)
    t = 0.5 * (d[1] / np.linalg.norm(d) + 1.0)
    assert np.allclose(img[0, 0], [(1 - t) + 0.5 * t, (1 - t) + 0.7 * t, 1.0], rtol=1e-6)


def check_jitter_footprint(be, spp=4096):
    # 9 x 1 pixels, vertical fov 90: the view is 2 high and 18 wide on the plane z = -1, a pixel is 2 x 2 and
    # pixel centres sit at x = -9 + (i + 0.5) * 2. Put the edge of the
    # white half-plane half a pixel to the right of the centre of pixel 4 (x = 0): direction x = (i + xi) px from
    # pixel00, xi in [0, 1), so pixel 4 is covered for xi < 1/2 only; a centred footprint would cover it fully.
    vp = Viewport.new_from_res(9, 1, spp, 2, 1.0, 90.0)
    px = 2.0
    img = vp._run(A.RTW_INT_ALBEDO, edge_scene(0.5 * px), A.RTW_MODE_RANDOM, be)[0, :, 0]
    assert np.all(img[:4] == 1.0) and np.all(img[5:] == 0.0)
    assert abs(img[4] - 0.5) < 4.0 * 0.5 / np.sqrt(spp)


def check_opacity_is_a_switch(be, exact=True):
    def render(opacity):
        glass = Material.new(1.0, opacity, 1.5)
        sc = Scene.new_sphere([Sphere.new_with_texture(Vec3(0.0, 0.0, -1.0), 0.5, None, glass, ImageTexture.from_color((0.9, 0.9, 0.9))),
                               Sphere.new_with_texture(Vec3(0.0, -100.5, -1.0), 100.0, None, SCATTER_M, ImageTexture.from_color((0.8, 0.8, 0.0)))])
        vp = Viewport.new_from_res(24, 16, 4, 8, 1.0, 90.0)
        return vp._run(A.RTW_INT_GRADIENT, sc, A.RTW_MODE_RANDOM, be)
    a, b, c = render(1.0), render(0.3), render(0.0)
    # materials.rs:106: only `opacity > 0` is read ... (on the device two runs may differ by the order of the f32 adds)
    assert np.array_equal(a, b) if exact else np.allclose(a, b, rtol=1e-5, atol=1e-6)
    assert not np.allclose(a, c, rtol=1e-3, atol=1e-3)              # ... and 0 makes it a mirror


# --------------------------------------------------------------------------------------------- CPU
def test_n7_estimator_and_n18_gamma(backend):
    check_estimator(backend)


def test_n18_rows_run_top_to_bottom(backend):
    check_rows_top_to_bottom(backend)


def test_n2_jitter_footprint_is_shifted_half_a_pixel(backend):
    check_jitter_footprint(backend)


def test_n9_opacity_value_is_unused(backend):
    check_opacity_is_a_switch(backend)


def test_n1_camera_basis_inherits_the_length_of_dir(oracle):
    """w = -dir is not normalised and v = w x u inherits |dir|; the viewport height is 2 tan(vfov/2) with no focal
    length; pixel00 is a direction. dir = (0, 0, -2), vfov 90, 4 x 2: by hand u = (1,0,0), v = (0,2,0),
    du = (1,0,0), dv = (0,-2,0), pixel00 = (-2,2,-2) + (du+dv)/2 = (-1.5, 1, -2). Product (host-only
    rtw_camera_init) and oracle."""
    lib = A.load_library()
    f3 = lambda v: (C.c_float * 3)(*v)
    cams = [A.rtw_camera(), oracle.camera_init(4, 2.0, 90.0, Vec3(5.0, 6.0, 7.0), Vec3(0.0, 0.0, -2.0), Vec3(0.0, 1.0, 0.0), 0.25)]
    A.check(lib.rtw_camera_init(C.byref(cams[0]), 4, 2.0, 90.0, f3((5, 6, 7)), f3((0, 0, -2)), f3((0, 1, 0)), 0.25))
    for cam in cams:
        assert (cam.width, cam.height) == (4, 2) and tuple(cam.origin) == (5.0, 6.0, 7.0) and cam.lens_radius == 0.25
        assert np.allclose(cam.u[:], (1, 0, 0), atol=1e-6) and np.allclose(cam.v[:], (0, 2, 0), atol=1e-6)
        assert np.allclose(cam.delta_u[:], (1, 0, 0), atol=1e-6) and np.allclose(cam.delta_v[:], (0, -2, 0), atol=1e-6)
        assert np.allclose(cam.pixel00[:], (-1.5, 1.0, -2.0), atol=1e-6)
    assert bytes(cams[0]) == bytes(cams[1])


def test_n10_refract_restated(oracle):
    """materials.rs:89-97 in float64: r_perp = ratio (uv + min(-uv.n, 1) n), r_par = -sqrt|1 - |r_perp|^2| n — the
    absolute value keeps it real past the critical angle (where the caller reflects instead, :125)."""
    def want(uv, n, ratio):
        uv, n = np.asarray(uv, np.float64), np.asarray(n, np.float64)
        cos = min(float(-uv @ n), 1.0)
        perp = ratio * (uv + cos * n)
        return perp - np.sqrt(abs(1.0 - perp @ perp)) * n
    out = (C.c_float * 3)()
    for uv, n, ratio in (((2 ** -0.5, -(2 ** -0.5), 0.0), (0, 1, 0), 1 / 1.5), ((0.6, -0.8, 0.0), (0, 1, 0), 1.5),
                         ((0.96, -0.28, 0.0), (0, 1, 0), 1.5), ((0.0, -1.0, 0.0), (0, 1, 0), 1 / 1.5)):
        oracle.lib.oracle_refract((C.c_float * 3)(*uv), (C.c_float * 3)(*n), C.c_float(ratio), out)
        assert np.allclose(out[:], want(uv, n, ratio), atol=2e-6), (uv, ratio)


def test_n19_uniforms_are_24_bit():
    assert u32_to_f32(np.array([0xFFFFFFFF], np.uint32))[0] == F(1.0) - F(2.0) ** -24     # never 1.0
    assert u32_to_f32(np.array([0x000000FF], np.uint32))[0] == 0.0                         # low 8 bits dropped
    assert u32_to_f32(np.array([0x80000000], np.uint32))[0] == 0.5


# --------------------------------------------------------------------------------------------- GPU
@pytest.mark.gpu
@pytest.mark.parametrize("flags", [A.RTW_FLAG_STRICT_FP | A.RTW_FLAG_MEGAKERNEL, A.RTW_FLAG_STRICT_FP | A.RTW_FLAG_WAVEFRONT, 0],
                         ids=["strict-megakernel", "strict-wavefront", "fast-auto"])
def test_gpu_known_answers(flags, cuda, monkeypatch):
    orig = Viewport.opts

    def opts(self, ray_color, mode):
        o = orig(self, ray_color, mode)
        o.flags |= flags
        return o
    monkeypatch.setattr(Viewport, "opts", opts)
    check_estimator(cuda)
    check_rows_top_to_bottom(cuda)
    check_jitter_footprint(cuda)
    if flags & A.RTW_FLAG_STRICT_FP:
        check_opacity_is_a_switch(cuda, exact=False)
