"""Triangles (Rust2/src/objects/triangle.rs) in the Rust/src object model: a quad record
whose inside test is `alfa >= 0, beta >= 0, alfa + beta <= 1` (triangle.rs:107-113) and whose leaf box
spans three corners (triangle.rs:50-89).

CPU: the oracle's restatement is pinned by an independent float64 point-in-triangle computation, by the
"two halves make the parallelogram" identity, and the kernels' own headers (tests/emu) are compared
with it. GPU (`-m gpu`): the CUDA path against the oracle through the C ABI; the `triangle` and
`triangle_mesh` scenes are also members of scenes.REFERENCE_TESTS, so every parametrised parity test of
test_gpu_parity.py / test_host_logic.py runs them too."""
import ctypes as C

import numpy as np
import pytest

import rtw_b200
from rtw_b200 import _abi as A, scenes
from rtw_b200.scene import METALLIC_M, SCATTER_M, ImageTexture, Instance, Quad, Scene, Triangle, Vec3
from rtw_b200.viewport import Viewport

Z = (0.0, 0.0, 0.0)


@pytest.fixture(scope="module")
def emu():
    from emu_backend import EmuBackend
    return EmuBackend()


def _centre_rays(vp, backend):
    """render_no_rand's rays (viewport.rs:479-516): origin, pixel00 + du*i + dv*j, in float64."""
    cam = vp.camera(backend)
    o = np.array(cam.origin[:], np.float64)
    p00, du, dv = (np.array(v[:], np.float64) for v in (cam.pixel00, cam.delta_u, cam.delta_v))
    jj, ii = np.mgrid[0:cam.height, 0:cam.width]
    return o, p00 + du * ii[..., None] + dv * jj[..., None]


def _bary(o, d, Q, u, v):
    """alfa, beta, t of triangle.rs:94-107 in float64."""
    n = np.cross(u, v)
    w = n / n.dot(n)
    nh = n / np.linalg.norm(n)
    t = (nh.dot(Q) - nh.dot(o)) / (d @ nh)
    planar = o + d * t[..., None] - Q
    return np.cross(planar, v) @ w, np.cross(u, planar) @ w, t


def test_oracle_triangle_matches_float64_point_in_triangle(oracle):
    """Pin of the restatement: which centre rays of Rust2's `triangle_test` scene (triangle.rs:149-199)
    hit, and with what colour, from first principles. Rays within 1e-4 of an edge are left out."""
    job = scenes.triangle_test(width=160, height=120)
    tri = job.scene.instances[0].quads[0]
    got = job.viewport._run(A.RTW_INT_ALBEDO, job.scene, A.RTW_MODE_NO_RAND, oracle)
    o, d = _centre_rays(job.viewport, oracle)
    al, be, t = _bary(o, d, *(x.astype(np.float64) for x in (tri.origin, tri.u, tri.v)))
    inside = (al >= 0) & (be >= 0) & (al + be <= 1) & (t > 1e-3)
    sure = np.minimum(np.minimum(np.abs(al), np.abs(be)), np.abs(al + be - 1)) > 1e-4
    assert 0.05 < inside.mean() < 0.5                       # the triangle is in view and is not a parallelogram
    upper_half = (al >= 0) & (be >= 0) & (al <= 1) & (be <= 1) & (al + be > 1)
    assert upper_half.sum() > 100                           # pixels a quad would cover and the triangle must not
    g = np.float32(1.0) ** np.float32(1.0)
    hit_col, bg = np.float32(0.5) * g, np.float32(0.6) * g
    assert np.all(got[inside & sure] == hit_col)
    assert np.all(got[~inside & sure] == bg)


def _split_scene(as_triangles, same_look=True):
    Q, u, v = np.array((-1.1, -0.7, -2.0), np.float32), np.array((2.3, 0.2, -0.4), np.float32), np.array((0.3, 1.6, 0.1), np.float32)
    tex = lambda c: ImageTexture.from_color(c)
    if not as_triangles:
        quads = [Quad.new(Q, u, v, SCATTER_M, Z, tex((0.4, 0.5, 0.6)))]
    else:
        c2 = (0.4, 0.5, 0.6) if same_look else (0.9, 0.1, 0.1)
        quads = [Triangle.new(Q, u, v, SCATTER_M, tex((0.4, 0.5, 0.6))),
                 Triangle.new(Q + u + v, -u, -v, SCATTER_M, tex(c2))]
    sc = Scene.new([], quads, [])
    sc.background_color = Vec3(0.1, 0.2, 0.3)
    return sc


def test_two_halves_make_the_parallelogram(oracle, emu):
    """A parallelogram and its two triangles (the second built from the opposite corner with reversed
    edges: same normal, triangle.rs:36-38) give the same first-hit images; only rays through the shared
    diagonal may fall in neither or both, and there both triangles look the same."""
    vp = Viewport.new_from_res(96, 72, 1, 4, 1.0, 90.0, Vec3(0.0, 0.0, 0.0), Vec3(0.0, 0.0, -1.0))
    for be in (oracle, emu):
        for ig in (A.RTW_INT_NORMAL, A.RTW_INT_ALBEDO):
            whole = vp._run(ig, _split_scene(False), A.RTW_MODE_NO_RAND, be)
            halves = vp._run(ig, _split_scene(True), A.RTW_MODE_NO_RAND, be)
            diff = (whole != halves).any(-1)
            assert diff.sum() <= 2, f"{diff.sum()} pixels differ"      # at most a diagonal ray that rounds out of both
    # and with different looks the two halves are told apart along alfa + beta = 1
    img = vp._run(A.RTW_INT_ALBEDO, _split_scene(True, same_look=False), A.RTW_MODE_NO_RAND, oracle)
    first = np.all(img == np.array((0.4, 0.5, 0.6), np.float32), axis=-1).sum()
    second = np.all(img == np.array((0.9, 0.1, 0.1), np.float32), axis=-1).sum()
    assert first > 300 and second > 300 and abs(int(first) - int(second)) < 0.35 * first   # perspective: not equal, but both there


def test_textured_triangle_uses_the_quad_texel_rule(oracle, emu):
    """Texel = floor(alfa*row), floor(beta*col) as for quads (quad.rs:66-76): with a two-colour texture
    split at alfa = 1/2 the triangle shows both colours, the second on the smaller part."""
    img = np.zeros((4, 8, 3), np.float32)
    img[:, :4] = (1.0, 0.0, 0.0)
    img[:, 4:] = (0.0, 0.0, 1.0)
    tri = Triangle.new(Vec3(-1.0, -1.0, -2.0), Vec3(2.0, 0.0, 0.0), Vec3(0.0, 2.0, 0.0), SCATTER_M,
                       ImageTexture.new(img.reshape(-1, 3), 8, 4))
    sc = Scene.new([], [tri], [])
    sc.background_color = Vec3(0.0, 1.0, 0.0)
    vp = Viewport.new_from_res(64, 64, 1, 4, 1.0, 90.0, Vec3(0.0, 0.0, 0.0), Vec3(0.0, 0.0, -1.0))
    a = vp._run(A.RTW_INT_ALBEDO, sc, A.RTW_MODE_NO_RAND, oracle)
    b = vp._run(A.RTW_INT_ALBEDO, sc, A.RTW_MODE_NO_RAND, emu)
    assert np.array_equal(a, b)
    red, blue = (a[..., 0] == 1.0).sum(), (a[..., 2] == 1.0).sum()
    # areas: alfa < 1/2 part of the triangle is 3/4 of it, alfa >= 1/2 the remaining 1/4
    assert red > 0 and blue > 0 and 2.5 < red / blue < 3.5


def test_thin_triangle_leaf_box_is_padded(oracle, emu):
    """triangle.rs:62-66: an axis-aligned triangle's flat axis gets a 0.005-wide interval, so the slab
    test in front of it (aabb.rs:109-126) passes and it is seen face-on; emulated kernels agree."""
    tri = Triangle.new(Vec3(-1.0, -1.0, -2.0), Vec3(2.0, 0.0, 0.0), Vec3(0.0, 2.0, 0.0), SCATTER_M, ImageTexture.from_color((0.5, 0.5, 0.5)))
    for sc in (Scene.new([], [tri], []), Scene.new([], [], [Instance.new_quads([tri])])):
        sc.background_color = Vec3(0.0, 1.0, 0.0)
        vp = Viewport.new_from_res(32, 32, 1, 4, 1.0, 90.0, Vec3(0.0, 0.0, 0.0), Vec3(0.0, 0.0, -1.0))
        a = vp._run(A.RTW_INT_ALBEDO, sc, A.RTW_MODE_NO_RAND, oracle)
        b = vp._run(A.RTW_INT_ALBEDO, sc, A.RTW_MODE_NO_RAND, emu)
        assert np.array_equal(a, b)
        # the square (-1..1)^2 at z = -2 fills a quarter of a 90-degree view; the triangle is half of it
        assert 0.09 < (a[..., 0] == 0.5).mean() < 0.16


def test_shape_field_is_validated_and_flattened():
    from test_host_logic import _desc_rc
    job = scenes.triangle_mesh_test()
    fs = job.scene.flatten()
    shapes = [fs.desc.quads[i].shape for i in range(fs.desc.n_quads)]
    assert shapes.count(A.RTW_SHAPE_TRIANGLE) == 3 + 4 + 1 + 4 and shapes.count(A.RTW_SHAPE_PARALLELOGRAM) == 6
    assert C.sizeof(A.rtw_quad) == 60                       # include/rtw.h: 12 floats + 3 words
    assert _desc_rc(job.scene) == 0

    def bad_shape(d): d.quads[0].shape = 2
    assert _desc_rc(job.scene, bad_shape) == A.RTW_E_INVALID


def test_path_radiance_with_triangles_matches_oracle(oracle, emu):
    """The whole estimator over the mesh scene (diffuse, mirror and glass triangles, a triangle-bounded
    constant-density volume): the emulated kernels follow the oracle path for path."""
    job = scenes.triangle_mesh_test(samples=4, size=72)
    job.viewport.gamma = 1.0
    ref = job.render(oracle)
    tie = oracle.scene(job.scene).last_tie.astype(bool)
    got = job.render(emu)
    so, se = oracle.scene(job.scene).stats(), emu.scene(job.scene).stats()
    if not tie.any():
        assert (se["paths"], se["rays"]) == (so.paths, so.rays)
    fin = np.isfinite(ref) & ~tie[..., None]
    assert np.array_equal(np.isfinite(got[~tie]), np.isfinite(ref[~tie]))
    scale = max(1.0, float(np.abs(ref[fin]).max()))
    assert np.all(np.abs(got[fin] - ref[fin]) <= 2e-6 * scale + 1e-5 * np.abs(ref[fin]))
    assert ref[fin].mean() > 0.01                           # the light reaches the camera


# ------------------------------------------------------------------------------------------- GPU
@pytest.mark.gpu
@pytest.mark.parametrize("pipe", [A.RTW_FLAG_MEGAKERNEL, A.RTW_FLAG_WAVEFRONT], ids=["megakernel", "wavefront"])
def test_gpu_triangles_first_hit_and_paths(pipe, oracle, cuda):
    """CUDA (strict build) against the oracle on the mesh scene: first-hit images bit for bit outside
    tie pixels, path radiance sample for sample."""
    job = scenes.triangle_mesh_test(samples=4, size=200)
    vp = job.viewport
    vp.gamma = 1.0
    vp.flags = A.RTW_FLAG_STRICT_FP | pipe
    for ig in (A.RTW_INT_NORMAL, A.RTW_INT_ALBEDO):
        ref = vp._run(ig, job.scene, A.RTW_MODE_NO_RAND, oracle)
        tie = oracle.scene(job.scene).last_tie.astype(bool)
        got = vp._run(ig, job.scene, A.RTW_MODE_NO_RAND, cuda)
        ok = ~tie
        assert np.allclose(got[ok], ref[ok], rtol=1e-5, atol=1e-6, equal_nan=True)
    ref = job.render(oracle)
    tie = oracle.scene(job.scene).last_tie.astype(bool)
    got = job.render(cuda)
    fin = np.isfinite(ref) & np.isfinite(got) & ~tie[..., None]
    scale = max(1.0, float(np.abs(ref[fin]).max()))
    close = np.abs(got - ref) <= 1e-4 * scale + 1e-4 * np.abs(ref)
    assert (~close[fin]).mean() < 5e-3


@pytest.mark.gpu
def test_gpu_fast_build_sees_the_same_triangles(cuda):
    """The fast build's branch-free quad test carries the same `alfa + beta <= lim` term: its first-hit
    image of the mesh scene equals the strict build's except at silhouette pixels."""
    job = scenes.triangle_mesh_test(samples=1, size=256)
    vp = job.viewport
    vp.gamma = 1.0
    vp.flags = A.RTW_FLAG_STRICT_FP
    strict = vp._run(A.RTW_INT_ALBEDO, job.scene, A.RTW_MODE_NO_RAND, cuda)
    vp.flags = 0
    fast = vp._run(A.RTW_INT_ALBEDO, job.scene, A.RTW_MODE_NO_RAND, cuda)
    assert (np.abs(fast - strict) > 1e-5).any(-1).mean() < 1e-2
    # and the half-parallelogram really is half: a triangle covers about half the pixels of its quad
    vp2 = Viewport.new_from_res(96, 72, 1, 4, 1.0, 90.0, Vec3(0.0, 0.0, 0.0), Vec3(0.0, 0.0, -1.0))
    whole = vp2._run(A.RTW_INT_ALBEDO, _split_scene(False), A.RTW_MODE_NO_RAND, cuda)
    halves = vp2._run(A.RTW_INT_ALBEDO, _split_scene(True), A.RTW_MODE_NO_RAND, cuda)
    assert ((whole != halves).any(-1)).sum() <= 4
