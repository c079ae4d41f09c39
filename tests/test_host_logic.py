"""Host logic of the product, on the CPU: scene flattening (Python mirror -> rtw_scene_desc), the host
scene builder (validation, derived quad/instance quantities, reference leaf boxes, two-level SAH BVH)
and — through tests/emu/ — the very traversal / shading headers the CUDA kernels are compiled from,
built for the host and compared with the oracle. The emulator is test infrastructure; the product
never loads it (test_abi.py checks the product refuses to run without a GPU)."""
import ctypes as C

import numpy as np
import pytest

import rtw_b200
from rtw_b200 import _abi as A, scenes
from rtw_b200.scene import SCATTER_M, ImageTexture, Instance, Quad, Scene, Sphere, Vec3, const_density
from rtw_b200.viewport import Viewport


@pytest.fixture(scope="module")
def emu():
    from emu_backend import EmuBackend
    return EmuBackend()


def _small(job, w, h, spp):
    vp = job.viewport
    nv = type(vp).new_from_res(w, h, spp, vp.depth, 1.0, vp.vfov, vp.origin, vp.direction, vp.vup, None, vp.lens_radius)
    nv.shutter_speed, nv.fps, nv.frame = vp.shutter_speed, vp.fps, vp.frame
    job.viewport = nv
    return job


ALL = dict(scenes.REFERENCE_TESTS)
ALL.update(scenes.CONFIGS)
ALL["C4_smoke"] = lambda **k: scenes.config_c4(smoke=True, **k)


@pytest.mark.parametrize("name", ["C3", "C4_smoke", "C5", "presentation", "texture", "triangle_mesh"])
def test_both_hit_flavours_are_the_same_arithmetic(name, emu):
    """rtw_shade.h evaluates the winning primitive's Hit in two shapes — one tail of texture lookup and material load for
    the whole warp (wf_shade), one per primitive kind (the megakernel): the same operations on the same values, so the
    images are equal bit for bit."""
    from emu_backend import EmuBackend
    if name not in ALL:
        pytest.skip(f"no scene {name}")
    job = _small(ALL[name](), 56, 40, 3)
    vp = job.viewport
    vp.flags = A.RTW_FLAG_STRICT_FP
    a = vp._run(job.ray_color, job.scene, job.mode, emu)
    b = vp._run(job.ray_color, job.scene, job.mode, EmuBackend(split_hit=True))
    assert np.array_equal(a, b, equal_nan=True)
    a = vp._run(A.RTW_INT_ALBEDO, job.scene, A.RTW_MODE_NO_RAND, emu)
    b = vp._run(A.RTW_INT_ALBEDO, job.scene, A.RTW_MODE_NO_RAND, EmuBackend(split_hit=True))
    assert np.array_equal(a, b, equal_nan=True)


@pytest.mark.parametrize("name", sorted(ALL))
def test_traversal_headers_match_oracle(name, oracle, emu):
    """Same closest hit (bit for bit outside tie pixels) and same path radiance as the oracle: the
    ordered, shrinking-interval SAH traversal is equivalent to the reference's visit-everything walk."""
    job = _small(ALL[name](), 64, 48, 3)
    vp = job.viewport
    for ig in (A.RTW_INT_NORMAL, A.RTW_INT_ALBEDO):
        ref = vp._run(ig, job.scene, A.RTW_MODE_NO_RAND, oracle)
        tie = oracle.scene(job.scene).last_tie.astype(bool)
        got = vp._run(ig, job.scene, A.RTW_MODE_NO_RAND, emu)
        assert np.array_equal(got[~tie], ref[~tie], equal_nan=True)
    ref = vp._run(job.ray_color, job.scene, job.mode, oracle)
    tie = oracle.scene(job.scene).last_tie.astype(bool)
    got = vp._run(job.ray_color, job.scene, job.mode, emu)
    so, se = oracle.scene(job.scene).stats(), emu.scene(job.scene).stats()
    if not tie.any():
        assert (se["paths"], se["rays"], se["nan_paths"]) == (so.paths, so.rays, so.nan_paths)
    ok = ~tie
    assert np.array_equal(np.isfinite(got[ok]), np.isfinite(ref[ok]))
    fin = np.isfinite(ref) & ok[..., None]
    scale = max(1.0, float(np.abs(ref[fin]).max()))
    assert np.all(np.abs(got[fin] - ref[fin]) <= 2e-6 * scale + 1e-5 * np.abs(ref[fin]))


def test_stratified_and_lens_and_frames(oracle, emu):
    job = scenes.camera_test("depth_of_field", width=40, height=30, samples=5)      # s_root = 3 -> 9 samples
    ref = job.render(oracle)
    got = job.render(emu)
    assert emu.scene(job.scene).stats()["paths"] == 40 * 30 * 9
    assert np.allclose(got, ref, rtol=1e-5, atol=1e-6)
    job = _small(scenes.ltr_scene(), 48, 27, 2)
    job.viewport.frame, job.viewport.fps, job.viewport.shutter_speed = 7, 15.0, 0.5
    assert np.allclose(job.render(emu), job.render(oracle), rtol=1e-5, atol=1e-6)


def test_sample_and_row_ranges_compose(emu):
    """RTW_OUT_SUM over sample sub-ranges adds up to the whole; row ranges are slices."""
    job = _small(scenes.config_c4(), 32, 32, 6)
    vp, be = job.viewport, emu
    h, cam = be.scene(job.scene), vp.camera(be)
    o = vp.opts(job.ray_color, job.mode)
    o.output = A.RTW_OUT_SUM
    full = h.render(cam, o)
    o.sample_begin, o.sample_count = 0, 4
    a = h.render(cam, o)
    o.sample_begin, o.sample_count = 4, 2
    b = h.render(cam, o)
    assert np.allclose(a + b, full, rtol=1e-6, atol=1e-6)
    o.sample_begin, o.sample_count, o.row_begin, o.row_count = 0, 0, 8, 5
    assert np.array_equal(h.render(cam, o), full[8:13])


def test_scene_info_and_identity_instance_flattening(emu):
    job = _small(scenes.config_c5(), 16, 16, 1)
    job.render(emu)
    inf = emu.scene(job.scene).info()
    assert inf.n_flattened_instances == 400                       # the untransformed ground boxes
    assert inf.n_tlas_leaves == 6 + 1 + 2400 + 3 - 1              # 5 static spheres + quad + box quads + 3 instances
    # 4-wide nodes: between (n-1)/3 (all full) and n-1 (all binary) nodes for n leaves
    assert (inf.n_tlas_leaves - 1) / 3 <= inf.n_tlas_nodes < (inf.n_tlas_leaves - 1) * 0.6
    assert 333 <= inf.n_blas_nodes < 600                          # the 1000-sphere cluster; 1-sphere volumes need no node
    assert 0 < inf.bvh_max_depth < 100                            # = stack words the deepest walk can push


def _desc_rc(scene, mutate=None):
    from emu_backend import load_emu
    fs = scene.flatten()
    if mutate:
        mutate(fs.desc)
    cam = A.rtw_camera(); cam.width = cam.height = 2
    o = A.rtw_render_opts(); o.samples, o.depth, o.gamma, o.fps = 1, 1, 1.0, 30.0
    out = np.zeros((2, 2, 3), dtype=np.float32)
    return load_emu().emu_render(C.byref(fs.desc), C.byref(cam), C.byref(o), out.ctypes.data_as(C.c_void_p), None, None)


def test_scene_validation_errors():
    sc = scenes.config_c4().scene
    assert _desc_rc(sc) == 0
    assert _desc_rc(sc, lambda d: setattr(d, "abi_version", 99)) == A.RTW_E_INVALID
    assert _desc_rc(sc, lambda d: setattr(d, "n_top_quads", d.n_quads + 1)) == A.RTW_E_INVALID

    def bad_mat(d): d.quads[0].mat_id = 1000
    assert _desc_rc(sc, bad_mat) == A.RTW_E_INVALID

    def bad_range(d): d.instances[0].n_quads = 10 ** 6
    assert _desc_rc(sc, bad_range) == A.RTW_E_INVALID

    def nan_sphere(d): d.spheres[0].radius = float("nan")
    assert _desc_rc(scenes.config_c1().scene, nan_sphere) == A.RTW_E_INVALID


def test_empty_and_single_primitive_scenes(oracle, emu):
    vp = Viewport.new_from_res(12, 9, 2, 5, 1.0)
    empty = Scene.new([], [], [])
    empty.background_color = Vec3(0.25, 0.5, 0.75)
    assert np.array_equal(vp.render_no_rand(rtw_b200.ray_color_albedo, empty, emu),
                          vp.render_no_rand(rtw_b200.ray_color_albedo, empty, oracle))
    one = Scene.new_sphere([Sphere.new(Vec3(0, 0, -1), 0.5)])
    assert np.array_equal(vp.render_no_rand(rtw_b200.ray_color_normal, one, emu),
                          vp.render_no_rand(rtw_b200.ray_color_normal, one, oracle))
    inst = Instance.new_sphere([])                               # an instance with nothing inside (box = [0,0])
    inst.translate(Vec3(0, 0, -3))
    sc = Scene.new([], [Quad.new(Vec3(-1, -1, -2), Vec3(2, 0, 0), Vec3(0, 2, 0), SCATTER_M, (0, 0, 0),
                                  ImageTexture.from_color((0.3, 0.6, 0.9)))], [inst])
    assert np.array_equal(vp.render_no_rand(rtw_b200.ray_color_albedo, sc, emu),
                          vp.render_no_rand(rtw_b200.ray_color_albedo, sc, oracle))


def test_axis_aligned_rays_use_the_literal_slab_rule(oracle, emu):
    """Direction components that are exactly 0 (objects.rs:126-135): a 1x1 image looking down -z."""
    vp = Viewport.new_from_res(1, 1, 1, 5, 1.0)
    sc = Scene.new([Sphere.new(Vec3(0, 0, -3), 1.0)], [Quad.new(Vec3(-1, -1, -5), Vec3(2, 0, 0), Vec3(0, 2, 0), SCATTER_M,
                   (0, 0, 0), ImageTexture.from_color((1, 0, 0)))], [])
    cam = vp.camera(oracle)
    assert cam.pixel00[0] == 0.0 and cam.pixel00[1] == 0.0
    for ig in (rtw_b200.ray_color_normal, rtw_b200.ray_color_albedo, rtw_b200.ray_color_gradient):
        assert np.array_equal(vp.render_no_rand(ig, sc, emu), vp.render_no_rand(ig, sc, oracle))


def test_deep_bvh_falls_back_to_a_balanced_split(emu):
    """2048 collinear, nested-size spheres would make a SAH tree deeper than the traversal stack; the
    builder's median-split guard keeps the binary tree below RTW_STACK levels, so the collapsed 4-wide
    tree can push at most 3 words per level pair."""
    sp = [Sphere.new(Vec3(0, 0, -float(2 ** (i / 64.0))), float(2 ** (i / 64.0)) * 0.9) for i in range(1024)]
    vp = Viewport.new_from_res(8, 8, 1, 3, 1.0)
    sc = Scene.new_sphere(sp)
    vp.render_no_rand(rtw_b200.ray_color_normal, sc, emu)
    assert emu.scene(sc).info().bvh_max_depth < 3 * 32


def test_dp_collapse_renders_identically(emu, tmp_path):
    """The DP-chosen 4-wide collapse (the default since the round-2 device A/B; RTW_BVH_COLLAPSE=greedy brings the
    largest-child-first one back) only regroups the same leaves into other wide nodes: strict renders are
    bit-identical under both. The variable is read once per process, hence the child."""
    import os
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    code = (
        "import sys; sys.path[:0] = [%r, %r]\n"
        "import numpy as np, rtw_b200\n"
        "from rtw_b200 import scenes\n"
        "from emu_backend import EmuBackend\n"
        "from test_host_logic import _small\n"
        "be = EmuBackend()\n"
        "out = {}\n"
        "for name in ('C2', 'C4_smoke', 'C5'):\n"
        "    f = scenes.CONFIGS.get(name) or (lambda: scenes.config_c4(smoke=True))\n"
        "    job = _small(f(), 40, 30, 2)\n"
        "    out[name] = job.render(be)\n"
        "    out[name + '_nodes'] = np.array(be.scene(job.scene).stats()['nodes'])\n"
        "np.savez(%r, **out)\n") % (root, os.path.join(root, "tests"), str(tmp_path / "dp.npz"))
    env = dict(os.environ, RTW_BVH_COLLAPSE="greedy")
    subprocess.check_call([sys.executable, "-c", code], env=env)
    got = np.load(str(tmp_path / "dp.npz"))
    fewer = 0
    for name in ("C2", "C4_smoke", "C5"):
        f = scenes.CONFIGS.get(name) or (lambda: scenes.config_c4(smoke=True))
        job = _small(f(), 40, 30, 2)
        assert np.array_equal(job.render(emu), got[name], equal_nan=True), name
        fewer += emu.scene(job.scene).stats()["nodes"] < int(got[name + "_nodes"])
    assert fewer >= 2                                   # and the default walks fewer nodes on most of them (the point of it)


BAKED_SCENES = ["instance_box_rot", "instance_box", "presentation", "triangle_mesh", "C4", "C4_smoke", "C5", "volume"]


@pytest.mark.parametrize("name", BAKED_SCENES)
def test_baked_instances_find_the_reference_hits(name, oracle):
    """rtw_host.h "baked instances": where an instance's frame change is a rigid motion, world-space copies of its
    primitives replace its leaf in the top level the fast kernels walk. The closest hit must be the reference's:
    same primitive (albedo bit for bit), same t up to the rounding of the copied centre / corners (normals within
    1e-5 of the oracle's outside tie pixels), and the Hit is still evaluated through the instance's frame. The
    emulated headers walk that top level here with the strict (f32-box) arithmetic; the GPU parity tests run it
    in the fast build."""
    from emu_backend import EmuBackend
    emu_b = EmuBackend(baked=True)
    job = _small(ALL[name](), 96, 72, 2)
    vp = job.viewport
    got_info = None
    for ig in (A.RTW_INT_ALBEDO, A.RTW_INT_NORMAL):
        ref = vp._run(ig, job.scene, A.RTW_MODE_NO_RAND, oracle)
        tie = oracle.scene(job.scene).last_tie.astype(bool)
        got = vp._run(ig, job.scene, A.RTW_MODE_NO_RAND, emu_b)
        got_info = emu_b.scene(job.scene).info()
        ok = ~tie
        assert np.array_equal(np.isfinite(got[ok]), np.isfinite(ref[ok]))
        fin = np.isfinite(ref) & ok[..., None]
        err = np.abs(got - ref)[fin] / np.maximum(np.abs(ref[fin]), 1e-6)
        if ig == A.RTW_INT_ALBEDO:
            assert (err > 0).mean() <= 2e-3, f"{name}: {(err > 0).sum()} albedo values differ"      # silhouette pixels only
        else:
            assert (err > 1e-5).mean() <= 4e-3, f"{name}: {(err > 1e-5).sum()} of {err.size} normal values off by > 1e-5"
    expect = {"instance_box_rot": 1, "instance_box": 1, "presentation": 1, "triangle_mesh": 2, "C4": 2, "C4_smoke": 0, "C5": 1, "volume": 0}[name]
    assert got_info.n_baked_instances == expect, (name, got_info.n_baked_instances)


def test_non_rigid_euler_maps_are_not_baked(oracle):
    """Vec3::rotated is a rotation only about y; an x- or z-'rotated' instance keeps its own level."""
    from emu_backend import EmuBackend
    emu_b = EmuBackend(baked=True)
    tex = ImageTexture.from_color((0.7, 0.4, 0.2))
    for rot, baked in (((0.3, 0.0, 0.0), 0), ((0.0, 0.0, 0.4), 0), ((0.0, 0.7, 0.0), 1), ((0.0, 0.0, 0.0), 1)):
        box = Instance.new_box(Vec3(-0.5, -0.5, -0.5), Vec3(0.5, 0.5, 0.5), tex, SCATTER_M)
        box.rotate(Vec3(*rot))
        box.translate(Vec3(0.1, 0.2, -3.0))
        sc = Scene.new([], [], [box])
        vp = Viewport.new_from_res(64, 48, 1, 5, 1.0)
        ref = vp._run(A.RTW_INT_NORMAL, sc, A.RTW_MODE_NO_RAND, oracle)
        tie = oracle.scene(sc).last_tie.astype(bool)
        got = vp._run(A.RTW_INT_NORMAL, sc, A.RTW_MODE_NO_RAND, emu_b)
        assert emu_b.scene(sc).info().n_baked_instances == baked, rot
        ok = ~tie
        assert np.allclose(got[ok], ref[ok], rtol=1e-5, atol=1e-6)


def test_interleaved_row_bands_are_those_rows_of_the_frame(oracle, emu):
    """rtw_render_opts.row_band / row_stride: one render of bands taken every row_stride rows (what a device of a
    tile split renders) returns exactly those rows of the whole frame, on the oracle and the kernels'
    headers; badly formed band descriptions are refused by the host logic both backends share."""
    job = _small(scenes.config_c4(smoke=True), 40, 37, 2)
    vp = job.viewport
    for be in (oracle, emu):
        h, cam = be.scene(job.scene), vp.camera(be)
        whole = h.render(cam, vp.opts(job.ray_color, job.mode))
        for band, world in ((4, 3), (5, 2), (8, 4)):
            got = np.zeros_like(whole)
            seen = np.zeros(cam.height, dtype=int)
            for rank in range(world):
                rows = np.concatenate([np.arange(r0, min(r0 + band, cam.height)) for r0 in range(rank * band, cam.height, world * band)] or [np.zeros(0, int)])
                if not rows.size:
                    continue
                o = vp.opts(job.ray_color, job.mode)
                o.row_begin, o.row_count, o.row_band, o.row_stride = rank * band, rows.size, band, world * band
                part = h.render(cam, o)
                assert part.shape == (rows.size, cam.width, 3)
                got[rows] = part
                seen[rows] += 1
            assert np.all(seen == 1)
            assert np.array_equal(got, whole, equal_nan=True)
    o = vp.opts(job.ray_color, job.mode)
    o.row_begin, o.row_count, o.row_band, o.row_stride = 0, 20, 4, 12          # the fifth band would start at row 48
    with pytest.raises(Exception):
        emu.scene(job.scene).render(vp.camera(emu), o)
