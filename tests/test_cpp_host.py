"""include/rtw.hpp — the C++17 host mirror of the reference's scene / render API (the reference
is compiled code and there is no Rust toolchain here, so the compiled caller's host side is C++).

tests/cpp/host_mirror.cpp writes two scenes against rtw.hpp the way a caller of the reference would
(`Sphere::new_(..)`, `Instance::new_box(..)`, `Viewport::new_from_res(..)`, `async_render(..)`); the same
scenes are built here with the Python mirror. CPU: both flatten to byte-identical rtw_scene_desc arrays,
cameras and render options, and the header needs nothing but librtw_b200.so. GPU: the images rendered
through rtw.hpp equal the Python mirror's up to f32 summation order (same description, same kernels)."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

import rtw_b200
from rtw_b200 import _abi as A, scenes
from rtw_b200.scene import (FUZZY3_M, GLASS_M, GLASSR_M, METALLIC_M, SCATTER_M, ImageTexture, Instance, Material,
                            PerlinNoise, Quad, Scene, Sphere, Triangle, Vec3, const_density)
from rtw_b200.viewport import Viewport, ray_color_bg_color

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CPP = os.path.join(ROOT, "tests", "cpp")
CSRC = os.path.join(ROOT, "raytracing-in-a-weekend_b200", "csrc")
F = np.float32


def build_host_mirror() -> str:
    path = os.path.join(CPP, "libhost_mirror.so")
    srcs = [os.path.join(CPP, "host_mirror.cpp")] + [os.path.join(ROOT, "include", f) for f in ("rtw.hpp", "rtw_json.hpp", "rtw.h")]
    if not os.path.exists(path) or any(os.path.getmtime(s) > os.path.getmtime(path) for s in srcs):
        subprocess.check_call(["g++", "-std=c++17", "-O1", "-Wall", "-Wextra", "-Werror", "-ffp-contract=off", "-fPIC", "-shared",
                               "-o", path, os.path.join(CPP, "host_mirror.cpp"), "-L" + CSRC, "-lrtw_b200",
                               "-Wl,-rpath,$ORIGIN/../../raytracing-in-a-weekend_b200/csrc"])
    return path


@pytest.fixture(scope="module")
def hm():
    A.load_library()                                   # librtw_b200.so first: fails loudly if it was not built
    lib = C.CDLL(build_host_mirror())
    lib.hm_flatten.restype = C.POINTER(A.rtw_scene_desc)
    lib.hm_flatten.argtypes = [C.c_int, C.c_uint32, C.c_uint32, C.c_uint32, C.POINTER(A.rtw_camera), C.POINTER(A.rtw_render_opts)]
    lib.hm_render.restype = C.c_int
    lib.hm_render.argtypes = [C.c_int, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32, C.c_void_p, C.c_void_p]
    lib.hm_json_dumps.restype = C.c_char_p
    lib.hm_json_flatten.restype = C.POINTER(A.rtw_scene_desc)
    lib.hm_json_flatten.argtypes = [C.c_char_p, C.POINTER(C.c_int)]
    return lib


def mixed_scene(width, height, samples):
    """The Python twin of host_mirror.cpp::mixed."""
    checker = np.zeros((8, 16, 3), F)
    for y in range(8):
        for x in range(16):
            checker[y, x] = (0.9, 0.9, 0.9) if ((x // 2) + (y // 2)) & 1 else (0.2, 0.3, 0.1)
    solid = ImageTexture.from_color
    noise = PerlinNoise.new(11)
    check_tex = ImageTexture.new(checker.reshape(-1, 3), 16, 8)
    check_tex2 = ImageTexture.new(checker.reshape(-1, 3), 16, 8)
    spheres = [
        Sphere.new(Vec3(-1.0, 0.0, -1.0), 0.5, Vec3(0.8, 0.8, 0.8), METALLIC_M),
        Sphere.new_with_texture(Vec3(0.0, -100.5, -1.0), 100.0, Vec3(1.0, 1.0, 1.0), SCATTER_M,
                                ImageTexture.from_color_noise(Vec3(1.0, 1.0, 1.0), 0.01, noise)),
        Sphere.new_moving(Vec3(1.0, 0.0, -1.0), 0.4, Vec3(0.9, 0.6, 0.2), FUZZY3_M, Vec3(0.5, 0.25, 0.0)),
        Sphere.new_with_texture(Vec3(0.0, 0.0, -1.2), 0.5, None, GLASS_M, solid((1.0, 1.0, 1.0))),
        Sphere.new_with_texture(Vec3(0.0, 0.0, -1.2), 0.4, None, GLASSR_M, solid((1.0, 1.0, 1.0))),
        Sphere.new_moving_with_texture(Vec3(0.0, 1.1, -1.5), 0.45, None, SCATTER_M, Vec3(0.0, 0.0, 0.0), check_tex),
        Sphere.new(Vec3(2.0, 0.3, -2.0), 0.6)]
    quads = [
        Triangle.new(Vec3(-2.5, -0.4, -2.5), Vec3(1.5, 0.1, 0.2), Vec3(0.2, 1.6, 0.1), SCATTER_M, check_tex2),
        Quad.new(Vec3(-0.5, 2.4, -2.0), Vec3(1.0, 0.0, 0.0), Vec3(0.0, 0.0, 1.0),
                 Material.new_emmiting(0.0, 0.0, 1.0, Vec3(4.0, 4.0, 4.0)), (0.0, 0.0, 0.0), solid((1.0, 1.0, 1.0)))]
    fog = Instance.new_sphere([Sphere.new_with_texture(Vec3(0.0, 0.0, 0.0), 0.7, None, SCATTER_M, solid((0.2, 0.4, 0.9)))])
    fog.translate(Vec3(-2.0, 0.4, -1.6))
    fog.dist_fn, fog.density = const_density, 2.0
    both = Instance.new([Sphere.new(Vec3(0.0, 0.0, 0.0), 0.25, Vec3(0.9, 0.1, 0.1), SCATTER_M)],
                        [Quad.new(Vec3(-0.5, -0.3, 0.0), Vec3(1.0, 0.0, 0.0), Vec3(0.0, 0.0, -1.0), METALLIC_M, (0.0, 0.0, 0.0),
                                  solid((0.7, 0.7, 0.7)))])
    both.rotate(Vec3(0.1, float(F(np.pi) / F(4.0)), -0.2))
    both.translate(Vec3(1.2, 0.9, -1.4))
    sc = Scene.new(spheres, quads, [fog, both])
    sc.background_color = Vec3(0.3, 0.4, 0.6)
    vp = Viewport.new(width, float(F(width) / F(height)), samples, 12, 2.0, 70.0, Vec3(0.2, 0.5, 1.5), Vec3(-0.05, -0.1, -1.0),
                      Vec3(0.0, 1.0, 0.0), "Rendering", 0.02)
    vp.shutter_speed, vp.fps, vp.frame = 0.5, 24.0, 3
    return scenes.Job("mixed", sc, vp, ray_color_bg_color, A.RTW_MODE_STRATIFIED)


def python_job(which, w, h, spp):
    return scenes.config_c4(width=w, height=h, samples=spp, smoke=True) if which == 0 else mixed_scene(w, h, spp)


def _bytes(ptr, n, T):
    return C.string_at(C.cast(ptr, C.c_void_p), n * C.sizeof(T))


@pytest.mark.parametrize("which", [0, 1], ids=["cornell_smoke", "mixed"])
def test_cpp_and_python_mirrors_flatten_identically(which, hm, oracle):
    w, h, spp = 60, 40, 5
    cam, opts = A.rtw_camera(), A.rtw_render_opts()
    dp = hm.hm_flatten(which, w, h, spp, C.byref(cam), C.byref(opts))
    assert dp, "rtw.hpp threw while flattening"
    d = dp.contents
    job = python_job(which, w, h, spp)
    fs = job.scene.flatten()
    p = fs.desc
    counts = ("abi_version", "n_spheres", "n_top_spheres", "n_quads", "n_top_quads", "n_instances", "n_materials",
              "n_textures", "n_noises")
    assert [getattr(d, k) for k in counts] == [getattr(p, k) for k in counts]
    assert tuple(d.background) == tuple(p.background)
    assert _bytes(d.spheres, d.n_spheres, A.rtw_sphere) == _bytes(p.spheres, p.n_spheres, A.rtw_sphere)
    assert _bytes(d.quads, d.n_quads, A.rtw_quad) == _bytes(p.quads, p.n_quads, A.rtw_quad)
    assert _bytes(d.instances, d.n_instances, A.rtw_instance) == _bytes(p.instances, p.n_instances, A.rtw_instance)
    assert _bytes(d.materials, d.n_materials, A.rtw_material) == _bytes(p.materials, p.n_materials, A.rtw_material)
    assert _bytes(d.noises, d.n_noises, A.rtw_noise) == _bytes(p.noises, p.n_noises, A.rtw_noise)      # same Philox tables
    for i in range(d.n_textures):
        a, b = d.textures[i], p.textures[i]
        assert (a.width, a.height, a.noise_id, a.noise_scale) == (b.width, b.height, b.noise_id, b.noise_scale)
        n = a.width * a.height * 3
        assert C.string_at(C.cast(a.rgb, C.c_void_p), 4 * n) == C.string_at(C.cast(b.rgb, C.c_void_p), 4 * n)
    # the camera (Viewport::new restated by rtw_camera_init) and the render options
    pc = job.viewport.camera(oracle)
    assert bytes(cam) == bytes(pc)
    po = job.viewport.opts(job.ray_color, job.mode)
    assert bytes(opts) == bytes(po)
    if which == 1:
        assert d.n_noises == 1 and d.n_instances == 2 and [d.quads[i].shape for i in range(d.n_quads)] == [1, 0, 0]


def _same_desc(d, p):
    assert (d.n_spheres, d.n_top_spheres, d.n_quads, d.n_instances, d.n_materials, d.n_textures, d.n_noises) == \
           (p.n_spheres, p.n_top_spheres, p.n_quads, p.n_instances, p.n_materials, p.n_textures, p.n_noises)
    assert _bytes(d.spheres, d.n_spheres, A.rtw_sphere) == _bytes(p.spheres, p.n_spheres, A.rtw_sphere)
    assert _bytes(d.materials, d.n_materials, A.rtw_material) == _bytes(p.materials, p.n_materials, A.rtw_material)
    for i in range(d.n_textures):
        a, b = d.textures[i], p.textures[i]
        assert (a.width, a.height, a.noise_id, a.noise_scale) == (b.width, b.height, b.noise_id, b.noise_scale)
        n = a.width * a.height * 3
        assert C.string_at(C.cast(a.rgb, C.c_void_p), 4 * n) == C.string_at(C.cast(b.rgb, C.c_void_p), 4 * n)


def test_json_wire_format_crosses_between_the_cpp_and_python_mirrors(hm):
    """include/rtw_json.hpp and scene_json.py restate the same Into / TryFrom<JsonValue> impls (sphere.rs:44-90,
    viewport.rs:174-205): each reads the other's text and arrives at the same flattened scene; what the format
    drops (noise, emission, quads, instances) is dropped on both sides; a missing field is a ParseError."""
    import json as pyjson
    from rtw_b200 import scene_json
    err = C.c_int(0)
    py_scene = Scene.new_sphere(mixed_scene(8, 8, 1).scene.spheres)
    # Python text -> C++
    text = scene_json.dumps(py_scene)
    dp = hm.hm_json_flatten(text.encode(), C.byref(err))
    assert dp and err.value == 0
    want = scene_json.loads(text).flatten()
    _same_desc(dp.contents, want.desc)
    assert dp.contents.n_noises == 0 and dp.contents.n_spheres == 7          # the Perlin table does not travel (texture.rs:28-59)
    # C++ text -> Python -> and back into C++
    ctext = hm.hm_json_dumps().decode()
    back = scene_json.loads(ctext)
    assert scene_json.scene_eq(back, py_scene) and len(back.spheres) == 7
    assert list(pyjson.loads(ctext)["spheres"][0]) == ["origin", "radius", "col_mod", "material", "velocity", "texture"]   # sphere.rs:46-53
    dp = hm.hm_json_flatten(ctext.encode(), C.byref(err))
    assert dp and err.value == 0
    _same_desc(dp.contents, back.flatten().desc)
    _same_desc(dp.contents, want.desc)                                          # nine digits round-trip every f32
    # errors: the reference's TryFrom fails as a whole
    obj = pyjson.loads(text)
    del obj["spheres"][2]["velocity"]
    for bad in (pyjson.dumps(obj), text[:-1], "{}", '{"spheres": 3}', text.replace('"radius":0.5', '"radius":"0.5"', 1)):
        assert not hm.hm_json_flatten(bad.encode(), C.byref(err)) and err.value == 1, bad[:60]
        with pytest.raises(scene_json.ParseError):
            scene_json.loads(bad)


def test_header_compiles_standalone_and_refuses_without_a_device(hm):
    """rtw.hpp is self-contained (g++ -fsyntax-only, -Wall -Wextra -Werror) and has no CPU path: without a
    device a render through it reports the library's RTW_E_CUDA."""
    subprocess.check_call(["g++", "-std=c++17", "-fsyntax-only", "-Wall", "-Wextra", "-Werror", "-x", "c++",
                           os.path.join(ROOT, "include", "rtw_json.hpp")])
    if rtw_b200.get_default_backend().device_count() == 0:
        out = np.zeros((8, 8, 3), F)
        assert hm.hm_render(1, 8, 8, 1, 0, out.ctypes.data_as(C.c_void_p), None) == A.RTW_E_CUDA


@pytest.mark.gpu
@pytest.mark.parametrize("which", [0, 1], ids=["cornell_smoke", "mixed"])
def test_gpu_render_through_cpp_mirror_equals_python_mirror(which, hm, cuda, oracle):
    w, h, spp = 96, 64, 9
    out = np.zeros((h, w, 3), F)
    out8 = np.zeros((h, w, 3), np.uint8)
    rc = hm.hm_render(which, w, h, spp, A.RTW_FLAG_STRICT_FP, out.ctypes.data_as(C.c_void_p), out8.ctypes.data_as(C.c_void_p))
    assert rc == 0
    job = python_job(which, w, h, spp)
    job.viewport.flags = A.RTW_FLAG_STRICT_FP
    want = job.render(cuda)
    # same description, same kernels, same Philox streams: two runs differ only by the order in which a
    # pixel's samples reach its f32 sum (atomic adds)
    assert np.array_equal(np.isfinite(out), np.isfinite(want))
    fin0 = np.isfinite(want)
    assert np.allclose(out[fin0], want[fin0], rtol=1e-5, atol=1e-6)
    assert np.array_equal(out8, oracle.quantize(out))      # rtw.hpp's quantize() = write_img.rs:12-14
    ref = job.render(oracle)                                # and the reference semantics behind both
    fin = np.isfinite(ref) & np.isfinite(out)
    close = np.abs(out - ref) <= 1e-4 * max(1.0, float(np.abs(ref[fin]).max())) + 1e-4 * np.abs(ref)
    assert (~close[fin]).mean() < 5e-3
