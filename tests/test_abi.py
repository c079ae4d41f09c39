"""The C-ABI shared library (raytracing-in-a-weekend_b200/csrc/librtw_b200.so) on a machine with no
GPU: it loads, exports every symbol include/rtw.h declares, its structs have the layout the ctypes
mirror assumes, the host-only helpers work, and every compute entry point FAILS LOUDLY (RTW_E_CUDA)
instead of falling back to a CPU path."""
import ctypes as C
import os
import re
import subprocess
import tempfile

import numpy as np
import pytest

import rtw_b200
from rtw_b200 import _abi as A, scenes

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HDR = os.path.join(ROOT, "include", "rtw.h")


@pytest.fixture(scope="module")
def lib():
    return rtw_b200.load_library()


def _declared_symbols():
    src = open(HDR).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(rtw_[a-z0-9_]+)\s*\(", src)))


def test_every_declared_symbol_is_exported_and_bound(lib):
    names = _declared_symbols()
    assert len(names) >= 16
    for n in names:
        assert hasattr(lib, n), f"librtw_b200.so does not export {n}"
        assert n in A.SYMBOLS, f"_abi.py does not bind {n}"
    assert sorted(A.SYMBOLS) == names
    assert lib.rtw_abi_version() == A.RTW_ABI_VERSION


def test_struct_layouts_match_the_header():
    """Compile a tiny C program against include/rtw.h and compare sizeof/offsetof with ctypes."""
    structs = ["rtw_material", "rtw_noise", "rtw_texture", "rtw_sphere", "rtw_quad", "rtw_instance", "rtw_scene_desc",
               "rtw_camera", "rtw_render_opts", "rtw_stats", "rtw_scene_info"]
    probes = [("rtw_render_opts", "seed"), ("rtw_render_opts", "flags"), ("rtw_scene_desc", "background"),
              ("rtw_texture", "noise_id"), ("rtw_stats", "hits"), ("rtw_camera", "lens_radius")]
    body = "#include <stdio.h>\n#include <stddef.h>\n#include \"rtw.h\"\nint main(void){\n"
    for s in structs:
        body += f'printf("%zu\\n", sizeof({s}));\n'
    for s, f in probes:
        body += f'printf("%zu\\n", offsetof({s}, {f}));\n'
    body += "return 0;}\n"
    with tempfile.TemporaryDirectory() as td:
        src, exe = os.path.join(td, "p.c"), os.path.join(td, "p")
        open(src, "w").write(body)
        subprocess.check_call(["gcc", "-std=c99", "-I", os.path.join(ROOT, "include"), src, "-o", exe])
        vals = [int(x) for x in subprocess.check_output([exe]).split()]
    for s, v in zip(structs, vals):
        assert C.sizeof(getattr(A, s)) == v, s
    for (s, f), v in zip(probes, vals[len(structs):]):
        assert getattr(getattr(A, s), f).offset == v, (s, f)


def test_no_gpu_means_loud_failure_not_a_cpu_fallback(lib):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    assert lib.rtw_device_count() == 0
    fs = scenes.config_c1().scene.flatten()
    h = C.c_void_p()
    rc = lib.rtw_scene_create(C.byref(fs.desc), -1, C.byref(h))
    assert rc == A.RTW_E_CUDA and not h.value
    assert b"no CPU path" in lib.rtw_last_error()
    with pytest.raises(rtw_b200.RtwError):
        scenes.config_c1(width=8, height=8, samples=1).render()      # the package's public call
    v = C.c_double()
    assert lib.rtw_microbench(-1, A.RTW_BENCH_FP32, 1, C.byref(v)) == A.RTW_E_CUDA


def test_null_arguments_are_errors(lib):
    assert lib.rtw_scene_create(None, 0, None) == A.RTW_E_INVALID
    assert lib.rtw_render(None, None, None, None) == A.RTW_E_INVALID
    assert lib.rtw_get_stats(None, None) == A.RTW_E_INVALID
    assert lib.rtw_resolve_device(None, 10, 1, 2.0, -1, None) == A.RTW_E_INVALID
    assert lib.rtw_last_error() != b""
    lib.rtw_scene_destroy(None)     # harmless, like free(NULL)


def test_camera_init_restates_viewport_new(lib, oracle):
    """rtw_camera_init (host code of the product) == the oracle's Viewport::new (viewport.rs:308-401)."""
    rs = np.random.RandomState(3)
    for _ in range(20):
        w = int(rs.randint(8, 900)); hgt = int(rs.randint(8, 700))
        vfov = float(rs.uniform(20, 120))
        origin, d, vup = rs.uniform(-5, 5, 3), rs.normal(size=3), (0.0, 1.0, 0.0)
        d = d / np.linalg.norm(d)
        got = A.rtw_camera()
        f3 = lambda v: (C.c_float * 3)(*[float(x) for x in v])
        assert lib.rtw_camera_init_res(C.byref(got), w, hgt, vfov, f3(origin), f3(d), f3(vup), 0.1) == 0
        want = oracle.camera_init(w, float(np.float32(w) / np.float32(hgt)), vfov, origin, d, vup, 0.1)
        assert bytes(got) == bytes(want)
    assert lib.rtw_camera_init(None, 1, 1.0, 90.0, None, None, None, 0.0) == A.RTW_E_INVALID


def test_render_opts_default(lib):
    o = A.rtw_render_opts()
    lib.rtw_render_opts_default(C.byref(o))
    assert (o.samples, o.depth, o.gamma, o.fps, o.shutter_speed, o.frame, o.seed) == (1, 10, 2.0, 30.0, 0.0, 0, 1)
