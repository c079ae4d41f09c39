"""rust_shim/ is the Rust binding a maintainer of the reference would add. The build image has
no cargo / rustc, so it cannot be compiled here; source-only code rots (round-1 review: build.rs had fallen behind
the Makefile and ffi.rs behind rtw.h). These checks keep it honest without a Rust toolchain:
  * build.rs compiles exactly the sources csrc/Makefile links into librtw_b200.so (same variants of rtw_kernels.cu);
  * ffi.rs declares every function include/rtw.h exports (via _abi.SYMBOLS, which test_abi.py holds against the built
    library) with the same number of arguments, and nothing else;
  * every #[repr(C)] struct has the fields of its ctypes mirror, in order, with matching scalar types;
  * the ABI version constant matches.
"""
import ctypes as C
import os
import re

import rtw_b200  # noqa: F401
from rtw_b200 import _abi as A

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SHIM = os.path.join(ROOT, "rust_shim")


def _read(*parts):
    return open(os.path.join(*parts)).read()


def test_build_rs_compiles_what_the_makefile_links():
    mk = _read(ROOT, "raytracing-in-a-weekend_b200", "csrc", "Makefile")
    objs = re.search(r"^OBJ := (.*)$", mk, re.M).group(1).split()
    rule_src = dict(re.findall(r"^\$\(BUILD\)/(\S+)\.o: (\S+)", mk, re.M))
    mk_sources = sorted(rule_src[os.path.basename(o)[:-2]] for o in objs)
    rs = _read(SHIM, "build.rs")
    rs_sources = re.findall(r'\("([\w.]+\.(?:cu|cpp))", (?:true|false)\)', rs)
    n_variants = len(re.findall(r'\("(fast|strict|count)", &\[', rs))
    rs_sources = sorted(rs_sources + ["rtw_kernels.cu"] * n_variants)
    assert rs_sources == mk_sources, (rs_sources, mk_sources)
    # same flags where they change results: the strict variant must not contract a*b+c, the count variant counts
    assert '"-fmad=false", "-DRTW_VARIANT=strict", "-DRTW_STRICT_FP=1"' in rs
    assert "-fmad=false -DRTW_VARIANT=strict -DRTW_STRICT_FP=1" in mk
    assert "arch=compute_100a,code=sm_100a" in rs and "arch=compute_100a,code=sm_100a" in mk
    assert "-ffp-contract=off" in rs and "-ffp-contract=off" in mk


def _rust_fns(text):
    block = text[text.index('extern "C" {'):]
    out = {}
    for m in re.finditer(r"pub fn (\w+)\s*\(([^)]*)\)", block, re.S):
        args = [a for a in m.group(2).split(",") if a.strip()]
        out[m.group(1)] = len(args)
    return out


def test_ffi_rs_declares_every_exported_function():
    fns = _rust_fns(_read(SHIM, "src", "ffi.rs"))
    assert sorted(fns) == sorted(A.SYMBOLS), (sorted(set(A.SYMBOLS) - set(fns)), sorted(set(fns) - set(A.SYMBOLS)))
    for name, (_, argtypes) in A.SYMBOLS.items():
        assert fns[name] == len(argtypes), name
    # ... and every one of them is declared in the header
    hdr = _read(ROOT, "include", "rtw.h")
    for name in fns:
        assert re.search(r"\b%s\s*\(" % name, hdr), name


_SCALAR = {C.c_float: "f32", C.c_uint32: "u32", C.c_int32: "i32", C.c_uint64: "u64"}


def _rust_type(ct):
    if ct in _SCALAR:
        return _SCALAR[ct]
    if hasattr(ct, "_length_"):
        return f"[{_rust_type(ct._type_)}; {ct._length_}]"
    if hasattr(ct, "contents") or "LP_" in ct.__name__:           # POINTER(x)
        return "*const " + (_SCALAR.get(ct._type_) or ct._type_.__name__)
    raise AssertionError(ct)


def test_ffi_rs_structs_match_the_ctypes_mirror():
    text = _read(SHIM, "src", "ffi.rs")
    structs = ["rtw_material", "rtw_noise", "rtw_texture", "rtw_sphere", "rtw_quad", "rtw_instance", "rtw_scene_desc",
               "rtw_camera", "rtw_render_opts", "rtw_stats", "rtw_scene_info"]
    for name in structs:
        m = re.search(r"pub struct %s\s*\{(.*?)\}" % name, text, re.S)
        assert m, name
        got = [(f.strip(), t.strip()) for f, t in re.findall(r"pub (\w+):\s*([^,}]+?)\s*(?:,|$)", m.group(1).replace("\n", " "))]
        want = [(f, _rust_type(t)) for f, t in getattr(A, name)._fields_]
        assert got == want, (name, got, want)


def test_versions_and_constants():
    text = _read(SHIM, "src", "ffi.rs")
    assert int(re.search(r"RTW_ABI_VERSION: u32 = (\d+)", text).group(1)) == A.RTW_ABI_VERSION
    assert f"(ABI version {A.RTW_ABI_VERSION})" in text
    hdr = _read(ROOT, "include", "rtw.h")
    assert int(re.search(r"#define RTW_ABI_VERSION (\d+)", hdr).group(1)) == A.RTW_ABI_VERSION
    for const in ("RTW_SPLIT_SAMPLES", "RTW_SPLIT_TILES", "RTW_FLAG_STRICT_FP", "RTW_FLAG_WAVEFRONT", "RTW_FLAG_COUNT_WORK",
                  "RTW_FLAG_MEGAKERNEL", "RTW_OUT_SUM", "RTW_SHAPE_TRIANGLE"):
        v = int(re.search(r"pub const %s: \w+ = (-?\d+);" % const, text).group(1))
        assert v == getattr(A, const), const
    # the entry points gpu.rs calls are declared
    gpu = _read(SHIM, "src", "gpu.rs")
    for fn in set(re.findall(r"\b(rtw_\w+)\s*\(", gpu)):
        assert fn in A.SYMBOLS, fn
