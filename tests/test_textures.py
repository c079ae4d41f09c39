"""The reference's own image assets on the texture path (texture.rs:215-226,259-267).

Rust/assets/earthmap.jpg, default.png and squares.png are decoded once in the build container by
tests/golden/make_textures.py and committed as lossless PNGs (tests/golden/assets/) with their pins
(tests/golden/ref_textures.npz). Here they go through ImageTexture.from_path, as in
viewport/texture_test.rs:63,114, and
  * the decoded texels are held against the pins (and, where /root/reference exists, against a fresh
    PIL decode of the reference's files);
  * a quad that fills the viewport exactly, textured with an asset and rendered with one centre ray per
    pixel (render_no_rand) and the albedo closure, must return the asset texel for texel (quad.rs:66-76:
    x = floor(alfa*row), y = floor(beta*col); texture.rs:265: img[y*row + x]) — bit for bit, on the
    oracle, the emulated kernel headers and the CUDA path (f32 texels -> float4 CUDA array -> tex2D,
    point filter), strict and fast builds, both pipelines.
"""
import hashlib
import os

import numpy as np
import pytest

import rtw_b200
from rtw_b200 import _abi as A, scenes
from rtw_b200.scene import SCATTER_M, ImageTexture, Quad, Scene, Vec3
from rtw_b200.viewport import Viewport, ray_color_albedo

HERE = os.path.dirname(os.path.abspath(__file__))
PINS = np.load(os.path.join(HERE, "golden", "ref_textures.npz"))
ASSETS = ("earthmap", "default", "squares")
REF_FILES = {"earthmap": "earthmap.jpg", "default": "default.png", "squares": "squares.png"}


def _u8(tex: ImageTexture) -> np.ndarray:
    a = np.round(tex.img.astype(np.float64) * 255.0).astype(np.uint8)
    assert np.array_equal((a.astype(np.float32) / np.float32(255.0)).astype(np.float32), tex.img)   # into_rgb32f: u8 / 255
    return a.reshape(tex.col, tex.row, 3)


@pytest.mark.parametrize("name", ASSETS)
def test_from_path_decodes_the_reference_assets(name):
    tex = ImageTexture.from_path(scenes.asset_path(name))
    h, w = (int(x) for x in PINS[name + "_shape"])
    assert (tex.row, tex.col) == (w, h)                       # texture.rs:219-221: row = width, col = height
    a = _u8(tex)
    assert hashlib.sha256(a.tobytes()).digest() == PINS[name + "_sha256"].tobytes()
    assert np.array_equal(a.reshape(-1, 3).astype(np.int64).sum(axis=0), PINS[name + "_sums"])
    assert np.array_equal(a.reshape(-1, 3)[PINS[name + "_idx"]], PINS[name + "_texels"])
    ref = os.path.join("/root/reference/Rust/assets", REF_FILES[name])
    if os.path.exists(ref):                                   # build container only
        from PIL import Image
        assert np.array_equal(a, np.asarray(Image.open(ref).convert("RGB"), dtype=np.uint8))


def test_configs_carry_the_reference_assets():
    """C3 and C5 as BASELINE.json names them: the earth sphere wears Rust/assets/earthmap.jpg, C3's fourth
    sphere Rust/assets/default.png."""
    earth = ImageTexture.from_path(scenes.asset_path("earthmap"))
    for job, idx in ((scenes.config_c3(samples=1), 2), (scenes.config_c5(samples=1), 4)):
        t = job.scene.spheres[idx].texture
        assert (t.row, t.col) == (1024, 512) and np.array_equal(t.img, earth.img)
    t = scenes.config_c3(samples=1).scene.spheres[3].texture
    assert (t.row, t.col) == (128, 64) and np.array_equal(t.img, ImageTexture.from_path(scenes.asset_path("default")).img)
    assert "reference texture assets" in scenes.texture_data_label()


def _billboard(name):
    """A quad that coincides with the viewport rectangle at distance 1 (vfov 90: height 2, width 2*aspect): the
    centre ray of pixel (i, j) meets it at alfa = (i + 0.5) / W, beta = (j + 0.5) / H."""
    tex = ImageTexture.from_path(scenes.asset_path(name))
    W, H = tex.row, tex.col
    asp = W / H
    q = Quad.new(Vec3(-asp, 1.0, -1.0), Vec3(2.0 * asp, 0.0, 0.0), Vec3(0.0, -2.0, 0.0), SCATTER_M, (0.0, 0.0, 0.0), tex)
    vp = Viewport.new_from_res(W, H, 1, 1, 1.0, 90.0, Vec3(0.0, 0.0, 0.0), Vec3(0.0, 0.0, -1.0), Vec3(0.0, 1.0, 0.0))
    return Scene.new_quad([q]), vp, tex.img.reshape(H, W, 3)


@pytest.mark.parametrize("name", ASSETS)
def test_oracle_and_headers_return_the_asset_texel_for_texel(name, oracle):
    from emu_backend import EmuBackend
    sc, vp, want = _billboard(name)
    for be in (oracle, EmuBackend()):
        got = vp._run(ray_color_albedo, sc, A.RTW_MODE_NO_RAND, be)
        assert np.array_equal(got, want)


@pytest.mark.gpu
@pytest.mark.parametrize("flags", [A.RTW_FLAG_STRICT_FP | A.RTW_FLAG_MEGAKERNEL, A.RTW_FLAG_STRICT_FP | A.RTW_FLAG_WAVEFRONT,
                                   A.RTW_FLAG_MEGAKERNEL, A.RTW_FLAG_WAVEFRONT],
                         ids=["strict-mega", "strict-wavefront", "fast-mega", "fast-wavefront"])
@pytest.mark.parametrize("name", ASSETS)
def test_cuda_texture_objects_return_the_asset_texel_for_texel(name, flags, cuda):
    """from_path -> rtw_scene_create (float4 cudaArray + texture object) -> tex2D in shade: every texel comes back."""
    sc, vp, want = _billboard(name)
    vp.flags = flags
    got = vp._run(ray_color_albedo, sc, A.RTW_MODE_NO_RAND, cuda)
    assert np.array_equal(got, want)
    assert cuda.scene(sc).stats().paths == want.shape[0] * want.shape[1]
