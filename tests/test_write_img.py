"""Output stage: write_img_f32's f32 -> u8 conversion (Rust/src/write_img.rs:6-19).

CPU: the oracle restatement against hand-derived known answers of `(x*255).clamp(0,255).round() as u8`
     (Rust: round half away from zero, `as u8` saturates, NaN -> 0) and the reference's own
     `test_write` gradient (write_img.rs:36-57).
GPU: rtw_quantize / rtw_render_rgb8 through the C ABI against the oracle, bit-exact.
"""
import os

import numpy as np
import pytest

import rtw_b200
from rtw_b200 import _abi as A, scenes


def test_oracle_quantize_known_answers(oracle):
    x = np.array([[[0.0, 1.0, 0.5], [-0.3, 1.7, np.nan], [np.inf, -np.inf, 0.25],
                   [0.5 / 255.0, 1.5 / 255.0, 2.5 / 255.0], [0.998, 0.9981, 1e-9]]], dtype=np.float32)
    q = oracle.quantize(x)
    # 0.5*255 = 127.5 -> 128 (half away from zero); NaN -> 0; +inf -> 255; 0.25*255 = 63.75 -> 64
    assert q.dtype == np.uint8 and q.shape == x.shape
    assert q[0, 0].tolist() == [0, 255, 128]
    assert q[0, 1].tolist() == [0, 255, 0]
    assert q[0, 2].tolist() == [255, 0, 64]
    # k + 0.5 computed in f32: (0.5/255)*255 may land a hair under .5 — take the f32 product as the truth
    prod = (x[0, 3] * np.float32(255.0)).astype(np.float32)
    want = np.where(prod - np.floor(prod) >= 0.5, np.floor(prod) + 1, np.floor(prod)).astype(np.uint8)
    assert q[0, 3].tolist() == want.tolist()
    assert q[0, 4].tolist() == [254, 255, 0]            # 254.49 -> 254, 254.5155 -> 255


def test_oracle_quantize_reference_gradient(oracle):
    """write_img.rs:36-57 (`test_write`): the 256x256 gradient the reference's own test writes."""
    w = h = 256
    i = np.arange(w, dtype=np.float32)[None, :].repeat(h, 0) / np.float32(w - 1)
    j = np.arange(h, dtype=np.float32)[:, None].repeat(w, 1) / np.float32(h - 1)
    img = np.stack([i, j, np.full_like(i, 0.25)], axis=-1)
    q = oracle.quantize(img)
    # (k/255)*255 rounds back to k for every k in 0..255
    assert np.array_equal(q[0, :, 0], np.arange(256, dtype=np.uint8))
    assert np.array_equal(q[:, 0, 1], np.arange(256, dtype=np.uint8))
    assert (q[..., 2] == 64).all()


def test_write_img_roundtrip_png(tmp_path, oracle):
    from PIL import Image
    rng = np.random.default_rng(3)
    img = rng.random((17, 23, 3), dtype=np.float32) * 1.2 - 0.1
    p = tmp_path / "o.png"
    rtw_b200.write_img_f32(img, str(p), backend=oracle)
    back = np.asarray(Image.open(p).convert("RGB"))
    assert np.array_equal(back, oracle.quantize(img))


@pytest.mark.gpu
def test_gpu_quantize_bit_exact(oracle, cuda):
    rng = np.random.default_rng(11)
    img = (rng.random((211, 307, 3), dtype=np.float32) * 1.5 - 0.25).astype(np.float32)
    img[3, 5, 1] = np.nan; img[4, 4, 0] = np.inf; img[9, 9, 2] = -np.inf
    k = np.arange(256, dtype=np.float32)
    img[0, :256, 0] = (k + 0.5) / 255.0                 # the rounding ties
    img[1, :256, 0] = k / 255.0
    assert np.array_equal(cuda.quantize(img), oracle.quantize(img))
    assert np.array_equal(rtw_b200.quantize(img, backend=cuda), oracle.quantize(img))


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["C3", "C5"])
def test_gpu_render_rgb8_matches_quantized_render(name, oracle, cuda):
    """rtw_render_rgb8 == quantise(rtw_render) for the same seed, and the strict first-hit image
    quantises to the oracle's 8-bit image except where the f32 values straddle a rounding tie."""
    job = scenes.CONFIGS[name]()
    vp = type(job.viewport).new_from_res(160, 120, 4, job.viewport.depth, job.viewport.gamma, job.viewport.vfov,
                                         job.viewport.origin, job.viewport.direction, job.viewport.vup, None,
                                         job.viewport.lens_radius)
    vp.flags = A.RTW_FLAG_STRICT_FP
    f32 = vp._run(job.ray_color, job.scene, A.RTW_MODE_NO_RAND, cuda)
    u8 = vp.render_rgb8(job.ray_color, job.scene, A.RTW_MODE_NO_RAND, cuda)
    assert u8.dtype == np.uint8 and u8.shape == f32.shape
    assert np.array_equal(u8, oracle.quantize(f32))
    ref = oracle.quantize(vp._run(rtw_b200.ray_color_albedo, job.scene, A.RTW_MODE_NO_RAND, oracle))
    got = vp.render_rgb8(rtw_b200.ray_color_albedo, job.scene, A.RTW_MODE_NO_RAND, cuda)
    assert (np.abs(got.astype(int) - ref.astype(int)) > 1).mean() < 2e-3
    assert (got != ref).mean() < 2e-2


def test_driver_test_run_and_frame_dump(tmp_path, oracle):
    """main.rs:61-87 `test_run` + the frame dump of main.rs:563-567, single process: the log lines, frame 0 in
    `f_name`, every frame as img%05d.png, all quantised by write_img.rs:12-14."""
    from PIL import Image
    from rtw_b200 import driver, scenes
    anim = scenes.ltr_scene(samples=2, width=32, height=18, frames=3)
    anim.viewport.start_frame = 2
    lines = []
    f0 = str(tmp_path / "First frame.png")
    video = driver.test_run(f0, anim.viewport, anim.ray_color, anim.scene, backend=oracle,
                            frames_dir=str(tmp_path / "video"), log=lines.append)
    assert len(video) == 3 and lines[0] == "Rendering %d samples" % (32 * 18 * 2 * 3)
    assert lines[1].startswith("Finished\nElapsed time: ") and lines[1].endswith("num of frames: 3")
    h = oracle.scene(anim.scene)
    want = h.render_frames(anim.viewport.camera(oracle), anim.viewport.opts(anim.ray_color, A.RTW_MODE_RANDOM), 2, 3)
    assert np.array_equal(np.stack(video), want, equal_nan=True)
    assert not np.array_equal(want[0], want[1])                                # the spheres move between frames
    assert sorted(os.listdir(tmp_path / "video")) == ["img00000.png", "img00001.png", "img00002.png"]
    for k in range(3):
        assert np.array_equal(np.asarray(Image.open(driver.frame_name(str(tmp_path / "video"), k))), oracle.quantize(want[k]))
    assert np.array_equal(np.asarray(Image.open(f0)), oracle.quantize(want[0]))


def test_driver_cli_needs_the_gpu_library():
    """`python rtw_b200.py` == `cargo run` (main.rs:421-423); without a device it fails loudly, no CPU fallback."""
    import subprocess, sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    r = subprocess.run([sys.executable, os.path.join(root, "rtw_b200.py"), "--help"], capture_output=True, text=True)
    assert r.returncode == 0 and "--frames-dir" in r.stdout
    import rtw_b200
    if rtw_b200.get_default_backend().device_count() == 0:
        r = subprocess.run([sys.executable, os.path.join(root, "rtw_b200.py"), "triangle", "--samples", "1", "--out", os.devnull],
                           capture_output=True, text=True)
        assert r.returncode != 0 and "Traceback" in r.stderr


@pytest.mark.gpu
def test_gpu_driver_frame_dump(tmp_path, oracle, cuda):
    """test_run on the device: frames = rtw_render_frames, files = rtw_quantize of the frames it returned
    (byte-exact against the oracle's write_img.rs:12-14)."""
    from PIL import Image
    from rtw_b200 import driver, scenes
    anim = scenes.ltr_scene(samples=4, width=64, height=36, frames=3)
    anim.viewport.flags = A.RTW_FLAG_STRICT_FP
    video = driver.test_run(str(tmp_path / "f0.png"), anim.viewport, anim.ray_color, anim.scene, backend=cuda,
                            frames_dir=str(tmp_path / "video"), log=lambda s: None)
    h = cuda.scene(anim.scene)
    want = h.render_frames(anim.viewport.camera(cuda), anim.viewport.opts(anim.ray_color, A.RTW_MODE_RANDOM), 0, 3)
    # a second render of the same frames differs only by the order of the f32 adds into each pixel
    assert np.allclose(np.stack(video), want, rtol=1e-5, atol=1e-6, equal_nan=True)
    assert not np.allclose(want[0], want[1], rtol=1e-3, atol=1e-3)              # the spheres move between frames
    for k in range(3):
        assert np.array_equal(np.asarray(Image.open(driver.frame_name(str(tmp_path / "video"), k))), oracle.quantize(video[k]))
    assert np.array_equal(np.asarray(Image.open(str(tmp_path / "f0.png"))), oracle.quantize(video[0]))
