"""Post-processing stage (SURVEY 8f-4 tail): the Rust2 tree's 8-bit bilateral filter
(Rust2/src/postprocessing.rs:11-135).

CPU: oracle/postprocessing.hpp against an independent pure-Python restatement of the reference (small
     images, both proximity types, sizes 0..4), the window quirks of Proximity::get_pixels, the flat-image
     behaviour, and the libm-expf / rounded-double-exp modes against each other.
GPU: rtw_bilateral_filter through the C ABI against the oracle, byte for byte.
The reference's own test of this function writes a PNG and asserts nothing (postprocessing.rs:396-449).
"""
import math

import numpy as np
import pytest

import rtw_b200
from rtw_b200 import Proximity, ProximityType

F = np.float32


def _as_u8(v) -> int:                      # Rust `f32 as u8`
    v = float(v)
    if math.isnan(v) or v <= 0.0:
        return 0
    return 255 if v >= 255.0 else int(v)


def _intensity(p):                         # postprocessing.rs:64-68, f32 step by step
    return (F(0.2989) * F(p[0]) / F(255.0) + F(0.5870) * F(p[1]) / F(255.0)) + F(0.1140) * F(p[2]) / F(255.0)


def py_bilateral(img: np.ndarray, prox: Proximity) -> np.ndarray:
    """The reference, statement by statement, with numpy f32 scalars (small images only)."""
    h, w, _ = img.shape
    with np.errstate(all="ignore"):
        spatial = F(np.ceil(F(0.02) * np.sqrt(F(w * w + h * h))))
        gsum = F(0.0)
        for y in range(1, h - 1):
            for x in range(1, w - 1):
                iu, idn = _intensity(img[y - 1, x]), _intensity(img[y + 1, x])
                ir, il = _intensity(img[y, x + 1]), _intensity(img[y, x - 1])
                gsum = F(gsum + np.sqrt(F((iu - idn) * (iu - idn) + (il - ir) * (il - ir))))
        avg = gsum / F((w - 2) * (h - 2))
        inv_range = F(0.5) / F(avg * avg)
        inv_spatial = F(0.5) / F(spatial * spatial)
        out = np.zeros_like(img)
        for y in range(h):
            for x in range(w):
                wsum = [F(0.0)] * 3
                csum = [F(0.0)] * 3
                for (xi, yi) in prox.get_pixels(x, y, w, h):
                    for c in range(3):
                        t = F(int(img[yi, xi, c]) - int(img[y, x, c])) / F(255.0)
                        arg = F(F(-inv_spatial) * F((xi - x) ** 2 + (yi - y) ** 2)) - F(inv_range * F(t * t))
                        wgt = F(np.exp(np.float64(arg)))       # correctly rounded f32 exp
                        csum[c] = F(csum[c] + F(F(F(img[yi, xi, c]) * wgt) / F(255.0)))
                        wsum[c] = F(wsum[c] + wgt)
                for c in range(3):
                    out[y, x, c] = _as_u8(F(F(csum[c] * F(255.0)) / wsum[c]))
    return out


def _noisy_steps(h, w, seed):
    """Piecewise-flat image plus noise: what the filter is for (edges to keep, noise to smooth)."""
    rng = np.random.default_rng(seed)
    base = np.zeros((h, w, 3), dtype=np.float64)
    base[:, : w // 2] = (60, 120, 200)
    base[:, w // 2:] = (220, 180, 40)
    base[h // 3: 2 * h // 3, w // 4: 3 * w // 4] = (30, 200, 90)
    return np.clip(base + rng.normal(0, 12, base.shape), 0, 255).astype(np.uint8)


def test_get_pixels_window_quirks():
    sq = Proximity.new(2, ProximityType.Square)
    # interior: [x-2, x+2) x [y-2, y+2), x outer
    px = sq.get_pixels(5, 5, 20, 20)
    assert px[0] == (3, 3) and px[1] == (3, 4) and px[-1] == (6, 6) and len(px) == 16
    assert (7, 5) not in px and (5, 7) not in px
    # last column: the pixel itself is outside its own window
    assert (19, 5) not in sq.get_pixels(19, 5, 20, 20)
    assert sq.get_pixels(19, 19, 20, 20) == [(17, 17), (17, 18), (18, 17), (18, 18)]
    # corner clips on the low side only
    assert sq.get_pixels(0, 0, 20, 20) == [(0, 0), (0, 1), (1, 0), (1, 1)]
    ed = Proximity.new(2, ProximityType.Edges)
    assert ed.get_pixels(5, 5, 20, 20) == [(4, 5), (5, 4), (5, 5), (5, 6), (6, 5)]
    assert Proximity.new(0, ProximityType.Square).get_pixels(5, 5, 20, 20) == []


@pytest.mark.parametrize("kind", [ProximityType.Square, ProximityType.Edges])
@pytest.mark.parametrize("size", [0, 1, 2, 4])
def test_oracle_matches_python_restatement(oracle, kind, size):
    img = _noisy_steps(9, 12, seed=size + 10 * int(kind))
    want = py_bilateral(img, Proximity.new(size, kind))
    got = oracle.bilateral_filter(img, size, int(kind), exp_mode=1)
    assert np.array_equal(got, want)


def test_oracle_params_and_smoothing(oracle):
    img = _noisy_steps(48, 64, seed=5)
    info = {}
    out = oracle.bilateral_filter(img, 4, 0, info)
    assert info["spatial"] == math.ceil(0.02 * math.sqrt(64 * 64 + 48 * 48))       # ceil(1.6) = 2
    assert info["inv_spatial"] == pytest.approx(0.5 / 4.0)
    assert info["avg_gradient"] > 0 and info["inv_range"] == pytest.approx(0.5 / info["avg_gradient"] ** 2, rel=1e-6)
    # noise goes down inside the flat regions, the vertical edge at w/2 stays an edge
    flat = (slice(2, 12), slice(2, 12))
    assert out[flat].astype(float).std(axis=(0, 1)).mean() < 0.8 * img[flat].astype(float).std(axis=(0, 1)).mean()
    assert abs(float(out[5, 28, 0]) - float(out[5, 36, 0])) > 100


def test_oracle_flat_image_comes_out_black(oracle):
    """avg_gradient = 0 -> inv_range = inf -> inf * 0 = NaN in every weight -> `as u8` gives 0 (reference behaviour)."""
    img = np.full((7, 9, 3), 77, dtype=np.uint8)
    assert not oracle.bilateral_filter(img, 2, 0).any()
    assert not oracle.bilateral_filter(img, 2, 1).any()
    # size 0: empty window, 0/0
    assert not oracle.bilateral_filter(_noisy_steps(7, 9, 1), 0, 0).any()


def test_oracle_exp_modes_agree_to_one_level(oracle):
    """libm expf (what f32::exp calls) against the rounded double exp the CUDA path uses: weights differ by at most
    an ulp now and then, which moves a truncated byte by at most one level, rarely."""
    img = _noisy_steps(60, 80, seed=9)
    a = oracle.bilateral_filter(img, 5, 0, exp_mode=0).astype(int)
    b = oracle.bilateral_filter(img, 5, 0, exp_mode=1).astype(int)
    assert np.abs(a - b).max() <= 1
    assert (a != b).mean() < 1e-3


def test_host_mirror_rejects_bad_images(oracle):
    with pytest.raises(ValueError):
        rtw_b200.bilateral_filter(np.zeros((4, 4, 3), dtype=np.float32), Proximity(1), backend=oracle)
    with pytest.raises(ValueError):
        rtw_b200.bilateral_filter(np.zeros((4, 4), dtype=np.uint8), Proximity(1), backend=oracle)
    out = rtw_b200.bilateral_filter(_noisy_steps(8, 8, 2), Proximity.new(2, ProximityType.Edges), backend=oracle)
    assert out.shape == (8, 8, 3) and out.dtype == np.uint8


# ----------------------------------------------------------------------------- GPU
@pytest.mark.gpu
@pytest.mark.parametrize("shape,size,kind", [
    ((3, 3), 1, 0), ((2, 2), 1, 0), ((2, 9), 2, 1), ((9, 12), 0, 0), ((9, 12), 4, 0), ((9, 12), 4, 1),
    ((37, 53), 3, 0), ((37, 53), 7, 1), ((64, 48), 64, 0), ((30, 40), 70, 0), ((30, 40), 70, 1),
    ((150, 200), 10, 0), ((150, 200), 10, 1),
])
def test_gpu_bilateral_matches_oracle(oracle, cuda, shape, size, kind):
    img = _noisy_steps(shape[0], shape[1], seed=shape[0] * 131 + size)
    oi, gi = {}, {}
    want = oracle.bilateral_filter(img, size, kind, oi)
    got = cuda.bilateral_filter(img, size, kind, gi)
    for k in ("spatial", "avg_gradient", "inv_range", "inv_spatial"):          # the ordered f32 sum, bit for bit
        assert np.float32(oi[k]).tobytes() == np.float32(gi[k]).tobytes() or (math.isnan(oi[k]) and math.isnan(gi[k])), k
    assert np.array_equal(got, want)


@pytest.mark.gpu
def test_gpu_bilateral_flat_and_random(oracle, cuda):
    flat = np.full((20, 31, 3), 200, dtype=np.uint8)
    assert not cuda.bilateral_filter(flat, 3, 0).any()
    rng = np.random.default_rng(4)
    img = rng.integers(0, 256, (90, 70, 3), dtype=np.uint8)
    assert np.array_equal(cuda.bilateral_filter(img, 6, 0), oracle.bilateral_filter(img, 6, 0))
    assert np.array_equal(cuda.bilateral_filter(img, 6, 1), oracle.bilateral_filter(img, 6, 1))


@pytest.mark.gpu
def test_gpu_bilateral_of_a_render(oracle, cuda):
    """The reference's use (postprocessing.rs:396-449): render, quantise, filter with Proximity::new(10, Square)."""
    from rtw_b200 import scenes
    job = scenes.config_c1(width=200, height=112, samples=8)
    img8 = job.viewport.render_rgb8(job.ray_color, job.scene, backend=cuda)
    info = {}
    got = rtw_b200.bilateral_filter(img8, Proximity.new(10, ProximityType.Square), backend=cuda, info=info)
    want = oracle.bilateral_filter(img8, 10, 0)
    assert np.array_equal(got, want)
    assert info["kernel_ms"] > 0
    # it does denoise: the sky gradient is smoother after the filter than before
    sky = (slice(2, 20), slice(60, 140))
    assert np.abs(np.diff(got[sky].astype(int), axis=1)).mean() <= np.abs(np.diff(img8[sky].astype(int), axis=1)).mean()


@pytest.mark.gpu
def test_gpu_bilateral_errors(cuda):
    from rtw_b200 import _abi as A
    img = np.zeros((1, 5, 3), dtype=np.uint8)
    with pytest.raises(A.RtwError):
        cuda.bilateral_filter(img, 1, 0)
    with pytest.raises(A.RtwError):
        cuda.bilateral_filter(np.zeros((4, 4, 3), dtype=np.uint8), 1, 7)


@pytest.mark.gpu
@pytest.mark.parametrize("pattern", ["binary_noise", "stripes", "checker", "sparse", "ramp", "tall", "wide"])
def test_gpu_ordered_gradient_sum_is_the_sequential_one(oracle, cuda, pattern):
    """sigma_range comes from one f32 accumulated over the interior pixels in row-major order (postprocessing.rs:75-87).
    The device takes that sum 1024 values at a time as integers of the running sum's binade (rtw_post.cu) and falls
    back to plain adds on ties and binade changes: images with few distinct gradient values (many exact ties), long
    runs of zeros, and extreme shapes must give the same bits as the sequential CPU sum."""
    rng = np.random.default_rng(17)
    h, w = 240, 320
    if pattern == "binary_noise":
        img = (rng.integers(0, 2, (h, w, 1), dtype=np.uint8) * 255).repeat(3, axis=2)
    elif pattern == "stripes":
        img = np.zeros((h, w, 3), dtype=np.uint8); img[:, ::4] = 128; img[:, 1::4] = 64
    elif pattern == "checker":
        yy, xx = np.mgrid[0:h, 0:w]; img = (((yy // 3 + xx // 5) % 2) * 200 + 20).astype(np.uint8)[..., None].repeat(3, axis=2)
    elif pattern == "sparse":
        img = np.full((h, w, 3), 10, dtype=np.uint8); img[rng.integers(0, h, 50), rng.integers(0, w, 50)] = 250
    elif pattern == "ramp":
        img = (np.arange(w, dtype=np.uint16)[None, :, None] * 255 // (w - 1)).astype(np.uint8).repeat(h, axis=0).repeat(3, axis=2)
    elif pattern == "tall":
        img = rng.integers(0, 256, (2000, 3, 3), dtype=np.uint8)
    else:
        img = rng.integers(0, 256, (3, 3000, 3), dtype=np.uint8)
    img = np.ascontiguousarray(img)
    oi, gi = {}, {}
    want = oracle.bilateral_filter(img, 2, 0, oi)
    got = cuda.bilateral_filter(img, 2, 0, gi)
    assert np.float32(oi["avg_gradient"]).tobytes() == np.float32(gi["avg_gradient"]).tobytes()
    assert np.array_equal(got, want)


# ----------------------------------------------------------------------------- the grouped ordered sum, modelled on the CPU
def _f32_bits(f):
    return int(np.float32(f).view(np.uint32))


def _grouped_sum_model(xs: np.ndarray, group: int):
    """rtw_post.cu post_params_kernel, statement for statement: the running f32 sum advances by a group's values as
    integers of its binade when no value ties / is out of range and the total stays in the binade; otherwise the
    group is added sequentially. Returns (sum, groups on the fast path, groups on the sequential path)."""
    s = F(0.0)
    fast = slow = 0
    for g0 in range(0, len(xs), group):
        grp = xs[g0:g0 + group]
        sb = _f32_bits(s)
        ex = (sb >> 23) & 0xFF
        ok = 23 < ex < 0xFE
        if ok:
            scale = np.uint32((277 - ex) << 23).view(F)
            unit = np.uint32((ex - 23) << 23).view(F)
            i0 = (sb & 0x7FFFFF) | 0x800000
            with np.errstate(all="ignore"):
                y = (grp * scale).astype(F)
                t = (y + F(8388608.0)).astype(F)
                d = (y - (t - F(8388608.0)).astype(F)).astype(F)
            good = bool(((y >= 0) & (y < F(2097152.0)) & (np.abs(d) != F(0.5))).all())
            total = int((t.view(np.uint32) & np.uint32(0x7FFFFF)).astype(np.uint64).sum())
            if good and i0 + total <= 0x1000000:
                s = F(F(i0 + total) * unit)
                fast += 1
                continue
        for x in grp:
            s = F(s + x)
        slow += 1
    return s, fast, slow


@pytest.mark.parametrize("kind", ["uniform", "ties", "zeros", "heavy_tail", "two_scales", "with_nan"])
def test_grouped_ordered_sum_model_equals_sequential_sum(kind):
    rng = np.random.default_rng(23)
    took_fast = 0
    with np.errstate(all="ignore"):
        for trial in range(6):
            n = int(rng.integers(1, 9000))
            if kind == "uniform":
                xs = (rng.random(n) * 1.5).astype(F)
            elif kind == "ties":
                xs = (rng.integers(0, 64, n) / 64.0).astype(F)
            elif kind == "zeros":
                xs = np.where(rng.random(n) < 0.7, 0.0, rng.random(n)).astype(F)
            elif kind == "heavy_tail":
                xs = (rng.random(n) ** 8 * 1e3).astype(F)
            elif kind == "two_scales":
                xs = (rng.random(n) * 1e-6).astype(F); xs[n // 2:] = rng.random(n - n // 2).astype(F)
            else:
                xs = rng.random(n).astype(F); xs[n // 3] = np.nan
            want = F(0.0)
            for x in xs:
                want = F(want + x)
            got, fast, slow = _grouped_sum_model(xs, int(rng.choice([32, 256, 1024])))
            assert _f32_bits(got) == _f32_bits(want) or (np.isnan(got) and np.isnan(want))
            took_fast += fast
    assert took_fast > 0 or kind == "with_nan"
