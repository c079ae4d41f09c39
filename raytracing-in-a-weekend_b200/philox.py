"""Philox4x32-10 (Salmon, Moraes, Dror, Shaw — SC'11) in numpy, for scene authoring only.

The reference builds its random scenes and Perlin tables from rand 0.8.5's unseeded thread_rng
(texture.rs:111-125, aabb.rs:42); the harness needs the same scene on every run and on every rank,
so scene-building randomness is a Philox stream with an explicit key (SURVEY §8d "Scene RNG").
The device kernels carry their own Philox implementation (csrc/rng.cuh); this one never runs on
the render path.
"""
from __future__ import annotations

import numpy as np

_M0, _M1 = np.uint64(0xD2511F53), np.uint64(0xCD9E8D57)
_W0, _W1 = 0x9E3779B9, 0xBB67AE85
_MASK = np.uint64(0xFFFFFFFF)


def philox4x32_10(ctr: np.ndarray, key) -> np.ndarray:
    """ctr: (..., 4) uint32 counters; key: (k0, k1). Returns (..., 4) uint32."""
    c = np.asarray(ctr, dtype=np.uint32).astype(np.uint64)
    c0, c1, c2, c3 = c[..., 0], c[..., 1], c[..., 2], c[..., 3]
    k0, k1 = int(key[0]) & 0xFFFFFFFF, int(key[1]) & 0xFFFFFFFF
    for _ in range(10):
        p0 = _M0 * c0
        p1 = _M1 * c2
        hi0, lo0 = p0 >> np.uint64(32), p0 & _MASK
        hi1, lo1 = p1 >> np.uint64(32), p1 & _MASK
        c0, c1, c2, c3 = hi1 ^ c1 ^ np.uint64(k0), lo1, hi0 ^ c3 ^ np.uint64(k1), lo0
        k0 = (k0 + _W0) & 0xFFFFFFFF
        k1 = (k1 + _W1) & 0xFFFFFFFF
    return np.stack([c0, c1, c2, c3], axis=-1).astype(np.uint32)


def u32_to_f32(w: np.ndarray) -> np.ndarray:
    """rand 0.8.5 Standard f32: 24 random bits, [0, 1)."""
    return ((np.asarray(w, dtype=np.uint32) >> np.uint32(8)).astype(np.float32) * np.float32(1.0 / 16777216.0))


class PhiloxStream:
    """Sequential stream of f32 uniforms: counter = (n, 0, 0, 0), key = (key, 0)."""

    def __init__(self, key: int):
        self.key = (int(key) & 0xFFFFFFFF, (int(key) >> 32) & 0xFFFFFFFF)
        self.n = 0            # next u32 index
        self._buf = np.zeros(0, dtype=np.uint32)

    def u32(self, count: int) -> np.ndarray:
        first_blk, last_blk = self.n // 4, (self.n + count + 3) // 4
        ctr = np.zeros((last_blk - first_blk, 4), dtype=np.uint32)
        ctr[:, 0] = np.arange(first_blk, last_blk, dtype=np.uint64).astype(np.uint32)
        words = philox4x32_10(ctr, self.key).reshape(-1)
        off = self.n - first_blk * 4
        self.n += count
        return words[off:off + count]

    def uniform_f32(self, count: int) -> np.ndarray:
        return u32_to_f32(self.u32(count))

    def random(self) -> float:
        """one `random::<f32>()`"""
        return float(self.uniform_f32(1)[0])

    def uniform(self, lo: float, hi: float) -> float:
        return float(np.float32(self.random()) * (np.float32(hi) - np.float32(lo)) + np.float32(lo))
