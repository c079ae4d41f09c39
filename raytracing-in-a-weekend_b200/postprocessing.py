"""Post-processing stage: the Rust2 tree's `postprocessing` module (Rust2/src/postprocessing.rs:1-135).

`bilateral_filter(img, Proximity::new(size, type))` takes and returns an 8-bit RGB image. The filter
itself runs on the GPU (librtw_b200.so: rtw_bilateral_filter, csrc/rtw_post.cu); this module mirrors
the reference's names and argument meaning. There is no CPU fallback.

The window is the reference's, quirks included (postprocessing.rs:30-61): `left..right` and
`up..down` are half-open, so the window of (x, y) is [x-size, x+size) x [y-size, y+size) clipped to the
image; ProximityType.Edges keeps |dx| + |dy| < size inside it.
"""
from __future__ import annotations

import enum
from typing import List, Tuple

import numpy as np

from .viewport import get_default_backend


class ProximityType(enum.IntEnum):       # postprocessing.rs:11-14
    Square = 0
    Edges = 1


class Proximity:                         # postprocessing.rs:16-28
    def __init__(self, size: int, proximity_type: ProximityType = ProximityType.Square):
        if size < 0:
            raise ValueError("size is a u32")
        self.size = int(size)
        self._type = ProximityType(proximity_type)

    @staticmethod
    def new(size: int, proximity_type: ProximityType) -> "Proximity":
        return Proximity(size, proximity_type)

    def get_pixels(self, x: int, y: int, width: int, height: int) -> List[Tuple[int, int]]:
        """postprocessing.rs:30-61, same order (x outer, y inner). Host-side mirror for callers and tests;
        the kernel walks the same ranges without materialising the list."""
        s = self.size
        left, up = x - min(x, s), y - min(y, s)
        right, down = x + min(width - x - 1, s), y + min(height - y - 1, s)
        out = []
        for i in range(left, right):
            for j in range(up, down):
                if self._type == ProximityType.Square or abs(i - x) + abs(j - y) < s:
                    out.append((i, j))
        return out


def bilateral_filter(img, proximity: Proximity, backend=None, info: dict | None = None) -> np.ndarray:
    """postprocessing.rs:70-135. `img`: H x W x 3 uint8 (what write_img.quantize / Viewport.render_rgb8 return).
    Returns a new H x W x 3 uint8 image."""
    be = backend or get_default_backend()
    a = np.asarray(img)
    if a.dtype != np.uint8 or a.ndim != 3 or a.shape[2] != 3:
        raise ValueError("expected an H x W x 3 uint8 image")
    return be.bilateral_filter(a, proximity.size, int(proximity._type), info)
