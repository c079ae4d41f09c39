"""Output stage: the reference's `img_writer` module (Rust/src/write_img.rs:1-29).

write_img_f32 converts an `Img` (H x W x RGB f32, gamma already applied by the render) to 8 bits per
channel with `(x * 255).clamp(0, 255).round() as u8` (NaN -> 0, write_img.rs:12-14) and saves it.
Here the conversion runs on the GPU (librtw_b200.so: rtw_quantize, or fused behind the render with
Viewport.render_rgb8 / rtw_render_rgb8 so that only a quarter of the bytes crosses PCIe); the PNG
encoding (`img.save`, the `image` crate) stays on the host, done by PIL.
"""
from __future__ import annotations

import numpy as np

from .viewport import get_default_backend


def quantize(arr, backend=None) -> np.ndarray:
    """write_img.rs:12-14 per channel. H x W x 3 f32 -> H x W x 3 uint8."""
    be = backend or get_default_backend()
    a = np.asarray(arr, dtype=np.float32)
    if a.ndim != 3 or a.shape[2] != 3:
        raise ValueError("expected an H x W x 3 image")
    return be.quantize(a)


def write_img(arr_u8, filename: str) -> None:
    """write_img (write_img.rs:21-29): save an 8-bit RGB image; the format follows the file extension."""
    from PIL import Image
    a = np.ascontiguousarray(arr_u8, dtype=np.uint8)
    if a.ndim != 3 or a.shape[2] != 3:
        raise ValueError("expected an H x W x 3 uint8 image")
    Image.fromarray(a, mode="RGB").save(filename)


def write_img_f32(arr, filename: str, backend=None) -> None:
    """write_img_f32 (write_img.rs:6-19)."""
    write_img(quantize(arr, backend), filename)
