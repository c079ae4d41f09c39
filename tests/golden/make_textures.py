"""Generates tests/golden/assets/*.png and tests/golden/ref_textures.npz from the reference's own
texture assets. Run in the build container (needs /root/reference):

    python tests/golden/make_textures.py

  Rust/assets/earthmap.jpg   1024x512  the earth map of texture_test.rs:63,114 and of configs C3 / C5
  Rust/assets/default.png     128x64   the magenta / black checker (texture_test.rs:53, C3)
  Rust/assets/squares.png     128x64   the colour grid

The reference decodes with `image 0.24.6` (`ImageReader::open(path)?.decode()?.into_rgb32f()`,
texture.rs:215-226: u8 / 255, no sRGB linearisation). Here PIL decodes once, in this container; the
decoded u8 texels are committed as LOSSLESS PNGs (tests/golden/assets/), so that
`ImageTexture.from_path` on the GPU box — where /root/reference does not exist — yields exactly the
array the oracle and the CUDA path were checked with. PNG decode is exact; the JPEG's IDCT may differ
from jpeg-decoder 0.3.0 by +-1/255 on some texels (decode parity unpinned).
ref_textures.npz pins the decoded arrays: shape, sha256 of the u8 bytes, channel sums and 64 sampled
texels per image.
"""
import hashlib
import os

import numpy as np
from PIL import Image

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = "/root/reference/Rust/assets"
FILES = {"earthmap": "earthmap.jpg", "default": "default.png", "squares": "squares.png"}


def main():
    os.makedirs(os.path.join(HERE, "assets"), exist_ok=True)
    pins = {}
    for name, fn in FILES.items():
        a = np.asarray(Image.open(os.path.join(SRC, fn)).convert("RGB"), dtype=np.uint8)
        Image.fromarray(a, "RGB").save(os.path.join(HERE, "assets", name + ".png"), optimize=True)
        back = np.asarray(Image.open(os.path.join(HERE, "assets", name + ".png")).convert("RGB"), dtype=np.uint8)
        assert np.array_equal(a, back)
        h, w, _ = a.shape
        idx = (np.arange(64, dtype=np.int64) * 2654435761 % (h * w)).astype(np.int64)
        pins[name + "_shape"] = np.array([h, w], dtype=np.int64)
        pins[name + "_sha256"] = np.frombuffer(hashlib.sha256(a.tobytes()).digest(), dtype=np.uint8)
        pins[name + "_sums"] = a.reshape(-1, 3).astype(np.int64).sum(axis=0)
        pins[name + "_idx"] = idx
        pins[name + "_texels"] = a.reshape(-1, 3)[idx]
        print(name, a.shape, hashlib.sha256(a.tobytes()).hexdigest()[:16])
    np.savez_compressed(os.path.join(HERE, "ref_textures.npz"), **pins)


if __name__ == "__main__":
    main()
