"""Generates tests/golden/ref_rust_*.npz from the two renders the reference ships with its Rust tree — the
only outputs of the reference's own Rust renderer that exist (no cargo/rustc here):

    Rust/Presentation.png   400x400, what `cargo run` writes: main.rs:89-419 `presentation_image()`
                            (room of quads, box instance, volume, glass / metal spheres; 2500 spp, gamma 2,
                            ray_color_bg_color)
    Rust/First frame.png    400x400: main.rs:427-561 through `test_run` (four spheres, ray_color_d, 100 spp)

Run in the build container (needs /root/reference):  python tests/golden/make_reference_images.py

The images are stochastic (unseeded thread_rng) and 8-bit, so the fixture keeps what a different sample
stream must reproduce: the 8x8 box-filtered image (50x50x3, f32 in [0,1]), the channel means, and the
fraction of exactly-black pixels (NaN-poisoned pixels quantise to 0: SURVEY N8). tests/test_oracle_pins.py
holds the oracle to them."""
import os

import numpy as np
from PIL import Image

HERE = os.path.dirname(os.path.abspath(__file__))
REF = "/root/reference/Rust"


def main():
    for name, png in (("presentation", "Presentation.png"), ("first_frame", "First frame.png")):
        a = np.asarray(Image.open(os.path.join(REF, png)).convert("RGB")).astype(np.float32) / np.float32(255.0)
        h, w, _ = a.shape
        assert (h, w) == (400, 400)
        box = a.reshape(h // 8, 8, w // 8, 8, 3).mean(axis=(1, 3)).astype(np.float32)
        np.savez_compressed(os.path.join(HERE, f"ref_rust_{name}.npz"), box8=box, mean=a.mean(axis=(0, 1)).astype(np.float32),
                            black_fraction=np.float32((a.sum(-1) == 0).mean()), shape=np.array([h, w]), source=np.array(f"Rust/{png}"))
        print("wrote", f"ref_rust_{name}.npz", box.shape, a.mean(axis=(0, 1)))


if __name__ == "__main__":
    main()
