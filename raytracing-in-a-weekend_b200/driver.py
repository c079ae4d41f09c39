"""Animation driver: the reference's `test_run` (Rust/src/main.rs:61-87), the frame dump
it keeps commented out (main.rs:563-567: `/tmp/video/img{:0>5}.png` per frame, then ffmpeg), and
`main()` itself (main.rs:421-423 -> `presentation_image`, main.rs:89-419 -> "Presentation.png").

`test_run` renders `viewport.number_of_frames` frames through `render_multi` — one process: frame
after frame on its GPU; under `torch.distributed` (one process per GPU): frame k on rank k mod n
(`sharding.render_multi_sharded`) — writes frame 0 to `f_name` and returns all frames. The 8-bit
conversion of every dumped frame runs on the device (`rtw_quantize`); PNG encoding stays on the host
(PIL), as in write_img.py. No ffmpeg on the box: `dump_frames` stops at the numbered PNGs.

    python rtw_b200.py                      # == `cargo run`: Presentation.png (400x400, 2500 spp)
    python rtw_b200.py ltr --frames 20 --frames-dir /tmp/video --samples 100
    torchrun --nproc-per-node 8 rtw_b200.py C5 --frames 8 --frames-dir /tmp/video   # one frame per GPU
"""
from __future__ import annotations

import os
import time
from typing import List, Optional

import numpy as np

from .viewport import render_multi
from .write_img import quantize, write_img, write_img_f32


def frame_name(directory: str, index: int) -> str:
    """main.rs:565: format!("/tmp/video/img{:0>5}.png", i)."""
    return os.path.join(directory, "img{:0>5}.png".format(index))


def dump_frames(frames: List[np.ndarray], directory: str, backend=None) -> List[str]:
    """main.rs:563-567 (`mkdir -p` + one write_img_f32 per frame); returns the file names."""
    os.makedirs(directory, exist_ok=True)
    names = []
    for i, img in enumerate(frames):
        names.append(frame_name(directory, i))
        write_img(quantize(img, backend), names[-1])
    return names


def _distributed():
    try:
        import torch.distributed as dist
    except ImportError:
        return None
    return dist if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1 else None


def test_run(f_name: str, viewport, ray_color, scene, backend=None, frames_dir: Optional[str] = None,
             log=print) -> Optional[List[np.ndarray]]:
    """main.rs:61-87. Returns the frames (`Vec<Img>`); under torch.distributed only rank 0 gets them
    (and writes the files), the other ranks return None."""
    log("Rendering {} samples".format(viewport.height * viewport.width * viewport.samples * viewport.number_of_frames))
    before = time.perf_counter()
    dist = _distributed()
    if dist is not None:
        from .sharding import render_multi_sharded
        img = render_multi_sharded(viewport, ray_color, scene, backend)
        if img is None:
            return None
    else:
        img = render_multi(viewport, ray_color, scene, backend)
    if img:
        write_img_f32(img[0], f_name, backend)
    if frames_dir is not None:
        dump_frames(img, frames_dir, backend)
    log("Finished\nElapsed time: {:.6f}s, num of frames: {}".format(time.perf_counter() - before, len(img)))
    return img


__test__ = False  # `test_run` is the reference's name, not a pytest test


def main(argv=None) -> int:
    """`cargo run` (main.rs:421-423) and, with a scene name, any job of scenes.py."""
    import argparse
    from . import scenes
    from .viewport import async_render
    jobs = dict(scenes.REFERENCE_TESTS)
    jobs.update(scenes.CONFIGS)
    jobs["ltr"] = scenes.ltr_scene
    ap = argparse.ArgumentParser(prog="rtw_b200.py", description=main.__doc__)
    ap.add_argument("scene", nargs="?", default="presentation", choices=sorted(jobs))
    ap.add_argument("--out", default=None, help="image of frame 0 (default: Presentation.png / <scene>.png)")
    ap.add_argument("--samples", type=int, default=None)
    ap.add_argument("--frames", type=int, default=None, help="number_of_frames: render through test_run / render_multi")
    ap.add_argument("--start-frame", type=int, default=0)
    ap.add_argument("--frames-dir", default=None, help="dump every frame as img%%05d.png (main.rs:563-567)")
    a = ap.parse_args(argv)
    job = jobs[a.scene]()
    vp = job.viewport
    if a.samples is not None:
        vp.samples = a.samples
    out = a.out or ("Presentation.png" if a.scene == "presentation" else a.scene + ".png")
    if "RANK" in os.environ and "WORLD_SIZE" in os.environ and int(os.environ["WORLD_SIZE"]) > 1:
        import torch
        import torch.distributed as dist
        torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", "0")))
        dist.init_process_group("nccl")
    if a.frames is None and a.frames_dir is None:
        print("Running")                                         # main.rs:412
        write_img_f32(async_render(vp, job.ray_color, job.scene), out)   # main.rs:413-418
        return 0
    vp.number_of_frames = a.frames if a.frames is not None else vp.number_of_frames
    vp.start_frame = a.start_frame
    test_run(out, vp, job.ray_color, job.scene, frames_dir=a.frames_dir)
    return 0
