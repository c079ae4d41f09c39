"""Multi-GPU sharding of a render: one process per GPU, the scene replicated.

The reference has no distributed mode: async_render gives one tokio task per row (viewport.rs:236-240)
and render_multi runs frames one after the other (viewport.rs:256-267). Samples, pixels and frames are
independent, so the three natural partitions are

  samples : rank r renders samples [r*spp/n, (r+1)*spp/n) of every pixel into its own f32 sum buffer
            (the Philox counter carries the GLOBAL sample index, so the image does not depend on n up
            to f32 summation order); one NCCL reduce(sum) of W*H*3 f32 to rank 0, then resolve
            (/spp, powf(1/gamma)) there. The only exchange of the whole frame: 7.68 MB for 800x800.
  tiles   : bands of `band` rows dealt round-robin; each rank resolves its own bands; gather to
            rank 0. No reduction across ranks: every pixel is summed on one GPU (a 1-GPU render's value up to the
            order of that GPU's own f32 adds; bit-identical on the deterministic CPU test backends).
  frames  : render_multi's frames dealt round-robin, frame k -> rank k mod n; no exchange at all.

`torch.distributed` is plumbing only (NCCL over NVLink on the GPU box, gloo in the CPU tests).
"""
from __future__ import annotations

from typing import List, Optional, Tuple

import numpy as np

from . import _abi as A


def sample_range(total: int, world: int, rank: int) -> Tuple[int, int]:
    """[begin, count) of `total` samples for `rank`; remainders go to the first ranks."""
    base, rem = divmod(total, world)
    begin = rank * base + min(rank, rem)
    return begin, base + (1 if rank < rem else 0)


def row_bands(height: int, world: int, rank: int, band: int = 8) -> List[Tuple[int, int]]:
    """(row_begin, row_count) bands owned by `rank`: bands of `band` rows, round-robin (path length
    varies across the image — the Cornell box's light vs its walls — so contiguous slabs would
    balance badly)."""
    out = []
    for b, r0 in enumerate(range(0, height, band)):
        if b % world == rank:
            out.append((r0, min(band, height - r0)))
    return out


def frames_for_rank(start_frame: int, n_frames: int, world: int, rank: int) -> List[int]:
    return [start_frame + k for k in range(n_frames) if k % world == rank]


def effective_samples(samples: int, mode: int) -> int:
    """Viewport::render rounds the sample count up to a square (viewport.rs:443)."""
    if mode == A.RTW_MODE_NO_RAND:
        return 1
    if mode == A.RTW_MODE_STRATIFIED:
        s_root = int(np.ceil(np.sqrt(np.float32(samples))))
        return s_root * s_root
    return int(samples)


def _dist():
    import torch.distributed as dist
    if not dist.is_available() or not dist.is_initialized():
        raise RuntimeError("torch.distributed is not initialised (launch with torchrun / init_process_group)")
    return dist


def render_samples_sharded(viewport, ray_color, scene, mode: int = A.RTW_MODE_RANDOM, backend=None, group=None,
                           device_tensor: bool = False):
    """Sample-range split + reduce(sum) to rank 0 + resolve there. Returns the Img on rank 0
    (numpy, or a CUDA tensor when device_tensor=True) and None elsewhere."""
    import torch
    from .viewport import get_default_backend
    dist = _dist()
    be = backend or get_default_backend()
    rank, world = dist.get_rank(group), dist.get_world_size(group)
    total = effective_samples(viewport.samples, mode)
    begin, count = sample_range(total, world, rank)
    cam = viewport.camera(be)
    o = viewport.opts(ray_color, mode)
    o.output = A.RTW_OUT_SUM
    h = be.scene(scene)
    shape = (cam.height, cam.width, 3)
    if getattr(be, "name", "") == "cuda":
        dev = torch.device("cuda", torch.cuda.current_device())
        buf = torch.empty(shape, dtype=torch.float32, device=dev)
        if count > 0:
            o.sample_begin, o.sample_count = begin, count
            h.render_device(cam, o, buf.data_ptr(), torch.cuda.current_stream().cuda_stream)
        else:
            buf.zero_()
        if world > 1:
            dist.reduce(buf, dst=0, op=dist.ReduceOp.SUM, group=group)
        if rank != 0:
            return None
        be.resolve_device(buf.data_ptr(), cam.height * cam.width, total, viewport.gamma,
                          torch.cuda.current_stream().cuda_stream)
        return buf if device_tensor else buf.cpu().numpy()
    # host path (CPU backends used by the gloo tests)
    if count > 0:
        o.sample_begin, o.sample_count = begin, count
        part = torch.from_numpy(np.ascontiguousarray(h.render(cam, o)))
    else:
        part = torch.zeros(shape, dtype=torch.float32)
    if world > 1:
        dist.reduce(part, dst=0, op=dist.ReduceOp.SUM, group=group)
    if rank != 0:
        return None
    img = part.numpy() / np.float32(total)
    with np.errstate(invalid="ignore"):
        return np.power(img, np.float32(1.0) / np.float32(viewport.gamma)).astype(np.float32)


def band_rows_for_rank(height: int, world: int, rank: int, band: int = 8) -> np.ndarray:
    """Image rows of `rank`'s bands, in the order an interleaved render returns them (rtw_render_opts.row_stride)."""
    rows = [np.arange(r0, r0 + rn) for r0, rn in row_bands(height, world, rank, band)]
    return np.concatenate(rows) if rows else np.zeros(0, dtype=np.int64)


def render_tiles_sharded(viewport, ray_color, scene, mode: int = A.RTW_MODE_RANDOM, backend=None, group=None,
                         band: int = 8) -> Optional[np.ndarray]:
    """Row-band split: rank r owns bands r, r + n, r + 2n, ... and renders them as ONE render of interleaved rows
    (rtw_render_opts.row_band / row_stride; a render per band would pay the pipeline's ramp and drain per band),
    resolves them and puts them in place; rank 0 gathers. No reduction across ranks: every pixel is summed on one
    GPU (bit-identical to one device on the deterministic CPU test backends)."""
    import torch
    from .viewport import get_default_backend
    dist = _dist()
    be = backend or get_default_backend()
    rank, world = dist.get_rank(group), dist.get_world_size(group)
    cam = viewport.camera(be)
    h = be.scene(scene)
    on_gpu = getattr(be, "name", "") == "cuda"
    dev = torch.device("cuda", torch.cuda.current_device()) if on_gpu else torch.device("cpu")
    img = torch.zeros((cam.height, cam.width, 3), dtype=torch.float32, device=dev)
    rows = band_rows_for_rank(cam.height, world, rank, band)
    if rows.size:
        o = viewport.opts(ray_color, mode)
        o.row_begin, o.row_count = rank * band, int(rows.size)
        o.row_band, o.row_stride = band, world * band
        idx = torch.from_numpy(rows.astype(np.int64)).to(dev)
        if on_gpu:
            part = torch.empty((rows.size, cam.width, 3), dtype=torch.float32, device=dev)
            h.render_device(cam, o, part.data_ptr(), torch.cuda.current_stream().cuda_stream)
        else:
            part = torch.from_numpy(np.ascontiguousarray(h.render(cam, o)))
        img[idx] = part
    if world > 1:
        # bands are disjoint and every other rank holds zeros there, so a sum is a gather; NaN
        # pixels survive because NaN + 0 = NaN.
        dist.reduce(img, dst=0, op=dist.ReduceOp.SUM, group=group)
    if rank != 0:
        return None
    return img.cpu().numpy()


def render_multi_sharded(viewport, ray_color, scene, backend=None, group=None) -> Optional[List[np.ndarray]]:
    """render_multi (viewport.rs:249-269) with frame k on rank k mod n. Rank 0 returns all frames in order."""
    import torch
    from .viewport import get_default_backend
    dist = _dist()
    be = backend or get_default_backend()
    rank, world = dist.get_rank(group), dist.get_world_size(group)
    cam = viewport.camera(be)
    h = be.scene(scene)
    mine = frames_for_rank(viewport.start_frame, viewport.number_of_frames, world, rank)
    o = viewport.opts(ray_color, A.RTW_MODE_RANDOM)
    imgs = {}
    for f in mine:
        imgs[f] = h.render_frames(cam, o, f, 1)[0]
    if world == 1:
        return [imgs[f] for f in sorted(imgs)]
    gathered = [None] * world
    dist.gather_object(imgs, gathered if rank == 0 else None, dst=0, group=group)
    if rank != 0:
        return None
    allf = {}
    for d in gathered:
        allf.update(d)
    return [allf[f] for f in sorted(allf)]
