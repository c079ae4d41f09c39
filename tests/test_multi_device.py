"""Several devices behind one call of the C ABI (include/rtw.h rtw_scene_multi_*, rtw_render_multi,
rtw_render_frames_multi; SURVEY 8b "device list, split {Samples, Tiles}", 8e A/B/C): the reference's caller is
one process (main.rs:76-80,412-414), so the partitions that `sharding.py` runs under torchrun also exist behind
a single call. A device may be listed more than once, which lets a one-GPU box run every code path (shares on
separate streams, the combine kernel, band placement, frame dealing); on a multi-GPU box the same tests run
over all devices and go through NVLink peer memory.

  samples split == one device up to f32 summation order (same Philox streams: the counter carries the global
                   sample index);
  tiles split   == one device, no reduction across devices;
  frames split  == rtw_render_frames.
"The same" is up to the order of the f32 adds into a pixel: several warps (sample chunks) add into one pixel with
atomics whose order is not fixed from launch to launch, so not even two identical renders on one device are
bit-identical; every path is (same Philox counter), which is what the 2e-5 bounds below pin.
"""


def _same(a, b, rtol=2e-5, atol=1e-6):
    fa, fb = np.isfinite(a), np.isfinite(b)
    return np.array_equal(fa, fb) and np.allclose(a[fa], b[fa], rtol=rtol, atol=atol)
import numpy as np
import pytest

import rtw_b200
from rtw_b200 import _abi as A, scenes

pytestmark = pytest.mark.gpu


def _device_lists(cuda):
    n = cuda.device_count()
    lists = [[0, 0, 0]]
    if n >= 2:
        lists.append(list(range(n)))
        lists.append([1, 0])
    return lists


def _job(name, w, h, spp):
    job = {"C4": scenes.config_c4, "C5": scenes.config_c5, "C2": scenes.config_c2}[name]()
    vp = job.viewport
    nv = type(vp).new_from_res(w, h, spp, vp.depth, vp.gamma, vp.vfov, vp.origin, vp.direction, vp.vup, None, vp.lens_radius)
    nv.shutter_speed, nv.fps, nv.frame = vp.shutter_speed, vp.fps, vp.frame
    job.viewport = nv
    return job


@pytest.mark.parametrize("flags", [A.RTW_FLAG_STRICT_FP, 0], ids=["strict", "fast"])
@pytest.mark.parametrize("name,w,h,spp,pipe", [("C5", 256, 256, 9, A.RTW_FLAG_WAVEFRONT), ("C4", 96, 96, 7, A.RTW_FLAG_MEGAKERNEL)])
def test_sample_split_equals_one_device(name, w, h, spp, pipe, flags, cuda):
    job = _job(name, w, h, spp)
    vp = job.viewport
    vp.flags = flags | pipe          # the pipeline is pinned: a share of the frame may fall below the size where the default switches
    cam, o = vp.camera(cuda), vp.opts(job.ray_color, job.mode)
    one = cuda.scene(job.scene).render(cam, o)
    st1 = cuda.scene(job.scene).stats()
    for devs in _device_lists(cuda):
        ms = cuda.scene_multi(job.scene, devs)
        got = ms.render(cam, o, A.RTW_SPLIT_SAMPLES)
        total, per, combine_ms = ms.stats()
        assert total.paths == st1.paths == w * h * spp and total.rays == st1.rays and total.nan_paths == st1.nan_paths
        assert sum(p.paths for p in per) == total.paths and combine_ms > 0.0
        assert np.array_equal(np.isfinite(got), np.isfinite(one))
        fin = np.isfinite(one)
        assert np.allclose(got[fin], one[fin], rtol=2e-5, atol=1e-6)            # f32 summation order only
        # raw sums come back too (RTW_OUT_SUM): the combine adds without resolving
        o2 = vp.opts(job.ray_color, job.mode)
        o2.output = A.RTW_OUT_SUM
        s_multi = ms.render(cam, o2, A.RTW_SPLIT_SAMPLES)
        s_one = cuda.scene(job.scene).render(cam, o2)
        fin = np.isfinite(s_one)
        assert np.allclose(s_multi[fin], s_one[fin], rtol=2e-5, atol=1e-5)
        ms.close()
        del job.scene._backend_cache[("cuda-multi", tuple(devs))]


def test_more_shares_than_samples(cuda):
    job = _job("C2", 64, 36, 2)
    vp = job.viewport
    cam, o = vp.camera(cuda), vp.opts(job.ray_color, job.mode)
    one = cuda.scene(job.scene).render(cam, o)
    ms = cuda.scene_multi(job.scene, [0, 0, 0])
    got = ms.render(cam, o, A.RTW_SPLIT_SAMPLES)
    assert np.allclose(got, one, rtol=2e-5, atol=1e-6, equal_nan=True)
    assert ms.stats()[0].paths == 64 * 36 * 2


@pytest.mark.parametrize("band", [0, 5, 64])
@pytest.mark.parametrize("name,w,h,spp", [("C4", 160, 123, 24), ("C5", 96, 64, 3)])
def test_tile_split_equals_one_device(name, w, h, spp, band, cuda):
    """Row bands dealt round-robin (C4 "tiles across 8 GPUs"): each band is rendered and resolved by
    one device and lands in its place in the host image: the frame equals the one-device frame path for path.
    (The default pipeline depends on the render's size, and a band is smaller than the frame: the comparison pins
    the pipeline.)"""
    job = _job(name, w, h, spp)
    vp = job.viewport
    vp.flags = A.RTW_FLAG_MEGAKERNEL
    cam, o = vp.camera(cuda), vp.opts(job.ray_color, job.mode)
    one = cuda.scene(job.scene).render(cam, o)
    for devs in _device_lists(cuda):
        ms = cuda.scene_multi(job.scene, devs)
        got = ms.render(cam, o, A.RTW_SPLIT_TILES, band)
        assert _same(got, one)
        total, per, combine_ms = ms.stats()
        assert total.paths == w * h * spp and combine_ms == 0.0
        assert sum(1 for p in per if p.paths) == min(len(devs), -(-h // (band or 8)))
        # a row sub-range of the frame, split again
        o3 = vp.opts(job.ray_color, job.mode)
        o3.row_begin, o3.row_count = 17, 40
        assert _same(ms.render(cam, o3, A.RTW_SPLIT_TILES, band), one[17:57])
        ms.close()
        del job.scene._backend_cache[("cuda-multi", tuple(devs))]


def test_frame_split_equals_render_frames(cuda):
    """render_multi (viewport.rs:249-269) with frame k on devices[k mod n] (C5's animation frames
    one per GPU): the frames equal rtw_render_frames' on one device."""
    anim = scenes.ltr_scene(samples=4, width=64, height=36, frames=5)
    vp = anim.viewport
    vp.shutter_speed = 0.5
    vp.flags = A.RTW_FLAG_MEGAKERNEL
    cam, o = vp.camera(cuda), vp.opts(anim.ray_color, A.RTW_MODE_RANDOM)
    one = cuda.scene(anim.scene).render_frames(cam, o, 2, 5)
    assert not np.array_equal(one[0], one[1])                      # the spheres move
    for devs in _device_lists(cuda):
        ms = cuda.scene_multi(anim.scene, devs)
        got = ms.render_frames(cam, o, 2, 5)
        assert _same(got, one)
        assert ms.stats()[0].paths == 5 * 64 * 36 * 4
        ms.close()
        del anim.scene._backend_cache[("cuda-multi", tuple(devs))]


def test_back_to_back_device_renders_on_one_scene(cuda):
    """rtw_render_device twice in a row on one scene with no host sync in between (what the row-band loops of
    sharding.py and rtw_render_multi do): the second wavefront render must not inherit the first one's counters
    (round-1 advisor finding: the counters were seeded from a pinned ring slot a pending copy could overwrite)."""
    torch = pytest.importorskip("torch")
    job = _job("C5", 512, 512, 2)                                   # 256 rows x 512 x 2 spp = 2^18 paths per band, through the wavefront pipeline
    vp = job.viewport
    cam = vp.camera(cuda)
    h = cuda.scene(job.scene)
    o = vp.opts(job.ray_color, job.mode)
    o.flags |= A.RTW_FLAG_WAVEFRONT                                 # the reference frame through the same pipeline as the bands
    whole = h.render(cam, o)
    dev = torch.device("cuda", torch.cuda.current_device())
    for rep in range(3):
        bufs = [torch.empty((256, 512, 3), dtype=torch.float32, device=dev) for _ in range(2)]
        st = torch.cuda.Stream()
        for k in range(2):
            ob = vp.opts(job.ray_color, job.mode)
            ob.row_begin, ob.row_count = 256 * k, 256
            ob.flags |= A.RTW_FLAG_WAVEFRONT
            h.render_device(cam, ob, bufs[k].data_ptr(), st.cuda_stream)
        st.synchronize()
        got = torch.cat(bufs).cpu().numpy()
        fin = np.isfinite(whole)
        assert np.array_equal(np.isfinite(got), fin)
        assert np.allclose(got[fin], whole[fin], rtol=1e-4, atol=1e-5), f"repeat {rep}"
