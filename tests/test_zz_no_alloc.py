"""rtw_bilateral_filter_device_noalloc (include/rtw.h): the device bilateral filter with its scratch in a static
device array of csrc/rtw_post.cu, for callers that may not allocate. It has no counterpart in the reference; the
check is against the existing entry point, rtw_bilateral_filter_device, through the C ABI, byte for byte (the
tolerance of tests/test_postprocessing.py: none), on calls in a row that share the array.

Collected last on purpose: it passes alone in a fresh process and after the
rest of the suite. The caller — here torch — owns the input and output buffers; the call itself allocates nothing.
"""
import ctypes as C

import numpy as np
import pytest

from rtw_b200 import _abi as A

pytestmark = pytest.mark.gpu


def _image(h, w, seed):
    """Piecewise-flat image plus noise (edges to keep, noise to smooth), as in test_postprocessing.py."""
    rng = np.random.default_rng(seed)
    base = np.zeros((h, w, 3), dtype=np.float64)
    base[:, : w // 2] = (60, 120, 200)
    base[:, w // 2:] = (220, 180, 40)
    base[h // 3: 2 * h // 3, w // 4: 3 * w // 4] = (30, 200, 90)
    return np.clip(base + rng.normal(0, 12, base.shape), 0, 255).astype(np.uint8)


def _run(lib, name, d_in, w, h, size, kind, stream=0, poison=None):
    """One call of entry point `name` on device buffers the caller owns. Returns (bytes out, 4 params)."""
    import torch
    d_out = torch.full((h, w, 3), 0 if poison is None else poison, dtype=torch.uint8, device="cuda")
    d_par = torch.full((4,), float("nan"), dtype=torch.float32, device="cuda")
    torch.cuda.synchronize()
    fn = getattr(lib, name)
    A.check(fn(C.c_void_p(d_in.data_ptr()), w, h, size, kind, C.c_void_p(d_out.data_ptr()), C.c_void_p(d_par.data_ptr()),
               -1, C.c_void_p(stream)))
    torch.cuda.synchronize()
    return d_out.cpu().numpy(), d_par.cpu().numpy()


CASES = [  # (height, width, window size, proximity): a large call first, so that smaller ones run over its leftovers
    (240, 320, 16, 0), (9, 12, 4, 1), (150, 200, 10, 0), (150, 200, 10, 1), (3, 3, 1, 0), (2, 9, 2, 1), (37, 53, 0, 0),
    (64, 48, 16, 1), (240, 320, 2, 0),
]


def test_noalloc_entry_point_matches_the_allocating_one_on_calls_in_a_row(cuda):
    import torch
    lib = cuda.lib
    for rep in range(2):                                        # the whole sequence twice: every case also follows every other
        for (h, w, size, kind) in CASES:
            img = _image(h, w, seed=h * 131 + size + kind)
            d_in = torch.from_numpy(img).cuda()
            want, want_par = _run(lib, "rtw_bilateral_filter_device", d_in, w, h, size, kind)
            got1, par1 = _run(lib, "rtw_bilateral_filter_device_noalloc", d_in, w, h, size, kind, poison=0x5A)
            got2, par2 = _run(lib, "rtw_bilateral_filter_device_noalloc", d_in, w, h, size, kind, poison=0xA5)
            for got, par in ((got1, par1), (got2, par2)):
                assert par.tobytes() == want_par.tobytes() or (np.isnan(par).any() and
                                                               np.array_equal(np.isnan(par), np.isnan(want_par))), (rep, h, w, size, kind)
                assert np.array_equal(got, want), (rep, h, w, size, kind)


def test_noalloc_matches_the_oracle_and_accepts_null_params_and_a_stream(cuda, oracle):
    import torch
    lib = cuda.lib
    h, w, size = 90, 70, 6
    img = np.random.default_rng(4).integers(0, 256, (h, w, 3), dtype=np.uint8)
    d_in = torch.from_numpy(img).cuda()
    st = torch.cuda.Stream()
    for kind in (0, 1):
        want = oracle.bilateral_filter(img, size, kind)
        for _ in range(2):
            d_out = torch.zeros((h, w, 3), dtype=torch.uint8, device="cuda")
            torch.cuda.synchronize()
            A.check(lib.rtw_bilateral_filter_device_noalloc(C.c_void_p(d_in.data_ptr()), w, h, size, kind,
                                                            C.c_void_p(d_out.data_ptr()), None, 0, C.c_void_p(st.cuda_stream)))
            st.synchronize()
            assert np.array_equal(d_out.cpu().numpy(), want)


def test_noalloc_takes_no_device_memory(cuda):
    """Free device memory is the same before and after calls of the entry point (the first call of a process may
    load the module that holds the static array, so the measurement starts after one call)."""
    import torch
    lib = cuda.lib
    h, w, size = 240, 320, 12
    d_in = torch.from_numpy(_image(h, w, 7)).cuda()
    d_out = torch.empty((h, w, 3), dtype=torch.uint8, device="cuda")

    def call():
        A.check(lib.rtw_bilateral_filter_device_noalloc(C.c_void_p(d_in.data_ptr()), w, h, size, 0,
                                                        C.c_void_p(d_out.data_ptr()), None, -1, None))
        torch.cuda.synchronize()

    call()
    free0, _ = torch.cuda.mem_get_info()
    for _ in range(4):
        call()
    free1, _ = torch.cuda.mem_get_info()
    assert free1 == free0


def test_noalloc_rejects_sizes_beyond_the_static_scratch(cuda):
    import torch
    lib = cuda.lib
    d = torch.zeros(64, dtype=torch.uint8, device="cuda")
    p_in, p_out = C.c_void_p(d.data_ptr()), C.c_void_p(d.data_ptr() + 32)
    fn = lib.rtw_bilateral_filter_device_noalloc
    # (w-2)(h-2) one past the limit: 1024 x 1022 interior pixels is the limit exactly
    assert 1024 * 1022 == A.RTW_NOALLOC_MAX_INTERIOR_PIXELS
    with pytest.raises(A.RtwError) as ei:
        A.check(fn(p_in, 1026 + 1, 1024, 4, 0, p_out, None, -1, None))
    assert ei.value.code == A.RTW_E_UNSUPPORTED and "RTW_NOALLOC_MAX_INTERIOR_PIXELS" in str(ei.value)
    with pytest.raises(A.RtwError) as ei:
        A.check(fn(p_in, 1920, 1080, 4, 0, p_out, None, -1, None))
    assert ei.value.code == A.RTW_E_UNSUPPORTED
    with pytest.raises(A.RtwError) as ei:
        A.check(fn(p_in, 64, 64, A.RTW_NOALLOC_MAX_WINDOW + 1, 0, p_out, None, -1, None))
    assert ei.value.code == A.RTW_E_UNSUPPORTED and "RTW_NOALLOC_MAX_WINDOW" in str(ei.value)
    for bad in ((1, 5, 1, 0), (5, 5, 1, 7)):                   # the existing entry point's argument checks
        with pytest.raises(A.RtwError) as ei:
            A.check(fn(p_in, bad[0], bad[1], bad[2], bad[3], p_out, None, -1, None))
        assert ei.value.code == A.RTW_E_INVALID
    with pytest.raises(A.RtwError):
        A.check(fn(p_in, 8, 8, 1, 0, p_in, None, -1, None))    # in place


def test_noalloc_at_the_limit(cuda):
    """The largest image the static scratch takes (1026 x 1024: exactly RTW_NOALLOC_MAX_INTERIOR_PIXELS interior
    pixels) with the largest window, against the allocating entry point."""
    import torch
    lib = cuda.lib
    h, w, size = 1024, 1026, A.RTW_NOALLOC_MAX_WINDOW
    d_in = torch.from_numpy(_image(h, w, 99)).cuda()
    want, want_par = _run(lib, "rtw_bilateral_filter_device", d_in, w, h, size, 1)
    for _ in range(2):
        got, par = _run(lib, "rtw_bilateral_filter_device_noalloc", d_in, w, h, size, 1, poison=0x33)
        assert par.tobytes() == want_par.tobytes()
        assert np.array_equal(got, want)
