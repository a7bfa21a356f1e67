"""plu_gpu_encode_bmp_device_noalloc: the .bmp encode on a fixed device array inside the library instead of a buffer
allocated per call (include/plugpu.h, no counterpart in the reference).

Through the C ABI, against the existing entry point plu_gpu_encode_bmp_device: the same bytes (this is byte work, so the
tolerance is zero, as in tests/test_post.py), on calls in a row that share the array — a large image first, then a
smaller one and an odd width (row padding), so stale bytes of the previous call would show — and a clear refusal beyond
the stated maximum.  The file is named so that pytest collects it last; it needs nothing from the other tests."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _encode(lib, h, name, dev, w, hgt, flags):
    n = lib.plu_bmp_size(w, hgt)
    out = np.zeros(n, np.uint8)
    import torch
    rc = getattr(lib, name)(h, w, hgt, C.c_void_p(dev.data_ptr()), flags, C.c_void_p(torch.cuda.current_stream().cuda_stream),
                            out.ctypes.data, n)
    return rc, out


def test_no_alloc_bmp_encode_equals_the_allocating_entry_point(lib):
    import torch
    from plutracer_b200 import GpuContext, _ffi
    from tests.orc import golden_scene
    ctx = GpuContext(0)
    try:
        h = ctx._h
        # a rendered image (values beyond [0, 1], exercised clamp + tone map), then synthetic ones of other shapes
        w0, h0, n0 = 640, 360, 2
        ctx.upload(golden_scene("cornell-box", w0, h0, n0))
        fb = torch.zeros(h0, w0, 3, dtype=torch.float32, device="cuda")
        ctx.render_device(GpuContext.params(w0, h0, n0, seed=5), fb.data_ptr(), torch.cuda.current_stream().cuda_stream)
        g = torch.Generator(device="cuda"); g.manual_seed(7)
        cases = [(fb, w0, h0)]
        for (w, hh) in ((1920, 1080), (333, 77), (5, 3), (1, 1)):   # the maximum that fits, odd widths (1 and 3 bytes of row padding), tiny
            cases.append((torch.rand(hh, w, 3, generator=g, device="cuda") * 3.0 - 0.5, w, hh))
        torch.cuda.synchronize()
        for flags in (0, _ffi.POST_TONEMAP):
            for img, w, hh in cases:   # consecutive calls share the array: each must be right whatever the previous one left
                rc_a, want = _encode(ctx._lib, h, "plu_gpu_encode_bmp_device", img, w, hh, flags)
                rc_b, got = _encode(ctx._lib, h, "plu_gpu_encode_bmp_device_noalloc", img, w, hh, flags)
                assert rc_a == 0 and rc_b == 0, ctx._lib.plu_gpu_last_error(h)
                assert np.array_equal(want, got), (w, hh, flags, int((want != got).sum()))
                assert want[:2].tobytes() == b"BM" and want[54:].any()
        # the same image twice in a row: identical both times
        img, w, hh = cases[0]
        first = _encode(ctx._lib, h, "plu_gpu_encode_bmp_device_noalloc", img, w, hh, 0)[1]
        second = _encode(ctx._lib, h, "plu_gpu_encode_bmp_device_noalloc", img, w, hh, 0)[1]
        assert np.array_equal(first, second)
        # beyond the stated maximum: refused with PLU_E_ARG and a message, nothing written
        big_w, big_h = 2048, 1400   # 8.6 MB body > 8 MiB
        assert ctx._lib.plu_bmp_size(big_w, big_h) - 54 > _ffi.NOALLOC_BMP_MAX_BODY
        big = torch.zeros(big_h, big_w, 3, dtype=torch.float32, device="cuda")
        rc, out = _encode(ctx._lib, h, "plu_gpu_encode_bmp_device_noalloc", big, big_w, big_h, 0)
        assert rc == -1 and not out.any()
        assert b"PLU_NOALLOC_BMP_MAX_BODY" in ctx._lib.plu_gpu_last_error(h)
        assert _encode(ctx._lib, h, "plu_gpu_encode_bmp_device", big, big_w, big_h, 0)[0] == 0   # the allocating call still takes it
        # and the entry point keeps working after a refusal
        rc, got = _encode(ctx._lib, h, "plu_gpu_encode_bmp_device_noalloc", img, w, hh, 0)
        assert rc == 0 and np.array_equal(got, first)
    finally:
        ctx.close()
