"""Post-process + BMP write: oracle pinned against the reference's own main.cpp
postprocesser and texture2d::write_bmp (tests/golden/post.npz, made by make_post_golden.py), CUDA path
against the oracle."""
import os

import numpy as np
import pytest

from tests.orc import GOLDEN


@pytest.fixture(scope="module")
def post():
    return np.load(os.path.join(GOLDEN, "post.npz"))


def _same_bits(a, b):
    return (a.view(np.uint32) == b.view(np.uint32)) | (np.isnan(a) & np.isnan(b))


def test_oracle_tonemap_matches_reference(oracle, post):
    got = oracle.postprocess(post["input"], pow_mode=0)
    want = post["tonemapped"]
    assert np.array_equal(np.isnan(got), np.isnan(want))
    ok = np.isfinite(want)
    # same formula, same libm family: bit-exact here; 1 ulp allowed in case the box's glibc powf differs
    assert np.allclose(got[ok], want[ok], rtol=1.2e-7, atol=0)
    assert _same_bits(got, want).mean() > 0.999
    assert np.array_equal(got[~ok & ~np.isnan(want)], want[~ok & ~np.isnan(want)])


def test_oracle_bmp_bytes_match_reference(oracle, post):
    assert np.array_equal(oracle.encode_bmp(post["input"]), post["bmp_raw"])
    assert np.array_equal(oracle.encode_bmp(post["tonemapped"]), post["bmp_post"])
    hdr = post["bmp_raw"][:54]
    assert bytes(hdr[:2]) == b"BM" and int.from_bytes(bytes(hdr[2:6]), "little") == post["bmp_raw"].size


def test_double_pow_is_within_one_ulp_of_libm_powf(oracle, post):
    a = oracle.postprocess(post["input"], pow_mode=0)
    b = oracle.postprocess(post["input"], pow_mode=1)
    ok = np.isfinite(a)
    assert np.array_equal(np.isnan(a), np.isnan(b))
    d = np.abs(a[ok].view(np.int32).astype(np.int64) - b[ok].view(np.int32).astype(np.int64))
    assert d.max() <= 1 and (d > 0).mean() < 1e-3


def test_bmp_size(lib):
    for w, h in ((1, 1), (37, 29), (640, 480), (1023, 3)):
        assert lib.plu_bmp_size(w, h) == 54 + (w * 3 + ((-w * 3) & 3)) * h
    assert lib.plu_bmp_size(0, 5) == 0


@pytest.mark.gpu
def test_gpu_tonemap_matches_oracle(gpu, oracle, post):
    want = oracle.postprocess(post["input"], pow_mode=1)
    got = gpu.postprocess(post["input"].copy())
    assert np.array_equal(np.isnan(got), np.isnan(want))
    ok = np.isfinite(want)
    d = np.abs(got[ok].view(np.int32).astype(np.int64) - want[ok].view(np.int32).astype(np.int64))
    assert d.max() <= 1 and (d > 0).mean() < 1e-3   # double pow on both sides, rounded once
    # and against the reference itself
    ref = post["tonemapped"]
    d = np.abs(got[ok].view(np.int32).astype(np.int64) - ref[ok].view(np.int32).astype(np.int64))
    assert d.max() <= 1 and (d > 0).mean() < 1e-3


@pytest.mark.gpu
def test_gpu_bmp_bytes_match_reference(gpu, oracle, post):
    assert np.array_equal(gpu.encode_bmp(post["input"]), post["bmp_raw"])          # bit-exact: no pow involved
    assert np.array_equal(gpu.encode_bmp(post["tonemapped"]), post["bmp_post"])
    fused = gpu.encode_bmp(post["input"], tonemap=True)                             # tone map fused in front
    diff = fused.astype(np.int16) - post["bmp_post"].astype(np.int16)
    assert np.array_equal(fused[:54], post["bmp_post"][:54])
    assert np.abs(diff).max() <= 1 and (diff != 0).mean() < 1e-3


@pytest.mark.gpu
def test_gpu_render_to_bmp_end_to_end(gpu, oracle):
    """render -> tone map -> BMP entirely on the device equals oracle render -> oracle post -> oracle BMP."""
    import torch
    from plutracer_b200 import GpuContext
    from tests.orc import golden_scene
    w, h, n = 44, 30, 3
    sc = golden_scene("cornell-box", w, h, n)
    gpu.upload(sc)
    p = GpuContext.params(w, h, n, seed=11)
    fb = torch.zeros(h, w, 3, dtype=torch.float32, device="cuda")
    gpu.render_device(p, fb.data_ptr(), torch.cuda.current_stream().cuda_stream)
    got = gpu.encode_bmp(None, tonemap=True, device_ptr=fb.data_ptr(), size=(w, h), stream=torch.cuda.current_stream().cuda_stream)
    img, _ = oracle.render(sc, p, trig=0)
    want = oracle.encode_bmp(oracle.postprocess(img, pow_mode=1))
    diff = got.astype(np.int16) - want.astype(np.int16)
    assert np.abs(diff).max() <= 1 and (diff != 0).mean() < 1e-3
    with pytest.raises(Exception):
        gpu.encode_bmp(np.zeros((2, 2, 3), np.float32)[:, :, :2])
