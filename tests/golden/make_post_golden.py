#!/usr/bin/env python
"""tests/golden/post.npz: the REFERENCE's post-process and BMP writer on a fixed float image
(oracle/_ref/plu_ref_post = src/main.cpp's postprocesser + texture2d::write_bmp, unmodified).
Run in the build container (needs /root/reference): python tests/golden/make_post_golden.py"""
import os
import subprocess
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)


def make_input(w=37, h=29, seed=5):
    """Linear radiance with everything write_bmp / the tone map can meet: a smooth ramp, values far above 1,
    exact zeros (0/0 in the tone map), negatives, +inf and NaN.  37 is not a multiple of 4: padded BMP rows."""
    rng = np.random.default_rng(seed)
    img = (rng.random((h, w, 3), dtype=np.float32) ** 3 * np.float32(6.0)).astype(np.float32)
    img[0, :, :] = np.linspace(0, 1.5, w, dtype=np.float32)[:, None]
    img[1, :5] = 0.0
    img[2, :5] = np.float32(-0.25)
    img[3, 0] = np.float32(np.inf)
    img[3, 1] = np.float32(np.nan)
    img[4, :8, 0] = 0.0   # one zero channel, non-zero luma
    return img


def main():
    ref = os.path.join(ROOT, "oracle", "_ref", "plu_ref_post")
    if not os.path.exists(ref):
        subprocess.run(["bash", os.path.join(ROOT, "oracle", "build_ref.sh")], check=True)
    img = make_input()
    h, w, _ = img.shape
    with tempfile.TemporaryDirectory() as tmp:
        fin, fbmp, ftm, fraw = (os.path.join(tmp, n) for n in ("in.f32", "post.bmp", "tm.f32", "raw.bmp"))
        img.tofile(fin)
        subprocess.run([ref, "post", str(w), str(h), fin, fbmp, ftm], check=True)
        subprocess.run([ref, "bmp", str(w), str(h), fin, fraw], check=True)
        tm = np.fromfile(ftm, np.float32).reshape(h, w, 3)
        bmp_post = np.fromfile(fbmp, np.uint8)
        bmp_raw = np.fromfile(fraw, np.uint8)
    out = os.path.join(ROOT, "tests", "golden", "post.npz")
    np.savez_compressed(out, input=img, tonemapped=tm, bmp_post=bmp_post, bmp_raw=bmp_raw)
    print(out, img.shape, bmp_post.size, bmp_raw.size)


if __name__ == "__main__":
    main()
