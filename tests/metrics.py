"""Image-parity metrics shared by the CPU (oracle vs reference) and GPU (CUDA vs reference) tests."""
import numpy as np


def converged_metrics(img, a, b):
    """north_star check 2 against the reference's own CPU renders (two independent seeds a, b).

    The named scenes are heavy-tailed: cornell-box has a 500-radiance light seen through a mirror and a glass
    sphere, and the two REFERENCE renders of it differ by 67 % in plain image mean at 1024 spp (fireflies), so
    "mean absolute error < 1 %" is only meaningful on tone-clipped radiance.  Metrics, all on finite pixels
    (the reference's NaN pixels - sphere poles, 0/0 - are masked and counted):
      mean_err       per channel |mean(clip(img)) - mean(clip(ref))| / mean(clip(ref)), clip at 4x the median
                     of the lit reference pixels; bound 1 %
      rmse_clipped   RMSE of the clipped images against a and b, next to the noise floor RMSE(clip a, clip b);
                     the largest 0.1 % of squared pixel errors are dropped on both sides: the estimator has
                     infinite variance (1/cos at the rim of the cosine-hemisphere warp under the skewed frames
                     of Q9), so a single firefly sample would otherwise decide the comparison
      rmse           the plain RMSE, reported only
    """
    ok = np.isfinite(a).all(axis=2) & np.isfinite(b).all(axis=2) & np.isfinite(img).all(axis=2)
    a, b, img = a[ok], b[ok], img[ok]
    ref = 0.5 * (a + b)
    lit = ref[ref > 0]
    cap = 4.0 * float(np.median(lit)) if lit.size else 1.0
    ca, cb, cg = np.minimum(a, cap), np.minimum(b, cap), np.minimum(img, cap)
    cr = [0.5 * float(ca[:, c].mean() + cb[:, c].mean()) for c in range(3)]
    rm = lambda x, y: float(np.sqrt(((x - y) ** 2).mean()))

    def rmt(x, y):
        e = np.sort(((x - y) ** 2).sum(axis=1))
        return float(np.sqrt(e[: max(1, int(len(e) * 0.999))].mean() / 3.0))
    return {
        "masked": int((~ok).sum()), "cap": cap,
        # clipping a noisy estimate is biased by its noise level, so the reference value is the mean of the two
        # clipped renders (same spp as the image under test), not the clipped mean of their average
        "mean_err": [abs(float(cg[:, c].mean() - cr[c])) / max(cr[c], 1e-9) for c in range(3)],
        "mean_err_ref_vs_ref": [abs(float(ca[:, c].mean() - cb[:, c].mean())) / max(cr[c], 1e-9) for c in range(3)],
        "rmse_floor_clipped": rmt(ca, cb), "rmse_gpu_clipped": float(np.sqrt(0.5 * (rmt(cg, ca) ** 2 + rmt(cg, cb) ** 2))),
        "rmse_floor": rm(a, b), "rmse_gpu": float(np.sqrt(0.5 * (rm(img, a) ** 2 + rm(img, b) ** 2))),
    }


_FLOOR = None


def noise_floor(scene):
    """tests/golden/noise_floor.json (make_noise_floor.py): distribution of these statistics over all triples of eight
    independent-seed renders of the UNMODIFIED REFERENCE of `scene` at 128x96 /smp 32."""
    global _FLOOR
    if _FLOOR is None:
        import json
        import os
        with open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "noise_floor.json")) as f:
            _FLOOR = json.load(f)
    return _FLOOR[scene]


def check_converged(m, npix, scene):
    """north_star check 2.  "RMSE within the Monte-Carlo noise floor": an image of the same estimator is exchangeable with
    a reference render, so its ratio rmse_gpu / rmse_floor must not exceed what the reference's own renders reach against
    each other — the recorded maximum over 168 reference triples (1.03 cornell-box / minimal0 / room, 1.14 refrac0, 1.26
    glass0; +2 % for the finite sample), not a constant.  "Per-channel mean error < 1 %": literal, except where the
    reference misses that against itself at this sample count (cornell-box at 1024 spp: up to 1.3 %), then its own maximum."""
    nf = noise_floor(scene)
    assert m["masked"] <= 0.01 * npix, m
    assert max(m["mean_err"]) < max(0.01, nf["mean_err_clipped"]["max"]), m
    ratio = m["rmse_gpu_clipped"] / max(m["rmse_floor_clipped"], 1e-30)
    assert ratio <= nf["ratio_clipped"]["max"] * 1.02, (ratio, nf["ratio_clipped"], m)
