#!/usr/bin/env python
"""Where the converged-image bound comes from (VERDICT r1 "weak" 1): the spread of the north_star statistics
among the REFERENCE's own renders.

    python tests/golden/make_noise_floor.py [--seeds 8]        (build container only: needs oracle/_ref/plu_ref)

For each of the reference's five scenes: K independent-seed renders of the unmodified reference (`plu_ref render-st`:
its camera::generate_ray + renderer::ray_color per sample, one mt19937 stream per pixel) at 128x96 /smp 32 — the
size of tests/golden/*.converged.npz.  If the GPU path is the same estimator, an image x of it is exchangeable
with a reference render, so the statistic the tests apply to (x; a, b)

    ratio = sqrt(0.5 * (RMSE(x, a)^2 + RMSE(x, b)^2)) / RMSE(a, b)          (tests/metrics.py converged_metrics)

must fall inside the distribution of the same statistic over all K * C(K-1, 2) triples (x; a, b) of DISTINCT
REFERENCE renders.  tests/golden/noise_floor.json records that distribution (clipped and plain RMSE, and the
per-channel mean error) per scene; tests/metrics.py takes its bound from the recorded maximum instead of a bare
constant.  The cornell-box renders themselves are kept (cornell-box.seeds.npz) so the GPU image can also be
compared against every pair."""
import argparse
import itertools
import json
import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
REF_BIN = os.path.join(ROOT, "oracle", "_ref", "plu_ref")
REF_SCENES = os.environ.get("PLU_REFERENCE", "/root/reference") + "/scenes"
SCENES = ["minimal0", "cornell-box", "glass0", "refrac0", "room"]
W, H, N = 128, 96, 32


def render(scene, seed, tmp):
    out = os.path.join(tmp, f"{scene}_{seed}.bin")
    if not os.path.exists(out):
        subprocess.run([REF_BIN, "render-st", f"{REF_SCENES}/{scene}.urn", str(W), str(H), str(N), str(seed), out], check=True, capture_output=True)
    return np.fromfile(out, np.float32).reshape(H, W, 3)


def dist(v):
    v = np.asarray(v, np.float64)
    return {"min": float(v.min()), "median": float(np.median(v)), "p95": float(np.percentile(v, 95)), "max": float(v.max()), "n": int(v.size)}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--seeds", type=int, default=8)
    a = ap.parse_args()
    from tests.metrics import converged_metrics
    tmp = "/tmp/plu_noise_floor"
    os.makedirs(tmp, exist_ok=True)
    seeds = [1000 + 37 * k for k in range(a.seeds)]
    out = {"config": {"width": W, "height": H, "smp": N, "seeds": seeds, "mode": "plu_ref render-st (unmodified reference, one mt19937 stream per pixel)"}}
    with ThreadPoolExecutor(os.cpu_count() or 8) as ex:
        imgs = {s: list(ex.map(lambda sd, s=s: render(s, sd, tmp), seeds)) for s in SCENES}
    for s in SCENES:
        im = imgs[s]
        rc, rp, me = [], [], []
        for x in range(len(im)):
            others = [k for k in range(len(im)) if k != x]
            for i, j in itertools.combinations(others, 2):
                m = converged_metrics(im[x], im[i], im[j])
                rc.append(m["rmse_gpu_clipped"] / max(m["rmse_floor_clipped"], 1e-30))
                rp.append(m["rmse_gpu"] / max(m["rmse_floor"], 1e-30))
                me.append(max(m["mean_err"]))
        plain_mean = [[float(np.nanmean(np.where(np.isfinite(i[..., c]), i[..., c], np.nan))) for c in range(3)] for i in im]
        out[s] = {"ratio_clipped": dist(rc), "ratio_plain": dist(rp), "mean_err_clipped": dist(me),
                  "plain_image_mean_per_seed": plain_mean,
                  "plain_mean_spread": float((np.max(plain_mean, axis=0) / np.maximum(np.min(plain_mean, axis=0), 1e-30)).max() - 1.0)}
        print(s, json.dumps(out[s]["ratio_clipped"]), json.dumps(out[s]["ratio_plain"]), json.dumps(out[s]["mean_err_clipped"]), flush=True)
    with open(os.path.join(HERE, "noise_floor.json"), "w") as f:
        json.dump(out, f, indent=1)
    np.savez_compressed(os.path.join(HERE, "cornell-box.seeds.npz"), cfg=np.array([W, H, N]), seeds=np.array(seeds), imgs=np.stack(imgs["cornell-box"]))


if __name__ == "__main__":
    main()
