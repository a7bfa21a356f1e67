#!/usr/bin/env python
"""Generate tests/golden/* from the UNMODIFIED reference compiled into oracle/_ref/plu_ref.

Run in the build container (needs /root/reference; the GPU box has neither, it only reads the
committed fixtures):    python tests/golden/make_golden.py [--converged]

Per named scene S in the reference's scenes/ directory:
  S.pluscene        flat tables of the loaded scene (integration/plu_gpu_renderer.h flattener run
                    inside the reference process; includes the real bvh_tree DFS leaf ranks)
  S.golden.npz      primary_{id,t,n}   pixel-centre pinhole rays at 96x72 through camera::generate_ray
                                       + bvh_tree::hit
                    rays_{o,d,id,t,n}  4096 seeded arbitrary rays (incl. axis-aligned, zero and
                                       non-unit directions, origins on surfaces) through bvh_tree::hit
                    st_img, st_samples `render-st` (single thread, mt19937 reseeded per pixel with
                                       seed 7): 32x24 at /smp 2, per-pixel image and per-sample radiance
                    x_first            evaluation order g++ gave vec2(next(), next()) (0: y drawn first)
With --converged (slow, ~25 min of CPU): two independent-seed `render` runs per scene at 128x96
/smp 32 and cornell-box at 256x256 /smp 64 -> S.converged.npz (float16-free: float32, compressed).
"""
import argparse
import json
import os
import subprocess
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
REF_BIN = os.path.join(ROOT, "oracle", "_ref", "plu_ref")
REF_SCENES = os.environ.get("PLU_REFERENCE", "/root/reference") + "/scenes"
SCENES = ["minimal0", "cornell-box", "glass0", "refrac0", "room"]   # the reference's own scenes/ (BASELINE.json configs)
OWN_SCENES = ["atrium", "colonnade", "shard", "gallery", "trilight"]   # this repo's scenes/: all material kinds, scripted objects + lens, triangle mesh
PW, PH = 96, 72
SW, SH, SN, SSEED = 32, 24, 2, 7


def _run(*args, env=None, cwd=None):
    e = dict(os.environ)
    e.update(env or {})
    res = subprocess.run([REF_BIN] + [str(a) for a in args], capture_output=True, text=True, env=e, cwd=cwd)
    if res.returncode != 0:
        raise RuntimeError(f"plu_ref {args}: {res.stderr}")
    return res.stdout


def read_hits(path, n):
    raw = np.fromfile(path, dtype=np.uint8)
    assert raw.size == n * 20, (raw.size, n)
    return raw[: 4 * n].view(np.int32).copy(), raw[4 * n: 8 * n].view(np.float32).copy(), raw[8 * n:].view(np.float32).reshape(n, 3).copy()


def make_rays(scene_npz_primary, rng, n=4096):
    """Seeded rays: random, axis-aligned, zero/non-unit directions, and origins ON surfaces
    (primary hit points re-shot along axes: the no-epsilon self-intersection cases)."""
    o = rng.uniform(-8, 8, (n, 3)).astype(np.float32)
    d = rng.normal(size=(n, 3)).astype(np.float32)
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    k = n // 8
    axes = np.eye(3, dtype=np.float32)
    d[:k] = axes[rng.integers(0, 3, k)] * rng.choice([-1.0, 1.0], (k, 1)).astype(np.float32)      # axis-aligned
    d[k:2 * k] *= rng.uniform(0.1, 3.0, (k, 1)).astype(np.float32)                                # non-unit
    d[2 * k:2 * k + 8] = 0.0                                                                         # zero direction
    o[3 * k:4 * k] = np.round(o[3 * k:4 * k])                                                     # lattice origins (slab-plane hits)
    cam_o, hit_p = scene_npz_primary
    m = min(len(hit_p), 2 * k)
    if m:
        sel = rng.choice(len(hit_p), m, replace=False)
        o[4 * k:4 * k + m] = hit_p[sel]                                                              # origins on surfaces
        d[4 * k:4 * k + m // 2] = axes[rng.integers(0, 3, m // 2)] * rng.choice([-1.0, 1.0], (m // 2, 1)).astype(np.float32)
    return o, d


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--converged", action="store_true")
    ap.add_argument("--only", default="", help="regenerate one scene only (own seeded ray set)")
    ap.add_argument("--big", default="", help="path of a pre-rendered cornell 256x256 /smp 64 pair prefix (a/b .bin)")
    ap.add_argument("--bands1024", default="", help="directory of pre-rendered cornell-box bands at the 1024x1024 camera, /smp 64 "
                    "(plu_ref render-st <urn> 1024 1024 64 <seed> s<seed>_y<y0>.bin 0 y0 1024 y0+32, seeds 101 and 202): "
                    "-> cornell-box.converged1024.npz")
    args = ap.parse_args()
    if not os.path.exists(REF_BIN):
        sys.exit("build oracle/_ref/plu_ref first (make -C oracle)")
    tmp = "/tmp/plu_golden"
    os.makedirs(tmp, exist_ok=True)
    rng = np.random.default_rng(20261104)
    todo = SCENES + OWN_SCENES
    if args.only:
        todo = [args.only]
        rng = np.random.default_rng(20261104 + sum(map(ord, args.only)))
    for s in todo:
        own = s in OWN_SCENES
        urn = f"{ROOT}/scenes/{s}.urn" if own else f"{REF_SCENES}/{s}.urn"
        cwd = f"{ROOT}/scenes" if own else None   # the reference opens triangle-mesh files relative to the cwd

        def run(*a, env=None, _cwd=cwd):
            return _run(*a, env=env, cwd=_cwd)
        run("dump-scene", urn, 0, 0, 0, f"{HERE}/{s}.pluscene")
        run("primary", urn, PW, PH, f"{tmp}/p.bin")
        pid, pt, pn = read_hits(f"{tmp}/p.bin", PW * PH)
        # hit points of the primary rays need the camera: re-derive from a flat dump at PWxPH
        run("dump-scene", urn, PW, PH, 0, f"{tmp}/s.pluscene")
        sys.path.insert(0, ROOT)
        from tests.orc import Oracle, load_scene
        sc = load_scene(f"{tmp}/s.pluscene")
        cam = sc.desc.cam
        # pixel-centre directions via the oracle's camera (bit-identical with the reference: checked by primary ids below)
        oid, ot, on, rays_o, rays_d = Oracle().primary_with_rays(sc, PW, PH)
        assert np.array_equal(oid, pid) and np.array_equal(ot.view(np.uint32), pt.view(np.uint32)), "oracle != reference on primary rays"
        hit = pid >= 0
        hit_p = (rays_o[hit] + rays_d[hit] * pt[hit, None]).astype(np.float32)
        o, d = make_rays((np.array(cam.pos[:]), hit_p), rng)
        np.concatenate([o, d], axis=1).astype(np.float32).tofile(f"{tmp}/rays.bin")
        run("rays", urn, f"{tmp}/rays.bin", f"{tmp}/r.bin")
        rid, rt, rn = read_hits(f"{tmp}/r.bin", len(o))
        run("render-st", urn, SW, SH, SN, SSEED, f"{tmp}/st.bin", 0, 0, SW, SH, f"{tmp}/ps.bin")
        st_img = np.fromfile(f"{tmp}/st.bin", np.float32).reshape(SH, SW, 3)
        st_smp = np.fromfile(f"{tmp}/ps.bin", np.float32).reshape(SH * SW, SN * SN, 3)
        np.savez_compressed(f"{HERE}/{s}.golden.npz", primary_id=pid, primary_t=pt, primary_n=pn, primary_wh=np.array([PW, PH]),
                            rays_o=o, rays_d=d, rays_id=rid, rays_t=rt, rays_n=rn,
                            st_img=st_img, st_samples=st_smp, st_cfg=np.array([SW, SH, SN, SSEED]), x_first=np.array(0))
        print(s, "hits", float(hit.mean()), "ray hits", float((rid >= 0).mean()))
        if args.converged and not own:
            out = {}
            for tag, seed in (("a", 101), ("b", 202)):
                js = _run("render", urn, 128, 96, 32, f"{tmp}/c.bin", env={"PLU_SEED": str(seed)})
                out[tag] = np.fromfile(f"{tmp}/c.bin", np.float32).reshape(96, 128, 3)
                out[tag + "_info"] = np.array(json.dumps(json.loads(js.strip().splitlines()[-1])))
            np.savez_compressed(f"{HERE}/{s}.converged.npz", cfg=np.array([128, 96, 32]), **out)
    if args.big:
        a = np.fromfile(args.big + "a.bin", np.float32).reshape(256, 256, 3)
        b = np.fromfile(args.big + "b.bin", np.float32).reshape(256, 256, 3)
        np.savez_compressed(f"{HERE}/cornell-box.converged256.npz", cfg=np.array([256, 256, 64]), a=a, b=b)


def pack_bands1024(directory):
    """BASELINE.json configs[1] at ITS size: the reference's own 4096-spp render of cornell-box with the 1024x1024 camera,
    as eight 32-row bands (a quarter of the image, evenly spaced), two independent seeds.  One band is 134 M paths =
    ~8 min of one core; `for seed in 101 202; for y0 in 96 224 352 480 608 736 864 992: plu_ref render-st
    cornell-box.urn 1024 1024 64 $seed s${seed}_y${y0}.bin 0 $y0 1024 $((y0+32))` (8 processes at a time)."""
    import glob
    import re
    y0s = sorted({int(re.search(r"_y(\d+)\.bin$", f).group(1)) for f in glob.glob(os.path.join(directory, "s101_y*.bin"))})
    out = {}
    for tag, seed in (("a", 101), ("b", 202)):
        out[tag] = np.stack([np.fromfile(os.path.join(directory, f"s{seed}_y{y0}.bin"), np.float32).reshape(32, 1024, 3) for y0 in y0s])
    np.savez_compressed(f"{HERE}/cornell-box.converged1024.npz", cfg=np.array([1024, 1024, 64]), y0=np.array(y0s), rows=np.array(32),
                        seeds=np.array([101, 202]), **out)
    print("bands", y0s, "mean a", float(np.nanmean(out["a"])), "mean b", float(np.nanmean(out["b"])))


if __name__ == "__main__":
    if "--bands1024" in sys.argv:
        pack_bands1024(sys.argv[sys.argv.index("--bands1024") + 1])
    else:
        main()
