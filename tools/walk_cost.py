#!/usr/bin/env python
"""Cost of the walk per ray class on the synthetic sphere lattice: rays that leave sphere surfaces with |d| = s."""
import argparse, json, os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    ap = argparse.ArgumentParser(); ap.add_argument("--n", type=int, default=200000); ap.add_argument("--rays", type=int, default=20000)
    a = ap.parse_args()
    from plutracer_b200 import GpuContext, _ffi
    from tools.gen_synth import make_scene, make_tables
    sc = make_scene(a.n, 1, "spheres", 64, 64, 1)
    t = make_tables(a.n, 1, "spheres")
    ctx = GpuContext(0); ctx.upload(sc)
    rng = np.random.default_rng(3)
    pick = rng.integers(0, a.n, a.rays)
    c, r = t["prim_data"][pick, 0:3], t["prim_data"][pick, 3:4]
    nrm = rng.normal(size=(a.rays, 3)); nrm /= np.linalg.norm(nrm, axis=1, keepdims=True)
    o = (c + nrm * r * 1.0001).astype(np.float32)
    d = rng.normal(size=(a.rays, 3)); d /= np.linalg.norm(d, axis=1, keepdims=True)
    d = np.where((d * nrm).sum(1, keepdims=True) < 0, -d, d)
    for s in (0.3, 0.7, 0.95, 1.0, 1.05, 1.2, 1.41):
        dd = (d * s).astype(np.float32)
        ids, tt, cost = ctx.trace_rays(o, dd, _ffi.RENDER_COUNT_BVH)
        nodes, prims = cost[:, 0], cost[:, 1]
        print(json.dumps({"s": s, "hit": float((ids >= 0).mean()), "nodes_mean": float(nodes.mean()), "nodes_p50": float(np.median(nodes)), "nodes_p99": float(np.percentile(nodes, 99)),
                          "nodes_max": float(nodes.max()), "prims_mean": float(prims.mean()), "prims_max": float(prims.max()), "t_med": float(np.median(tt[ids >= 0])) if (ids >= 0).any() else None}), flush=True)
    # camera-like rays from outside
    o2 = np.tile(np.array([[0, 0, -140]], np.float32), (a.rays, 1)); d2 = rng.normal(size=(a.rays, 3)); d2[:, 2] = np.abs(d2[:, 2]) * 4 + 2; d2 /= np.linalg.norm(d2, axis=1, keepdims=True)
    ids, tt, cost = ctx.trace_rays(o2, d2.astype(np.float32), _ffi.RENDER_COUNT_BVH)
    print(json.dumps({"camera": True, "hit": float((ids >= 0).mean()), "nodes_mean": float(cost[:, 0].mean()), "nodes_max": float(cost[:, 0].max()), "prims_mean": float(cost[:, 1].mean())}))
    ctx.close()


if __name__ == "__main__":
    main()
