#!/usr/bin/env python
"""BVH-traversal stress: the synthetic n-primitive scene (tools/gen_synth.py) on one B200.
Reports Msamples/s, Mrays/s, node/primitive fetch rates and the fraction of the measured HBM peak the
fast-BVH traversal kernels achieve (algorithmic bytes: 64 B per interior node visited + 32 / 48 B per box /
triangle tested + 48 B of ray/hit rows)."""
import argparse
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--n", type=int, default=1000000)
    ap.add_argument("--kind", default="boxes")
    ap.add_argument("--res", type=int, default=1024)
    ap.add_argument("--smp", type=int, default=8)
    ap.add_argument("--steps", type=int, default=3)
    a = ap.parse_args()
    import numpy as np
    from plutracer_b200 import GpuContext, _ffi
    from tools.gen_synth import make_scene
    t0 = time.perf_counter()
    sc = make_scene(a.n, 1, a.kind, a.res, a.res, a.smp)
    t1 = time.perf_counter()
    ctx = GpuContext(0)
    ctx.upload(sc)
    t2 = time.perf_counter()
    img = np.zeros((a.res, a.res, 3), np.float32)
    p = GpuContext.params(a.res, a.res, a.smp, seed=20261104)
    ctx.render(p, img)
    ms = []
    for _ in range(a.steps):
        img[:] = 0
        st = ctx.render(p, img)
        ms.append(st.ms_device)
    pc = GpuContext.params(a.res, a.res, a.smp, seed=20261104, flags=_ffi.RENDER_COUNT_BVH | _ffi.RENDER_PROFILE, sample_count=min(8, a.smp * a.smp), sample_stride=1)
    cs = ctx.render(pc, img)
    rays = cs.rays_extend + cs.rays_shadow
    prim_bytes = 40.0 if a.kind == "mixed" else 32.0
    bvh_bytes = 64.0 * cs.bvh_nodes_visited + prim_bytes * cs.prims_tested
    trav_ms = cs.ms_extend + cs.ms_connect
    peak = json.load(open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")))["hbm_gbs"] if os.path.exists("MEASURED_PEAKS.json") else 6650.0
    dt = min(ms) * 1e-3
    out = {"scene": f"synthetic {a.kind} n={sc.n_prims}", "res": a.res, "spp": a.smp ** 2, "msamples_per_s": st.samples / dt / 1e6,
           "mrays_per_s": (st.rays_extend + st.rays_shadow) / dt / 1e6, "ms_render": min(ms), "scene_build_s": t1 - t0, "upload_bvh_build_s": t2 - t1,
           "nodes_per_ray": cs.bvh_nodes_visited / rays, "prims_per_ray": cs.prims_tested / rays,
           "traversal_gbs_counted_pass": (bvh_bytes + 48.0 * rays) / (trav_ms * 1e-3) / 1e9 if trav_ms > 0 else None,
           "traversal_frac_of_measured_hbm": (bvh_bytes + 48.0 * rays) / (trav_ms * 1e-3) / 1e9 / peak if trav_ms > 0 else None,
           "ms_by_kernel_counted_pass": {"extend": cs.ms_extend, "shade": cs.ms_shade, "connect": cs.ms_connect}, "mean_radiance": float(np.nanmean(img))}
    print(json.dumps(out))


if __name__ == "__main__":
    main()
