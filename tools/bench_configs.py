#!/usr/bin/env python
"""Throughput of the render path on every BASELINE.json configuration, one B200 (device-timed, inputs resident).

    python tools/bench_configs.py [--max-spp 256] [--synth-n 1000000]

One JSON line per configuration: Msamples/s, Mrays/s, rays per sample, per-kernel split.  A configuration
whose full spp would take long is rendered on its first --max-spp strata (stated in the line); throughput
does not depend on which strata are rendered."""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

CONFIGS = [   # (scene, W, H, /smp)  — BASELINE.json configs[0..3], sizes from SURVEY.md 8(d)
    ("minimal0", 512, 512, 8),
    ("cornell-box", 1024, 1024, 64),
    ("glass0", 1280, 960, 8),
    ("refrac0", 1280, 960, 8),
    ("room", 1920, 1080, 32),
]


def run(ctx, sc, name, w, h, n, max_spp, steps=3, flags=0, spw=0):
    import numpy as np
    import torch
    from plutracer_b200 import GpuContext, _ffi
    spp = min(n * n, max_spp)
    ctx.upload(sc)
    fb = torch.zeros(h, w, 3, dtype=torch.float32, device="cuda")
    st_ = torch.cuda.current_stream().cuda_stream
    p = GpuContext.params(w, h, n, seed=20261104, flags=flags, sample_begin=0, sample_stride=1, sample_count=spp, strata_per_wave=spw)
    ctx.render_device(p, fb.data_ptr(), st_)
    best = None
    for _ in range(steps):
        fb.zero_()
        st = ctx.render_device(p, fb.data_ptr(), st_)
        best = st.ms_device if best is None else min(best, st.ms_device)
    pp = GpuContext.params(w, h, n, seed=20261104, flags=flags | _ffi.RENDER_PROFILE | _ffi.RENDER_COUNT_BVH, sample_begin=0, sample_stride=1, sample_count=min(spp, 16))
    ps = ctx.render_device(pp, fb.data_ptr(), st_)
    rays = st.rays_extend + st.rays_shadow
    prays = max(ps.rays_extend + ps.rays_shadow, 1)
    img = fb.cpu().numpy()
    return {"scene": name, "width": w, "height": h, "spp_config": n * n, "spp_rendered": spp, "samples": int(st.samples), "ms_device": best,
            "msamples_per_s": st.samples / best / 1e3, "mrays_per_s": rays / best / 1e3, "rays_per_sample": rays / st.samples,
            "nodes_per_ray": ps.bvh_nodes_visited / prays, "prims_per_ray": ps.prims_tested / prays,
            "ms_by_kernel_profiled_pass": {"generate": ps.ms_generate, "extend": ps.ms_extend, "shade": ps.ms_shade, "connect": ps.ms_connect, "accumulate": ps.ms_accumulate},
            "mean_radiance": float(np.nanmean(img)), "nonfinite_pixels": int((~np.isfinite(img)).any(axis=2).sum())}


def run_cold(sc, name, w, h, n, max_spp):
    """First render of a scene on a FRESH context (nothing learnt, nothing allocated) against the second one:
    wall clock around plu_gpu_render_device and the device-timed span."""
    import time
    import torch
    from plutracer_b200 import GpuContext
    spp = min(n * n, max_spp)
    ctx = GpuContext(0)
    ctx.upload(sc)
    fb = torch.zeros(h, w, 3, dtype=torch.float32, device="cuda")
    torch.cuda.synchronize()
    st_ = torch.cuda.current_stream().cuda_stream
    p = GpuContext.params(w, h, n, seed=20261104, sample_begin=0, sample_stride=1, sample_count=spp)
    out = {"scene": name, "samples": w * h * spp}
    for tag in ("first", "second", "third"):
        t0 = time.perf_counter()
        st = ctx.render_device(p, fb.data_ptr(), st_)
        dt = time.perf_counter() - t0
        out[tag] = {"wall_ms": dt * 1e3, "ms_device": st.ms_device, "msamples_per_s_device": st.samples / st.ms_device / 1e3,
                    "msamples_per_s_wall": st.samples / dt / 1e6, "waves": int(st.waves), "launches": int(st.kernel_launches)}
    ctx.close()
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--cold", action="store_true", help="first vs second render on a fresh context per scene")
    ap.add_argument("--max-spp", type=int, default=256)
    ap.add_argument("--synth-n", type=int, default=1000000)
    ap.add_argument("--only", default="")
    ap.add_argument("--strata-per-wave", type=int, default=0, help="paths in flight per wave = W*H*this (0: library default, 2^22 paths)")
    ap.add_argument("--no-refill", action="store_true", help="PLU_RENDER_NO_LANE_REFILL: the one-ray-per-lane schedule on fast-BVH scenes")
    a = ap.parse_args()
    from plutracer_b200 import GpuContext, Scene, _ffi
    from tools.gen_synth import make_scene
    ctx = GpuContext(0)
    flags = _ffi.RENDER_NO_LANE_REFILL if a.no_refill else 0
    for name, w, h, n in CONFIGS:
        if a.only and name not in a.only.split(","):
            continue
        sc = Scene.load_flat(os.path.join(ROOT, "tests", "golden", name + ".pluscene")).override(w, h, n)
        if a.cold:
            print(json.dumps(run_cold(sc, name, w, h, n, a.max_spp)), flush=True)
            continue
        print(json.dumps(run(ctx, sc, name, w, h, n, a.max_spp, flags=flags, spw=a.strata_per_wave)), flush=True)
    if a.synth_n > 0 and (not a.only or "synth" in a.only):
        for kind in ("boxes", "mixed"):
            sc = make_scene(a.synth_n, 1, kind, 1024, 1024, 8)
            print(json.dumps(run(ctx, sc, f"synthetic {kind} n={sc.n_prims}", 1024, 1024, 8, a.max_spp, flags=flags)), flush=True)
    if a.synth_n > 0 and a.only and "obj" in a.only.split(","):
        # second flavour of config 5: the triangles as ONE triangle-mesh OBJ through the scene loader
        import tempfile
        import time
        from tools.gen_synth import write_obj_scene
        d = tempfile.mkdtemp(prefix="plu_synth_obj")
        t0 = time.time(); urn = write_obj_scene(d, a.synth_n); t1 = time.time()
        sc = Scene.load_urn(urn); t2 = time.time()
        line = run(ctx, sc, f"synthetic triangle-mesh OBJ n={sc.n_prims}", 1024, 1024, 8, a.max_spp, flags=flags)
        line["s_generate_obj"], line["s_load_urn_obj"] = round(t1 - t0, 2), round(t2 - t1, 2)
        print(json.dumps(line), flush=True)
    if a.synth_n > 0 and a.only and "spheres" in a.only.split(","):
        # config 5 as SURVEY.md words it (spheres).  sphere::hit assumes unit directions and the reference feeds it
        # non-unit ones, for which it "hits" spheres far off the ray: those rays must walk the reference's own tree
        # (DESIGN.md 5.4), so this is the slow, exact mode — small image, few samples, bounded time.
        sc = make_scene(a.synth_n, 1, "spheres", 256, 256, 8)
        print(json.dumps(run(ctx, sc, f"synthetic spheres n={sc.n_prims}", 256, 256, 8, min(a.max_spp, 2), steps=1, flags=flags)), flush=True)
    ctx.close()


if __name__ == "__main__":
    main()
