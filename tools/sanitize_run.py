#!/usr/bin/env python
"""Small inputs through every kernel family of libplugpu.so in a few seconds — the command to put under
`compute-sanitizer --tool memcheck --error-exitcode 9` where that tool is available (it is closed on this round's GPU pool:
the run there was plain, rc 0):
staged kernels (cornell-box: k_primary / k_extend / k_shade / k_connect / k_accumulate, graph replay), the same scene
through the global-memory kernels (PLU_RENDER_BVH_FAST), a SAH-tree scene (k_generate / k_extend_pt / k_connect_pt with the
seeded area-light walk), a scene with spheres (pruned walk of the reference tree), both device builds, the parity hooks
(trace_primary / trace_rays / sample_radiance), tone map + BMP encoders incl. the no-allocation one.  Prints one line per
step; any CUDA error raises."""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    from plutracer_b200 import GpuContext, Scene, _ffi
    from tools.gen_synth import make_scene
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

    def golden_scene(name, w, h, n):
        return Scene.load_flat(os.path.join(root, "tests", "golden", name + ".pluscene")).override(w, h, n)
    REF, FAST = _ffi.RENDER_BVH_REFERENCE, _ffi.RENDER_BVH_FAST
    ctx = GpuContext(0)
    steps = 0

    def render(sc, w, h, n, flags=0, seed=7):
        p = GpuContext.params(w, h, n, seed=seed, flags=flags)
        img = np.zeros((h, w, 3), np.float32)
        st = ctx.render(p, img)
        assert np.isfinite(img).any()
        return img, st

    sc = golden_scene("cornell-box", 48, 40, 3)
    ctx.upload(sc)
    a, st = render(sc, 48, 40, 3)
    b, _ = render(sc, 48, 40, 3)               # second call replays the wave graph
    assert np.array_equal(a.view(np.uint32), b.view(np.uint32))
    print("staged", st.samples, "samples", st.kernel_launches, "launches"); steps += 1
    g, st = render(sc, 48, 40, 3, FAST)        # global-memory kernels on the same scene
    print("global", st.samples, "samples", st.kernel_launches, "launches"); steps += 1
    rgb = a.copy()
    ctx.postprocess(rgb)
    bmp = ctx.encode_bmp(a, tonemap=True)
    print("post + bmp", len(bmp), "bytes"); steps += 1
    ids, t, nrm = ctx.trace_primary(48, 40, REF)
    s = ctx.sample_radiance(GpuContext.params(48, 40, 3, seed=7), 0, 48 * 4)
    print("hooks", int((ids >= 0).sum()), "primary hits", s.shape); steps += 1

    rng = np.random.default_rng(3)
    for kind in ("mixed", "spheres"):
        sc = make_scene(6000, seed=2, kind=kind, width=64, height=48, n_strata=2)
        for build in ("device", "host"):
            ctx.upload(sc, build=build)
        img, st = render(sc, 64, 48, 2)
        ro = rng.uniform(-55, 55, (2000, 3)).astype(np.float32)
        rd = rng.normal(size=(2000, 3)).astype(np.float32)
        x = ctx.trace_rays(ro, rd, REF)
        y = ctx.trace_rays(ro, rd, 0)
        print(kind, st.samples, "samples", st.kernel_launches, "launches;", int((x[0] >= 0).sum()), "/", int((y[0] >= 0).sum()), "ray hits"); steps += 1
    ctx.close()
    print("sanitize_run: ok,", steps, "steps")


if __name__ == "__main__":
    main()
