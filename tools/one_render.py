#!/usr/bin/env python
"""One render of a BASELINE.json configuration (for ncu captures): python tools/one_render.py glass0 [--spp K] [--reps R]"""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("config")
    ap.add_argument("--spp", type=int, default=0)
    ap.add_argument("--reps", type=int, default=1)
    ap.add_argument("--flags", type=int, default=0)
    a = ap.parse_args()
    import torch
    import bench
    from plutracer_b200 import GpuContext
    c = bench.CONFIGS[a.config]
    sc = bench.load_scene(a.config)
    ctx = GpuContext(0)
    ctx.upload(sc)
    w, h, n = c["w"], c["h"], c["n"]
    spp = a.spp if a.spp > 0 else n * n
    fb = torch.zeros(h, w, 3, dtype=torch.float32, device="cuda")
    p = GpuContext.params(w, h, n, seed=bench.SEED, flags=a.flags, sample_begin=0, sample_stride=1, sample_count=spp)
    for _ in range(a.reps):
        st = ctx.render_device(p, fb.data_ptr(), torch.cuda.current_stream().cuda_stream)
    print(a.config, "samples", st.samples, "ms_device", round(st.ms_device, 3), "Msamples/s", round(st.samples / st.ms_device / 1e3, 1), "waves", st.waves, "launches", st.kernel_launches)
    ctx.close()


if __name__ == "__main__":
    main()
