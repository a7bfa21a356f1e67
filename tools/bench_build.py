#!/usr/bin/env python
"""Build time of the traversal structures, host restatement vs device kernels.

    python tools/bench_build.py [--n 1000000] [--kind boxes] [--repeat 3]

One JSON line per (n, where): plu_build_stats of the best of `repeat` uploads (the upload of the tables
themselves is outside ms_total)."""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--n", default="1000,100000,1000000")
    ap.add_argument("--kind", default="boxes")
    ap.add_argument("--repeat", type=int, default=3)
    a = ap.parse_args()
    from plutracer_b200 import GpuContext
    from tools.gen_synth import make_scene
    ctx = GpuContext(0)
    for n in [int(x) for x in a.n.split(",")]:
        sc = make_scene(n, 1, a.kind, 64, 64, 2)
        for where in ("host", "device"):
            best = None
            for _ in range(a.repeat):
                bs = ctx.upload(sc, build=where)
                if best is None or bs.ms_total < best.ms_total:
                    best = bs
            print(json.dumps({"n_surfaces": sc.n_prims, "kind": a.kind, "build": where, "ms_total": round(best.ms_total, 3),
                              "ms_ref_order": round(best.ms_ref_order, 3), "ms_ref_nodes": round(best.ms_ref_nodes, 3),
                              "ms_fast_bvh": round(best.ms_fast_bvh, 3), "levels_ref": best.levels_ref, "levels_fast": best.levels_fast,
                              "host_tie_segments": best.host_tie_segments, "n_fast_nodes": best.n_fast_nodes}), flush=True)
    ctx.close()


if __name__ == "__main__":
    main()
