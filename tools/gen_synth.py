#!/usr/bin/env python
"""Synthetic BVH-stress scene (BASELINE.json configs[4]): a jittered lattice of n small primitives.

    python tools/gen_synth.py --n 1000000 --seed 1 --out synth1m.pluscene [--urn synth.urn] [--kind boxes|mixed]

Primitives are boxes (kind=boxes) or boxes and triangles (kind=mixed) — no spheres: sphere::hit assumes unit
directions and the reference feeds it non-unit ones, which no geometric BVH can follow, so a
sphere-free scene is the one on which the fast BVH is exact for every ray.  m^3 cells in [-50, 50]^3 with
m = ceil(n^(1/3)), jitter U(-0.15, 0.15) of a cell, half extents U(0.15, 0.35) of a cell, eight diffuse
materials (albedo from splitmix64), a floor box, one point light and one box area light; camera
[0 0 -140] -> [0 0 0], 1024x1024, /smp 8.  Surfaces are emitted in lattice order; the tie ranks of the
reference tree come from libplugpu's restatement of the reference build.

The same tables can be written as .urn text (--urn; exact decimal floats) to go through the scene loader —
use a small n for that: the reference's loader is quadratic-ish in scene size.
"""
import argparse
import ctypes as C
import os
import sys
import tempfile

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

MASK = (1 << 64) - 1


def splitmix64(x):
    x = (x + 0x9E3779B97F4A7C15) & MASK
    z = x
    z = ((z ^ (z >> 30)) * 0xBF58476D1CE4E5B9) & MASK
    z = ((z ^ (z >> 27)) * 0x94D049BB133111EB) & MASK
    return z ^ (z >> 31)


def camera_for(pos, target, width, height):
    """camera::camera (camera.h:17-23) through the scene loader, so the bits are the loader's."""
    from plutracer_b200 import Scene
    txt = (f"[\n resolution: [{width} {height}]\n antialiasing-samples: 1\n camera: [ position: [{pos[0]} {pos[1]} {pos[2]}] "
           f"target: [{target[0]} {target[1]} {target[2]}] ]\n materials: [ ]\n objects: [ ]\n]\n")
    with tempfile.NamedTemporaryFile("w", suffix=".urn", delete=False) as f:
        f.write(txt)
    try:
        sc = Scene.load_urn(f.name)
        cam = type(sc.desc.cam).from_buffer_copy(sc.desc.cam)
    finally:
        os.unlink(f.name)
    return cam


def make_tables(n, seed=1, kind="boxes"):
    from plutracer_b200 import _ffi
    rng = np.random.default_rng(seed)
    m = int(np.ceil(n ** (1.0 / 3.0)))
    cell = 100.0 / m
    idx = np.arange(n)
    ix, iy, iz = idx % m, (idx // m) % m, idx // (m * m)
    c = (np.stack([ix, iy, iz], 1) + 0.5) * cell - 50.0 + rng.uniform(-0.15, 0.15, (n, 3)) * cell
    c = c.astype(np.float32)
    e = (rng.uniform(0.15, 0.35, (n, 3)) * cell).astype(np.float32)
    data = np.zeros((n + 2, 12), np.float32)
    tag = np.full(n + 2, _ffi.PRIM_BOX, np.int32)
    data[:n, 0:3] = c - e
    data[:n, 4:7] = c + e
    uv = nr = None
    if kind == "spheres":   # SURVEY.md 8(d) config 5 as written: a lattice of spheres (radius = the x half extent)
        data[:n] = 0
        data[:n, 0:3] = c
        data[:n, 3] = e[:, 0]
        tag[:n] = _ffi.PRIM_SPHERE
    if kind == "mixed":
        tri = rng.random(n) < 0.5
        k = int(tri.sum())
        v = c[tri, None, :] + (rng.uniform(-1, 1, (k, 3, 3)) * e[tri, None, :] * 1.5).astype(np.float32)
        data[:n][tri] = 0
        blk = np.zeros((k, 12), np.float32)
        blk[:, 0:3], blk[:, 4:7], blk[:, 8:11] = v[:, 0], v[:, 1], v[:, 2]
        data[:n][tri] = blk
        tag[:n][tri] = _ffi.PRIM_TRIANGLE
        uv = np.zeros((n + 2, 6), np.float32)
        uv[:, 2] = 1; uv[:, 5] = 1
        nr = np.zeros((n + 2, 9), np.float32)
        g = np.cross(v[:, 1] - v[:, 0], v[:, 2] - v[:, 0])
        g /= np.maximum(np.linalg.norm(g, axis=1, keepdims=True), 1e-20)
        nr[:n][tri] = np.repeat(g, 3, axis=0).reshape(k, 9).astype(np.float32)
    mat = np.zeros(n + 2, np.int32)
    mat[:n] = idx % 8
    # floor (material 8) and the area light's box (emission, material 9)
    data[n, 0:3], data[n, 4:7] = (-60, -52.2, -60), (60, -52, 60)
    mat[n] = 8
    data[n + 1, 0:3], data[n + 1, 4:7] = (-10, 69.9, -10), (10, 70.1, 10)
    mat[n + 1] = 9
    materials = []
    for k in range(8):
        h = splitmix64(seed * 8 + k)
        md = _ffi.MaterialDesc(); md.tag, md.texture, md.light = _ffi.MAT_DIFFUSE, -1, -1
        md.color[:] = [0.15 + 0.7 * ((h >> s) & 0xFFFF) / 65535.0 for s in (0, 16, 32)]
        materials.append(md)
    fl = _ffi.MaterialDesc(); fl.tag, fl.texture, fl.light = _ffi.MAT_DIFFUSE, -1, -1; fl.color[:] = [0.5, 0.5, 0.5]
    em = _ffi.MaterialDesc(); em.tag, em.texture, em.light = _ffi.MAT_EMISSION, -1, 1
    materials += [fl, em]
    pl = _ffi.LightDesc(); pl.tag, pl.prim = _ffi.LIGHT_POINT, -1; pl.a[:] = [0, 80, 0]; pl.b[:] = [20000, 20000, 20000]
    al = _ffi.LightDesc(); al.tag, al.prim = _ffi.LIGHT_DIFFUSE_AREA, n + 1; al.b[:] = [50, 50, 50]
    d = (data[n + 1, 4:7] - data[n + 1, 0:3]).astype(np.float32)
    al.area = float(np.float32(2) * (d[0] * d[1] + d[0] * d[2] + d[1] * d[2]))   # box::area, box.h:19-21
    return dict(prim_tag=tag, prim_material=mat, prim_data=data, materials=materials, lights=[pl, al], tri_uv=uv, tri_normal=nr)


def make_scene(n, seed=1, kind="boxes", width=1024, height=1024, n_strata=8):
    from plutracer_b200.scene import scene_from_tables
    t = make_tables(n, seed, kind)
    cam = camera_for((0, 0, -140), (0, 0, 0), width, height)
    return scene_from_tables(cam, width=width, height=height, n_strata=n_strata, **t)


def write_urn(path, n, seed=1, width=1024, height=1024, n_strata=8):
    """boxes only; box [c] [e] is re-derived from min/max as exact decimals of the float32 midpoint / half extent"""
    from plutracer_b200 import _ffi
    t = make_tables(n, seed, "boxes")
    f32 = np.float32

    def dec(x):
        return np.format_float_positional(f32(x), unique=True, trim="0")

    with open(path, "w") as f:
        f.write(f"[\n\tresolution: [{width} {height}]\n\tantialiasing-samples: {n_strata}\n\tcamera: [ position: [0 0 -140] target: [0 0 0] up: [0 1 0] ]\n\tmaterials: [\n")
        for k, m in enumerate(t["materials"][:9]):
            f.write(f"\t\tm{k}: [ diffuse [{dec(m.color[0])} {dec(m.color[1])} {dec(m.color[2])}] ]\n")
        f.write("\t]\n\tobjects: [\n")
        data, mat = t["prim_data"], t["prim_material"]
        for i in range(n + 1):
            lo, hi = data[i, 0:3], data[i, 4:7]
            c, e = (lo + hi) * f32(0.5), (hi - lo) * f32(0.5)
            f.write(f"\t\tbox [{dec(c[0])} {dec(c[1])} {dec(c[2])}] [{dec(e[0])} {dec(e[1])} {dec(e[2])}] 'm{mat[i]}\n")
        f.write("\t\tpoint-light [0 80 0] [20000 20000 20000]\n")
        f.write("\t\tdiffuse-area-light [ box [0 70 0] [10 0.1 10] ] [50 50 50]\n\t]\n]\n")


def write_obj_scene(directory, n, seed=1, width=1024, height=1024, n_strata=8):
    """SURVEY.md 8(d) config 5, second flavour: n triangles as ONE triangle-mesh OBJ with full v/vt/vn corners
    (scene.h:138-201) next to the floor box and the two lights; returns the path of the .urn.  Vertices are the
    lattice triangles of kind=mixed (every cell a triangle), written with 9 significant digits."""
    rng = np.random.default_rng(seed)
    m = int(np.ceil(n ** (1.0 / 3.0)))
    cell = 100.0 / m
    idx = np.arange(n)
    c = (np.stack([idx % m, (idx // m) % m, idx // (m * m)], 1) + 0.5) * cell - 50.0 + rng.uniform(-0.15, 0.15, (n, 3)) * cell
    e = rng.uniform(0.15, 0.35, (n, 3)) * cell
    v = (c[:, None, :] + rng.uniform(-1, 1, (n, 3, 3)) * e[:, None, :] * 1.5).astype(np.float32)
    g = np.cross(v[:, 1] - v[:, 0], v[:, 2] - v[:, 0])
    g = (g / np.maximum(np.linalg.norm(g, axis=1, keepdims=True), 1e-20)).astype(np.float32)
    os.makedirs(directory, exist_ok=True)
    obj = os.path.join(directory, "synth_tri.obj")
    with open(obj, "w") as f:
        f.write("# tools/gen_synth.py write_obj_scene\nvt 0 0\nvt 1 0\nvt 0 1\n")
        step = 50000
        for b in range(0, n, step):
            vv, gg = v[b:b + step].reshape(-1, 3), g[b:b + step]
            f.write("".join("v %.9g %.9g %.9g\n" % tuple(r) for r in vv.tolist()))
            f.write("".join("vn %.9g %.9g %.9g\n" % tuple(r) for r in gg.tolist()))
        for b in range(0, n, step):
            k = np.arange(b, min(b + step, n))
            f.write("".join("f %d/1/%d %d/2/%d %d/3/%d\n" % (3 * i + 1, i + 1, 3 * i + 2, i + 1, 3 * i + 3, i + 1) for i in k.tolist()))
    urn = os.path.join(directory, "synth_tri.urn")
    with open(urn, "w") as f:
        f.write(f"[\n\tresolution: [{width} {height}]\n\tantialiasing-samples: {n_strata}\n\tcamera: [ position: [0 0 -140] target: [0 0 0] up: [0 1 0] ]\n"
                "\tmaterials: [\n\t\tfloor: [ diffuse [0.5 0.5 0.5] ]\n\t]\n\tobjects: [\n"
                "\t\ttriangle-mesh \"synth_tri.obj\" [ diffuse [0.7 0.6 0.4] ]\n"
                "\t\tbox [0 -52.1 0] [60 0.1 60] 'floor\n"
                "\t\tpoint-light [0 80 0] [20000 20000 20000]\n"
                "\t\tdiffuse-area-light [ box [0 70 0] [10 0.1 10] ] [50 50 50]\n\t]\n]\n")
    return urn


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--n", type=int, default=1000000)
    ap.add_argument("--seed", type=int, default=1)
    ap.add_argument("--kind", default="boxes")
    ap.add_argument("--out", default="")
    ap.add_argument("--urn", default="")
    ap.add_argument("--obj-dir", default="", help="write synth_tri.obj + synth_tri.urn (n triangles as one triangle-mesh) here")
    a = ap.parse_args()
    if a.obj_dir:
        print(write_obj_scene(a.obj_dir, a.n, a.seed))
    if a.urn:
        write_urn(a.urn, a.n, a.seed)
    if a.out:
        make_scene(a.n, a.seed, a.kind).save_flat(a.out)
