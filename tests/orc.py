"""ctypes binding of the CPU oracle (oracle/_build/liborc.so).  TEST INFRASTRUCTURE: only tests/,
__graft_entry__.smoke() and bench.py's CPU-baseline legs import this."""
import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORC_PATH = os.path.join(ROOT, "oracle", "_build", "liborc.so")
GOLDEN = os.path.join(ROOT, "tests", "golden")


def build_oracle():
    src = [os.path.join(ROOT, "oracle", f) for f in ("oracle_pt.cpp", "oracle_pt.h", "oracle_math.h")]
    if not os.path.exists(ORC_PATH) or any(os.path.getmtime(s) > os.path.getmtime(ORC_PATH) for s in src):
        os.makedirs(os.path.dirname(ORC_PATH), exist_ok=True)
        subprocess.run(["g++", "-std=c++17", "-O2", "-ffp-contract=off", "-fPIC", "-shared", "-pthread", "-o", ORC_PATH, src[0]], check=True)
    return ORC_PATH


def load_scene(path):
    from plutracer_b200 import Scene
    return Scene.load_flat(path)


def golden_scene(name, width=0, height=0, n_strata=0):
    sc = load_scene(os.path.join(GOLDEN, name + ".pluscene"))
    return sc.override(width, height, n_strata)


class Oracle:
    def __init__(self):
        self.lib = C.CDLL(build_oracle())
        self.lib.orc_bvh_nodes.restype = C.c_int

    @staticmethod
    def _p(a):
        return a.ctypes.data_as(C.c_void_p)

    def trace_rays(self, scene, o, d, trig=0):
        o = np.ascontiguousarray(o, np.float32); d = np.ascontiguousarray(d, np.float32)
        n = o.shape[0]
        ids, t, nrm = np.zeros(n, np.int32), np.zeros(n, np.float32), np.zeros((n, 3), np.float32)
        rc = self.lib.orc_trace_rays(scene.desc_ptr, C.c_int64(n), self._p(o), self._p(d), trig, self._p(ids), self._p(t), self._p(nrm))
        assert rc == 0
        return ids, t, nrm

    def trace_primary(self, scene, w, h, trig=0):
        n = w * h
        ids, t, nrm = np.zeros(n, np.int32), np.zeros(n, np.float32), np.zeros((n, 3), np.float32)
        rc = self.lib.orc_trace_primary(scene.desc_ptr, w, h, trig, self._p(ids), self._p(t), self._p(nrm))
        assert rc == 0
        return ids, t, nrm

    def primary_with_rays(self, scene, w, h):
        """primary hits + the rays themselves (numpy restatement of camera::generate_ray, float32)."""
        ids, t, nrm = self.trace_primary(scene, w, h, 1)
        cam = scene.desc.cam
        f = np.float32
        pos, look = np.array(cam.pos[:], f), np.array(cam.look[:], f)
        up, right = np.array(cam.up[:], f), np.array(cam.right[:], f)
        xs, ys = np.meshgrid(np.arange(w, dtype=f) + f(0.5), np.arange(h, dtype=f) + f(0.5))
        ux = (xs * f(cam.inv_image_size[0])) * f(2) - f(1)
        uy = ((ys * f(cam.inv_image_size[1])) * f(2) - f(1)) * f(-1)
        d = (f(cam.w) * look)[None, None, :] + ux[..., None] * right + uy[..., None] * up
        dd = (d * d)
        inv = f(1) / np.sqrt((dd[..., 0] + dd[..., 1]) + dd[..., 2], dtype=f)
        d = (d * inv[..., None]).reshape(-1, 3).astype(f)
        o = np.broadcast_to(pos, d.shape).copy()
        return ids, t, nrm, o, d

    def leaf_order(self, scene):
        rank = np.zeros(max(scene.n_prims, 1), np.int32)
        assert self.lib.orc_leaf_order(scene.desc_ptr, self._p(rank)) == 0
        return rank[: scene.n_prims]

    def bvh_nodes(self, scene):
        n = 2 * max(scene.n_prims, 1)
        lro, b6 = np.zeros((n, 3), np.int32), np.zeros((n, 6), np.float32)
        cnt = self.lib.orc_bvh_nodes(scene.desc_ptr, n, self._p(lro), self._p(b6))
        return lro[:cnt], b6[:cnt]

    def nan_slab_rays(self, scene, o, d):
        """Rays for which an interior node of the reference tree evaluates 0 * inf (d_k == 0 with the
        origin exactly on a slab plane): the reference's outcome then hinges on NaN comparisons."""
        lro, b6 = self.bvh_nodes(scene)
        inner = lro[:, 2] < 0
        planes = b6[inner]
        bad = np.zeros(len(o), bool)
        for k in range(3):
            zero = np.flatnonzero(d[:, k] == 0)   # only these rays can evaluate 0 * inf on axis k
            if zero.size == 0:
                continue
            vals = np.unique(np.concatenate([planes[:, k], planes[:, 3 + k]]))
            bad[zero] |= np.isin(o[zero, k], vals)
        return bad

    def sample_radiance(self, scene, params, pixel_begin, pixel_count, trig=0):
        count = params.sample_count if params.sample_count > 0 else params.n_strata ** 2
        out = np.zeros((pixel_count, count, 3), np.float32)
        rc = self.lib.orc_sample_radiance(scene.desc_ptr, C.byref(params), C.c_int64(pixel_begin), C.c_int64(pixel_count), trig, self._p(out))
        assert rc == 0
        return out

    def render(self, scene, params, out=None, trig=0, n_threads=0):
        from plutracer_b200 import _ffi
        if out is None:
            out = np.zeros((params.height, params.width, 3), np.float32)
        st = _ffi.RenderStats()
        rc = self.lib.orc_render(scene.desc_ptr, C.byref(params), trig, n_threads, self._p(out), C.byref(st))
        assert rc == 0
        return out, st

    def render_mt(self, scene, w, h, n, seed, x0, y0, x1, y1, trig=1, x_first=0, per_sample=True):
        img = np.zeros((y1 - y0, x1 - x0, 3), np.float32)
        ps = np.zeros(((y1 - y0) * (x1 - x0), n * n, 3), np.float32) if per_sample else None
        rc = self.lib.orc_render_mt(scene.desc_ptr, w, h, n, C.c_uint32(seed), x0, y0, x1, y1, trig, x_first, self._p(img),
                                    self._p(ps) if per_sample else None)
        assert rc == 0
        return img, ps

    def postprocess(self, rgb, pow_mode=0):
        out = np.ascontiguousarray(rgb, np.float32).copy()
        self.lib.orc_postprocess(C.c_int64(out.shape[0] * out.shape[1]), self._p(out), int(pow_mode))
        return out

    def encode_bmp(self, rgb):
        rgb = np.ascontiguousarray(rgb, np.float32)
        h, w = rgb.shape[:2]
        self.lib.orc_encode_bmp.restype = C.c_int64
        cap = 54 + (w * 3 + ((-w * 3) & 3)) * h
        out = np.zeros(cap, np.uint8)
        n = self.lib.orc_encode_bmp(w, h, self._p(rgb), self._p(out), C.c_int64(cap))
        assert n == cap
        return out

    def philox(self, ctr, key):
        c = (C.c_uint32 * 4)(*ctr); k = (C.c_uint32 * 2)(*key); o = (C.c_uint32 * 4)()
        self.lib.orc_philox4x32_10(c, k, o)
        return list(o)

    def det_trig(self, which, x):
        x = np.ascontiguousarray(x, np.float32); y = np.zeros_like(x)
        self.lib.orc_det_trig(which, C.c_int64(x.size), self._p(x), self._p(y))
        return y
