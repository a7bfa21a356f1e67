"""Host-side mirror of the reference's scene surface (inc/scene.h, inc/camera.h, inc/sampler.h,
inc/texture.h) on top of the C ABI: a loaded scene is a `plu_scene` handle owning flat tables."""
import ctypes as C

import numpy as np

from . import _ffi


class Scene:
    """plu::scenes::scene (scene.h:64-299): resolution, samples, camera, surfaces, lights — flattened.

    Scene.load_urn(path, args) takes the reference's CLI overrides (`/res WxH`, `/smp N`,
    scene.h:232-238,251-256) as `args`, e.g. ["/res", "512x512", "/smp", "8"].
    """

    def __init__(self, handle):
        self._h = C.c_void_p(handle) if not isinstance(handle, C.c_void_p) else handle
        self._lib = _ffi.lib()

    @classmethod
    def load_urn(cls, path, args=()):
        lib = _ffi.lib()
        h = C.c_void_p()
        err = C.create_string_buffer(512)
        argv = (C.c_char_p * max(len(args), 1))(*[str(a).encode() for a in args])
        rc = lib.plu_scene_load_urn(str(path).encode(), argv, len(args), C.byref(h), err, len(err))
        if rc != _ffi.PLU_OK:
            raise _ffi.PluError(rc, err.value.decode(errors="replace"))
        return cls(h)

    @classmethod
    def load_flat(cls, path):
        lib = _ffi.lib()
        h = C.c_void_p()
        err = C.create_string_buffer(512)
        rc = lib.plu_scene_load_flat(str(path).encode(), C.byref(h), err, len(err))
        if rc != _ffi.PLU_OK:
            raise _ffi.PluError(rc, err.value.decode(errors="replace"))
        return cls(h)

    @classmethod
    def from_desc(cls, desc, width, height, n_strata):
        lib = _ffi.lib()
        h = C.c_void_p()
        rc = lib.plu_scene_from_desc(C.byref(desc), width, height, n_strata, C.byref(h))
        if rc != _ffi.PLU_OK:
            raise _ffi.PluError(rc, "invalid scene description")
        return cls(h)

    def save_flat(self, path):
        w, h, n = self.defaults
        rc = self._lib.plu_scene_save_flat(self.desc_ptr, w, h, n, str(path).encode())
        if rc != _ffi.PLU_OK:
            raise _ffi.PluError(rc, f"cannot write {path}")

    def override(self, width=0, height=0, n_strata=0):
        """`/res WxH` and `/smp N`."""
        rc = self._lib.plu_scene_override(self._h, width, height, n_strata)
        if rc != _ffi.PLU_OK:
            raise _ffi.PluError(rc, "override failed")
        return self

    @property
    def desc_ptr(self):
        return self._lib.plu_scene_get_desc(self._h)

    @property
    def desc(self):
        return self.desc_ptr.contents

    @property
    def defaults(self):
        w, h, n = C.c_int32(), C.c_int32(), C.c_int32()
        self._lib.plu_scene_get_defaults(self._h, C.byref(w), C.byref(h), C.byref(n))
        return w.value, h.value, n.value

    @property
    def resolution(self):
        return self.defaults[:2]

    @property
    def samples(self):
        return self.defaults[2]

    @property
    def n_prims(self):
        return self.desc.n_prims

    @property
    def n_lights(self):
        return self.desc.n_lights

    def reference_leaf_order(self):
        rank = np.zeros(max(self.n_prims, 1), np.int32)
        rc = self._lib.plu_scene_reference_leaf_order(self.desc_ptr, rank.ctypes.data)
        if rc != _ffi.PLU_OK:
            raise _ffi.PluError(rc, "leaf order failed")
        return rank[: self.n_prims]

    def __del__(self):
        try:
            if self._h:
                self._lib.plu_scene_free(self._h)
                self._h = None
        except Exception:
            pass


class stratified_sampler:
    """samplers::stratified_sampler (sampler.h:52-93): dims = image size, sample_count = (n, n).
    The GPU path replaces its mt19937 draws by the Philox stream; only sample_count is read."""

    def __init__(self, dims, samples, jitter=True, start=(0, 0)):
        self.dims = tuple(int(v) for v in dims)
        self.sample_count = (int(samples[0]), int(samples[1])) if hasattr(samples, "__len__") else (int(samples), int(samples))
        self.jitter_samples = bool(jitter)
        self.start_pos = tuple(start)
        if self.sample_count[0] != self.sample_count[1]:
            raise ValueError("the reference always passes uvec2(sc.samples): sample_count.x == sample_count.y")


def pinned_zeros(shape, dtype=np.float32):
    """A zero-filled numpy array in page-locked host memory (plu_host_alloc), freed with the array.  Raises PluError
    when there is no CUDA device."""
    import weakref
    lib = _ffi.lib()
    n = int(np.prod(shape)) * np.dtype(dtype).itemsize
    p = C.c_void_p()
    rc = lib.plu_host_alloc(n, C.byref(p))
    if rc != _ffi.PLU_OK:
        raise _ffi.PluError(rc, "plu_host_alloc: no page-locked host memory (no CUDA device?)")
    buf = (C.c_char * n).from_address(p.value)
    a = np.frombuffer(buf, dtype=dtype).reshape(shape)
    weakref.finalize(buf, lib.plu_host_free, C.c_void_p(p.value))   # `a` keeps `buf` alive through its .base chain
    return a


class texture2d:
    """texture2d as a render target (texture.h:26-47): float RGB, row-major, y down, size = (W, H).
    pinned=True puts the pixels in page-locked memory (plu_host_alloc): the two PCIe copies of renderer.render then run
    at the link rate instead of through the driver's staging buffer; pinned=None does so when a device is present."""

    def __init__(self, size, pinned=False):
        self.size = (int(size[0]), int(size[1]))
        shape = (self.size[1], self.size[0], 3)
        self.pinned = False
        if pinned or pinned is None:
            try:
                self.pixels = pinned_zeros(shape)
                self.pinned = True
            except (_ffi.PluError, RuntimeError):
                if pinned:
                    raise
        if not self.pinned:
            self.pixels = np.zeros(shape, np.float32)   # GLM vec3() zero-inits

    def pixel(self, c):
        return self.pixels[c[1], c[0]]


def scene_from_tables(cam, prim_tag, prim_material, prim_data, materials, lights, width, height, n_strata,
                      textures=(), tri_uv=None, tri_normal=None, tie_rank=None):
    """Build a Scene from numpy tables (the same tables the reference-side flattener fills):
    prim_tag int32 [n], prim_material int32 [n], prim_data float32 [n, 12], materials / textures / lights
    sequences of the _ffi *Desc structs; triangles need tri_uv [n, 6] and tri_normal [n, 9]."""
    import numpy as np
    n = len(prim_tag)
    tag = np.ascontiguousarray(prim_tag, np.int32)
    mat = np.ascontiguousarray(prim_material, np.int32)
    data = np.ascontiguousarray(prim_data, np.float32).reshape(n, 12)
    d = _ffi.SceneDesc()
    d.abi_version = _ffi.ABI_VERSION
    d.cam = cam
    d.n_prims, d.n_materials, d.n_textures, d.n_lights = n, len(materials), len(textures), len(lights)
    fp, ip = C.POINTER(C.c_float), C.POINTER(C.c_int32)
    d.prim_tag, d.prim_material, d.prim_data = tag.ctypes.data_as(ip), mat.ctypes.data_as(ip), data.ctypes.data_as(fp)
    keep = [tag, mat, data]
    if tri_uv is not None:
        uv = np.ascontiguousarray(tri_uv, np.float32).reshape(n, 6)
        nr = np.ascontiguousarray(tri_normal, np.float32).reshape(n, 9)
        d.prim_tri_uv, d.prim_tri_normal = uv.ctypes.data_as(fp), nr.ctypes.data_as(fp)
        keep += [uv, nr]
    if tie_rank is not None:
        tr = np.ascontiguousarray(tie_rank, np.int32)
        d.prim_tie_rank = tr.ctypes.data_as(ip)
        keep.append(tr)
    marr = (_ffi.MaterialDesc * max(len(materials), 1))(*materials)
    tarr = (_ffi.TextureDesc * max(len(textures), 1))(*textures)
    larr = (_ffi.LightDesc * max(len(lights), 1))(*lights)
    d.materials, d.textures, d.lights = marr, tarr, larr
    sc = Scene.from_desc(d, width, height, n_strata)   # deep copy: the numpy arrays may go away afterwards
    del keep
    return sc
