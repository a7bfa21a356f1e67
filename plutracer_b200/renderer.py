"""plu::renderer (reference inc/renderer.h:11-31) on the B200 through libplugpu's C ABI.

    r = renderer(scene, None, (32, 32), stratified_sampler(tx.size, (n, n)), None)
    r.render(tx)

keeps the reference's constructor shape `renderer(surface* S, const camera& C, uvec2 tile_size,
sampler* smp, const vector<shared_ptr<light>>& l)`.  On this side of the boundary the scene handle
already carries camera, surfaces and lights (the flat tables), so `C` and `l` may be None; a
CameraDesc passed as `C` overrides the scene's camera.  `tile_size` is accepted and ignored (the
GPU path has no tiles).  There is no CPU fallback.
"""
import ctypes as C

import numpy as np

from . import _ffi
from .scene import Scene, stratified_sampler, texture2d  # noqa: F401  (re-exported)

DEFAULT_SEED = 20261104


def devices_from_env(default=0):
    """PLU_DEVICES="0,1,2,3" | "all" | unset (-> `default`): the device list of the drop-in, as
    integration/plu_gpu_renderer.h reads it."""
    import os
    v = os.environ.get("PLU_DEVICES", "").strip()
    if not v:
        return default
    if v.lower() == "all":
        return "all"
    return [int(x) for x in v.split(",") if x.strip()]


class GpuContext:
    """One libplugpu context: one CUDA device (`device` an int: one process per GPU), or several GPUs of
    the box driven from this process (`device` a list of ints, or "all"): plu_gpu_create_multi — the scene is
    replicated, samples are split k = g (mod n) and the framebuffers are combined by one NCCL reduce."""

    def __init__(self, device=0):
        self._lib = _ffi.lib()
        self._h = C.c_void_p()
        if isinstance(device, (list, tuple)) or device == "all":
            if device == "all":
                rc = self._lib.plu_gpu_create_multi(None, 0, C.byref(self._h))
            else:
                arr = (C.c_int * len(device))(*[int(d) for d in device])
                rc = self._lib.plu_gpu_create_multi(arr, len(device), C.byref(self._h))
        else:
            rc = self._lib.plu_gpu_create(int(device), C.byref(self._h))
        if rc != _ffi.PLU_OK:
            raise _ffi.PluError(rc, self._lib.plu_gpu_last_error(None).decode(errors="replace"))
        self.n_devices = int(self._lib.plu_gpu_device_count(self._h))
        self.device = device if isinstance(device, int) else (0 if device == "all" else int(device[0]))
        self._scene = None

    def _check(self, rc):
        if rc != _ffi.PLU_OK:
            raise _ffi.PluError(rc, self._lib.plu_gpu_last_error(self._h).decode(errors="replace"))

    def upload(self, scene, build="device"):
        """Copy the scene tables to the GPU and build both traversal structures there (build="device", the
        default) or with the host restatement of the reference's builder (build="host").  Returns BuildStats."""
        bs = _ffi.BuildStats()
        where = {"device": _ffi.BUILD_DEVICE, "host": _ffi.BUILD_HOST}[build]
        self._check(self._lib.plu_gpu_upload_scene_ex(self._h, scene.desc_ptr, where, C.byref(bs)))
        self._scene = scene
        return bs

    def download_table(self, which):
        """Raw bytes (numpy uint8) of a device table of the uploaded scene: _ffi.TABLE_*."""
        n = C.c_int64(0)
        self._check(self._lib.plu_gpu_download_table(self._h, int(which), None, 0, C.byref(n)))
        out = np.zeros(n.value, dtype=np.uint8)
        if n.value:
            self._check(self._lib.plu_gpu_download_table(self._h, int(which), out.ctypes.data, n.value, C.byref(n)))
        return out

    @staticmethod
    def params(width, height, n_strata, seed=DEFAULT_SEED, flags=0, max_vertices=8, sample_begin=0, sample_stride=1,
               sample_count=0, strata_per_wave=0):
        p = _ffi.RenderParams()
        p.width, p.height, p.n_strata, p.max_vertices = int(width), int(height), int(n_strata), int(max_vertices)
        p.seed, p.flags = int(seed), int(flags)
        p.sample_begin, p.sample_stride, p.sample_count = int(sample_begin), int(sample_stride), int(sample_count)
        p.strata_per_wave = int(strata_per_wave)
        return p

    def render(self, params, out_rgb):
        """renderer::render into a host float32 (H, W, 3) array, accumulating (+=)."""
        assert out_rgb.dtype == np.float32 and out_rgb.flags.c_contiguous and out_rgb.size == params.width * params.height * 3
        st = _ffi.RenderStats()
        self._check(self._lib.plu_gpu_render(self._h, C.byref(params), out_rgb.ctypes.data, C.byref(st)))
        return st

    def render_device(self, params, device_ptr, stream=None):
        """Same, into a device float buffer (e.g. a torch tensor's data_ptr()) on `stream`."""
        st = _ffi.RenderStats()
        self._check(self._lib.plu_gpu_render_device(self._h, C.byref(params), C.c_void_p(int(device_ptr)),
                                                    C.c_void_p(int(stream) if stream else None), C.byref(st)))
        return st

    def trace_primary(self, width, height, flags=0):
        n = width * height
        ids, t, nrm = np.zeros(n, np.int32), np.zeros(n, np.float32), np.zeros((n, 3), np.float32)
        self._check(self._lib.plu_gpu_trace_primary(self._h, width, height, flags, ids.ctypes.data, t.ctypes.data, nrm.ctypes.data))
        return ids, t, nrm

    def trace_rays(self, origins, dirs, flags=0):
        o = np.ascontiguousarray(origins, np.float32).reshape(-1, 3)
        d = np.ascontiguousarray(dirs, np.float32).reshape(-1, 3)
        n = o.shape[0]
        ids, t, nrm = np.zeros(n, np.int32), np.zeros(n, np.float32), np.zeros((n, 3), np.float32)
        self._check(self._lib.plu_gpu_trace_rays(self._h, n, o.ctypes.data, d.ctypes.data, flags, ids.ctypes.data, t.ctypes.data, nrm.ctypes.data))
        return ids, t, nrm

    def sample_radiance(self, params, pixel_begin, pixel_count):
        count = params.sample_count if params.sample_count > 0 else params.n_strata ** 2
        out = np.zeros((pixel_count, count, 3), np.float32)
        self._check(self._lib.plu_gpu_sample_radiance(self._h, C.byref(params), pixel_begin, pixel_count, out.ctypes.data))
        return out

    def postprocess(self, rgb):
        """postprocesser::postprocess (main.cpp:88-111) on a host float32 (H, W, 3) image, in place."""
        assert rgb.dtype == np.float32 and rgb.flags.c_contiguous and rgb.ndim == 3 and rgb.shape[2] == 3
        self._check(self._lib.plu_gpu_postprocess(self._h, rgb.shape[1], rgb.shape[0], rgb.ctypes.data))
        return rgb

    def postprocess_resident(self, out=None):
        """Tone-map the image the last render() left on the device, in place there; `out` (host float32) receives it."""
        self._check(self._lib.plu_gpu_postprocess_resident(self._h, out.ctypes.data if out is not None else None))
        return out

    def encode_bmp_resident(self, size, tonemap=False):
        """.bmp bytes of the resident image (size = (W, H) of the last render)."""
        n = self._lib.plu_bmp_size(int(size[0]), int(size[1]))
        out = np.zeros(n, np.uint8)
        self._check(self._lib.plu_gpu_encode_bmp_resident(self._h, _ffi.POST_TONEMAP if tonemap else 0, out.ctypes.data, n))
        return out

    def encode_bmp(self, rgb, tonemap=False, device_ptr=None, size=None, stream=None):
        """The bytes texture2d::write_bmp (texture.cpp:178-194) writes, from a host float32 (H, W, 3) image or,
        with device_ptr + size=(W, H), from a device framebuffer; tonemap=True fuses the post-process in front."""
        flags = _ffi.POST_TONEMAP if tonemap else 0
        if device_ptr is None:
            assert rgb.dtype == np.float32 and rgb.flags.c_contiguous and rgb.ndim == 3 and rgb.shape[2] == 3
            W, H = rgb.shape[1], rgb.shape[0]
        else:
            W, H = int(size[0]), int(size[1])
        n = self._lib.plu_bmp_size(W, H)
        out = np.zeros(n, np.uint8)
        if device_ptr is None:
            self._check(self._lib.plu_gpu_encode_bmp(self._h, W, H, rgb.ctypes.data, flags, out.ctypes.data, n))
        else:
            self._check(self._lib.plu_gpu_encode_bmp_device(self._h, W, H, C.c_void_p(int(device_ptr)), flags,
                                                            C.c_void_p(int(stream) if stream else None), out.ctypes.data, n))
        return out

    def close(self):
        if self._h:
            self._lib.plu_gpu_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class renderer:
    """Drop-in for plu::renderer: same constructor shape, same render(target) call."""

    def __init__(self, S, C_=None, tile_size=(32, 32), smp=None, l=None, device=0, seed=DEFAULT_SEED, flags=0):
        if not isinstance(S, Scene):
            raise TypeError("S must be a plutracer_b200.Scene (the flattened surface graph)")
        if smp is None:
            raise ValueError("renderer needs a sampler (sample_count gives the strata per pixel)")
        self.scene, self.cam, self.tile_size, self.main_sampler, self.lights = S, C_, tuple(tile_size), smp, l
        self.seed, self.flags = seed, flags
        self.ctx = GpuContext(device)
        self.last_stats = None

    def render(self, target):
        W, H = target.size
        sc = self.scene
        if self.cam is not None:   # a camera passed explicitly replaces the scene's
            d = _ffi.SceneDesc.from_buffer_copy(sc.desc)
            d.cam = self.cam
            sc = Scene.from_desc(d, W, H, self.main_sampler.sample_count[0])
        self.ctx.upload(sc)
        p = GpuContext.params(W, H, self.main_sampler.sample_count[0], self.seed, self.flags)
        self.last_stats = self.ctx.render(p, target.pixels)
        return self.last_stats


class postprocesser:
    """Drop-in for plu::postprocesser (main.cpp:76-112): postprocess(rt) tone-maps the render target in place."""

    def __init__(self, ctx=None, device=0):
        self.ctx = ctx if ctx is not None else GpuContext(device)

    def postprocess(self, rt):
        self.ctx.postprocess(rt.pixels)


def write_bmp(rt, filename, ctx=None, device=0):
    """texture2d::write_bmp(filename) (texture.cpp:178-194) with the clamp / quantise / layout on the GPU."""
    ctx = ctx if ctx is not None else GpuContext(device)
    data = ctx.encode_bmp(rt.pixels)
    with open(filename, "wb") as f:
        f.write(data.tobytes())
    return len(data)
