"""ctypes binding of include/plugpu.h (libplugpu.so).  Fails loudly when the library is missing."""
import ctypes as C
import os

from ._build import LIB_PATH  # noqa: F401

PLU_OK = 0
PLU_E_ARG, PLU_E_CUDA, PLU_E_NOSCENE, PLU_E_IO, PLU_E_PARSE, PLU_E_NOMEM = -1, -2, -3, -4, -5, -6
ERRORS = {-1: "PLU_E_ARG", -2: "PLU_E_CUDA", -3: "PLU_E_NOSCENE", -4: "PLU_E_IO", -5: "PLU_E_PARSE", -6: "PLU_E_NOMEM"}

PRIM_SPHERE, PRIM_BOX, PRIM_TRIANGLE = 0, 1, 2
MAT_DIFFUSE, MAT_PERFECT_REFLECTION, MAT_PERFECT_REFRACTION, MAT_GLASS, MAT_EMISSION = range(5)
TEX_CHECKERBOARD, TEX_GRID, TEX_IMAGE = range(3)
LIGHT_POINT, LIGHT_DIFFUSE_AREA = 0, 1

RENDER_COMPAT_OVERFLOW_COLUMN = 1
RENDER_BVH_REFERENCE = 2
RENDER_COUNT_BVH = 4
RENDER_BVH_FAST = 8
RENDER_PROFILE = 16
RENDER_NO_LANE_REFILL = 32
POST_TONEMAP = 1
ABI_VERSION = 2
NOALLOC_BMP_MAX_BODY = 8 << 20


class CameraDesc(C.Structure):
    _fields_ = [("pos", C.c_float * 3), ("look", C.c_float * 3), ("up", C.c_float * 3), ("right", C.c_float * 3),
                ("inv_image_size", C.c_float * 2), ("w", C.c_float), ("lens_radius", C.c_float), ("focal_distance", C.c_float)]


class MaterialDesc(C.Structure):
    _fields_ = [("tag", C.c_int32), ("texture", C.c_int32), ("light", C.c_int32), ("reserved", C.c_int32),
                ("color", C.c_float * 3), ("p0", C.c_float * 3), ("p1", C.c_float * 3)]


class TextureDesc(C.Structure):
    _fields_ = [("tag", C.c_int32), ("img_w", C.c_int32), ("img_h", C.c_int32), ("img_offset", C.c_int32),
                ("c0", C.c_float * 3), ("c1", C.c_float * 3), ("scale", C.c_float), ("line", C.c_float)]


class LightDesc(C.Structure):
    _fields_ = [("tag", C.c_int32), ("prim", C.c_int32), ("a", C.c_float * 3), ("b", C.c_float * 3),
                ("area", C.c_float), ("reserved", C.c_float)]


class SceneDesc(C.Structure):
    _fields_ = [("abi_version", C.c_uint32), ("reserved0", C.c_uint32), ("cam", CameraDesc),
                ("n_prims", C.c_int32), ("n_materials", C.c_int32), ("n_textures", C.c_int32), ("n_lights", C.c_int32),
                ("prim_tag", C.POINTER(C.c_int32)), ("prim_material", C.POINTER(C.c_int32)),
                ("prim_data", C.POINTER(C.c_float)), ("prim_tri_uv", C.POINTER(C.c_float)),
                ("prim_tri_normal", C.POINTER(C.c_float)), ("materials", C.POINTER(MaterialDesc)),
                ("textures", C.POINTER(TextureDesc)), ("lights", C.POINTER(LightDesc)),
                ("image_texels", C.POINTER(C.c_float)), ("n_image_texels", C.c_int64),
                ("prim_tie_rank", C.POINTER(C.c_int32))]


class RenderParams(C.Structure):
    _fields_ = [("width", C.c_int32), ("height", C.c_int32), ("n_strata", C.c_int32), ("max_vertices", C.c_int32),
                ("seed", C.c_uint64), ("flags", C.c_uint32), ("sample_begin", C.c_int32), ("sample_stride", C.c_int32),
                ("sample_count", C.c_int32), ("strata_per_wave", C.c_int32), ("reserved", C.c_int32)]


class RenderStats(C.Structure):
    _fields_ = [("samples", C.c_uint64), ("rays_extend", C.c_uint64), ("rays_shadow", C.c_uint64),
                ("bvh_nodes_visited", C.c_uint64), ("prims_tested", C.c_uint64), ("ms_device", C.c_double),
                ("ms_extend", C.c_double), ("ms_shade", C.c_double), ("ms_connect", C.c_double),
                ("kernel_launches", C.c_uint64), ("ms_generate", C.c_double), ("ms_accumulate", C.c_double),
                ("shade_vertices", C.c_uint64), ("shadow_entries", C.c_uint64), ("waves", C.c_uint64),
                ("ms_reduce", C.c_double), ("n_devices", C.c_int32), ("reserved", C.c_int32)]


class BuildStats(C.Structure):
    _fields_ = [("ms_total", C.c_double), ("ms_ref_order", C.c_double), ("ms_ref_nodes", C.c_double),
                ("ms_fast_bvh", C.c_double), ("levels_ref", C.c_int32), ("levels_fast", C.c_int32),
                ("host_tie_segments", C.c_int32), ("n_fast_nodes", C.c_int32), ("built_on_device", C.c_int32),
                ("reserved", C.c_int32)]


BUILD_DEVICE, BUILD_HOST = 0, 1
TABLE_REF_NODES, TABLE_PRIM_META, TABLE_FAST_NODES, TABLE_FAST_INDEX, TABLE_FAST_LEAF = range(5)

# every symbol include/plugpu.h declares: name -> (restype, argtypes)
_P = C.c_void_p
SYMBOLS = {
    "plu_gpu_create": (C.c_int, [C.c_int, C.POINTER(_P)]),
    "plu_gpu_create_multi": (C.c_int, [C.POINTER(C.c_int), C.c_int, C.POINTER(_P)]),
    "plu_gpu_device_count": (C.c_int, [_P]),
    "plu_gpu_destroy": (None, [_P]),
    "plu_gpu_last_error": (C.c_char_p, [_P]),
    "plu_gpu_upload_scene": (C.c_int, [_P, C.POINTER(SceneDesc)]),
    "plu_gpu_upload_scene_ex": (C.c_int, [_P, C.POINTER(SceneDesc), C.c_uint32, C.POINTER(BuildStats)]),
    "plu_gpu_download_table": (C.c_int, [_P, C.c_int, _P, C.c_int64, C.POINTER(C.c_int64)]),
    "plu_gpu_render": (C.c_int, [_P, C.POINTER(RenderParams), _P, C.POINTER(RenderStats)]),
    "plu_gpu_render_device": (C.c_int, [_P, C.POINTER(RenderParams), _P, _P, C.POINTER(RenderStats)]),
    "plu_gpu_trace_primary": (C.c_int, [_P, C.c_int32, C.c_int32, C.c_uint32, _P, _P, _P]),
    "plu_gpu_trace_rays": (C.c_int, [_P, C.c_int64, _P, _P, C.c_uint32, _P, _P, _P]),
    "plu_gpu_sample_radiance": (C.c_int, [_P, C.POINTER(RenderParams), C.c_int64, C.c_int64, _P]),
    "plu_gpu_postprocess": (C.c_int, [_P, C.c_int32, C.c_int32, _P]),
    "plu_bmp_size": (C.c_int64, [C.c_int32, C.c_int32]),
    "plu_gpu_encode_bmp": (C.c_int, [_P, C.c_int32, C.c_int32, _P, C.c_uint32, _P, C.c_int64]),
    "plu_gpu_encode_bmp_device": (C.c_int, [_P, C.c_int32, C.c_int32, _P, C.c_uint32, _P, _P, C.c_int64]),
    "plu_gpu_encode_bmp_device_noalloc": (C.c_int, [_P, C.c_int32, C.c_int32, _P, C.c_uint32, _P, _P, C.c_int64]),
    "plu_gpu_postprocess_resident": (C.c_int, [_P, _P]),
    "plu_gpu_encode_bmp_resident": (C.c_int, [_P, C.c_uint32, _P, C.c_int64]),
    "plu_host_alloc": (C.c_int, [C.c_size_t, C.POINTER(_P)]),
    "plu_host_free": (C.c_int, [_P]),
    "plu_host_register": (C.c_int, [_P, C.c_size_t]),
    "plu_host_unregister": (C.c_int, [_P]),
    "plu_scene_load_urn": (C.c_int, [C.c_char_p, C.POINTER(C.c_char_p), C.c_int, C.POINTER(_P), C.c_char_p, C.c_int]),
    "plu_scene_load_flat": (C.c_int, [C.c_char_p, C.POINTER(_P), C.c_char_p, C.c_int]),
    "plu_scene_save_flat": (C.c_int, [C.POINTER(SceneDesc), C.c_int32, C.c_int32, C.c_int32, C.c_char_p]),
    "plu_scene_from_desc": (C.c_int, [C.POINTER(SceneDesc), C.c_int32, C.c_int32, C.c_int32, C.POINTER(_P)]),
    "plu_scene_get_desc": (C.POINTER(SceneDesc), [_P]),
    "plu_scene_get_defaults": (None, [_P, C.POINTER(C.c_int32), C.POINTER(C.c_int32), C.POINTER(C.c_int32)]),
    "plu_scene_override": (C.c_int, [_P, C.c_int32, C.c_int32, C.c_int32]),
    "plu_scene_free": (None, [_P]),
    "plu_scene_reference_leaf_order": (C.c_int, [C.POINTER(SceneDesc), _P]),
    "plu_abi_version": (C.c_uint32, []),
}

_lib = None


def lib():
    """The loaded libplugpu.so.  No fallback: a missing library is an error."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                f"{LIB_PATH} is missing: build it with `python -m plutracer_b200._build` "
                "(or __graft_entry__.build()).  plutracer_b200 has no CPU or PyTorch fallback.")
        L = C.CDLL(LIB_PATH)
        for name, (res, args) in SYMBOLS.items():
            fn = getattr(L, name)
            fn.restype = res
            fn.argtypes = args
        _lib = L
    return _lib


class PluError(RuntimeError):
    def __init__(self, code, message):
        super().__init__(f"{ERRORS.get(code, code)}: {message}")
        self.code = code
