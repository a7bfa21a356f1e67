/*
 * plugpu.h — C ABI of libplugpu.so, the B200-native render path for andrew-pa/plutracer.
 *
 * This is the drop-in boundary: everything the reference does between
 *     plu::renderer r(scene, cam, tile, sampler, lights);  r.render(target);
 * (reference src/main.cpp:171,177  ->  src/renderer.cpp:98-150 render, :59-96 ray_color,
 *  :5-57 estimate_direct_light / uniform_sample_one_light) happens behind these entry points.
 * Host code (the reference's own C++ via integration/plu_gpu_renderer.h, or this repo's .urn
 * loader) flattens the loaded scene into the POD tables below and calls in; there are no C++
 * or torch types in any signature.  No exceptions cross the boundary: every call returns
 * PLU_OK (0) or a negative PLU_E_* code, and plu_gpu_last_error() gives the message
 * (the reference throws runtime_error / bare `throw;`, surface.h:17,21, scene.h:90-123).
 *
 * There is NO CPU fallback: calls that need the GPU return PLU_E_CUDA when no device is usable.
 */
#ifndef PLUGPU_H
#define PLUGPU_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define PLUGPU_ABI_VERSION 2

/* ---- error codes ------------------------------------------------------------------------- */
enum {
	PLU_OK          =  0,
	PLU_E_ARG       = -1,  /* bad argument / malformed table                                   */
	PLU_E_CUDA      = -2,  /* CUDA runtime error (message names the call)                      */
	PLU_E_NOSCENE   = -3,  /* render / trace before plu_gpu_upload_scene                       */
	PLU_E_IO        = -4,  /* file could not be read / written                                 */
	PLU_E_PARSE     = -5,  /* .urn / OBJ / BMP syntax or semantic error                        */
	PLU_E_NOMEM     = -6
};

/* ---- tags (virtual dispatch in the reference -> tag-indexed device tables) ---------------- */
enum { /* surface kinds: reference inc/surfaces/{sphere,box,triangle}.h                        */
	PLU_PRIM_SPHERE   = 0, /* data: cx cy cz r                                                 */
	PLU_PRIM_BOX      = 1, /* data: min.xyz 0 | max.xyz 0          (box_bounds, box.h:7)       */
	PLU_PRIM_TRIANGLE = 2  /* data: v0.xyz 0 | v1.xyz 0 | v2.xyz 0 (triangle.h:7-9)            */
};
enum { /* material kinds: reference inc/material.h:213-254, inc/lights/area_light.h:45-56       */
	PLU_MAT_DIFFUSE            = 0, /* color                                                   */
	PLU_MAT_PERFECT_REFLECTION = 1, /* color, p0 = eta (vec3), p1 = k (vec3)                    */
	PLU_MAT_PERFECT_REFRACTION = 2, /* color, p0[0] = eta_t, p0[1] = eta_i (ctor order, :236)   */
	PLU_MAT_GLASS              = 3, /* color, p0[0] = ior                                      */
	PLU_MAT_EMISSION           = 4  /* light = index of its diffuse_area_light                  */
};
enum { /* texture kinds: reference inc/textures/textures.h, inc/texture.h:53-60                  */
	PLU_TEX_CHECKERBOARD = 0,  /* c0 = colors[0], c1 = colors[1], scale                          */
	PLU_TEX_GRID         = 1,  /* c0 = fg_color, c1 = bg_color (ctor order :25), scale, line     */
	PLU_TEX_IMAGE        = 2   /* img_w, img_h, img_offset into image_texels (rgb float)         */
};
enum { /* light kinds: reference inc/light.h:15-35, inc/lights/area_light.h:12-43                */
	PLU_LIGHT_POINT        = 0, /* a = pos, b = intensity                                       */
	PLU_LIGHT_DIFFUSE_AREA = 1  /* b = Lemit, prim = index of its surface, area = surf->area()  */
};

/* ---- flat scene tables -------------------------------------------------------------------- */
typedef struct plu_camera_desc {   /* public fields of plu::camera, reference inc/camera.h:9-14   */
	float pos[3];
	float look[3];
	float up[3];
	float right[3];
	float inv_image_size[2];       /* 1/(W,H) the camera was BUILT for (scene resolution)        */
	float w;
	float lens_radius;
	float focal_distance;
} plu_camera_desc;

typedef struct plu_material_desc {
	int32_t tag;          /* PLU_MAT_*                                                          */
	int32_t texture;      /* index into textures, or -1: constant `color` (props.h:24-30)       */
	int32_t light;        /* PLU_MAT_EMISSION: index into lights; else -1                       */
	int32_t reserved;
	float   color[3];
	float   p0[3];
	float   p1[3];
} plu_material_desc;

typedef struct plu_texture_desc {
	int32_t tag;          /* PLU_TEX_*                                                          */
	int32_t img_w, img_h; /* image only                                                         */
	int32_t img_offset;   /* first texel (in texels, not floats) inside image_texels            */
	float   c0[3];
	float   c1[3];
	float   scale;
	float   line;
} plu_texture_desc;

typedef struct plu_light_desc {
	int32_t tag;          /* PLU_LIGHT_*                                                        */
	int32_t prim;         /* area light: index of the emitting surface in the prim table        */
	float   a[3];         /* point: pos                                                         */
	float   b[3];         /* point: intensity; area: Lemit                                      */
	float   area;         /* area light: surf->area() as the reference computes it (quirks in)  */
	float   reserved;
} plu_light_desc;

/*
 * The whole scene.  All pointers are HOST pointers borrowed for the duration of the call.
 * Surface ids everywhere in this ABI = index into the prim table = position in
 * plu::scenes::scene::surfs in LOAD order, captured before bvh_tree sorts it in place
 * (reference src/surfaces/bvh_tree.cpp:23).
 */
typedef struct plu_scene_desc {
	uint32_t abi_version;               /* PLUGPU_ABI_VERSION                                   */
	uint32_t reserved0;
	plu_camera_desc cam;
	int32_t  n_prims;
	int32_t  n_materials;
	int32_t  n_textures;
	int32_t  n_lights;
	const int32_t* prim_tag;            /* [n_prims] PLU_PRIM_*                                 */
	const int32_t* prim_material;       /* [n_prims] index into materials                       */
	const float*   prim_data;           /* [n_prims*12] three float4 per prim, see PLU_PRIM_*   */
	const float*   prim_tri_uv;         /* [n_prims*6]  triangle texture_coords[0..2], or NULL  */
	const float*   prim_tri_normal;     /* [n_prims*9]  triangle normals[0..2], or NULL         */
	const plu_material_desc* materials;
	const plu_texture_desc*  textures;
	const plu_light_desc*    lights;    /* order = plu::scenes::scene::lights                   */
	const float*   image_texels;        /* rgb float, all image textures back to back, or NULL  */
	int64_t  n_image_texels;
	const int32_t* prim_tie_rank;       /* [n_prims] or NULL.  Rank of each surface among the
	                                       leaves of the reference bvh_tree in DFS (left first)
	                                       order; equal-t hits resolve to the HIGHER rank
	                                       (bvh_tree.cpp:73).  NULL: the library restates the
	                                       reference build (bvh_tree.cpp:7-36) to derive it.    */
} plu_scene_desc;

/* ---- render call -------------------------------------------------------------------------- */
enum { /* plu_render_params.flags                                                                */
	PLU_RENDER_COMPAT_OVERFLOW_COLUMN = 1u << 0, /* reproduce sampler.h:74-89: the x==W strip is
	                                                rendered and lands in pixel (0,y+1)          */
	PLU_RENDER_BVH_REFERENCE          = 1u << 1, /* traverse the reference's own tree literally
	                                                (origin-inflated bounds, both children, no
	                                                early-out) — verification mode               */
	PLU_RENDER_COUNT_BVH              = 1u << 2, /* fill bvh_nodes_visited / prims_tested         */
	PLU_RENDER_BVH_FAST               = 1u << 3, /* use the tight SAH BVH even for a small scene
	                                                (default: scenes with > 64 surfaces)         */
	PLU_RENDER_PROFILE                = 1u << 4, /* CUDA events between all launches: fills
	                                                ms_generate/extend/shade/connect/accumulate   */
	PLU_RENDER_NO_LANE_REFILL         = 1u << 5  /* fast-BVH scenes: one ray per lane per iteration
	                                                instead of persistent warps that refill idle
	                                                lanes (A/B and parity of the two schedules)   */
};

typedef struct plu_render_params {
	int32_t  width, height;        /* target->size (texture.h:32)                               */
	int32_t  n_strata;             /* sampler->sample_count.x (== .y); spp = n_strata^2         */
	int32_t  max_vertices;         /* 8 = the reference (renderer.cpp:83 `bounces > 6`)         */
	uint64_t seed;                 /* Philox key                                                */
	uint32_t flags;
	int32_t  sample_begin;         /* this call renders sample indices                          */
	int32_t  sample_stride;        /*   s = sample_begin + k*sample_stride, k < sample_count    */
	int32_t  sample_count;         /*   (0 -> all n_strata^2: begin 0, stride 1).  Every s must
	                                  be < n_strata^2 (else PLU_E_ARG); a subset is still
	                                  weighted by 1/n_strata^2 per sample (renderer.cpp:100),
	                                  so the parts of a split add up to the whole image         */
	int32_t  strata_per_wave;      /* paths in flight = W*H*this; 0 -> library picks             */
	int32_t  reserved;
} plu_render_params;

typedef struct plu_render_stats {
	uint64_t samples;              /* ray_color calls (paths)                                   */
	uint64_t rays_extend;          /* closest-hit queries from the path loop                    */
	uint64_t rays_shadow;          /* light-sample + BSDF-MIS connection rays                   */
	uint64_t bvh_nodes_visited;    /* PLU_RENDER_COUNT_BVH only                                 */
	uint64_t prims_tested;         /* PLU_RENDER_COUNT_BVH only                                 */
	double   ms_device;            /* CUDA-event time of the whole render on its stream         */
	double   ms_extend;            /* per-stage CUDA-event sums (PLU_RENDER_PROFILE only)       */
	double   ms_shade;
	double   ms_connect;
	uint64_t kernel_launches;      /* kernels this call launched                                */
	double   ms_generate;
	double   ms_accumulate;
	uint64_t shade_vertices;       /* path vertices shaded (= extension rays that hit)          */
	uint64_t shadow_entries;       /* shaded vertices that needed a connection                  */
	uint64_t waves;                /* waves (batches of paths in flight) this call ran          */
	/* ABI 2: multi-device contexts (plu_gpu_create_multi).  Counters above are sums over the
	 * devices, ms_device is the slowest device's render time.                                  */
	double   ms_reduce;            /* the NCCL reduce of the per-device framebuffers (0: one GPU) */
	int32_t  n_devices;            /* devices that took part in this call                       */
	int32_t  reserved;
} plu_render_stats;

typedef struct plu_gpu_ctx plu_gpu_ctx;

/* Create a context on CUDA device `device` (one process / context per GPU).                     */
int  plu_gpu_create(int device, plu_gpu_ctx** out);
/* Create ONE context that drives `n_devices` GPUs of this box from a single process (SURVEY.md
 * 8b/8e): the scene is deep-copied to every device; plu_gpu_render / plu_gpu_render_device give
 * device g the samples k = g (mod n_devices) of the call's sample set (per-pixel stratification
 * is kept, sampler.h:58-69; nothing is exchanged while rendering), one host thread per device
 * enqueues its waves, and the per-device float framebuffers are combined by ONE
 * ncclReduce(sum, float32, W*H*3, root = devices[0]) over NVLink on communicators the context
 * owns (ncclCommInitAll; libnccl.so.2 is loaded at run time, only by this call).  All other
 * entry points (parity hooks, post-process, table download) run on devices[0].  devices == NULL:
 * the first n_devices visible devices; n_devices == 0 with devices == NULL: every visible one.
 * n_devices == 1 is exactly plu_gpu_create(devices[0]).  The image is independent of the device
 * count up to the float summation order of the reduce (Philox counters carry the global sample
 * index).                                                                                        */
int  plu_gpu_create_multi(const int* devices, int n_devices, plu_gpu_ctx** out);
/* Number of devices the context drives (1 for plu_gpu_create).                                  */
int  plu_gpu_device_count(const plu_gpu_ctx* ctx);
void plu_gpu_destroy(plu_gpu_ctx* ctx);
/* Message of the last failing call on ctx (ctx may be NULL: last plu_gpu_create failure).       */
const char* plu_gpu_last_error(const plu_gpu_ctx* ctx);

/* Deep-copy the tables to the device, build the traversal structure.  Replaces scene->hit,
 * the material/light/texture object graphs and bvh_tree construction (main.cpp:168-171).        */
int plu_gpu_upload_scene(plu_gpu_ctx* ctx, const plu_scene_desc* scene);
/* The context remembers the tables it was given: uploading a scene whose surfaces, materials,
 * textures, lights and tie ranks are byte for byte those already on the device (only the camera
 * may differ) returns at once and keeps the traversal structures — a caller that flattens and
 * uploads before every render (the C++ drop-in: the object graph may have changed) pays for a
 * rebuild only when something did.  plu_build_stats.reserved == 1 then, ms_total == 0.           */

/* Where the traversal structures are built.  The reference builds its tree on the host with a
 * recursive median split (bvh_tree.cpp:7-36: round-robin axis, std::sort by box centre, halves);
 * PLU_BUILD_DEVICE builds the same tree — same leaf order, hence same tie ranks — and the fast
 * BVH with CUDA kernels, level by level; PLU_BUILD_HOST keeps the host
 * restatement (host/bvh_build.cpp).  Both produce byte-identical device tables.                */
enum { PLU_BUILD_DEVICE = 0u, PLU_BUILD_HOST = 1u };

typedef struct plu_build_stats {
	double  ms_total;            /* wall clock of the structure build inside the upload          */
	double  ms_ref_order;        /* reference leaf order (tie ranks)                             */
	double  ms_ref_nodes;        /* reference tree, pre-order with skip links                    */
	double  ms_fast_bvh;         /* binned-SAH BVH + leaf records                                */
	int32_t levels_ref;          /* sort levels of the median split                              */
	int32_t levels_fast;         /* levels of the SAH build                                      */
	int32_t host_tie_segments;   /* segments (> 16 surfaces, equal box centres) whose order the
	                                host's std::sort had to decide (bvh_tree.cpp:23 is unstable) */
	int32_t n_fast_nodes;
	int32_t built_on_device;     /* 1 / 0                                                        */
	int32_t reserved;
} plu_build_stats;

/* plu_gpu_upload_scene with the build location chosen (PLU_BUILD_*); stats may be NULL.
 * plu_gpu_upload_scene(ctx, s) == plu_gpu_upload_scene_ex(ctx, s, PLU_BUILD_DEVICE, NULL).     */
int plu_gpu_upload_scene_ex(plu_gpu_ctx* ctx, const plu_scene_desc* scene, uint32_t build,
                            plu_build_stats* stats);

/* Copy a device table of the uploaded scene back (diagnostics and the build-parity tests).
 * out may be NULL to query *n_bytes.                                                            */
enum { PLU_TABLE_REF_NODES = 0, PLU_TABLE_PRIM_META = 1, PLU_TABLE_FAST_NODES = 2,
       PLU_TABLE_FAST_INDEX = 3, PLU_TABLE_FAST_LEAF = 4 };
int plu_gpu_download_table(plu_gpu_ctx* ctx, int which, void* out, int64_t cap_bytes, int64_t* n_bytes);

/* renderer::render (renderer.cpp:98-150): out_rgb is target->_pixels, HOST, W*H*3 float,
 * row-major y down (texture.h:40-47), ACCUMULATED INTO (+=) like the reference.  Blocks.      */
int plu_gpu_render(plu_gpu_ctx* ctx, const plu_render_params* p, float* out_rgb,
                   plu_render_stats* stats);

/* Same, but d_rgb is a DEVICE pointer on ctx's (first) device and the work is enqueued on
 * `stream` (a cudaStream_t; NULL = stream 0, the legacy default stream, as everywhere in CUDA —
 * so work the caller enqueued on stream 0, e.g. the framebuffer initialisation, is ordered
 * before the render) — for callers that keep the framebuffer on the device (one process per
 * GPU + their own reduce, or a device-side post-process).  Synchronises `stream` before
 * returning.                                                                                   */
int plu_gpu_render_device(plu_gpu_ctx* ctx, const plu_render_params* p, float* d_rgb,
                          void* stream, plu_render_stats* stats);

/* Parity hooks.  Pixel-centre pinhole rays through camera::generate_ray (camera.h:25-37) then
 * scene->hit: per pixel the hit surface id (-1 = miss), t, and hit_record::norm.               */
int plu_gpu_trace_primary(plu_gpu_ctx* ctx, int32_t width, int32_t height, uint32_t flags,
                          int32_t* surf_id, float* t, float* normal3);
/* Arbitrary rays (o,d each n*3 float) through scene->hit.  With PLU_RENDER_COUNT_BVH in flags
 * normal3 receives the cost of each ray's walk instead: (interior nodes visited, surfaces
 * tested, 0) — diagnostics of the traversal structures.                                         */
int plu_gpu_trace_rays(plu_gpu_ctx* ctx, int64_t n, const float* origin3, const float* dir3,
                       uint32_t flags, int32_t* surf_id, float* t, float* normal3);
/* Per-sample radiance of ray_color for samples [sample_begin, +sample_count) of the pixels
 * [pixel_begin, +pixel_count) (row-major index): out_rgb[(p*sample_count+s)*3].  Lets tests
 * compare single paths with the oracle driven by the same Philox stream.                       */
int plu_gpu_sample_radiance(plu_gpu_ctx* ctx, const plu_render_params* p, int64_t pixel_begin,
                            int64_t pixel_count, float* out_rgb);

/* ---- post-process + image write: the step right after the path (main.cpp:183, :209) -------- */
enum {
	PLU_POST_TONEMAP = 1u << 0   /* apply postprocesser::postprocess before quantising             */
};
/* postprocesser::postprocess (main.cpp:76-111): white-preserving luma Reinhard (white = 2) and
 * gamma 1/2.2 over a HOST float RGB image, in place.  pow is evaluated in double and rounded to
 * float (host powf differs from it in the last bit for a vanishing fraction of inputs).        */
int plu_gpu_postprocess(plu_gpu_ctx* ctx, int32_t width, int32_t height, float* rgb_inout);
/* Size of the .bmp file texture2d::write_bmp produces for a W x H image (texture.cpp:167-173).  */
int64_t plu_bmp_size(int32_t width, int32_t height);
/* texture2d::write_bmp (texture.cpp:178-194) without the file: clamp to [0,1], byte =
 * (unsigned char)(c * 255), 24-bit BGR bottom-up rows padded to 4 bytes behind the 54-byte
 * header.  rgb: HOST float image; with PLU_POST_TONEMAP the tone map is fused in front (the
 * float image itself is not modified).  out_bmp: HOST, at least plu_bmp_size bytes.             */
int plu_gpu_encode_bmp(plu_gpu_ctx* ctx, int32_t width, int32_t height, const float* rgb,
                       uint32_t flags, unsigned char* out_bmp, int64_t cap);
/* Same from a DEVICE framebuffer (e.g. the one plu_gpu_render_device filled): tone map +
 * quantise + BMP layout in one kernel, only the bytes travel to the host (3 instead of 12 per
 * pixel).  Work is enqueued on `stream` (NULL = stream 0, the legacy default stream) and
 * synchronised.                                                                                 */
int plu_gpu_encode_bmp_device(plu_gpu_ctx* ctx, int32_t width, int32_t height, const float* d_rgb,
                              uint32_t flags, void* stream, unsigned char* out_bmp, int64_t cap);

/* plu_gpu_encode_bmp_device WITHOUT ANY ALLOCATION.  This entry point has no counterpart in the
 * reference: it was added for callers that cannot call
 * cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * Same arguments, same bytes as plu_gpu_encode_bmp_device, but the .bmp body is staged in a
 * fixed device array inside the library instead of a buffer allocated and freed per call: the
 * call makes no cudaMalloc / cudaFree and no allocating library call.
 *   limit     the body (plu_bmp_size(w, h) - 54 bytes) must fit PLU_NOALLOC_BMP_MAX_BODY (8 MiB:
 *             up to 1920 x 1080); larger images are rejected with PLU_E_ARG and a message.
 *   cost      every GPU that loads libplugpu holds those 8 MiB until the process exits.
 *   sharing   ALL calls on a device share the one array, which keeps the last call's bytes (each
 *             call overwrites every byte it then reads back, so nothing needs clearing): calls
 *             that overlap on the GPU (different streams or contexts on one device) RACE — keep
 *             them on one stream or serialise them.  The call synchronises `stream` before
 *             returning, as plu_gpu_encode_bmp_device does.                                     */
#define PLU_NOALLOC_BMP_MAX_BODY (8u << 20)
int plu_gpu_encode_bmp_device_noalloc(plu_gpu_ctx* ctx, int32_t width, int32_t height,
                                      const float* d_rgb, uint32_t flags, void* stream,
                                      unsigned char* out_bmp, int64_t cap);

/* The context keeps the image of its last plu_gpu_render on its (first) device.  These two run
 * the steps after the path on that RESIDENT framebuffer, so the float pixels cross PCIe once
 * (the render's own download) instead of once per step: tone map in place on the device
 * (rgb_out, HOST, W*H*3 float, may be NULL: nothing is copied back), and the .bmp bytes
 * (PLU_POST_TONEMAP fuses the tone map in front without modifying the resident image).
 * PLU_E_NOSCENE when no render has left an image.  The caller vouches that the host copy has not
 * been edited in between (the reference draws its watermark on the host between main.cpp:183 and
 * :209 — use plu_gpu_encode_bmp for that flow).                                                 */
int plu_gpu_postprocess_resident(plu_gpu_ctx* ctx, float* rgb_out);
int plu_gpu_encode_bmp_resident(plu_gpu_ctx* ctx, uint32_t flags, unsigned char* out_bmp, int64_t cap);

/* Page-locked host images.  texture2d::_pixels is `new vec3[w*h]` (texture.h:29,36): pageable
 * memory, which the driver copies through its own staging buffer at a fraction of the PCIe rate —
 * for an image that renders in a few milliseconds (glass0, refrac0, minimal0) the two copies of
 * plu_gpu_render then cost as much as the render.  plu_host_alloc returns ZERO-FILLED memory
 * (vec3() zero-initialises) that is pinned for every device of the box; plu_host_register pins
 * memory the caller already owns (an existing texture2d's pixels) until plu_host_unregister,
 * which must run before that memory is freed.  Any pointer into such memory may be handed to the
 * entry points that take HOST images; nothing else changes (pageable pointers keep working).
 * PLU_E_ARG for NULL / 0 bytes, PLU_E_CUDA when there is no device or the driver refuses.      */
int plu_host_alloc(size_t bytes, void** out);
int plu_host_free(void* p);
int plu_host_register(void* p, size_t bytes);
int plu_host_unregister(void* p);

/* ---- host-side scene plumbing (no GPU needed) --------------------------------------------- */
typedef struct plu_scene plu_scene;   /* owns the arrays a plu_scene_desc points at              */

/* Load a .urn scene the way plu::scenes::scene does (scene.h:229-298).  args = the reference
 * CLI overrides ("/res", "640x480", "/smp", "4").  Replaces urn::value + scenes::scene.       */
int plu_scene_load_urn(const char* path, const char* const* args, int n_args,
                       plu_scene** out, char* err, int err_len);
/* Flat binary form of a plu_scene_desc (".pluscene"), for fixtures and caching.                */
int plu_scene_load_flat(const char* path, plu_scene** out, char* err, int err_len);
int plu_scene_save_flat(const plu_scene_desc* desc, int32_t width, int32_t height,
                        int32_t n_strata, const char* path);
/* Wrap/deep-copy an existing desc (e.g. produced by the integration adapter).                  */
int plu_scene_from_desc(const plu_scene_desc* desc, int32_t width, int32_t height,
                        int32_t n_strata, plu_scene** out);
const plu_scene_desc* plu_scene_get_desc(const plu_scene* s);
void plu_scene_get_defaults(const plu_scene* s, int32_t* width, int32_t* height,
                            int32_t* n_strata);  /* resolution / antialiasing-samples           */
/* The reference CLI overrides applied after loading: "/res WxH" rebuilds the camera's
 * inv_image_size = 1/(W,H) (scene.h:232-248, camera.h:18) and "/smp N" replaces
 * antialiasing-samples (scene.h:251-256).  A value <= 0 leaves that setting alone.              */
int plu_scene_override(plu_scene* s, int32_t width, int32_t height, int32_t n_strata);
void plu_scene_free(plu_scene* s);

/* Restatement of the reference BVH build order (bvh_tree.cpp:7-36): fills rank[n_prims] with
 * each surface's position among the leaves in DFS order.  Exposed for tests.                   */
int plu_scene_reference_leaf_order(const plu_scene_desc* desc, int32_t* rank);

uint32_t plu_abi_version(void);

#ifdef __cplusplus
}
#endif
#endif /* PLUGPU_H */
