// plugpu.cu — C ABI of libplugpu.so (include/plugpu.h): context, scene upload, wave scheduling.
// Replaces plu::renderer::render (reference src/renderer.cpp:98-150) and its callees behind the
// drop-in boundary of SURVEY.md 8b.  There is no CPU fallback: every entry point that needs the GPU
// fails with PLU_E_CUDA when no device is usable.
#include "device/wavefront.cuh"
#include "device/post.cuh"
#include "shade_tu.h"
#include "host/bvh_build.h"
#include "host/bvh_gpu.h"
#include "host/scene_store.h"
#include <nccl.h>    // types only: libnccl.so.2 is loaded at run time by plu_gpu_create_multi (no link-time dependency)
#include <dlfcn.h>
#include <algorithm>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cmath>
#include <cstring>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

using namespace plud;

static_assert(sizeof(pluhost::ref_node_h) == sizeof(ref_node), "ref node layout");
static_assert(sizeof(pluhost::fast_node_h) == sizeof(fast_node), "fast node layout");

static thread_local std::string g_create_error;

#define PLU_PIPES 4

// The input tables of the scene that is on the device, kept on the host so that plu_gpu_upload_scene can tell an unchanged
// scene from a new one by comparing bytes (not hashes): the C++ drop-in flattens and uploads on every render() because the
// reference's object graph may have changed between frames; when it has not — or only the camera has — the upload returns
// at once and the traversal structures stay (10^6 surfaces: 0.38 s of upload + GPU build per call and per device).
struct upload_signature {
	bool valid = false; uint32_t build = 0; bool has_rank = false, has_tri = false;
	int32_t n_prims = 0, n_materials = 0, n_textures = 0, n_lights = 0; int64_t n_texels = 0;
	std::vector<unsigned char> bytes;   // prim_tag | prim_material | prim_data | tri_uv | tri_normal | tie_rank | materials | textures | lights | texels
	template <class F> static void each(const plu_scene_desc* d, F f) {
		const size_t n = (size_t)d->n_prims;
		f(d->prim_tag, n * 4); f(d->prim_material, n * 4); f(d->prim_data, n * 48);
		if (d->prim_tri_uv && d->prim_tri_normal) { f(d->prim_tri_uv, n * 24); f(d->prim_tri_normal, n * 36); }
		if (d->prim_tie_rank) f(d->prim_tie_rank, n * 4);
		f(d->materials, (size_t)d->n_materials * sizeof(plu_material_desc)); f(d->textures, (size_t)d->n_textures * sizeof(plu_texture_desc));
		f(d->lights, (size_t)d->n_lights * sizeof(plu_light_desc));
		if (d->image_texels && d->n_image_texels > 0) f(d->image_texels, (size_t)d->n_image_texels * 12);
	}
	bool head_equals(const plu_scene_desc* d, uint32_t b) const {
		return valid && build == b && has_rank == (d->prim_tie_rank != nullptr) && has_tri == (d->prim_tri_uv && d->prim_tri_normal) &&
		       n_prims == d->n_prims && n_materials == d->n_materials && n_textures == d->n_textures && n_lights == d->n_lights &&
		       n_texels == ((d->image_texels && d->n_image_texels > 0) ? d->n_image_texels : 0);
	}
	bool equals(const plu_scene_desc* d, uint32_t b) const {
		if (!head_equals(d, b)) return false;
		size_t at = 0; bool same = true;
		each(d, [&](const void* p, size_t len) { if (same && (at + len > bytes.size() || std::memcmp(bytes.data() + at, p, len) != 0)) same = false; at += len; });
		return same && at == bytes.size();
	}
	void store(const plu_scene_desc* d, uint32_t b) {
		build = b; has_rank = d->prim_tie_rank != nullptr; has_tri = d->prim_tri_uv && d->prim_tri_normal;
		n_prims = d->n_prims; n_materials = d->n_materials; n_textures = d->n_textures; n_lights = d->n_lights;
		n_texels = (d->image_texels && d->n_image_texels > 0) ? d->n_image_texels : 0;
		size_t total = 0; each(d, [&](const void*, size_t len) { total += len; });
		bytes.resize(total);
		size_t at = 0; each(d, [&](const void* p, size_t len) { if (len) std::memcpy(bytes.data() + at, p, len); at += len; });
		valid = true;
	}
};

struct plu_gpu_ctx {
	int device = 0;
	int sm_count = 0;
	cudaStream_t stream = nullptr;
	std::string err;
	// scene
	bool have_scene = false;
	upload_signature sig;          // what is on the device (plu_gpu_upload_scene skips an identical scene)
	plu_build_stats last_build{};  // stats of the build that produced it
	scene_ref sc{};
	unsigned char* d_blob = nullptr;
	float *d_tri_uv = nullptr, *d_tri_normal = nullptr, *d_texels = nullptr;
	fast_node* d_fast = nullptr; int32_t* d_fast_index = nullptr; float4* d_fast_leaf = nullptr;
	walk_node* d_walk = nullptr;   // scenes with spheres on the fast path: the reference tree with tight boxes (trace_ref_pruned)
	bool blob_in_smem = false;
	bool use_fast = false;
	bool has_specular = false;     // any perfect-reflection / perfect-refraction / glass material
	int32_t fast_threshold = 64;   // scenes with more surfaces than this use the fast BVH by default
	uint32_t pt_refill = PT_REFILL; // k_extend_pt: idle lanes of a warp that trigger a retire + refill round
	// wave buffers
	// wave buffers: two sets, so that two waves can be in flight on two streams (render_device)
	uint32_t cap[PLU_PIPES] = {};      // path slots
	wave_buffers wbs[PLU_PIPES] = {};
	void* d_wave[PLU_PIPES] = {};
	// One CUDA graph per pipe = the whole launch sequence of a wave (generate -> [extend, shade, shade, connect] x depth);
	// valid while its key (scene tables, kernel variants, grids, buffers) is unchanged, see wave_graph_for()
	struct wave_graph { cudaGraphExec_t exec = nullptr; std::vector<unsigned char> key; int launches = 0, body_launches = 0; bool looped = false; };
	wave_graph graphs[PLU_PIPES];
	wave_desc* h_descs = nullptr; size_t h_descs_cap = 0;   // pinned: the descriptors of a render's waves, copied to wb.desc wave by wave
	int n_pipes = 4;                   // waves in flight (PLU_PIPES=1..4 overrides); 2 suffice for cornell-box, launch-bound scenes gain up to 4
	float* d_fb = nullptr; size_t fb_floats = 0;
	int32_t res_w = 0, res_h = 0;      // size of the image plu_gpu_render left in d_fb (0: none) — plu_gpu_*_resident
	// multi-device group (plu_gpu_create_multi): this context is devices[0]; it owns one context per further device
	// and the NCCL communicators of the group (comms[0] is its own)
	std::vector<plu_gpu_ctx*> peers;
	std::vector<ncclComm_t> comms;
	cudaEvent_t ev_r0 = nullptr, ev_r1 = nullptr;   // around the framebuffer reduce
	cudaEvent_t ev0 = nullptr, ev1 = nullptr;
	cudaStream_t pipe[PLU_PIPES] = {};             // wave i of a render runs on pipe[i % n_pipes]: always the context's own streams (the caller's may be the legacy default stream, which cannot be captured into a graph); ordered against the caller's stream by events
	cudaStream_t aux[PLU_PIPES] = {};              // per pipe: the specular shade kernel of a bounce runs beside the diffuse one
	cudaEvent_t ev_fork[PLU_PIPES] = {}, ev_join[PLU_PIPES] = {};
	cudaEvent_t ev_acc[PLU_PIPES] = {}, ev_start = nullptr;   // framebuffer adds stay in wave order
	int grid_generate = 0, grid_extend = 0, grid_shade = 0, grid_connect = 0;
};

#define CU(call)                                                                                         \
	do {                                                                                                 \
		cudaError_t e_ = (call);                                                                         \
		if (e_ != cudaSuccess) {                                                                         \
			ctx->err = std::string(#call) + ": " + cudaGetErrorString(e_);                               \
			return PLU_E_CUDA;                                                                           \
		}                                                                                                \
	} while (0)

static int fail(plu_gpu_ctx* ctx, int code, const std::string& msg) { ctx->err = msg; return code; }

extern "C" int plu_gpu_create(int device, plu_gpu_ctx** out) {
	if (!out) return PLU_E_ARG;
	*out = nullptr;
	int count = 0;
	cudaError_t e = cudaGetDeviceCount(&count);
	if (e != cudaSuccess || count <= 0) {
		g_create_error = std::string("no usable CUDA device (libplugpu has no CPU fallback): ") + cudaGetErrorString(e);
		return PLU_E_CUDA;
	}
	if (device < 0 || device >= count) { g_create_error = "device index out of range"; return PLU_E_ARG; }
	plu_gpu_ctx* ctx = new plu_gpu_ctx;
	ctx->device = device;
	cudaDeviceProp prop;
	if ((e = cudaSetDevice(device)) != cudaSuccess || (e = cudaGetDeviceProperties(&prop, device)) != cudaSuccess ||
	    (e = cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking)) != cudaSuccess ||
	    (e = cudaEventCreate(&ctx->ev0)) != cudaSuccess || (e = cudaEventCreate(&ctx->ev1)) != cudaSuccess ||
	    (e = cudaEventCreateWithFlags(&ctx->ev_start, cudaEventDisableTiming)) != cudaSuccess) {
		g_create_error = std::string("CUDA context set-up failed: ") + cudaGetErrorString(e);
		delete ctx;
		return PLU_E_CUDA;
	}
	for (int k = 0; k < PLU_PIPES; ++k) {
		if ((e = cudaStreamCreateWithFlags(&ctx->pipe[k], cudaStreamNonBlocking)) != cudaSuccess ||
		    (e = cudaStreamCreateWithFlags(&ctx->aux[k], cudaStreamNonBlocking)) != cudaSuccess ||
		    (e = cudaEventCreateWithFlags(&ctx->ev_fork[k], cudaEventDisableTiming)) != cudaSuccess ||
		    (e = cudaEventCreateWithFlags(&ctx->ev_join[k], cudaEventDisableTiming)) != cudaSuccess ||
		    (e = cudaEventCreateWithFlags(&ctx->ev_acc[k], cudaEventDisableTiming)) != cudaSuccess) {
			g_create_error = std::string("CUDA context set-up failed: ") + cudaGetErrorString(e);
			plu_gpu_destroy(ctx);
			return PLU_E_CUDA;
		}
	}
	ctx->sm_count = prop.multiProcessorCount;
	{   // scene tables and build scratch are stream-ordered allocations: keep the pool's memory between uploads
		cudaMemPool_t pool = nullptr; uint64_t keep = ~0ull;
		if (cudaDeviceGetDefaultMemPool(&pool, device) == cudaSuccess) cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &keep);
	}
	if (const char* e = std::getenv("PLU_PIPES")) { int v = std::atoi(e); if (v >= 1 && v <= PLU_PIPES) ctx->n_pipes = v; }   // tuning knob
	if (const char* e = std::getenv("PLU_PT_REFILL")) { int v = std::atoi(e); if (v >= 1 && v <= 32) ctx->pt_refill = (uint32_t)v; }   // tuning knob
	*out = ctx;
	return PLU_OK;
}

// ---- NCCL, loaded at run time (only multi-device contexts need it) ---------------------------------------------
struct nccl_api {
	void* lib = nullptr;
	ncclResult_t (*CommInitAll)(ncclComm_t*, int, const int*) = nullptr;
	ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
	ncclResult_t (*Reduce)(const void*, void*, size_t, ncclDataType_t, ncclRedOp_t, int, ncclComm_t, cudaStream_t) = nullptr;
	ncclResult_t (*GroupStart)() = nullptr;
	ncclResult_t (*GroupEnd)() = nullptr;
	const char* (*GetErrorString)(ncclResult_t) = nullptr;
	std::string err;
};
static nccl_api* load_nccl() {
	static nccl_api api; static std::once_flag once;
	std::call_once(once, [] {
		const char* names[] = {std::getenv("PLU_NCCL_LIB"), "libnccl.so.2", "libnccl.so"};
		for (const char* n : names) { if (n && *n && (api.lib = dlopen(n, RTLD_NOW | RTLD_GLOBAL))) break; }
		if (!api.lib) { api.err = std::string("cannot load libnccl.so.2: ") + dlerror(); return; }
#define PLU_SYM(field, name) api.field = (decltype(api.field))dlsym(api.lib, name); if (!api.field) api.err = std::string("libnccl lacks ") + name;
		PLU_SYM(CommInitAll, "ncclCommInitAll") PLU_SYM(CommDestroy, "ncclCommDestroy") PLU_SYM(Reduce, "ncclReduce")
		PLU_SYM(GroupStart, "ncclGroupStart") PLU_SYM(GroupEnd, "ncclGroupEnd") PLU_SYM(GetErrorString, "ncclGetErrorString")
#undef PLU_SYM
	});
	return &api;
}

extern "C" int plu_gpu_create_multi(const int* devices, int n_devices, plu_gpu_ctx** out) {
	if (!out) return PLU_E_ARG;
	*out = nullptr;
	int count = 0;
	cudaError_t e = cudaGetDeviceCount(&count);
	if (e != cudaSuccess || count <= 0) {
		g_create_error = std::string("no usable CUDA device (libplugpu has no CPU fallback): ") + cudaGetErrorString(e);
		return PLU_E_CUDA;
	}
	if (n_devices < 0 || (n_devices == 0 && devices)) { g_create_error = "n_devices must be positive"; return PLU_E_ARG; }
	std::vector<int> devs;
	if (!devices) { if (n_devices == 0) n_devices = count; for (int i = 0; i < n_devices; ++i) devs.push_back(i); }
	else devs.assign(devices, devices + n_devices);
	for (size_t i = 0; i < devs.size(); ++i) {
		if (devs[i] < 0 || devs[i] >= count) { g_create_error = "device index out of range"; return PLU_E_ARG; }
		for (size_t j = 0; j < i; ++j) if (devs[j] == devs[i]) { g_create_error = "device listed twice"; return PLU_E_ARG; }
	}
	plu_gpu_ctx* ctx = nullptr;
	int rc = plu_gpu_create(devs[0], &ctx);
	if (rc != PLU_OK || devs.size() == 1) { *out = ctx; return rc; }
	for (size_t i = 1; i < devs.size(); ++i) {
		plu_gpu_ctx* peer = nullptr;
		rc = plu_gpu_create(devs[i], &peer);
		if (rc != PLU_OK) { plu_gpu_destroy(ctx); return rc; }
		ctx->peers.push_back(peer);
	}
	nccl_api* nc = load_nccl();
	if (!nc->err.empty()) { g_create_error = nc->err; plu_gpu_destroy(ctx); return PLU_E_CUDA; }
	ctx->comms.assign(devs.size(), nullptr);
	ncclResult_t nr = nc->CommInitAll(ctx->comms.data(), (int)devs.size(), devs.data());
	if (nr != ncclSuccess) {
		g_create_error = std::string("ncclCommInitAll: ") + nc->GetErrorString(nr);
		ctx->comms.clear(); plu_gpu_destroy(ctx); return PLU_E_CUDA;
	}
	cudaSetDevice(devs[0]);
	if ((e = cudaEventCreate(&ctx->ev_r0)) != cudaSuccess || (e = cudaEventCreate(&ctx->ev_r1)) != cudaSuccess) {
		g_create_error = std::string("CUDA context set-up failed: ") + cudaGetErrorString(e);
		plu_gpu_destroy(ctx); return PLU_E_CUDA;
	}
	*out = ctx;
	return PLU_OK;
}

extern "C" int plu_gpu_device_count(const plu_gpu_ctx* ctx) { return ctx ? 1 + (int)ctx->peers.size() : 0; }

static void free_scene(plu_gpu_ctx* ctx) {
	// scene tables are stream-ordered allocations (see bvh_gpu.cu: dev_mem): no device-wide synchronisation per table
	void* tables[] = {ctx->d_blob, ctx->d_tri_uv, ctx->d_tri_normal, ctx->d_texels, ctx->d_walk, ctx->d_fast, ctx->d_fast_index, ctx->d_fast_leaf};
	for (void* t : tables) if (t) cudaFreeAsync(t, ctx->stream);
	ctx->d_walk = nullptr; ctx->d_fast_leaf = nullptr;
	ctx->d_blob = nullptr; ctx->d_tri_uv = ctx->d_tri_normal = ctx->d_texels = nullptr; ctx->d_fast = nullptr; ctx->d_fast_index = nullptr;
	ctx->have_scene = false;
}

extern "C" void plu_gpu_destroy(plu_gpu_ctx* ctx) {
	if (!ctx) return;
	if (!ctx->comms.empty()) { nccl_api* nc = load_nccl(); for (ncclComm_t c : ctx->comms) if (c) nc->CommDestroy(c); ctx->comms.clear(); }
	for (plu_gpu_ctx* peer : ctx->peers) plu_gpu_destroy(peer);
	ctx->peers.clear();
	cudaSetDevice(ctx->device);
	free_scene(ctx);
	if (ctx->ev_r0) cudaEventDestroy(ctx->ev_r0);
	if (ctx->ev_r1) cudaEventDestroy(ctx->ev_r1);
	for (int k = 0; k < PLU_PIPES; ++k) { cudaFree(ctx->d_wave[k]); if (ctx->graphs[k].exec) cudaGraphExecDestroy(ctx->graphs[k].exec); }
	if (ctx->h_descs) cudaFreeHost(ctx->h_descs);
	cudaFree(ctx->d_fb);
	if (ctx->ev0) cudaEventDestroy(ctx->ev0);
	if (ctx->ev1) cudaEventDestroy(ctx->ev1);
	if (ctx->stream) cudaStreamDestroy(ctx->stream);
	if (ctx->ev_start) cudaEventDestroy(ctx->ev_start);
	for (int k = 0; k < PLU_PIPES; ++k) {
		if (ctx->pipe[k]) cudaStreamDestroy(ctx->pipe[k]);
		if (ctx->aux[k]) cudaStreamDestroy(ctx->aux[k]);
		if (ctx->ev_fork[k]) cudaEventDestroy(ctx->ev_fork[k]);
		if (ctx->ev_join[k]) cudaEventDestroy(ctx->ev_join[k]);
		if (ctx->ev_acc[k]) cudaEventDestroy(ctx->ev_acc[k]);
	}
	delete ctx;
}

extern "C" const char* plu_gpu_last_error(const plu_gpu_ctx* ctx) { return ctx ? ctx->err.c_str() : g_create_error.c_str(); }

// ---- scene upload ---------------------------------------------------------------------------------------------
template <class T>
static uint32_t blob_put(std::vector<unsigned char>& blob, const T* p, size_t n) {
	size_t off = (blob.size() + 15) & ~(size_t)15;
	blob.resize(off + std::max<size_t>(n * sizeof(T), 16));
	if (n) std::memcpy(blob.data() + off, p, n * sizeof(T));
	return (uint32_t)off;
}

static double wall_ms() { return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now().time_since_epoch()).count(); }

extern "C" int plu_gpu_upload_scene(plu_gpu_ctx* ctx, const plu_scene_desc* d) { return plu_gpu_upload_scene_ex(ctx, d, PLU_BUILD_DEVICE, nullptr); }

// ---- walk_node table of trace_ref_pruned (scene_dev.cuh): the reference tree + the tight box of every subtree ----
// float <-> unsigned images that keep the order, so that boxes merge with atomicMin / atomicMax
__device__ __forceinline__ uint32_t f2ord(float f) { uint32_t u = __float_as_uint(f); return (u & 0x80000000u) ? ~u : (u | 0x80000000u); }
__device__ __forceinline__ float ord2f(uint32_t u) { return __uint_as_float((u & 0x80000000u) ? (u & 0x7fffffffu) : ~u); }

__global__ void k_walk_init(int32_t n_nodes, const ref_node* ref, walk_node* out) {
	const int32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n_nodes) return;
	const uint32_t pinf = f2ord(INFINITY), ninf = f2ord(-INFINITY);
	walk_node w; w.a = ref[i].lo; w.b = ref[i].hi;
	w.c = make_float4(__uint_as_float(pinf), __uint_as_float(pinf), __uint_as_float(pinf), __int_as_float(1));
	w.d = make_float4(__uint_as_float(ninf), __uint_as_float(ninf), __uint_as_float(ninf), __int_as_float(0));
	out[i] = w;
}
// one thread per surface: its padded tight box joins every ancestor on the way from the root to its leaf (the path
// follows from the leaf's DFS rank: left subtree = the first k / 2 leaves)
__global__ void k_walk_merge(int32_t n, const float* prim, const int32_t* meta, walk_node* out) {
	const int32_t s = blockIdx.x * blockDim.x + threadIdx.x;
	if (s >= n) return;
	const float* f = prim + (size_t)s * 12;
	const int32_t tag = meta[4 * s], rank = meta[4 * s + 3];
	float mn[3], mx[3];
	for (int i = 0; i < 3; ++i) {
		if (tag == PLU_PRIM_SPHERE) { mn[i] = f[i] - fabsf(f[3]); mx[i] = f[i] + fabsf(f[3]); }
		else if (tag == PLU_PRIM_BOX) { mn[i] = fminf(f[i], f[4 + i]); mx[i] = fmaxf(f[i], f[4 + i]); }
		else { mn[i] = fminf(f[i], fminf(f[4 + i], f[8 + i])); mx[i] = fmaxf(f[i], fmaxf(f[4 + i], f[8 + i])); }
	}
	float big = 0.f, diag = 0.f;
	for (int i = 0; i < 3; ++i) { big = fmaxf(big, fmaxf(fabsf(mn[i]), fabsf(mx[i]))); diag += mx[i] - mn[i]; }
	const float pad = 1e-5f * big + 1e-5f * diag + 1e-30f;   // guard band >> rounding of the slab / hit arithmetic (as the fast BVH)
	int32_t node = 0, lo = 0, k = n;
	while (k > 1) {
		uint32_t* c = (uint32_t*)&out[node].c; uint32_t* d = (uint32_t*)&out[node].d;
		for (int i = 0; i < 3; ++i) { atomicMin(c + i, f2ord(mn[i] - pad)); atomicMax(d + i, f2ord(mx[i] + pad)); }
		((int32_t*)c)[3] = k;   // every surface below writes the same value
		const int32_t half = k / 2;
		if (rank < lo + half) { node += 1; k = half; } else { node += 2 * half; lo += half; k -= half; }
	}
	((int32_t*)&out[node].c)[3] = 1;
	((int32_t*)&out[node].d)[3] = rank;
}
__global__ void k_walk_finish(int32_t n_nodes, walk_node* out) {
	const int32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n_nodes) return;
	if (__float_as_int(out[i].b.w) >= 0) return;   // leaf
	float4 c = out[i].c, d = out[i].d;
	c.x = ord2f(__float_as_uint(c.x)); c.y = ord2f(__float_as_uint(c.y)); c.z = ord2f(__float_as_uint(c.z));
	d.x = ord2f(__float_as_uint(d.x)); d.y = ord2f(__float_as_uint(d.y)); d.z = ord2f(__float_as_uint(d.z));
	const float ex = d.x - c.x, ey = d.y - c.y, ez = d.z - c.z;
	d.w = 0.5f * sqrtf(ex * ex + ey * ey + ez * ez) * 1.0001f;   // radius of the box's bounding ball (trace_ref_pruned)
	out[i].c = c; out[i].d = d;
}

static int upload_single(plu_gpu_ctx* ctx, const plu_scene_desc* d, uint32_t build, plu_build_stats* stats);

// A multi-device context deep-copies the scene to every device and builds the traversal structures on each (one host
// thread per device; the builds are deterministic, so all devices hold byte-identical tables).
extern "C" int plu_gpu_upload_scene_ex(plu_gpu_ctx* ctx, const plu_scene_desc* d, uint32_t build, plu_build_stats* stats) {
	if (!ctx) return PLU_E_ARG;
	if (ctx->peers.empty()) return upload_single(ctx, d, build, stats);
	const size_t n = ctx->peers.size();
	std::vector<int> rc(n, PLU_OK);
	std::vector<std::thread> th;
	for (size_t g = 0; g < n; ++g) th.emplace_back([&, g] { rc[g] = upload_single(ctx->peers[g], d, build, nullptr); });
	int rc0 = upload_single(ctx, d, build, stats);
	for (auto& t : th) t.join();
	if (rc0 != PLU_OK) return rc0;
	for (size_t g = 0; g < n; ++g) if (rc[g] != PLU_OK) return fail(ctx, rc[g], "device " + std::to_string(ctx->peers[g]->device) + ": " + ctx->peers[g]->err);
	return PLU_OK;
}

static int upload_single_impl(plu_gpu_ctx* ctx, const plu_scene_desc* d, uint32_t build, plu_build_stats* stats);
static int upload_single(plu_gpu_ctx* ctx, const plu_scene_desc* d, uint32_t build, plu_build_stats* stats) {
	try { return upload_single_impl(ctx, d, build, stats); }   // no exception crosses the C ABI (or leaves a worker thread)
	catch (const std::bad_alloc&) { free_scene(ctx); return fail(ctx, PLU_E_NOMEM, "out of host memory while packing the scene"); }
	catch (const std::exception& e) { free_scene(ctx); return fail(ctx, PLU_E_ARG, std::string("scene upload: ") + e.what()); }
}

static int upload_single_impl(plu_gpu_ctx* ctx, const plu_scene_desc* d, uint32_t build, plu_build_stats* stats) {
	if (build != PLU_BUILD_DEVICE && build != PLU_BUILD_HOST) return fail(ctx, PLU_E_ARG, "build must be PLU_BUILD_DEVICE or PLU_BUILD_HOST");
	std::string v = pluhost::validate_desc(d);
	if (!v.empty()) return fail(ctx, PLU_E_ARG, v);
	CU(cudaSetDevice(ctx->device));
	static const bool no_cache = std::getenv("PLU_NO_UPLOAD_CACHE") != nullptr;
	if (!no_cache && ctx->have_scene && ctx->sig.equals(d, build)) {
		// the same surfaces, materials, textures and lights as the scene on the device: only the camera may differ
		ctx->sc.cam = d->cam;
		if (stats) { *stats = ctx->last_build; stats->ms_total = 0.0; stats->reserved = 1; }   // reserved = 1: tables reused, nothing rebuilt
		return PLU_OK;
	}
	ctx->sig.valid = false;
	free_scene(ctx);
	size_t n = (size_t)d->n_prims;
	const bool on_device = build == PLU_BUILD_DEVICE;
	plu_build_stats bs; std::memset(&bs, 0, sizeof(bs));
	double t_build = 0.0, t0 = wall_ms();

	std::vector<int32_t> rank(n, 0);
	if (d->prim_tie_rank) rank.assign(d->prim_tie_rank, d->prim_tie_rank + n);
	else if (!on_device) pluhost::reference_leaf_order(d, rank);
	bs.ms_ref_order = wall_ms() - t0;
	if (d->prim_tie_rank || !on_device) {   // the tie ranks must be a permutation
		std::vector<char> seen(n, 0);
		for (size_t i = 0; i < n; ++i) { if (rank[i] < 0 || (size_t)rank[i] >= n || seen[(size_t)rank[i]]) return fail(ctx, PLU_E_ARG, "prim_tie_rank is not a permutation"); seen[(size_t)rank[i]] = 1; }
	}
	std::vector<pluhost::ref_node_h> ref_nodes;
	bool use_fast = d->n_prims > ctx->fast_threshold;
	// the reference tree of a huge scene is only needed in verification mode; it is still built (O(n))
	t0 = wall_ms();
	if (on_device) ref_nodes.assign(n ? 2 * n - 1 : 0, pluhost::ref_node_h{});   // written by the device build below
	else pluhost::build_ref_nodes(d, rank, ref_nodes);
	bs.ms_ref_nodes = wall_ms() - t0;
	t_build = bs.ms_ref_order + bs.ms_ref_nodes;

	std::vector<int32_t> meta(4 * n);
	for (size_t s = 0; s < n; ++s) {
		const plu_material_desc& m = d->materials[d->prim_material[s]];
		meta[4 * s] = d->prim_tag[s]; meta[4 * s + 1] = d->prim_material[s];
		meta[4 * s + 2] = m.tag == PLU_MAT_EMISSION ? m.light : -1;   // material::area_light()
		meta[4 * s + 3] = rank[s];
	}
	std::vector<unsigned char> blob;
	scene_ref sc{};
	sc.off_ref_nodes = blob_put(blob, ref_nodes.data(), ref_nodes.size());
	sc.off_prim = blob_put(blob, d->prim_data, n * 12);
	sc.off_prim_meta = blob_put(blob, meta.data(), meta.size());
	sc.off_material = blob_put(blob, d->materials, (size_t)d->n_materials);
	sc.off_texture = blob_put(blob, d->textures, (size_t)d->n_textures);
	sc.off_light = blob_put(blob, d->lights, (size_t)d->n_lights);
	blob.resize((blob.size() + 15) & ~(size_t)15);
	sc.blob_bytes = (uint32_t)blob.size();
	sc.n_ref_nodes = (int32_t)ref_nodes.size();
	sc.n_prims = d->n_prims; sc.n_materials = d->n_materials; sc.n_textures = d->n_textures; sc.n_lights = d->n_lights;
	sc.cam = d->cam;
	sc.sphere_rmin = 3.0e38f;
	for (size_t s = 0; s < n; ++s) if (d->prim_tag[s] == PLU_PRIM_SPHERE) {
		const float r = std::fabs(d->prim_data[12 * s + 3]);
		sc.has_spheres = 1; sc.sphere_rmax = std::max(sc.sphere_rmax, r); sc.sphere_rmin = std::min(sc.sphere_rmin, r);
	}
	if (!sc.has_spheres) sc.sphere_rmin = 0.f;
	CU(cudaMallocAsync(&ctx->d_blob, blob.size(), ctx->stream));
	CU(cudaMemcpyAsync(ctx->d_blob, blob.data(), blob.size(), cudaMemcpyHostToDevice, ctx->stream));
	sc.blob = ctx->d_blob;
	if (d->prim_tri_uv && d->prim_tri_normal && n) {
		CU(cudaMallocAsync(&ctx->d_tri_uv, n * 6 * sizeof(float), ctx->stream)); CU(cudaMemcpyAsync(ctx->d_tri_uv, d->prim_tri_uv, n * 6 * sizeof(float), cudaMemcpyHostToDevice, ctx->stream));
		CU(cudaMallocAsync(&ctx->d_tri_normal, n * 9 * sizeof(float), ctx->stream)); CU(cudaMemcpyAsync(ctx->d_tri_normal, d->prim_tri_normal, n * 9 * sizeof(float), cudaMemcpyHostToDevice, ctx->stream));
	}
	if (d->image_texels && d->n_image_texels > 0) {
		size_t bytes = (size_t)d->n_image_texels * 3 * sizeof(float);
		CU(cudaMallocAsync(&ctx->d_texels, bytes, ctx->stream)); CU(cudaMemcpyAsync(ctx->d_texels, d->image_texels, bytes, cudaMemcpyHostToDevice, ctx->stream));
	}
	sc.tri_uv = ctx->d_tri_uv; sc.tri_normal = ctx->d_tri_normal; sc.texels = ctx->d_texels;
	if (on_device) {
		plubuild::device_tables tb;
		tb.d_prim = (const float*)(ctx->d_blob + sc.off_prim);
		tb.d_prim_meta = (int32_t*)(ctx->d_blob + sc.off_prim_meta);
		tb.d_ref_nodes = ctx->d_blob + sc.off_ref_nodes;
		std::string berr;
		CU(cudaStreamSynchronize(ctx->stream));   // the table uploads above are staged from host vectors
		t0 = wall_ms();
		bool ok = plubuild::build_on_device(d, tb, ctx->stream, &bs, berr);
		ctx->d_fast = (fast_node*)tb.d_fast; ctx->d_fast_index = tb.d_fast_index; ctx->d_fast_leaf = (float4*)tb.d_fast_leaf;
		if (!ok) { free_scene(ctx); return fail(ctx, PLU_E_CUDA, "device BVH build: " + berr); }
		sc.n_fast_nodes = n ? tb.n_fast_nodes : 0;
		t_build = wall_ms() - t0;
	} else {
		t0 = wall_ms();
		std::vector<pluhost::fast_node_h> fn; std::vector<int32_t> fi;
		pluhost::build_fast_bvh(d, fn, fi);
		if (!fn.empty()) {
			// leaf references carry their surface count; the surfaces are copied into leaf order (scene_dev.cuh)
			for (auto& nd : fn) {
				for (int k = 0; k < 2; ++k)
					if (nd.child[k] < 0) nd.child[k] = ~(((~nd.child[k]) << 3) | (nd.count[k] & 7));
				// builder order: c0.min, c0.max, c1.min, c1.max  ->  device order: (child 0, child 1) pairs per plane
				float f[12]; std::memcpy(f, nd.f, sizeof(f));
				for (int ax = 0; ax < 3; ++ax) {
					nd.f[2 * ax] = f[ax]; nd.f[2 * ax + 1] = f[6 + ax];             // min planes
					nd.f[6 + 2 * ax] = f[3 + ax]; nd.f[6 + 2 * ax + 1] = f[9 + ax];   // max planes
				}
			}
			std::vector<float> leaf(12 * fi.size(), 0.f);
			for (size_t k = 0; k < fi.size(); ++k) {
				int32_t s = fi[k], kind = d->prim_tag[s];
				float* o = leaf.data() + 12 * k;
				std::memcpy(o, d->prim_data + (size_t)s * 12, 12 * sizeof(float));
				int32_t kr = (int32_t)(((uint32_t)kind << 30) | (uint32_t)rank[(size_t)s]);
				std::memcpy(o + 7, &kr, 4);
				std::memcpy(kind == PLU_PRIM_SPHERE ? o + 4 : o + 3, &s, 4);
			}
			CU(cudaMallocAsync(&ctx->d_fast_leaf, leaf.size() * sizeof(float), ctx->stream));
			CU(cudaMemcpyAsync(ctx->d_fast_leaf, leaf.data(), leaf.size() * sizeof(float), cudaMemcpyHostToDevice, ctx->stream));
			CU(cudaMallocAsync(&ctx->d_fast, fn.size() * sizeof(fast_node), ctx->stream)); CU(cudaMemcpyAsync(ctx->d_fast, fn.data(), fn.size() * sizeof(fast_node), cudaMemcpyHostToDevice, ctx->stream));
			CU(cudaMallocAsync(&ctx->d_fast_index, fi.size() * sizeof(int32_t), ctx->stream)); CU(cudaMemcpyAsync(ctx->d_fast_index, fi.data(), fi.size() * sizeof(int32_t), cudaMemcpyHostToDevice, ctx->stream));
			CU(cudaStreamSynchronize(ctx->stream));   // the host vectors go out of scope
		}
		sc.n_fast_nodes = (int32_t)fn.size();
		bs.ms_fast_bvh = wall_ms() - t0; bs.n_fast_nodes = sc.n_fast_nodes;
		t_build += bs.ms_fast_bvh;
	}
	bs.ms_total = t_build;
	if (stats) *stats = bs;
	{   // the traversal stacks hold PLU_WALK_DEPTH_MAX entries (isect.cuh: walk_stack): refuse a tree that could overflow them
		int lg = 0; while (((size_t)1 << lg) < std::max<size_t>(n, 2)) ++lg;
		if (41 + lg > PLU_WALK_DEPTH_MAX || bs.levels_fast > PLU_WALK_DEPTH_MAX) { free_scene(ctx); return fail(ctx, PLU_E_ARG, "fast BVH deeper than the traversal stack"); }
	}
	sc.fast_nodes = use_fast ? ctx->d_fast : nullptr;
	ctx->use_fast = use_fast;
	ctx->has_specular = false;
	for (int32_t i = 0; i < d->n_materials; ++i) if (d->materials[i].tag >= PLU_MAT_PERFECT_REFLECTION && d->materials[i].tag <= PLU_MAT_GLASS) ctx->has_specular = true;
	sc.fast_prim_index = ctx->d_fast_index; sc.fast_leaf = ctx->d_fast_leaf;
	sc.walk_nodes = nullptr;
	if (sc.has_spheres && sc.n_ref_nodes > 0) {   // (small scenes only reach it with PLU_RENDER_BVH_FAST)
		// sphere::hit is fed non-unit directions, which no walk of the SAH tree can follow: those scenes walk the reference
		// tree with tight-box pruning instead (isect.cuh: trace_ref_pruned)
		CU(cudaMallocAsync(&ctx->d_walk, (size_t)sc.n_ref_nodes * sizeof(walk_node), ctx->stream));
		const int nb = (sc.n_ref_nodes + 255) / 256;
		k_walk_init<<<nb, 256, 0, ctx->stream>>>(sc.n_ref_nodes, (const ref_node*)(ctx->d_blob + sc.off_ref_nodes), ctx->d_walk);
		k_walk_merge<<<(sc.n_prims + 255) / 256, 256, 0, ctx->stream>>>(sc.n_prims, (const float*)(ctx->d_blob + sc.off_prim), (const int32_t*)(ctx->d_blob + sc.off_prim_meta), ctx->d_walk);
		k_walk_finish<<<nb, 256, 0, ctx->stream>>>(sc.n_ref_nodes, ctx->d_walk);
		CU(cudaGetLastError());
		CU(cudaStreamSynchronize(ctx->stream));
		sc.walk_nodes = ctx->d_walk;
	}
	ctx->sc = sc;
	ctx->blob_in_smem = sc.blob_bytes <= PLU_SMEM_BLOB_MAX;
	CU(cudaStreamSynchronize(ctx->stream));   // uploads and builds have landed before any other stream renders from them
	ctx->have_scene = true;
	ctx->last_build = bs;
	if (!no_cache) ctx->sig.store(d, build);
	return PLU_OK;
}

extern "C" int plu_gpu_download_table(plu_gpu_ctx* ctx, int which, void* out, int64_t cap_bytes, int64_t* n_bytes) {
	if (!ctx || !n_bytes) return PLU_E_ARG;
	if (!ctx->have_scene) return fail(ctx, PLU_E_NOSCENE, "no scene uploaded");
	CU(cudaSetDevice(ctx->device));
	const scene_ref& sc = ctx->sc;
	const void* src = nullptr; int64_t bytes = 0;
	const bool fast = ctx->d_fast != nullptr;
	switch (which) {
	case PLU_TABLE_REF_NODES: src = ctx->d_blob + sc.off_ref_nodes; bytes = (int64_t)sc.n_ref_nodes * (int64_t)sizeof(ref_node); break;
	case PLU_TABLE_PRIM_META: src = ctx->d_blob + sc.off_prim_meta; bytes = (int64_t)sc.n_prims * 16; break;
	case PLU_TABLE_FAST_NODES: src = ctx->d_fast; bytes = fast ? (int64_t)sc.n_fast_nodes * (int64_t)sizeof(fast_node) : 0; break;
	case PLU_TABLE_FAST_INDEX: src = ctx->d_fast_index; bytes = fast ? (int64_t)sc.n_prims * 4 : 0; break;
	case PLU_TABLE_FAST_LEAF: src = ctx->d_fast_leaf; bytes = fast ? (int64_t)sc.n_prims * 48 : 0; break;
	default: return fail(ctx, PLU_E_ARG, "unknown table");
	}
	*n_bytes = bytes;
	if (!out || bytes == 0) return PLU_OK;
	if (cap_bytes < bytes) return fail(ctx, PLU_E_ARG, "buffer too small for the table");
	CU(cudaMemcpy(out, src, (size_t)bytes, cudaMemcpyDeviceToHost));
	return PLU_OK;
}

static void select_bvh(const plu_gpu_ctx* ctx, scene_ref& sc, uint32_t flags) {
	if (flags & PLU_RENDER_BVH_REFERENCE) sc.fast_nodes = nullptr;
	else if (flags & PLU_RENDER_BVH_FAST) sc.fast_nodes = ctx->d_fast;
}

// ---- wave buffers ---------------------------------------------------------------------------------------------
#define PLU_MAX_VERTICES_CAP 64
#ifndef PLU_WAVE_PATHS
#define PLU_WAVE_PATHS (1u << 22)
#endif
#define PLU_WAVE_PATHS_MAX (1u << 24)

static int ensure_wave(plu_gpu_ctx* ctx, uint32_t slots, int P = 0) {
	if (ctx->cap[P] >= slots && ctx->d_wave[P]) return PLU_OK;
	CU(cudaFree(ctx->d_wave[P])); ctx->d_wave[P] = nullptr; ctx->cap[P] = 0;
	size_t N = slots;
	auto al = [](size_t b) { return (b + 255) & ~(size_t)255; };
	size_t bytes = al(N * 64) + 8 * al(N * 16) + 2 * al(N * 4) + al((PLU_MAX_VERTICES_CAP + 2) * CNT_STRIDE * sizeof(uint32_t)) + al(8 * sizeof(unsigned long long)) + al(sizeof(wave_desc)) + al((N + 31) / 32 * 4);
	CU(cudaMalloc(&ctx->d_wave[P], bytes));
	unsigned char* p = (unsigned char*)ctx->d_wave[P];
	auto take = [&](size_t b) { unsigned char* r = p; p += al(b); return r; };
	wave_buffers& wb = ctx->wbs[P];
	// 16-byte rows first so every array is 256-byte aligned
	wb.rec = (float4*)take(N * 64); wb.rad = (float4*)take(N * 16);
	wb.sh_p = (float4*)take(N * 16); wb.sh_w1 = (float4*)take(N * 16); wb.sh_c1 = (float4*)take(N * 16);
	wb.sh_w2 = (float4*)take(N * 16); wb.sh_c2 = (float4*)take(N * 16);
	wb.queue[0] = (uint32_t*)take(N * 4); wb.queue[1] = (uint32_t*)take(N * 4);
	wb.q_hit[0] = (uint4*)take(N * 16); wb.q_hit[1] = (uint4*)take(N * 16);
	wb.counters = (uint32_t*)take((PLU_MAX_VERTICES_CAP + 2) * CNT_STRIDE * sizeof(uint32_t));
	wb.stats = (unsigned long long*)take(8 * sizeof(unsigned long long));
	wb.desc = (const wave_desc*)take(sizeof(wave_desc));
	wb.dirty = (uint32_t*)take((N + 31) / 32 * 4);
	if ((size_t)(p - (unsigned char*)ctx->d_wave[P]) != bytes) return fail(ctx, PLU_E_NOMEM, "wave buffer carve-out mismatch");
	ctx->cap[P] = slots;
	return PLU_OK;
}

template <class K>
static int grid_for(plu_gpu_ctx* ctx, K kernel, size_t smem) {
	int occ = 0;
	cudaFuncAttributes fa; size_t fixed = 0;   // static shared memory (push counters, PLU_SSTK stack levels) counts towards the 48 KB default limit
	if (cudaFuncGetAttributes(&fa, kernel) == cudaSuccess) fixed = fa.sharedSizeBytes;
	if (smem + fixed > 48 * 1024) cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
	if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kernel, PLU_CTA, smem) != cudaSuccess || occ < 1) occ = 1;
	return ctx->sm_count * occ;
}

// dynamic shared memory of a kernel = the staged scene blob (small scenes) + its two cp.async stages
struct launch_cfg { bool smem, count, no_pt, no_fork; size_t sm_ext, sm_shade, sm_conn; int g_gen, g_ext, g_prim, g_ext_pt, g_conn_pt, g_shade[2], g_conn, g_acc; };

// Waves of a large scene (fast BVH / pruned reference tree) extend with persistent warps (k_extend_pt) and start with
// k_generate; all other waves start with k_primary (camera ray generation fused into the first extension).
static bool wave_uses_pt(const scene_ref& sc, const launch_cfg& c) { return !c.smem && sc.fast_nodes != nullptr && !c.no_pt; }

// force_global: PLU_RENDER_BVH_FAST on a scene that would be staged in shared memory — it then runs through the
// global-memory kernels (the only ones that hold the large-scene walks)
static launch_cfg make_cfg(plu_gpu_ctx* ctx, bool count, bool force_global = false) {
	launch_cfg c; std::memset(&c, 0, sizeof(c));   // (the struct is part of the graph key: no indeterminate padding)
	c.smem = ctx->blob_in_smem && !force_global; c.count = count; c.no_pt = false; c.no_fork = std::getenv("PLU_NO_FORK") != nullptr;
	if (ctx->sc.has_spheres) c.g_ext_pt = count ? grid_for(ctx, k_extend_pt<true, pruned_walker>, 0) : grid_for(ctx, k_extend_pt<false, pruned_walker>, 0);
	else c.g_ext_pt = count ? grid_for(ctx, k_extend_pt<true>, 0) : grid_for(ctx, k_extend_pt<false>, 0);
	if (ctx->sc.has_spheres) c.g_conn_pt = count ? grid_for(ctx, k_connect_pt<true, pruned_walker>, 0) : grid_for(ctx, k_connect_pt<false, pruned_walker>, 0);
	else c.g_conn_pt = count ? grid_for(ctx, k_connect_pt<true, fast_walker>, 0) : grid_for(ctx, k_connect_pt<false, fast_walker>, 0);
	size_t blob = c.smem ? ((size_t)ctx->sc.blob_bytes + 127) & ~(size_t)127 : 0;
	c.sm_ext = blob + PLU_STAGE_EXTEND; c.sm_shade = blob + PLU_STAGE_SHADE; c.sm_conn = blob + PLU_STAGE_CONNECT;
	c.g_gen = grid_for(ctx, k_generate, 0);
	c.g_acc = grid_for(ctx, k_accumulate, 0);
#define PICK(var, K, SM) \
	var = c.smem ? (count ? grid_for(ctx, K<true, true>, SM) : grid_for(ctx, K<true, false>, SM)) \
	             : (count ? grid_for(ctx, K<false, true>, SM) : grid_for(ctx, K<false, false>, SM))
	PICK(c.g_ext, k_extend, c.sm_ext); PICK(c.g_conn, k_connect, c.sm_conn);
#undef PICK
	c.g_prim = c.smem ? (count ? grid_for(ctx, k_extend<true, true, true>, c.sm_ext) : grid_for(ctx, k_extend<true, false, true>, c.sm_ext))
	                  : (count ? grid_for(ctx, k_extend<false, true, true>, c.sm_ext) : grid_for(ctx, k_extend<false, false, true>, c.sm_ext));
	c.g_shade[0] = ctx->sm_count * plushade::max_blocks_per_sm(c.smem, CLS_DIFFUSE, c.sm_shade);   // (k_shade lives in shade_tu.cu)
	c.g_shade[1] = ctx->sm_count * plushade::max_blocks_per_sm(c.smem, CLS_SPECULAR, c.sm_shade);
	return c;
}

// PLU_RENDER_PROFILE: one event after every launch of a wave; the gaps are summed per kernel class.
struct wave_profiler {
	bool on = false;
	std::vector<cudaEvent_t> ev;
	std::vector<int> cls;       // class of the launch that ENDS at event i (0 gen, 1 extend, 2 shade, 3 connect, 4 accumulate, -1 start)
	double ms[5] = {0, 0, 0, 0, 0};
	size_t used = 0;
	void mark(int c, cudaStream_t st) {
		if (!on) return;
		if (used == ev.size()) { cudaEvent_t e; cudaEventCreate(&e); ev.push_back(e); cls.push_back(0); }
		cls[used] = c; cudaEventRecord(ev[used], st); used++;
	}
	void collect(cudaStream_t st) {
		if (!on || used == 0) return;
		cudaStreamSynchronize(st);
		for (size_t i = 1; i < used; ++i) if (cls[i] >= 0) { float m = 0.f; cudaEventElapsedTime(&m, ev[i - 1], ev[i]); ms[cls[i]] += m; }
		used = 0;
	}
	~wave_profiler() { for (auto e : ev) cudaEventDestroy(e); }
};

// Enqueue one wave (generate -> [extend, shade, connect] x max_vertices); returns the launch count.
// ---- launch sequence of a wave -----------------------------------------------------------------------------------
// The wave's descriptor must already be on its way to wb.desc on `st` (put_desc); nothing below depends on the wave, so
// the same sequence serves every wave of a render (and is what wave_graph_for records).  `b` is the path vertex, or
// PLU_B_LOOP / PLU_B_LOOP_NEXT inside the graph's while loop (the kernels then read the vertex from device memory).
// persistent warps with lane refill over the scene's walk: the SAH tree, or (scenes with spheres) the pruned reference tree
static void launch_extend_pt(plu_gpu_ctx* ctx, const scene_ref& sc, const launch_cfg& c, cudaStream_t st, const wave_buffers& wb, int b) {
	if (sc.has_spheres) {
		if (c.count) k_extend_pt<true, pruned_walker><<<c.g_ext_pt, PLU_CTA, 0, st>>>(sc, wb, b, ctx->pt_refill);
		else k_extend_pt<false, pruned_walker><<<c.g_ext_pt, PLU_CTA, 0, st>>>(sc, wb, b, ctx->pt_refill);
	} else {
		if (c.count) k_extend_pt<true><<<c.g_ext_pt, PLU_CTA, 0, st>>>(sc, wb, b, ctx->pt_refill);
		else k_extend_pt<false><<<c.g_ext_pt, PLU_CTA, 0, st>>>(sc, wb, b, ctx->pt_refill);
	}
}

static void launch_head(plu_gpu_ctx* ctx, const scene_ref& sc, int max_vertices, const launch_cfg& c, cudaStream_t st, wave_profiler* pf, int P, bool pt) {
	wave_buffers& wb = ctx->wbs[P];
	cudaMemsetAsync(wb.counters, 0, (size_t)(max_vertices + 2) * CNT_STRIDE * sizeof(uint32_t), st);   // incl. the loop word, row max_vertices + 1
	pf->mark(-1, st);
	if (pt) {
		k_generate<<<c.g_gen, 256, 0, st>>>(sc, wb); pf->mark(0, st);
		launch_extend_pt(ctx, sc, c, st, wb, 0);
	} else if (c.smem) {   // k_primary: generate + extend of vertex 0 in one kernel (its time is booked under "extend")
		if (c.count) k_extend<true, true, true><<<c.g_prim, PLU_CTA, c.sm_ext, st>>>(sc, wb, 0); else k_extend<true, false, true><<<c.g_prim, PLU_CTA, c.sm_ext, st>>>(sc, wb, 0);
	} else {
		if (c.count) k_extend<false, true, true><<<c.g_prim, PLU_CTA, c.sm_ext, st>>>(sc, wb, 0); else k_extend<false, false, true><<<c.g_prim, PLU_CTA, c.sm_ext, st>>>(sc, wb, 0);
	}
	pf->mark(1, st);
}

static void launch_extend(plu_gpu_ctx* ctx, const scene_ref& sc, const launch_cfg& c, cudaStream_t st, wave_profiler* pf, int P, bool pt, int b) {
	wave_buffers& wb = ctx->wbs[P];
	if (pt) {   // fast BVH, no sphere (whose non-unit-direction rays need another walk): persistent warps with lane refill
		launch_extend_pt(ctx, sc, c, st, wb, b);
	} else if (c.smem) {
		if (c.count) k_extend<true, true><<<c.g_ext, PLU_CTA, c.sm_ext, st>>>(sc, wb, b);
		else k_extend<true, false><<<c.g_ext, PLU_CTA, c.sm_ext, st>>>(sc, wb, b);
	} else {
		if (c.count) k_extend<false, true><<<c.g_ext, PLU_CTA, c.sm_ext, st>>>(sc, wb, b);
		else k_extend<false, false><<<c.g_ext, PLU_CTA, c.sm_ext, st>>>(sc, wb, b);
	}
	pf->mark(1, st);
}

// shade (one kernel per material class) + connect of one vertex; returns the launch count
static int launch_shade_connect(plu_gpu_ctx* ctx, const scene_ref& sc, const launch_cfg& c, cudaStream_t st, wave_profiler* pf, int P, int b) {
	wave_buffers& wb = ctx->wbs[P];
	// the two shade kernels of a vertex touch disjoint slots and push through atomics: the specular one (few vertices,
	// mostly prologue and tail) runs on a second stream and fills the SMs the diffuse one leaves idle
	const bool fork = ctx->has_specular && !pf->on && !c.no_fork;
	int launches = 2;
	if (fork) cudaEventRecord(ctx->ev_fork[P], st);
	plushade::launch(c.smem, CLS_DIFFUSE, c.g_shade[0], c.sm_shade, st, sc, wb, b);
	if (ctx->has_specular) {
		cudaStream_t s2 = fork ? ctx->aux[P] : st;
		if (fork) cudaStreamWaitEvent(s2, ctx->ev_fork[P], 0);
		plushade::launch(c.smem, CLS_SPECULAR, c.g_shade[1], c.sm_shade, s2, sc, wb, b);
		launches++;
		if (fork) { cudaEventRecord(ctx->ev_join[P], s2); cudaStreamWaitEvent(st, ctx->ev_join[P], 0); }
	}
	pf->mark(2, st);
	if (wave_uses_pt(sc, c) && std::getenv("PLU_NO_CONNECT_PT") == nullptr) {   // large-scene walks: persistent warps with lane refill
		if (sc.has_spheres) { if (c.count) k_connect_pt<true, pruned_walker><<<c.g_conn_pt, PLU_CTA, 0, st>>>(sc, wb, b, ctx->pt_refill); else k_connect_pt<false, pruned_walker><<<c.g_conn_pt, PLU_CTA, 0, st>>>(sc, wb, b, ctx->pt_refill); }
		else { if (c.count) k_connect_pt<true, fast_walker><<<c.g_conn_pt, PLU_CTA, 0, st>>>(sc, wb, b, ctx->pt_refill); else k_connect_pt<false, fast_walker><<<c.g_conn_pt, PLU_CTA, 0, st>>>(sc, wb, b, ctx->pt_refill); }
	}
	else if (c.smem) { if (c.count) k_connect<true, true><<<c.g_conn, PLU_CTA, c.sm_conn, st>>>(sc, wb, b); else k_connect<true, false><<<c.g_conn, PLU_CTA, c.sm_conn, st>>>(sc, wb, b); }
	else if (sc.fast_nodes == nullptr) { if (c.count) k_connect<false, true><<<c.g_conn, PLU_CTA, c.sm_conn, st>>>(sc, wb, b); else k_connect<false, false><<<c.g_conn, PLU_CTA, c.sm_conn, st>>>(sc, wb, b); }
	else if (sc.has_spheres) { if (c.count) k_connect<false, true, PLU_WALKS_PRUNED><<<c.g_conn, PLU_CTA, c.sm_conn, st>>>(sc, wb, b); else k_connect<false, false, PLU_WALKS_PRUNED><<<c.g_conn, PLU_CTA, c.sm_conn, st>>>(sc, wb, b); }
	else { if (c.count) k_connect<false, true, PLU_WALKS_FAST><<<c.g_conn, PLU_CTA, c.sm_conn, st>>>(sc, wb, b); else k_connect<false, false, PLU_WALKS_FAST><<<c.g_conn, PLU_CTA, c.sm_conn, st>>>(sc, wb, b); }
	pf->mark(3, st);
	return launches;
}

// The whole chain, vertex by vertex (profiled renders, the parity hooks, and the fallback when graphs are unavailable).
static int enqueue_wave(plu_gpu_ctx* ctx, const scene_ref& sc, int max_vertices, const launch_cfg& c, cudaStream_t st, wave_profiler* pf = nullptr, int P = 0) {
	wave_profiler none; if (!pf) pf = &none;
	const bool pt = wave_uses_pt(sc, c);
	launch_head(ctx, sc, max_vertices, c, st, pf, P, pt);
	int launches = pt ? 2 : 1;
	for (int b = 0; b < max_vertices; ++b) {
		if (b > 0) { launch_extend(ctx, sc, c, st, pf, P, pt, b); launches++; }
		launches += launch_shade_connect(ctx, sc, c, st, pf, P, b);
	}
	return launches;
}

// The pinned descriptor array of a render (one entry per wave; an entry is read by its async copy when the stream gets
// there, so entries are never reused inside a render).
static int ensure_descs(plu_gpu_ctx* ctx, size_t n) {
	if (ctx->h_descs_cap >= n) return PLU_OK;
	if (ctx->h_descs) { CU(cudaFreeHost(ctx->h_descs)); ctx->h_descs = nullptr; ctx->h_descs_cap = 0; }
	size_t cap = std::max<size_t>(n, 1024);
	CU(cudaHostAlloc(&ctx->h_descs, cap * sizeof(wave_desc), cudaHostAllocDefault));
	ctx->h_descs_cap = cap;
	return PLU_OK;
}

// The graph of one wave on pipe P (launch sequence of enqueue_wave, captured once and replayed for every wave).  It is
// rebuilt when anything a kernel parameter was captured from changes: the scene tables, the kernel variants and grids,
// the path-length limit, the pipe's buffers or streams.  Returns nullptr when the capture fails (the caller then
// enqueues the launches directly — same kernels either way).
static plu_gpu_ctx::wave_graph* wave_graph_for(plu_gpu_ctx* ctx, const scene_ref& sc, int max_vertices, const launch_cfg& c, cudaStream_t st, int P) {
	std::vector<unsigned char> key;
	auto put = [&](const void* p, size_t n) { const unsigned char* b = (const unsigned char*)p; key.insert(key.end(), b, b + n); };
	put(&sc, sizeof(sc)); put(&c, sizeof(c)); put(&max_vertices, sizeof(max_vertices)); put(&ctx->wbs[P], sizeof(wave_buffers));
	put(&st, sizeof(st)); put(&ctx->has_specular, sizeof(bool)); put(&ctx->pt_refill, sizeof(uint32_t));
	plu_gpu_ctx::wave_graph& g = ctx->graphs[P];
	if (g.exec && g.key == key) return &g;
	if (g.exec) { cudaGraphExecDestroy(g.exec); g.exec = nullptr; }
	wave_profiler none;
	const bool pt = wave_uses_pt(sc, c);
	cudaGraph_t graph = nullptr;
	cudaError_t e = cudaSuccess;
	g.launches = 0; g.looped = false;
	// measured (profiles/r02_summary.md): a round of the while loop costs more than the ~4 launches on empty queues it saves
	// (glass0 16.8 looped vs 18.2 unrolled Gsamples/s, room 8.3 vs 8.5), so the loop is opt-in
	const bool no_loop = std::getenv("PLU_GRAPH_LOOP") == nullptr;
	if (!no_loop) {
		// memset -> first extension (k_primary | k_generate + k_extend_pt) -> WHILE { shade, connect of vertex v; extension of
		// vertex v + 1; k_loop_next: ++v, go on iff v < max_vertices and the extension found a surface }: a wave stops at the
		// depth its longest path reaches instead of running all 4 x max_vertices launches on empty queues
		const cudaGraphNode_t* deps = nullptr; size_t ndeps = 0;
		cudaStreamCaptureStatus cs; unsigned long long id; cudaGraph_t cg = nullptr;
		cudaGraphConditionalHandle handle;
		cudaGraphNode_t loop_node;
		cudaGraphNodeParams cp = {};
		const char* step = "cudaGraphCreate";
		e = cudaGraphCreate(&graph, 0);
		if (e == cudaSuccess) { step = "begin capture (head)"; e = cudaStreamBeginCaptureToGraph(st, graph, nullptr, nullptr, 0, cudaStreamCaptureModeThreadLocal); }
		if (e == cudaSuccess) {
			launch_head(ctx, sc, max_vertices, c, st, &none, P, pt);
			step = "capture info"; e = cudaStreamGetCaptureInfo_v2(st, &cs, &id, &cg, &deps, &ndeps);
			std::vector<cudaGraphNode_t> tail(deps, deps + (e == cudaSuccess ? ndeps : 0));
			cudaGraph_t ended = nullptr;
			cudaError_t e2 = cudaStreamEndCapture(st, &ended);
			if (e == cudaSuccess && e2 != cudaSuccess) { step = "end capture (head)"; e = e2; }
			if (e == cudaSuccess) { step = "conditional handle"; e = cudaGraphConditionalHandleCreate(&handle, graph, 1, cudaGraphCondAssignDefault); }
			if (e == cudaSuccess) {
				cp.type = cudaGraphNodeTypeConditional; cp.conditional.handle = handle; cp.conditional.type = cudaGraphCondTypeWhile; cp.conditional.size = 1;
				step = "add while node"; e = cudaGraphAddNode(&loop_node, graph, tail.data(), tail.size(), &cp);
			}
			if (e == cudaSuccess) { step = "begin capture (body)"; e = cudaStreamBeginCaptureToGraph(st, cp.conditional.phGraph_out[0], nullptr, nullptr, 0, cudaStreamCaptureModeThreadLocal); }
			if (e == cudaSuccess) {
				int per_vertex = launch_shade_connect(ctx, sc, c, st, &none, P, PLU_B_LOOP);
				launch_extend(ctx, sc, c, st, &none, P, pt, PLU_B_LOOP_NEXT);
				k_loop_next<<<1, 32, 0, st>>>(ctx->wbs[P], handle);
				step = "end capture (body)"; e = cudaStreamEndCapture(st, &ended);
				g.launches = pt ? 2 : 1;            // the head; the rounds of the loop are counted on the device (stats[6])
				g.body_launches = per_vertex + 2;
			}
		}
		if (e == cudaSuccess) { step = "instantiate"; e = cudaGraphInstantiate(&g.exec, graph, 0); }
		if (graph) cudaGraphDestroy(graph);
		graph = nullptr;
		if (e == cudaSuccess) { g.looped = true; g.key.swap(key); return &g; }
		if (std::getenv("PLU_GRAPH_DEBUG")) std::fprintf(stderr, "libplugpu: looped wave graph unavailable (%s: %s), using the unrolled one\n", step, cudaGetErrorString(e));
		cudaGetLastError(); g.exec = nullptr; g.body_launches = 0;   // fall through to the unrolled graph
	}
	if (cudaStreamBeginCapture(st, cudaStreamCaptureModeThreadLocal) != cudaSuccess) { cudaGetLastError(); return nullptr; }
	g.launches = enqueue_wave(ctx, sc, max_vertices, c, st, nullptr, P);
	e = cudaStreamEndCapture(st, &graph);
	if (e == cudaSuccess) e = cudaGraphInstantiate(&g.exec, graph, 0);
	if (graph) cudaGraphDestroy(graph);
	if (e != cudaSuccess) { cudaGetLastError(); g.exec = nullptr; return nullptr; }
	g.key.swap(key);
	return &g;
}

static int put_desc(plu_gpu_ctx* ctx, const wave_desc* host, cudaStream_t st, int P) {
	CU(cudaMemcpyAsync((void*)ctx->wbs[P].desc, host, sizeof(wave_desc), cudaMemcpyHostToDevice, st));
	return PLU_OK;
}

static int check_params(plu_gpu_ctx* ctx, const plu_render_params* p) {
	if (!p) return fail(ctx, PLU_E_ARG, "render params are NULL");
	if (p->width <= 0 || p->height <= 0 || p->n_strata <= 0) return fail(ctx, PLU_E_ARG, "width, height and n_strata must be positive");
	if ((int64_t)p->width * p->height > (int64_t)1 << 30) return fail(ctx, PLU_E_ARG, "image too large");
	if (p->n_strata > 32768) return fail(ctx, PLU_E_ARG, "n_strata too large");
	if (p->max_vertices < 0 || p->max_vertices > PLU_MAX_VERTICES_CAP) return fail(ctx, PLU_E_ARG, "max_vertices out of range");
	if (p->sample_count < 0 || (p->sample_count > 0 && (p->sample_stride <= 0 || p->sample_begin < 0))) return fail(ctx, PLU_E_ARG, "bad sample range");
	if (p->sample_count > 0 && (int64_t)p->sample_begin + (int64_t)(p->sample_count - 1) * p->sample_stride >= (int64_t)p->n_strata * p->n_strata)
		return fail(ctx, PLU_E_ARG, "sample range exceeds n_strata^2 (a sample index beyond the strata grid would land in the next pixel row)");
	if (!ctx->have_scene) return fail(ctx, PLU_E_NOSCENE, "no scene uploaded");
	return PLU_OK;
}

// udiv32 for divisor d >= 1 (wavefront.cuh: div_by), checked on the numerators around the multiples of d and at 2^31 - 1
static udiv32 set_div(uint32_t d) {
	udiv32 v{d, 0u, 0u};
	if (d <= 1u) { v.d = 1u; return v; }
	uint32_t l = 0; while ((1ull << l) < d) ++l;
	v.m = (uint32_t)(((1ull << (31 + l)) / d) + 1ull); v.sh = l - 1u;
	const uint32_t probe[] = {0u, d - 1u, d, d + 1u, 2u * d - 1u, 0x7fffffffu / d * d - 1u, 0x7fffffffu / d * d, 0x7fffffffu};
	for (uint32_t n : probe)
		if (n <= 0x7fffffffu && (uint32_t)(((unsigned long long)n * v.m) >> 32) >> v.sh != n / d) { std::fprintf(stderr, "plugpu: set_div(%u) fails at %u\n", d, n); std::abort(); }
	return v;
}

static wave_desc base_desc(const plu_scene_desc*, const plu_render_params* p, const scene_ref& sc) {
	wave_desc w{};
	w.width = p->width; w.height = p->height; w.n_strata = p->n_strata;
	w.div_width = set_div((uint32_t)p->width); w.div_strata = set_div((uint32_t)p->n_strata);
	{ volatile float one = 1.f, nf = (float)(uint32_t)p->n_strata; w.inv_strata = one / nf; }   // (volatile: a division at run time, not a folded constant)
	w.sample_begin = p->sample_count > 0 ? p->sample_begin : 0;
	w.sample_stride = p->sample_count > 0 ? p->sample_stride : 1;
	w.max_vertices = p->max_vertices > 0 ? p->max_vertices : 8;
	w.seed_lo = (uint32_t)p->seed; w.seed_hi = (uint32_t)(p->seed >> 32);
	w.use_lens = sc.cam.lens_radius > 0.f ? 1 : 0;
	return w;
}

// One device: the waves of the call's sample set on `st` (and the context's pipes), synchronised before returning.
static int render_device_single(plu_gpu_ctx* ctx, const plu_render_params* p, float* d_rgb, cudaStream_t st, plu_render_stats* stats) {
	int rc = check_params(ctx, p);
	if (rc != PLU_OK) return rc;
	if (!d_rgb) return fail(ctx, PLU_E_ARG, "framebuffer pointer is NULL");
	CU(cudaSetDevice(ctx->device));
	scene_ref sc = ctx->sc;
	select_bvh(ctx, sc, p->flags);
	wave_desc w = base_desc(nullptr, p, sc);
	const int64_t wh = (int64_t)p->width * p->height;
	const int32_t spp = p->n_strata * p->n_strata;
	const int32_t count = p->sample_count > 0 ? p->sample_count : spp;
	const float smp_wgt = 1.f / (float)(uint32_t)spp;   // renderer.cpp:100
	// Paths per wave: the caller's strata_per_wave, else 2^22 for the first waves and then a size DERIVED IN THIS CALL from
	// what the first wave did (no state is kept between renders): a wave is a chain of ~4 dependent launches per path vertex
	// whatever its size, so scenes whose paths are short (few rays per path: most launches of a wave nearly empty) run
	// larger waves — glass0 18.2 -> 22.5, room 8.5 -> 9.2 Gsamples/s at 2^24; cornell-box (5.9 rays per path) stays at 2^22.
	const bool adaptive = p->strata_per_wave <= 0 && std::getenv("PLU_NO_ADAPTIVE_WAVES") == nullptr;
	uint32_t want = p->strata_per_wave > 0 ? (uint32_t)std::min<int64_t>(wh * p->strata_per_wave, (int64_t)1 << 26) : PLU_WAVE_PATHS;
	want = (uint32_t)std::min<int64_t>(want, wh * (int64_t)count);
	want = std::max<uint32_t>(want, (uint32_t)std::min<int64_t>(wh, 1 << 22));
	// buffers are sized for the largest wave this call may choose, so that growing the waves never allocates mid-render
	uint32_t want_max = adaptive ? (uint32_t)std::max<int64_t>(want, std::min<int64_t>((int64_t)PLU_WAVE_PATHS_MAX, wh * (int64_t)count / 4)) : want;
	rc = ensure_wave(ctx, want_max);
	if (rc != PLU_OK) return rc;
	launch_cfg cfg = make_cfg(ctx, (p->flags & PLU_RENDER_COUNT_BVH) != 0, sc.fast_nodes != nullptr);
	cfg.no_pt = (p->flags & PLU_RENDER_NO_LANE_REFILL) != 0;
	w.use_dirty = wave_uses_pt(sc, cfg) ? 0 : 1;
	wave_profiler prof; prof.on = (p->flags & PLU_RENDER_PROFILE) != 0;
	const bool use_graph = !prof.on && std::getenv("PLU_NO_GRAPH") == nullptr;
	// Two waves in flight: even waves on the caller's stream, odd waves on ctx->pipe1 with the second buffer set.
	// Every kernel of a wave is a full-occupancy persistent grid whose CTAs end at different times; with two
	// independent waves the SMs one wave's kernel leaves idle at its tail (and between its launches) run the other
	// wave's kernels.  The framebuffer adds stay in wave order (k_accumulate waits for its predecessor's event),
	// so the image does not depend on the overlap.
	int64_t wcap = want;   // paths per wave (the buffers may be larger)
	const int64_t k_per_wave = std::max<int64_t>(1, wcap / std::min<int64_t>(wh, wcap));
	const int64_t waves_total = ((wh + wcap - 1) / wcap) * ((count + k_per_wave - 1) / k_per_wave);
	int NP = (prof.on || waves_total < 2) ? 1 : (int)std::min<int64_t>(ctx->n_pipes, waves_total);
	for (int k = 1; k < NP; ++k)
		if (ensure_wave(ctx, want_max, k) != PLU_OK) { cudaGetLastError(); NP = k; break; }   // no memory for another set: fewer waves in flight
	const bool two = NP > 1;
	rc = ensure_descs(ctx, (size_t)waves_total + (size_t)((p->flags & PLU_RENDER_COMPAT_OVERFLOW_COLUMN) ? (int64_t)count * ((p->height + wcap - 1) / wcap) : 0));
	if (rc != PLU_OK) return rc;
	for (int k = 0; k < NP; ++k) CU(cudaMemsetAsync(ctx->wbs[k].stats, 0, 8 * sizeof(unsigned long long), st));
	CU(cudaEventRecord(ctx->ev0, st));
	CU(cudaEventRecord(ctx->ev_start, st));
	for (int k = 0; k < NP; ++k) CU(cudaStreamWaitEvent(ctx->pipe[k], ctx->ev_start, 0));
	uint64_t launches = 0, samples = 0, waves = 0, first_wave_samples = 1;
	bool probe = adaptive && !prof.on && want_max > want && wh <= (int64_t)want;   // (images larger than a wave keep one size)
	int body_launches = 0;   // launches per round of the graph loop (rounds are counted on the device)
	// main image, then (compat) the overflow-column strip: sampler.h:74-78 renders x == W for rows 0..H-1
	for (int pass = 0; pass < ((p->flags & PLU_RENDER_COMPAT_OVERFLOW_COLUMN) ? 2 : 1); ++pass) {
		int64_t pbeg = pass == 0 ? 0 : wh, pend = pass == 0 ? wh : wh + p->height - 1;
		for (int64_t p0 = pbeg; p0 < pend;) {
			int64_t np = std::min<int64_t>(pend - p0, wcap);
			int32_t K = 0;
			for (int32_t k0 = 0; k0 < count; k0 += K) {
				if (probe && waves == (uint64_t)NP) {   // NP waves keep the device busy while the first one's counters are read
					probe = false;
					unsigned long long h0[8];
					CU(cudaEventSynchronize(ctx->ev_acc[0]));
					CU(cudaMemcpyAsync(h0, ctx->wbs[0].stats, sizeof(h0), cudaMemcpyDeviceToHost, ctx->stream));
					CU(cudaStreamSynchronize(ctx->stream));
					const double r = (double)(h0[0] + h0[1]) / (double)first_wave_samples;   // rays per path
					wcap = std::min<int64_t>(want_max, r < 2.0 ? (int64_t)want << 2 : (r < 3.5 ? (int64_t)want << 1 : (int64_t)want));
				}
				K = (int32_t)std::max<int64_t>(1, std::min<int64_t>(count, wcap / np));
				w.pix0 = (int32_t)p0; w.npix = (int32_t)np; w.k0 = k0; w.kcount = std::min(K, count - k0);
				w.n_slots = (uint32_t)(np * w.kcount); w.div_npix = set_div((uint32_t)np);
				const int P = (int)(waves % (uint64_t)NP);
				cudaStream_t sp = ctx->pipe[P];
				if (waves >= ctx->h_descs_cap) return fail(ctx, PLU_E_NOMEM, "wave descriptor array too small");
				ctx->h_descs[waves] = w;
				rc = put_desc(ctx, ctx->h_descs + waves, sp, P);
				if (rc != PLU_OK) return rc;
				plu_gpu_ctx::wave_graph* wg = use_graph ? wave_graph_for(ctx, sc, w.max_vertices, cfg, sp, P) : nullptr;
				if (wg) { CU(cudaGraphLaunch(wg->exec, sp)); launches += (uint64_t)wg->launches; body_launches = wg->body_launches; }
				else launches += (uint64_t)enqueue_wave(ctx, sc, w.max_vertices, cfg, sp, &prof, P);
				if (two && waves > 0) CU(cudaStreamWaitEvent(sp, ctx->ev_acc[(P + NP - 1) % NP], 0));   // the previous wave's add
				k_accumulate<<<cfg.g_acc, 256, 0, sp>>>(w, ctx->wbs[P], d_rgb, smp_wgt);
				CU(cudaEventRecord(ctx->ev_acc[P], sp));
				prof.mark(4, sp); prof.collect(sp);
				launches++;
				if (waves == 0) first_wave_samples = w.n_slots;
				samples += w.n_slots; waves++;
			}
			p0 += np;
		}
	}
	if (waves > 0) CU(cudaStreamWaitEvent(st, ctx->ev_acc[(waves - 1) % (uint64_t)NP], 0));   // the caller's stream resumes after the last add
	CU(cudaEventRecord(ctx->ev1, st));
	CU(cudaGetLastError());
	CU(cudaStreamSynchronize(st));
	{
		plu_render_stats local; if (!stats) stats = &local;
		std::memset(stats, 0, sizeof(*stats));
		unsigned long long h[8], h1[8];
		CU(cudaMemcpy(h, ctx->wbs[0].stats, sizeof(h), cudaMemcpyDeviceToHost));
		for (int q = 1; q < NP; ++q) { CU(cudaMemcpy(h1, ctx->wbs[q].stats, sizeof(h1), cudaMemcpyDeviceToHost)); for (int k = 0; k < 8; ++k) h[k] += h1[k]; }
		float ms = 0.f; CU(cudaEventElapsedTime(&ms, ctx->ev0, ctx->ev1));
		stats->samples = samples; stats->rays_extend = h[0]; stats->rays_shadow = h[1];
		stats->bvh_nodes_visited = h[2]; stats->prims_tested = h[3];
		stats->ms_device = ms; stats->kernel_launches = launches + h[6] * (uint64_t)body_launches;
		stats->shade_vertices = h[4]; stats->shadow_entries = h[5]; stats->waves = waves;
		stats->ms_generate = prof.ms[0]; stats->ms_extend = prof.ms[1]; stats->ms_shade = prof.ms[2];
		stats->ms_connect = prof.ms[3]; stats->ms_accumulate = prof.ms[4];
	}
	return PLU_OK;
}

static int ensure_fb(plu_gpu_ctx* ctx, size_t floats) {
	if (ctx->fb_floats >= floats) return PLU_OK;
	CU(cudaFree(ctx->d_fb)); ctx->d_fb = nullptr; ctx->fb_floats = 0; ctx->res_w = ctx->res_h = 0;
	CU(cudaMalloc(&ctx->d_fb, floats * sizeof(float))); ctx->fb_floats = floats;
	return PLU_OK;
}

// All devices of a multi-device context: device g renders the samples k = g (mod n) of the call's sample set into its
// own framebuffer (device 0: the caller's, which already holds the pixels to add to; the others: zeroed), one host
// thread per device enqueues that device's waves, then ONE ncclReduce(sum) to device 0 on the caller's stream.
static int render_device_multi(plu_gpu_ctx* ctx, const plu_render_params* p, float* d_rgb, cudaStream_t st, plu_render_stats* stats) {
	int rc = check_params(ctx, p);
	if (rc != PLU_OK) return rc;
	if (!d_rgb) return fail(ctx, PLU_E_ARG, "framebuffer pointer is NULL");
	const int n = 1 + (int)ctx->peers.size();
	const size_t floats = (size_t)p->width * p->height * 3;
	const int32_t spp = p->n_strata * p->n_strata;
	const int32_t begin = p->sample_count > 0 ? p->sample_begin : 0, stride = p->sample_count > 0 ? p->sample_stride : 1;
	const int32_t count = p->sample_count > 0 ? p->sample_count : spp;
	std::vector<plu_render_params> pg((size_t)n, *p);
	std::vector<plu_render_stats> sg((size_t)n);
	std::vector<int> rcs((size_t)n, PLU_OK);
	for (int g = 0; g < n; ++g) {
		pg[g].sample_begin = begin + g * stride; pg[g].sample_stride = stride * n;
		pg[g].sample_count = g < count ? (count - g + n - 1) / n : 0;
		std::memset(&sg[g], 0, sizeof(plu_render_stats));
	}
	auto run_peer = [&](int g) {
		plu_gpu_ctx* pc = ctx->peers[(size_t)g - 1];
		cudaError_t e = cudaSetDevice(pc->device);
		if (e == cudaSuccess && ensure_fb(pc, floats) != PLU_OK) { rcs[g] = PLU_E_CUDA; return; }
		if (e == cudaSuccess) e = cudaMemsetAsync(pc->d_fb, 0, floats * sizeof(float), pc->stream);
		if (e != cudaSuccess) { pc->err = std::string("framebuffer set-up: ") + cudaGetErrorString(e); rcs[g] = PLU_E_CUDA; return; }
		if (pg[g].sample_count > 0) rcs[g] = render_device_single(pc, &pg[g], pc->d_fb, pc->stream, &sg[g]);
		else if (cudaStreamSynchronize(pc->stream) != cudaSuccess) rcs[g] = PLU_E_CUDA;
	};
	std::vector<std::thread> th;
	for (int g = 1; g < n; ++g) th.emplace_back(run_peer, g);
	rcs[0] = render_device_single(ctx, &pg[0], d_rgb, st, &sg[0]);   // count >= 1, so device 0 always has samples
	for (auto& t : th) t.join();
	for (int g = 0; g < n; ++g)
		if (rcs[g] != PLU_OK) return g == 0 ? rcs[0] : fail(ctx, rcs[g], "device " + std::to_string(ctx->peers[(size_t)g - 1]->device) + ": " + ctx->peers[(size_t)g - 1]->err);
	// the single collective of the path: per-device framebuffers -> device 0
	nccl_api* nc = load_nccl();
	CU(cudaSetDevice(ctx->device));
	CU(cudaEventRecord(ctx->ev_r0, st));
	ncclResult_t nr = nc->GroupStart();
	for (int g = 0; g < n && nr == ncclSuccess; ++g) {
		plu_gpu_ctx* c = g == 0 ? ctx : ctx->peers[(size_t)g - 1];
		float* buf = g == 0 ? d_rgb : c->d_fb;
		nr = nc->Reduce(buf, buf, floats, ncclFloat32, ncclSum, 0, ctx->comms[(size_t)g], g == 0 ? st : c->stream);
	}
	ncclResult_t ne = nc->GroupEnd();
	if (nr == ncclSuccess) nr = ne;
	if (nr != ncclSuccess) return fail(ctx, PLU_E_CUDA, std::string("ncclReduce: ") + nc->GetErrorString(nr));
	CU(cudaSetDevice(ctx->device));
	CU(cudaEventRecord(ctx->ev_r1, st));
	CU(cudaStreamSynchronize(st));
	for (plu_gpu_ctx* pc : ctx->peers) { CU(cudaSetDevice(pc->device)); CU(cudaStreamSynchronize(pc->stream)); }
	CU(cudaSetDevice(ctx->device));
	float ms_r = 0.f; CU(cudaEventElapsedTime(&ms_r, ctx->ev_r0, ctx->ev_r1));
	if (stats) {
		plu_render_stats t = sg[0];
		for (int g = 1; g < n; ++g) {
			const plu_render_stats& a = sg[g];
			t.samples += a.samples; t.rays_extend += a.rays_extend; t.rays_shadow += a.rays_shadow;
			t.bvh_nodes_visited += a.bvh_nodes_visited; t.prims_tested += a.prims_tested;
			t.kernel_launches += a.kernel_launches; t.shade_vertices += a.shade_vertices; t.shadow_entries += a.shadow_entries;
			t.waves += a.waves;
			t.ms_device = std::max(t.ms_device, a.ms_device);
			t.ms_generate = std::max(t.ms_generate, a.ms_generate); t.ms_extend = std::max(t.ms_extend, a.ms_extend);
			t.ms_shade = std::max(t.ms_shade, a.ms_shade); t.ms_connect = std::max(t.ms_connect, a.ms_connect);
			t.ms_accumulate = std::max(t.ms_accumulate, a.ms_accumulate);
		}
		t.ms_reduce = ms_r; t.n_devices = n;
		*stats = t;
	}
	return PLU_OK;
}

extern "C" int plu_gpu_render_device(plu_gpu_ctx* ctx, const plu_render_params* p, float* d_rgb, void* stream, plu_render_stats* stats) {
	if (!ctx) return PLU_E_ARG;
	// NULL is stream 0, the legacy default stream (not the context's own non-blocking stream): whatever the caller
	// enqueued on stream 0 before this call — typically the framebuffer initialisation — is ordered before the render
	cudaStream_t st = stream ? (cudaStream_t)stream : cudaStreamLegacy;
	if (!ctx->peers.empty()) return render_device_multi(ctx, p, d_rgb, st, stats);
	int rc = render_device_single(ctx, p, d_rgb, st, stats);
	if (rc == PLU_OK && stats) { stats->ms_reduce = 0.0; stats->n_devices = 1; }
	return rc;
}

extern "C" int plu_gpu_render(plu_gpu_ctx* ctx, const plu_render_params* p, float* out_rgb, plu_render_stats* stats) {
	if (!ctx) return PLU_E_ARG;
	int rc = check_params(ctx, p);
	if (rc != PLU_OK) return rc;
	if (!out_rgb) return fail(ctx, PLU_E_ARG, "out_rgb is NULL");
	CU(cudaSetDevice(ctx->device));
	size_t floats = (size_t)p->width * p->height * 3;
	rc = ensure_fb(ctx, floats);
	if (rc != PLU_OK) return rc;
	ctx->res_w = ctx->res_h = 0;
	// `+=` semantics (renderer.cpp:132): the device buffer starts from the caller's pixels
	CU(cudaMemcpyAsync(ctx->d_fb, out_rgb, floats * sizeof(float), cudaMemcpyHostToDevice, ctx->stream));
	rc = plu_gpu_render_device(ctx, p, ctx->d_fb, ctx->stream, stats);
	if (rc != PLU_OK) return rc;
	CU(cudaMemcpyAsync(out_rgb, ctx->d_fb, floats * sizeof(float), cudaMemcpyDeviceToHost, ctx->stream));
	CU(cudaStreamSynchronize(ctx->stream));
	ctx->res_w = p->width; ctx->res_h = p->height;   // the image stays resident for plu_gpu_*_resident
	return PLU_OK;
}

// ---- parity hooks ---------------------------------------------------------------------------------------------
static int trace_common(plu_gpu_ctx* ctx, int64_t n, const float* o3, const float* d3, int pw, int ph, uint32_t flags,
                        int32_t* surf_id, float* t, float* normal3) {
	if (!ctx->have_scene) return fail(ctx, PLU_E_NOSCENE, "no scene uploaded");
	if (n < 0 || !surf_id || !t || !normal3) return fail(ctx, PLU_E_ARG, "bad trace arguments");
	if (n == 0) return PLU_OK;
	CU(cudaSetDevice(ctx->device));
	scene_ref sc = ctx->sc;
	select_bvh(ctx, sc, flags);
	float *d_o = nullptr, *d_d = nullptr, *d_t = nullptr, *d_n = nullptr; int32_t* d_id = nullptr;
	int rc = PLU_OK;
	auto cleanup = [&]() { void* b[] = {d_o, d_d, d_t, d_n, d_id}; for (void* q : b) if (q) cudaFreeAsync(q, ctx->stream); };
#define CUX(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) { ctx->err = std::string(#call) + ": " + cudaGetErrorString(e_); cleanup(); return PLU_E_CUDA; } } while (0)
	if (pw <= 0) {
		CUX(cudaMallocAsync(&d_o, (size_t)n * 12, ctx->stream)); CUX(cudaMallocAsync(&d_d, (size_t)n * 12, ctx->stream));
		CUX(cudaMemcpyAsync(d_o, o3, (size_t)n * 12, cudaMemcpyHostToDevice, ctx->stream));
		CUX(cudaMemcpyAsync(d_d, d3, (size_t)n * 12, cudaMemcpyHostToDevice, ctx->stream));
	}
	CUX(cudaMallocAsync(&d_id, (size_t)n * 4, ctx->stream)); CUX(cudaMallocAsync(&d_t, (size_t)n * 4, ctx->stream)); CUX(cudaMallocAsync(&d_n, (size_t)n * 12, ctx->stream));
	int grid = (int)std::min<int64_t>((n + 255) / 256, (int64_t)ctx->sm_count * 8);
	const bool staged = ctx->blob_in_smem && sc.fast_nodes == nullptr;   // the large-scene walks live in the global-memory kernels
	if (flags & PLU_RENDER_COUNT_BVH) {
		if (staged) k_trace_rays<true, true><<<grid, 256, sc.blob_bytes, ctx->stream>>>(sc, n, d_o, d_d, pw, ph, d_id, d_t, d_n);
		else k_trace_rays<false, true><<<grid, 256, 0, ctx->stream>>>(sc, n, d_o, d_d, pw, ph, d_id, d_t, d_n);
	} else if (staged) k_trace_rays<true><<<grid, 256, sc.blob_bytes, ctx->stream>>>(sc, n, d_o, d_d, pw, ph, d_id, d_t, d_n);
	else k_trace_rays<false><<<grid, 256, 0, ctx->stream>>>(sc, n, d_o, d_d, pw, ph, d_id, d_t, d_n);
	CUX(cudaGetLastError());
	CUX(cudaMemcpyAsync(surf_id, d_id, (size_t)n * 4, cudaMemcpyDeviceToHost, ctx->stream));
	CUX(cudaMemcpyAsync(t, d_t, (size_t)n * 4, cudaMemcpyDeviceToHost, ctx->stream));
	CUX(cudaMemcpyAsync(normal3, d_n, (size_t)n * 12, cudaMemcpyDeviceToHost, ctx->stream));
	CUX(cudaStreamSynchronize(ctx->stream));
#undef CUX
	cleanup();
	return rc;
}

extern "C" int plu_gpu_trace_primary(plu_gpu_ctx* ctx, int32_t width, int32_t height, uint32_t flags, int32_t* surf_id, float* t, float* normal3) {
	if (!ctx) return PLU_E_ARG;
	if (width <= 0 || height <= 0) return fail(ctx, PLU_E_ARG, "width and height must be positive");
	return trace_common(ctx, (int64_t)width * height, nullptr, nullptr, width, height, flags, surf_id, t, normal3);
}

extern "C" int plu_gpu_trace_rays(plu_gpu_ctx* ctx, int64_t n, const float* origin3, const float* dir3, uint32_t flags, int32_t* surf_id, float* t, float* normal3) {
	if (!ctx) return PLU_E_ARG;
	if (n > 0 && (!origin3 || !dir3)) return fail(ctx, PLU_E_ARG, "ray arrays are NULL");
	return trace_common(ctx, n, origin3, dir3, 0, 0, flags, surf_id, t, normal3);
}

extern "C" int plu_gpu_sample_radiance(plu_gpu_ctx* ctx, const plu_render_params* p, int64_t pixel_begin, int64_t pixel_count, float* out_rgb) {
	if (!ctx) return PLU_E_ARG;
	int rc = check_params(ctx, p);
	if (rc != PLU_OK) return rc;
	const int64_t wh = (int64_t)p->width * p->height;
	if (!out_rgb || pixel_begin < 0 || pixel_count < 0 || pixel_begin + pixel_count > wh + p->height) return fail(ctx, PLU_E_ARG, "bad pixel range");
	if (pixel_count == 0) return PLU_OK;
	CU(cudaSetDevice(ctx->device));
	scene_ref sc = ctx->sc;
	select_bvh(ctx, sc, p->flags);
	wave_desc w = base_desc(nullptr, p, sc);
	const int32_t count = p->sample_count > 0 ? p->sample_count : p->n_strata * p->n_strata;
	rc = ensure_wave(ctx, (uint32_t)std::min<int64_t>((int64_t)1 << 22, std::max<int64_t>(pixel_count, 1) * count));
	if (rc != PLU_OK) return rc;
	launch_cfg cfg = make_cfg(ctx, false, sc.fast_nodes != nullptr);
	w.use_dirty = wave_uses_pt(sc, cfg) ? 0 : 1;
	std::vector<float4> host; std::vector<uint32_t> dirty;
	for (int64_t p0 = pixel_begin; p0 < pixel_begin + pixel_count;) {
		int64_t np = std::min<int64_t>(pixel_begin + pixel_count - p0, ctx->cap[0]);
		int32_t K = (int32_t)std::max<int64_t>(1, std::min<int64_t>(count, ctx->cap[0] / np));
		for (int32_t k0 = 0; k0 < count; k0 += K) {
			w.pix0 = (int32_t)p0; w.npix = (int32_t)np; w.k0 = k0; w.kcount = std::min(K, count - k0);
			w.n_slots = (uint32_t)(np * w.kcount); w.div_npix = set_div((uint32_t)np);
			rc = put_desc(ctx, &w, ctx->stream, 0);   // pageable source: staged before the call returns
			if (rc != PLU_OK) return rc;
			enqueue_wave(ctx, sc, w.max_vertices, cfg, ctx->stream);
			CU(cudaGetLastError());
			host.resize(w.n_slots);
			CU(cudaMemcpyAsync(host.data(), ctx->wbs[0].rad, (size_t)w.n_slots * sizeof(float4), cudaMemcpyDeviceToHost, ctx->stream));
			if (w.use_dirty) {   // slots whose camera ray missed were never written: L == 0
				dirty.resize(((size_t)w.n_slots + 31) / 32);
				CU(cudaMemcpyAsync(dirty.data(), ctx->wbs[0].dirty, dirty.size() * 4, cudaMemcpyDeviceToHost, ctx->stream));
			}
			CU(cudaStreamSynchronize(ctx->stream));
			if (w.use_dirty) for (size_t i = 0; i < (size_t)w.n_slots; ++i) if (!((dirty[i >> 5] >> (i & 31)) & 1u)) host[i] = make_float4(0.f, 0.f, 0.f, 0.f);
			for (int32_t k = 0; k < w.kcount; ++k)
				for (int64_t q = 0; q < np; ++q) {
					const float4& c = host[(size_t)k * (size_t)np + (size_t)q];
					float* o = out_rgb + ((size_t)(p0 - pixel_begin + q) * (size_t)count + (size_t)(k0 + k)) * 3;
					o[0] = c.x; o[1] = c.y; o[2] = c.z;
				}
		}
		p0 += np;
	}
	return PLU_OK;
}

// ---- post-process + image write (main.cpp:183, :209) -------------------------------------------------------------
extern "C" int64_t plu_bmp_size(int32_t w, int32_t h) {
	if (w <= 0 || h <= 0) return 0;
	return 54 + ((int64_t)w * 3 + ((-w * 3) & 3)) * h;
}

static void bmp_header(int32_t W, int32_t H, unsigned char* o) {   // stbi_write_bmp, texture.cpp:167-173
	std::memset(o, 0, 54);
	auto put32 = [&](int at, uint32_t v) { o[at] = v & 255; o[at + 1] = (v >> 8) & 255; o[at + 2] = (v >> 16) & 255; o[at + 3] = (v >> 24) & 255; };
	o[0] = 'B'; o[1] = 'M'; put32(2, (uint32_t)plu_bmp_size(W, H)); put32(10, 14 + 40);
	put32(14, 40); put32(18, (uint32_t)W); put32(22, (uint32_t)H); o[26] = 1; o[28] = 24;
}

extern "C" int plu_gpu_postprocess(plu_gpu_ctx* ctx, int32_t w, int32_t h, float* rgb) {
	if (!ctx) return PLU_E_ARG;
	if (w <= 0 || h <= 0 || !rgb) return fail(ctx, PLU_E_ARG, "bad image arguments");
	CU(cudaSetDevice(ctx->device));
	size_t floats = (size_t)w * h * 3;
	int rc = ensure_fb(ctx, floats);
	if (rc != PLU_OK) return rc;
	ctx->res_w = ctx->res_h = 0;
	CU(cudaMemcpyAsync(ctx->d_fb, rgb, floats * sizeof(float), cudaMemcpyHostToDevice, ctx->stream));
	ctx->res_w = w; ctx->res_h = h;
	return plu_gpu_postprocess_resident(ctx, rgb);
}

extern "C" int plu_gpu_postprocess_resident(plu_gpu_ctx* ctx, float* rgb_out) {
	if (!ctx) return PLU_E_ARG;
	if (ctx->res_w <= 0 || ctx->res_h <= 0 || !ctx->d_fb) return fail(ctx, PLU_E_NOSCENE, "no resident image: plu_gpu_render has not run on this context");
	CU(cudaSetDevice(ctx->device));
	const int64_t n = (int64_t)ctx->res_w * ctx->res_h;
	int grid = (int)std::min<int64_t>((n + 255) / 256, (int64_t)ctx->sm_count * 8);
	k_tonemap<<<grid, 256, 0, ctx->stream>>>(n, ctx->d_fb);
	CU(cudaGetLastError());
	if (rgb_out) CU(cudaMemcpyAsync(rgb_out, ctx->d_fb, (size_t)n * 3 * sizeof(float), cudaMemcpyDeviceToHost, ctx->stream));
	CU(cudaStreamSynchronize(ctx->stream));
	return PLU_OK;
}

extern "C" int plu_gpu_encode_bmp_resident(plu_gpu_ctx* ctx, uint32_t flags, unsigned char* out, int64_t cap) {
	if (!ctx) return PLU_E_ARG;
	if (ctx->res_w <= 0 || ctx->res_h <= 0 || !ctx->d_fb) return fail(ctx, PLU_E_NOSCENE, "no resident image: plu_gpu_render has not run on this context");
	return plu_gpu_encode_bmp_device(ctx, ctx->res_w, ctx->res_h, ctx->d_fb, flags, ctx->stream, out, cap);
}

extern "C" int plu_gpu_encode_bmp_device(plu_gpu_ctx* ctx, int32_t w, int32_t h, const float* d_rgb, uint32_t flags, void* stream,
                                         unsigned char* out, int64_t cap) {
	if (!ctx) return PLU_E_ARG;
	if (w <= 0 || h <= 0 || !d_rgb || !out) return fail(ctx, PLU_E_ARG, "bad image arguments");
	const int64_t total = plu_bmp_size(w, h);
	if (cap < total) return fail(ctx, PLU_E_ARG, "output buffer smaller than plu_bmp_size");
	CU(cudaSetDevice(ctx->device));
	cudaStream_t st = stream ? (cudaStream_t)stream : cudaStreamLegacy;
	unsigned char* d_body = nullptr;
	CU(cudaMalloc(&d_body, (size_t)(total - 54)));
	int64_t n = (int64_t)w * h;
	int grid = (int)std::min<int64_t>((n + 255) / 256, (int64_t)ctx->sm_count * 8);
	k_encode_bmp<<<grid, 256, 0, st>>>(w, h, d_rgb, (flags & PLU_POST_TONEMAP) ? 1 : 0, d_body);
	cudaError_t e = cudaGetLastError();
	if (e == cudaSuccess) e = cudaMemcpyAsync(out + 54, d_body, (size_t)(total - 54), cudaMemcpyDeviceToHost, st);
	if (e == cudaSuccess) e = cudaStreamSynchronize(st);
	cudaFree(d_body);
	if (e != cudaSuccess) return fail(ctx, PLU_E_CUDA, std::string("bmp encode: ") + cudaGetErrorString(e));
	bmp_header(w, h, out);
	return PLU_OK;
}

// ---- plu_gpu_encode_bmp_device without any allocation (include/plugpu.h: PLU_NOALLOC_BMP_MAX_BODY) ----------------------
// The .bmp body of plu_gpu_encode_bmp_device goes through a scratch buffer that call cudaMallocs and frees; callers that
// must not allocate get the same kernel on a FIXED scratch: a file-scope device array of a stated maximum size.  Every GPU
// that loads the library holds these bytes until the process exits.  `g_noalloc_bmp_body` below is, in host code, only a
// host-side stand-in for the array (not a device address): its device address is taken ONCE PER DEVICE, with that device
// current, by cudaGetSymbolAddress, and that pointer is what the kernel receives.
static __device__ __align__(256) unsigned char g_noalloc_bmp_body[PLU_NOALLOC_BMP_MAX_BODY];

#define PLU_NOALLOC_MAX_DEVICES 64
static int noalloc_bmp_scratch(plu_gpu_ctx* ctx, unsigned char** out) {   // caller has made ctx->device current
	static std::mutex mu;
	static unsigned char* ptr[PLU_NOALLOC_MAX_DEVICES] = {};   // per device: the array is one object per GPU
	if (ctx->device < 0 || ctx->device >= PLU_NOALLOC_MAX_DEVICES) return fail(ctx, PLU_E_ARG, "device index beyond the no-allocation scratch table");
	std::lock_guard<std::mutex> lock(mu);
	if (!ptr[ctx->device]) {
		void* q = nullptr;
		cudaError_t e = cudaGetSymbolAddress(&q, g_noalloc_bmp_body);
		if (e != cudaSuccess || !q) return fail(ctx, PLU_E_CUDA, std::string("cudaGetSymbolAddress(g_noalloc_bmp_body): ") + cudaGetErrorString(e));
		ptr[ctx->device] = (unsigned char*)q;
	}
	*out = ptr[ctx->device];
	return PLU_OK;
}

extern "C" int plu_gpu_encode_bmp_device_noalloc(plu_gpu_ctx* ctx, int32_t w, int32_t h, const float* d_rgb, uint32_t flags, void* stream,
                                                 unsigned char* out, int64_t cap) {
	if (!ctx) return PLU_E_ARG;
	if (w <= 0 || h <= 0 || !d_rgb || !out) return fail(ctx, PLU_E_ARG, "bad image arguments");
	const int64_t total = plu_bmp_size(w, h);
	if (cap < total) return fail(ctx, PLU_E_ARG, "output buffer smaller than plu_bmp_size");
	if (total - 54 > (int64_t)PLU_NOALLOC_BMP_MAX_BODY)
		return fail(ctx, PLU_E_ARG, "image too large for the no-allocation entry point: its .bmp body (" + std::to_string(total - 54) +
		            " bytes) exceeds PLU_NOALLOC_BMP_MAX_BODY (" + std::to_string((long long)PLU_NOALLOC_BMP_MAX_BODY) + "); use plu_gpu_encode_bmp_device");
	CU(cudaSetDevice(ctx->device));
	unsigned char* d_body = nullptr;
	int rc = noalloc_bmp_scratch(ctx, &d_body);
	if (rc != PLU_OK) return rc;
	cudaStream_t st = stream ? (cudaStream_t)stream : cudaStreamLegacy;
	const int64_t n = (int64_t)w * h;
	const int grid = (int)std::min<int64_t>((n + 255) / 256, (int64_t)ctx->sm_count * 8);
	// the array keeps the previous call's bytes; nothing has to start at zero: k_encode_bmp writes EVERY byte of the
	// total - 54 it is given (pixels and the zero padding of each row), and only those are copied out
	k_encode_bmp<<<grid, 256, 0, st>>>(w, h, d_rgb, (flags & PLU_POST_TONEMAP) ? 1 : 0, d_body);
	cudaError_t e = cudaGetLastError();
	if (e == cudaSuccess) e = cudaMemcpyAsync(out + 54, d_body, (size_t)(total - 54), cudaMemcpyDeviceToHost, st);
	if (e == cudaSuccess) e = cudaStreamSynchronize(st);
	if (e != cudaSuccess) return fail(ctx, PLU_E_CUDA, std::string("bmp encode (no allocation): ") + cudaGetErrorString(e));
	bmp_header(w, h, out);
	return PLU_OK;
}

extern "C" int plu_gpu_encode_bmp(plu_gpu_ctx* ctx, int32_t w, int32_t h, const float* rgb, uint32_t flags, unsigned char* out, int64_t cap) {
	if (!ctx) return PLU_E_ARG;
	if (w <= 0 || h <= 0 || !rgb || !out) return fail(ctx, PLU_E_ARG, "bad image arguments");
	CU(cudaSetDevice(ctx->device));
	size_t floats = (size_t)w * h * 3;
	int rc = ensure_fb(ctx, floats);
	if (rc != PLU_OK) return rc;
	ctx->res_w = ctx->res_h = 0;
	CU(cudaMemcpyAsync(ctx->d_fb, rgb, floats * sizeof(float), cudaMemcpyHostToDevice, ctx->stream));
	ctx->res_w = w; ctx->res_h = h;
	return plu_gpu_encode_bmp_device(ctx, w, h, ctx->d_fb, flags, ctx->stream, out, cap);
}

// ---- host-side helpers ------------------------------------------------------------------------------------------
extern "C" int plu_scene_reference_leaf_order(const plu_scene_desc* desc, int32_t* rank) {
	if (!rank || !pluhost::validate_desc(desc).empty()) return PLU_E_ARG;
	std::vector<int32_t> r;
	pluhost::reference_leaf_order(desc, r);
	std::copy(r.begin(), r.end(), rank);
	return PLU_OK;
}

// ---- page-locked host images (plugpu.h: plu_host_*) ---------------------------------------------------------------
// No context: the memory is pinned for every device (cudaHostAllocPortable / cudaHostRegisterPortable), so the same image can
// be handed to a single- or a multi-device context.  cudaGetLastError() clears the sticky error a refused call leaves.
extern "C" int plu_host_alloc(size_t bytes, void** out) {
	if (!out) return PLU_E_ARG;
	*out = nullptr;
	if (bytes == 0) return PLU_E_ARG;
	void* p = nullptr;
	if (cudaHostAlloc(&p, bytes, cudaHostAllocPortable) != cudaSuccess) { cudaGetLastError(); return PLU_E_CUDA; }
	std::memset(p, 0, bytes);
	*out = p;
	return PLU_OK;
}
extern "C" int plu_host_free(void* p) {
	if (!p) return PLU_OK;
	if (cudaFreeHost(p) != cudaSuccess) { cudaGetLastError(); return PLU_E_CUDA; }
	return PLU_OK;
}
extern "C" int plu_host_register(void* p, size_t bytes) {
	if (!p || bytes == 0) return PLU_E_ARG;
	if (cudaHostRegister(p, bytes, cudaHostRegisterPortable) != cudaSuccess) { cudaGetLastError(); return PLU_E_CUDA; }
	return PLU_OK;
}
extern "C" int plu_host_unregister(void* p) {
	if (!p) return PLU_E_ARG;
	if (cudaHostUnregister(p) != cudaSuccess) { cudaGetLastError(); return PLU_E_CUDA; }
	return PLU_OK;
}
