// bvh_gpu.cu — both traversal structures built on the device, level by level.
//
// 1. Reference leaf order.  bvh_tree's constructor (reference src/surfaces/bvh_tree.cpp:7-36) sorts its
//    surfaces by box centre along a round-robin axis, splits them n/2 | n - n/2 and recurses until one or
//    two are left.  All nodes of a depth share the axis and the segment boundaries depend on n alone, so a
//    depth is ONE stable radix sort of (segment path | ordered centre bits) over all surfaces.  std::sort
//    is an insertion sort (stable) up to 16 elements and unique without ties, so the result is the host's
//    permutation except where a segment of more than 16 surfaces holds equal centres; those segments are
//    detected here and the host's std::sort decides them (counted in plu_build_stats.host_tie_segments).
// 2. Reference nodes.  The tree's shape is a function of n: a thread finds a node's pre-order index by
//    walking the halving recursion, leaves are written first, interior bounds (origin-inflated unions,
//    cmmn.h:78-85) bottom-up, one launch per depth.
// 3. Fast BVH.  Binned SAH (16 bins x 3 axes) restating host/bvh_build.cpp decision for decision: segments
//    above 1024 surfaces are processed element-parallel (CTA-private bins in shared memory, one global
//    scan for the stable partition), smaller ones by one warp each entirely in shared memory.  Nodes are
//    numbered breadth-first while building and put into the host builder's pre-order by one sort on
//    (first leaf slot, depth).
#include "host/bvh_gpu.h"
#include "host/bvh_build.h"
#include "device/scene_dev.cuh"
#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_scan.cuh>
#include <algorithm>
#include <atomic>
#include <chrono>
#include <thread>
#include <vector>

namespace plubuild {
namespace {

using plud::fast_node;
using plud::ref_node;

struct box3 { float mn[3], mx[3]; };

// float <-> unsigned with the same order (so atomicMin / atomicMax and radix sort apply)
__host__ __device__ inline uint32_t f2o(float f) {
	uint32_t u;
#ifdef __CUDA_ARCH__
	u = __float_as_uint(f);
#else
	memcpy(&u, &f, 4);
#endif
	return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__device__ __forceinline__ float o2f(uint32_t o) { return __uint_as_float((o & 0x80000000u) ? (o ^ 0x80000000u) : ~o); }
#define O_POS_INF 0xff800000u   // f2o(+inf)
#define O_NEG_INF 0x007fffffu   // f2o(-inf)

__device__ __forceinline__ void add_point(box3& b, const float* p) {   // aabb::add_point, cmmn.h:88-97
	for (int i = 0; i < 3; ++i) { if (p[i] > b.mx[i]) b.mx[i] = p[i]; if (p[i] < b.mn[i]) b.mn[i] = p[i]; }
}
__device__ __forceinline__ box3 origin_box() { box3 b; for (int i = 0; i < 3; ++i) b.mn[i] = b.mx[i] = 0.f; return b; }
// surface::bounds(): sphere.h:11-13, box.h:11-13, triangle.h:20-24 (a box that starts at the origin)
__device__ __forceinline__ box3 surface_bounds(const float* f, int tag) {
	box3 b;
	if (tag == PLU_PRIM_SPHERE) { for (int i = 0; i < 3; ++i) { b.mn[i] = f[i] - f[3]; b.mx[i] = f[i] + f[3]; } }
	else if (tag == PLU_PRIM_BOX) { for (int i = 0; i < 3; ++i) { b.mn[i] = f[i]; b.mx[i] = f[4 + i]; } }
	else { b = origin_box(); add_point(b, f); add_point(b, f + 4); add_point(b, f + 8); }
	return b;
}
__device__ __forceinline__ box3 tight_bounds(const float* f, int tag) {
	if (tag != PLU_PRIM_TRIANGLE) return surface_bounds(f, tag);
	box3 b;
	for (int i = 0; i < 3; ++i) { b.mn[i] = fminf(f[i], fminf(f[4 + i], f[8 + i])); b.mx[i] = fmaxf(f[i], fmaxf(f[4 + i], f[8 + i])); }
	return b;
}

// ---- the shape of the median-split tree ---------------------------------------------------------------
// Segment of position i at depth L: walk n -> n/2 | n - n/2 (bvh_tree.cpp:19-32); recursion ends at <= 2.
struct ref_seg { uint32_t begin, size, path; };
__host__ __device__ inline ref_seg ref_segment(uint32_t n, uint32_t i, int L) {
	ref_seg s{0u, n, 0u};
	int l = 0;
	for (; l < L && s.size > 2; ++l) {
		uint32_t mid = s.size / 2;
		if (i - s.begin < mid) { s.size = mid; s.path <<= 1; }
		else { s.begin += mid; s.size -= mid; s.path = (s.path << 1) | 1u; }
	}
	s.path <<= (L - l);
	return s;
}

__global__ void k_ref_centres(int32_t n, const float* prim, const int32_t* meta, float* cen) {
	int32_t s = blockIdx.x * blockDim.x + threadIdx.x;
	if (s >= n) return;
	box3 b = surface_bounds(prim + (size_t)s * 12, meta[4 * s]);
	for (int a = 0; a < 3; ++a) {
		float c = (b.mn[a] + b.mx[a]) * 1.f / 2.f;   // aabb::center, cmmn.h:204-207
		cen[(size_t)a * n + s] = c == 0.f ? 0.f : c;   // -0 and +0 compare equal on the host
	}
}

__global__ void k_iota(int32_t n, int32_t* v) { int32_t i = blockIdx.x * blockDim.x + threadIdx.x; if (i < n) v[i] = i; }

__global__ void k_ref_keys(int32_t n, int L, const float* cen_axis, const int32_t* order, unsigned long long* keys) {
	int32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n) return;
	ref_seg s = ref_segment((uint32_t)n, (uint32_t)i, L);
	uint32_t lo = s.size > 2 ? f2o(cen_axis[order[i]]) : 0u;   // finished segments keep their order
	keys[i] = ((unsigned long long)s.path << 32) | lo;
}

// equal neighbours inside a segment of more than 16: std::sort's introsort decides their order
__global__ void k_ref_ties(int32_t n, int L, const unsigned long long* keys, uint32_t* count, uint32_t* list, uint32_t cap) {
	int32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i + 1 >= n || keys[i] != keys[i + 1]) return;
	ref_seg s = ref_segment((uint32_t)n, (uint32_t)i, L);
	if (s.size <= 16) return;
	uint32_t k = atomicAdd(count, 1u);
	if (k < cap) list[k] = s.begin;
}

__global__ void k_rank_from_order(int32_t n, const int32_t* order, int32_t* meta) {
	int32_t k = blockIdx.x * blockDim.x + threadIdx.x;
	if (k < n) meta[4 * order[k] + 3] = k;
}
__global__ void k_order_from_rank(int32_t n, const int32_t* meta, int32_t* order) {
	int32_t s = blockIdx.x * blockDim.x + threadIdx.x;
	if (s < n) order[meta[4 * s + 3]] = s;
}

// ---- reference nodes ------------------------------------------------------------------------------------
// pre-order: a subtree over `size` leaves has 2*size-1 nodes; left child = me+1, right = me + 2*mid
__global__ void k_ref_leaves(int32_t n, const int32_t* order, const float* prim, const int32_t* meta, ref_node* nodes) {
	int32_t k = blockIdx.x * blockDim.x + threadIdx.x;
	if (k >= n) return;
	uint32_t b = 0, sz = (uint32_t)n, me = 0;
	while (sz > 1) {
		uint32_t mid = sz / 2;
		if ((uint32_t)k - b < mid) { me += 1; sz = mid; } else { me += 2 * mid; b += mid; sz -= mid; }
	}
	int32_t s = order[k], tag = meta[4 * s];
	const float* f = prim + (size_t)s * 12;
	box3 bb = surface_bounds(f, tag);
	// leaf bounds are never tested (bvh_tree.cpp:40-42): a sphere leaf carries (centre | radius) instead
	if (tag == PLU_PRIM_SPHERE) for (int i = 0; i < 3; ++i) { bb.mn[i] = f[i]; bb.mx[i] = f[3]; }
	nodes[me].lo = make_float4(bb.mn[0], bb.mn[1], bb.mn[2], __int_as_float(tag));
	nodes[me].hi = make_float4(bb.mx[0], bb.mx[1], bb.mx[2], __int_as_float(s));
}

__device__ __forceinline__ box3 ref_child_box(const ref_node* nodes, uint32_t node, uint32_t pos, uint32_t size,
                                               const int32_t* order, const float* prim, const int32_t* meta) {
	if (size == 1) { int32_t s = order[pos]; return surface_bounds(prim + (size_t)s * 12, meta[4 * s]); }
	box3 b; float4 lo = nodes[node].lo, hi = nodes[node].hi;
	b.mn[0] = lo.x; b.mn[1] = lo.y; b.mn[2] = lo.z; b.mx[0] = hi.x; b.mx[1] = hi.y; b.mx[2] = hi.z;
	return b;
}

__global__ void k_ref_interior(int32_t n, int L, const int32_t* order, const float* prim, const int32_t* meta, ref_node* nodes) {
	uint32_t p = blockIdx.x * blockDim.x + threadIdx.x;
	if (p >= (1u << L)) return;
	uint32_t b = 0, sz = (uint32_t)n, me = 0;
	for (int l = L - 1; l >= 0; --l) {
		if (sz < 2) return;
		uint32_t mid = sz / 2;
		if ((p >> l) & 1u) { me += 2 * mid; b += mid; sz -= mid; } else { me += 1; sz = mid; }
	}
	if (sz < 2) return;
	uint32_t mid = sz / 2;
	box3 l = ref_child_box(nodes, me + 1, b, mid, order, prim, meta);
	box3 r = ref_child_box(nodes, me + 2 * mid, b + mid, sz - mid, order, prim, meta);
	box3 u = origin_box();   // aabb(const aabb&, const aabb&), cmmn.h:78-85
	add_point(u, l.mn); add_point(u, l.mx); add_point(u, r.mn); add_point(u, r.mx);
	nodes[me].lo = make_float4(u.mn[0], u.mn[1], u.mn[2], __int_as_float((int32_t)(me + 2 * sz - 1)));
	nodes[me].hi = make_float4(u.mx[0], u.mx[1], u.mx[2], __int_as_float(-1));
}

inline int blocks(size_t n, int t = 256) { return (int)((n + (size_t)t - 1) / (size_t)t); }

struct kv { float k; int32_t v; };

// Scratch of the build: stream-ordered allocations (cudaMallocAsync on the build's stream).  cudaMalloc / cudaFree are
// device-wide synchronisation points and, in a process that drives several GPUs with peer access enabled (the
// multi-device context: NCCL), each one has to be mapped on every peer — dozens of them per build cost more than the
// build itself.  The pool keeps the memory across builds (plugpu.cu sets the release threshold).
struct dev_mem {   // frees everything it handed out
	std::vector<void*> p; cudaStream_t st;
	explicit dev_mem(cudaStream_t s) : st(s) {}
	template <class T> T* get(size_t count) { void* q = nullptr; if (cudaMallocAsync(&q, std::max<size_t>(count, 1) * sizeof(T), st) != cudaSuccess) return nullptr; p.push_back(q); return (T*)q; }
	~dev_mem() { for (void* q : p) cudaFreeAsync(q, st); }
};

// host comparator of bvh_tree.cpp:23-27 for the tie fix-up
struct host_centre {
	const plu_scene_desc* d; int axis;
	float operator()(int32_t s) const {
		const float* f = d->prim_data + (size_t)s * 12;
		float mn, mx;
		if (d->prim_tag[s] == PLU_PRIM_SPHERE) { mn = f[axis] - f[3]; mx = f[axis] + f[3]; }
		else if (d->prim_tag[s] == PLU_PRIM_BOX) { mn = f[axis]; mx = f[4 + axis]; }
		else { mn = mx = 0.f; for (int v = 0; v < 3; ++v) { float p = f[4 * v + axis]; if (p > mx) mx = p; if (p < mn) mn = p; } }
		return (mn + mx) * 1.f / 2.f;
	}
};

#define CK(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) { err = std::string(#call) + ": " + cudaGetErrorString(e_); return false; } } while (0)

double now_ms() { return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now().time_since_epoch()).count(); }

// Part 1: the leaf order of the reference tree, into meta[].w (tie rank) and `order` (rank -> surface)
bool ref_order_on_device(const plu_scene_desc* d, device_tables& t, int32_t* d_order, cudaStream_t st, plu_build_stats* stats, std::string& err) {
	const int32_t n = d->n_prims;
	dev_mem mem(st);
	float* cen = mem.get<float>(3 * (size_t)n);
	unsigned long long* keys = mem.get<unsigned long long>(n); unsigned long long* keys2 = mem.get<unsigned long long>(n);
	int32_t* order2 = mem.get<int32_t>(n);
	const uint32_t cap = 1u << 16;
	uint32_t* ties = mem.get<uint32_t>(cap + 1);
	if (!cen || !keys || !keys2 || !order2 || !ties) { err = "cudaMalloc (reference order)"; return false; }
	size_t tmp_bytes = 0;
	CK(cub::DeviceRadixSort::SortPairs(nullptr, tmp_bytes, keys, keys2, d_order, order2, n, 0, 64, st));
	unsigned char* tmp = mem.get<unsigned char>(tmp_bytes);
	if (!tmp) { err = "cudaMalloc (sort scratch)"; return false; }
	k_ref_centres<<<blocks(n), 256, 0, st>>>(n, t.d_prim, t.d_prim_meta, cen);
	k_iota<<<blocks(n), 256, 0, st>>>(n, d_order);
	int32_t* cur = d_order; int32_t* nxt = order2;
	std::vector<float> hc[3]; std::vector<int32_t> h_cur, h_nxt;   // host side of the tie fix-up, filled on demand
	int L = 0;
	for (uint32_t widest = (uint32_t)n; widest > 2; widest -= widest / 2, ++L) {
		const int axis = L % 3;
		k_ref_keys<<<blocks(n), 256, 0, st>>>(n, L, cen + (size_t)axis * n, cur, keys);
		CK(cub::DeviceRadixSort::SortPairs(tmp, tmp_bytes, keys, keys2, cur, nxt, n, 0, 32 + L, st));
		if (widest > 16) {   // only then can a segment of this depth exceed 16
			CK(cudaMemsetAsync(ties, 0, sizeof(uint32_t), st));
			k_ref_ties<<<blocks(n), 256, 0, st>>>(n, L, keys2, ties, ties + 1, cap);
			uint32_t n_ties = 0;
			CK(cudaMemcpyAsync(&n_ties, ties, sizeof(uint32_t), cudaMemcpyDeviceToHost, st));
			CK(cudaStreamSynchronize(st));
			if (n_ties) {   // the host's std::sort on the segment as it was BEFORE this depth's sort
				std::vector<uint32_t> begins;
				if (n_ties <= cap) {
					begins.resize(n_ties);
					CK(cudaMemcpyAsync(begins.data(), ties + 1, n_ties * sizeof(uint32_t), cudaMemcpyDeviceToHost, st));
					CK(cudaStreamSynchronize(st));
					std::sort(begins.begin(), begins.end());
					begins.erase(std::unique(begins.begin(), begins.end()), begins.end());
				} else {   // too many to list: every segment of this depth
					for (uint32_t i = 0; i < (uint32_t)n;) { ref_seg s = ref_segment((uint32_t)n, i, L); if (s.size > 16) begins.push_back(s.begin); i = s.begin + s.size; }
				}
				// same std::sort, same comparator outcomes, same input order as bvh_tree.cpp:23-27 => same permutation;
				// (centre, id) pairs instead of ids keep the comparisons in cache, segments are sorted in parallel
				if (hc[axis].empty()) { hc[axis].resize((size_t)n); host_centre centre{d, axis}; for (int32_t s = 0; s < n; ++s) hc[axis][(size_t)s] = centre(s); }
				h_cur.resize((size_t)n); h_nxt.resize((size_t)n);
				// every copy is ordered on `st` (a non-blocking stream does not wait for the legacy default stream)
				CK(cudaMemcpyAsync(h_cur.data(), cur, (size_t)n * sizeof(int32_t), cudaMemcpyDeviceToHost, st));
				CK(cudaMemcpyAsync(h_nxt.data(), nxt, (size_t)n * sizeof(int32_t), cudaMemcpyDeviceToHost, st));
				CK(cudaStreamSynchronize(st));
				const std::vector<float>& key = hc[axis];
				std::atomic<size_t> next_seg{0};
				auto worker = [&]() {
					std::vector<kv> seg;
					for (size_t w = next_seg++; w < begins.size(); w = next_seg++) {
						ref_seg s = ref_segment((uint32_t)n, begins[w], L);
						seg.resize(s.size);
						for (uint32_t i = 0; i < s.size; ++i) { int32_t id = h_cur[s.begin + i]; seg[i] = kv{key[(size_t)id], id}; }
						std::sort(seg.begin(), seg.end(), [](const kv& a, const kv& c) { return a.k < c.k; });
						for (uint32_t i = 0; i < s.size; ++i) h_nxt[s.begin + i] = seg[i].v;
					}
				};
				size_t n_threads = std::min<size_t>(begins.size(), std::max(1u, std::min(32u, std::thread::hardware_concurrency())));
				std::vector<std::thread> pool;
				for (size_t k = 1; k < n_threads; ++k) pool.emplace_back(worker);
				worker();
				for (auto& th : pool) th.join();
				CK(cudaMemcpyAsync(nxt, h_nxt.data(), (size_t)n * sizeof(int32_t), cudaMemcpyHostToDevice, st));
				CK(cudaStreamSynchronize(st));
				if (stats) stats->host_tie_segments += (int32_t)begins.size();
			}
		}
		std::swap(cur, nxt);
	}
	if (cur != d_order) CK(cudaMemcpyAsync(d_order, cur, (size_t)n * sizeof(int32_t), cudaMemcpyDeviceToDevice, st));
	k_rank_from_order<<<blocks(n), 256, 0, st>>>(n, d_order, t.d_prim_meta);
	CK(cudaStreamSynchronize(st));
	if (stats) stats->levels_ref = L;
	return true;
}

// Part 2: the reference tree in pre-order
bool ref_nodes_on_device(int32_t n, device_tables& t, const int32_t* d_order, cudaStream_t st, std::string& err) {
	ref_node* nodes = (ref_node*)t.d_ref_nodes;
	k_ref_leaves<<<blocks(n), 256, 0, st>>>(n, d_order, t.d_prim, t.d_prim_meta, nodes);
	int deepest = -1;   // last depth that still holds a segment of two or more
	for (uint32_t widest = (uint32_t)n; widest >= 2; widest -= widest / 2) deepest++;
	for (int L = deepest; L >= 0; --L)
		k_ref_interior<<<blocks((size_t)1 << L), 256, 0, st>>>(n, L, d_order, t.d_prim, t.d_prim_meta, nodes);
	CK(cudaGetLastError());
	return true;
}

// ---- fast BVH: binned SAH, level by level ---------------------------------------------------------------
#define FB_SMALL 1024              // segments up to this many surfaces are split by one warp in shared memory
#define FB_NB 16
#define FB_BINW (3 * FB_NB * 7)    // per segment: axis x bin x (min.xyz, max.xyz, count)
#define FB_FULL 0xffffffffu

struct fseg { int32_t lo, hi, parent, info; };                        // info = slot | depth << 1; parent = node id, -1: root
struct bnode { float f[12]; int32_t child[2]; int32_t count[2]; };     // builder layout == pluhost::fast_node_h
struct fsplit { int32_t axis, bin, mid, pad; };                        // axis < 0: halve in current order

__device__ __forceinline__ bool seg_interior(int32_t n, int32_t depth) { return !(n <= PLU_LEAF_SMALL || (n <= 4 && depth > 8)); }
__device__ __forceinline__ int bin_of(float c, float base, float scale) { int k = (int)((c - base) * scale); return min(FB_NB - 1, max(0, k)); }
__device__ __forceinline__ float half_area(const box3& b) { float x = b.mx[0] - b.mn[0], y = b.mx[1] - b.mn[1], z = b.mx[2] - b.mn[2]; return x * y + y * z + z * x; }

// padded tight box and centre of every surface: pa = (min | c.x), pb = (max | c.y), cz = c.z
__global__ void k_fast_prims(int32_t n, const float* prim, const int32_t* meta, float4* pa, float4* pb, float* cz) {
	int32_t s = blockIdx.x * blockDim.x + threadIdx.x;
	if (s >= n) return;
	box3 b = tight_bounds(prim + (size_t)s * 12, meta[4 * s]);
	float big = 0.f, diag = 0.f, c[3];
	for (int i = 0; i < 3; ++i) { big = fmaxf(big, fmaxf(fabsf(b.mn[i]), fabsf(b.mx[i]))); diag += b.mx[i] - b.mn[i]; }
	float pad = 1e-5f * big + 1e-5f * diag + 1e-30f;
	for (int i = 0; i < 3; ++i) { c[i] = 0.5f * (b.mn[i] + b.mx[i]); b.mn[i] -= pad; b.mx[i] += pad; }
	pa[s] = make_float4(b.mn[0], b.mn[1], b.mn[2], c[0]);
	pb[s] = make_float4(b.mx[0], b.mx[1], b.mx[2], c[1]);
	cz[s] = c[2];
}

__global__ void k_fill_i32(int32_t n, int32_t* v, int32_t x) { int32_t i = blockIdx.x * blockDim.x + threadIdx.x; if (i < n) v[i] = x; }

// the 12 ordered words of a surface: box min, box max, centre (as min), centre (as max)
__device__ __forceinline__ void surf_words(int32_t s, const float4* pa, const float4* pb, const float* cz, uint32_t v[12]) {
	float4 a = pa[s], b = pb[s]; float z = cz[s];
	v[0] = f2o(a.x); v[1] = f2o(a.y); v[2] = f2o(a.z); v[3] = f2o(b.x); v[4] = f2o(b.y); v[5] = f2o(b.z);
	v[6] = v[9] = f2o(a.w); v[7] = v[10] = f2o(b.w); v[8] = v[11] = f2o(z);
}
__device__ __forceinline__ bool word_is_min(int k) { return k < 3 || (k >= 6 && k < 9); }

// per level: which segments split, which are big; neutral boxes
__global__ void k_lvl_prep(int32_t m, const fseg* segs, unsigned long long* packed, uint32_t* seg_box) {
	int32_t s = blockIdx.x * blockDim.x + threadIdx.x;
	if (s >= m) return;
	fseg g = segs[s];
	int32_t n = g.hi - g.lo;
	bool interior = seg_interior(n, g.info >> 1);
	packed[s] = (unsigned long long)(interior ? 1u : 0u) | ((unsigned long long)(interior && n > FB_SMALL ? 1u : 0u) << 32);
	for (int k = 0; k < 12; ++k) seg_box[(size_t)s * 12 + k] = word_is_min(k) ? O_POS_INF : O_NEG_INF;
}
__global__ void k_lvl_totals(int32_t m, const unsigned long long* packed, const unsigned long long* ranks, uint32_t* totals) {
	unsigned long long t = ranks[m - 1] + packed[m - 1];
	totals[0] = (uint32_t)t; totals[1] = (uint32_t)(t >> 32);
}

__global__ void k_big_bounds(int32_t n, const int32_t* idx, const int32_t* seg_of, const float4* pa, const float4* pb, const float* cz, uint32_t* seg_box) {
	int32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	int32_t s = i < n ? seg_of[i] : -1;
	uint32_t v[12];
	for (int k = 0; k < 12; ++k) v[k] = word_is_min(k) ? O_POS_INF : O_NEG_INF;
	if (s >= 0) surf_words(idx[i], pa, pb, cz, v);
	int32_t s0 = __shfl_sync(FB_FULL, s, 0);
	if (__all_sync(FB_FULL, s == s0)) {   // one atomic per warp and word
		if (s0 < 0) return;
		for (int k = 0; k < 12; ++k) {
			uint32_t r = word_is_min(k) ? __reduce_min_sync(FB_FULL, v[k]) : __reduce_max_sync(FB_FULL, v[k]);
			if ((threadIdx.x & 31) == 0) { if (word_is_min(k)) atomicMin(&seg_box[(size_t)s0 * 12 + k], r); else atomicMax(&seg_box[(size_t)s0 * 12 + k], r); }
		}
	} else if (s >= 0) {
		for (int k = 0; k < 12; ++k) { if (word_is_min(k)) atomicMin(&seg_box[(size_t)s * 12 + k], v[k]); else atomicMax(&seg_box[(size_t)s * 12 + k], v[k]); }
	}
}

// one warp per segment that is not big (leaves included: the parent stores their box)
__global__ void k_small_bounds(int32_t m, const fseg* segs, const unsigned long long* packed, const int32_t* idx,
                               const float4* pa, const float4* pb, const float* cz, uint32_t* seg_box) {
	int32_t s = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
	if (s >= m || (packed[s] >> 32)) return;
	fseg g = segs[s];
	uint32_t acc[12];
	for (int k = 0; k < 12; ++k) acc[k] = word_is_min(k) ? O_POS_INF : O_NEG_INF;
	for (int32_t i = g.lo + lane; i < g.hi; i += 32) {
		uint32_t v[12]; surf_words(idx[i], pa, pb, cz, v);
		for (int k = 0; k < 12; ++k) acc[k] = word_is_min(k) ? min(acc[k], v[k]) : max(acc[k], v[k]);
	}
	for (int k = 0; k < 12; ++k) {
		uint32_t r = word_is_min(k) ? __reduce_min_sync(FB_FULL, acc[k]) : __reduce_max_sync(FB_FULL, acc[k]);
		if (lane == 0) seg_box[(size_t)s * 12 + k] = r;
	}
}

// a segment reports (box, reference, count) into its slot of the parent node; a splitting segment gets its node
__global__ void k_lvl_parent(int32_t m, const fseg* segs, const unsigned long long* packed, const unsigned long long* ranks,
                             int32_t node_base, const uint32_t* seg_box, bnode* nodes, unsigned long long* node_key, float* root_box) {
	int32_t s = blockIdx.x * blockDim.x + threadIdx.x;
	if (s >= m) return;
	fseg g = segs[s];
	bool interior = (uint32_t)packed[s] != 0u;
	int32_t me = node_base + (int32_t)(uint32_t)ranks[s];
	float b[6];
	for (int k = 0; k < 6; ++k) b[k] = o2f(seg_box[(size_t)s * 12 + k]);
	if (g.parent >= 0) {
		bnode& p = nodes[g.parent];
		int slot = g.info & 1;
		for (int k = 0; k < 6; ++k) p.f[6 * slot + k] = b[k];
		p.child[slot] = interior ? me : ~g.lo;
		p.count[slot] = interior ? 0 : g.hi - g.lo;
	} else {
		for (int k = 0; k < 6; ++k) root_box[k] = b[k];
	}
	if (interior) node_key[me] = ((unsigned long long)(uint32_t)g.lo << 8) | (unsigned long long)(g.info >> 1);
}

// SAH decision of host/bvh_build.cpp:build_rec on a segment's bins (ordered words)
__device__ void sah_pick(const uint32_t* bins, const uint32_t* box_words, int depth, int& axis, int& bin, int& nleft) {
	axis = -1; bin = -1; nleft = 0;
	if (depth >= 40) return;
	float best = INFINITY;
	for (int a = 0; a < 3; ++a) {
		float ext = o2f(box_words[9 + a]) - o2f(box_words[6 + a]);
		if (!(ext > 0.f)) continue;
		const uint32_t* b = bins + a * FB_NB * 7;
		float la[FB_NB]; int lc[FB_NB];
		box3 acc; int ac = 0;
		for (int i = 0; i < 3; ++i) { acc.mn[i] = INFINITY; acc.mx[i] = -INFINITY; }
		for (int k = 0; k < FB_NB; ++k) {
			for (int i = 0; i < 3; ++i) { acc.mn[i] = fminf(acc.mn[i], o2f(b[k * 7 + i])); acc.mx[i] = fmaxf(acc.mx[i], o2f(b[k * 7 + 3 + i])); }
			ac += (int)b[k * 7 + 6];
			la[k] = ac ? half_area(acc) : 0.f; lc[k] = ac;
		}
		ac = 0;
		for (int i = 0; i < 3; ++i) { acc.mn[i] = INFINITY; acc.mx[i] = -INFINITY; }
		for (int k = FB_NB - 1; k > 0; --k) {
			for (int i = 0; i < 3; ++i) { acc.mn[i] = fminf(acc.mn[i], o2f(b[k * 7 + i])); acc.mx[i] = fmaxf(acc.mx[i], o2f(b[k * 7 + 3 + i])); }
			ac += (int)b[k * 7 + 6];
			if (ac == 0 || lc[k - 1] == 0) continue;
			float cost = la[k - 1] * (float)lc[k - 1] + half_area(acc) * (float)ac;
			if (cost < best) { best = cost; axis = a; bin = k; nleft = lc[k - 1]; }
		}
	}
}

__device__ __forceinline__ void emit_children(fseg* next, int32_t irank, const fseg& g, int32_t mid, int32_t node) {
	int32_t d1 = ((g.info >> 1) + 1) << 1;
	next[2 * irank] = fseg{g.lo, mid, node, d1};
	next[2 * irank + 1] = fseg{mid, g.hi, node, d1 | 1};
}
__device__ __forceinline__ float centre_of(int32_t s, int axis, const float4* pa, const float4* pb, const float* cz) {
	return axis == 0 ? pa[s].w : (axis == 1 ? pb[s].w : cz[s]);
}

__global__ void k_bins_init(int32_t count, uint32_t* bins) {
	int32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= count) return;
	int w = i % 7;
	bins[i] = w < 3 ? O_POS_INF : (w < 6 ? O_NEG_INF : 0u);
}
__device__ __forceinline__ void bin_update(uint32_t* slot7, const uint32_t v[12]) {
	for (int i = 0; i < 3; ++i) { atomicMin(slot7 + i, v[i]); atomicMax(slot7 + 3 + i, v[3 + i]); }
	atomicAdd(slot7 + 6, 1u);
}

// big segments: CTA-private bins in shared memory when the whole CTA sits in one segment
__global__ void __launch_bounds__(256) k_big_bin(int32_t n, const int32_t* idx, const int32_t* seg_of, const fseg* segs, const unsigned long long* ranks,
                                                  const float4* pa, const float4* pb, const float* cz, const uint32_t* seg_box, uint32_t* bins) {
	__shared__ uint32_t sb[FB_BINW];
	__shared__ int32_t s_first;
	int32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	int32_t s = i < n ? seg_of[i] : -1;
	if (threadIdx.x == 0) s_first = s;
	__syncthreads();
	const bool uni = __syncthreads_and(s == s_first) && s_first >= 0;
	if (uni) for (int k = threadIdx.x; k < FB_BINW; k += blockDim.x) { int w = k % 7; sb[k] = w < 3 ? O_POS_INF : (w < 6 ? O_NEG_INF : 0u); }
	__syncthreads();
	if (s >= 0 && (segs[s].info >> 1) < 40) {
		uint32_t v[12]; surf_words(idx[i], pa, pb, cz, v);
		uint32_t* dst = uni ? sb : bins + (size_t)(ranks[s] >> 32) * FB_BINW;
		for (int a = 0; a < 3; ++a) {
			float base = o2f(seg_box[(size_t)s * 12 + 6 + a]), ext = o2f(seg_box[(size_t)s * 12 + 9 + a]) - base;
			if (!(ext > 0.f)) continue;
			int k = bin_of(o2f(v[6 + a]), base, (float)FB_NB / ext);
			bin_update(dst + (a * FB_NB + k) * 7, v);
		}
	}
	__syncthreads();
	if (uni) {
		uint32_t* dst = bins + (size_t)(ranks[s_first] >> 32) * FB_BINW;
		for (int k = threadIdx.x; k < FB_BINW; k += blockDim.x) {
			int w = k % 7; uint32_t x = sb[k];
			if (w < 3) { if (x != O_POS_INF) atomicMin(dst + k, x); } else if (w < 6) { if (x != O_NEG_INF) atomicMax(dst + k, x); } else if (x) atomicAdd(dst + k, x);
		}
	}
}

__global__ void k_big_sah(int32_t m, const fseg* segs, const unsigned long long* packed, const unsigned long long* ranks, int32_t node_base,
                          const uint32_t* seg_box, const uint32_t* bins, fsplit* splits, fseg* next) {
	int32_t s = blockIdx.x * blockDim.x + threadIdx.x;
	if (s >= m || !(packed[s] >> 32)) return;
	fseg g = segs[s];
	int32_t brank = (int32_t)(ranks[s] >> 32), irank = (int32_t)(uint32_t)ranks[s];
	int axis, bin, nleft;
	sah_pick(bins + (size_t)brank * FB_BINW, seg_box + (size_t)s * 12, g.info >> 1, axis, bin, nleft);
	int32_t mid = axis >= 0 ? g.lo + nleft : g.lo + (g.hi - g.lo) / 2;
	splits[brank] = fsplit{axis, bin, mid, 0};
	emit_children(next, irank, g, mid, node_base + irank);
}

__global__ void k_big_flags(int32_t n, const int32_t* idx, const int32_t* seg_of, const fseg* segs, const unsigned long long* ranks, const fsplit* splits,
                            const float4* pa, const float4* pb, const float* cz, const uint32_t* seg_box, int32_t* flag) {
	int32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n) return;
	int32_t s = seg_of[i], f = 0;
	if (s >= 0) {
		fsplit sp = splits[ranks[s] >> 32];
		if (sp.axis < 0) f = i < sp.mid;
		else {
			float base = o2f(seg_box[(size_t)s * 12 + 6 + sp.axis]), ext = o2f(seg_box[(size_t)s * 12 + 9 + sp.axis]) - base;
			f = bin_of(centre_of(idx[i], sp.axis, pa, pb, cz), base, 16.f / ext) < sp.bin;
		}
	}
	flag[i] = f;
}
__global__ void k_big_scatter(int32_t n, const int32_t* idx, const int32_t* seg_of, const fseg* segs, const unsigned long long* ranks, const fsplit* splits,
                              const int32_t* flag, const int32_t* pre, int32_t* idx_tmp, int32_t* seg_tmp) {
	int32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n) return;
	int32_t s = seg_of[i];
	if (s < 0) return;
	fseg g = segs[s];
	int32_t mid = splits[ranks[s] >> 32].mid, irank = (int32_t)(uint32_t)ranks[s];
	int32_t lr = pre[i] - pre[g.lo];
	bool left = flag[i] != 0;
	int32_t pos = left ? g.lo + lr : mid + (i - g.lo - lr);
	int32_t child_n = left ? mid - g.lo : g.hi - mid;
	idx_tmp[pos] = idx[i];
	seg_tmp[pos] = child_n > FB_SMALL ? 2 * irank + (left ? 0 : 1) : -1;
}
__global__ void k_big_copyback(int32_t n, int32_t* idx, int32_t* seg_of, const int32_t* idx_tmp, const int32_t* seg_tmp) {
	int32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i < n && seg_of[i] >= 0) { idx[i] = idx_tmp[i]; seg_of[i] = seg_tmp[i]; }
}

// one warp per splitting segment of up to FB_SMALL surfaces: bins, decision and stable partition in shared memory
#define FS_WARPS 4
__global__ void __launch_bounds__(FS_WARPS * 32) k_small_split(int32_t m, const fseg* segs, const unsigned long long* packed, const unsigned long long* ranks,
                                                                 int32_t node_base, const uint32_t* seg_box, int32_t* idx,
                                                                 const float4* pa, const float4* pb, const float* cz, fseg* next) {
	__shared__ uint32_t s_bins[FS_WARPS][FB_BINW];
	__shared__ int32_t s_stage[FS_WARPS][FB_SMALL];
	__shared__ int32_t s_pick[FS_WARPS][4];
	const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
	int32_t s = blockIdx.x * FS_WARPS + w;
	if (s >= m) return;
	unsigned long long pk = packed[s];
	if (!(uint32_t)pk || (pk >> 32)) return;   // leaf, or big
	fseg g = segs[s];
	const int32_t n = g.hi - g.lo, depth = g.info >> 1, irank = (int32_t)(uint32_t)ranks[s];
	uint32_t* bins = s_bins[w]; int32_t* stage = s_stage[w];
	const uint32_t* bw = seg_box + (size_t)s * 12;
	for (int k = lane; k < FB_BINW; k += 32) { int q = k % 7; bins[k] = q < 3 ? O_POS_INF : (q < 6 ? O_NEG_INF : 0u); }
	for (int32_t i = lane; i < n; i += 32) stage[i] = idx[g.lo + i];
	__syncwarp();
	float base[3], scale[3];
	for (int a = 0; a < 3; ++a) { base[a] = o2f(bw[6 + a]); float ext = o2f(bw[9 + a]) - base[a]; scale[a] = ext > 0.f ? (float)FB_NB / ext : 0.f; }
	if (depth < 40)
		for (int32_t i = lane; i < n; i += 32) {
			uint32_t v[12]; surf_words(stage[i], pa, pb, cz, v);
			for (int a = 0; a < 3; ++a) if (scale[a] != 0.f) bin_update(bins + (a * FB_NB + bin_of(o2f(v[6 + a]), base[a], scale[a])) * 7, v);
		}
	__syncwarp();
	if (lane == 0) {
		int axis, bin, nleft;
		sah_pick(bins, bw, depth, axis, bin, nleft);
		s_pick[w][0] = axis; s_pick[w][1] = bin; s_pick[w][2] = axis >= 0 ? g.lo + nleft : g.lo + n / 2;
		emit_children(next, irank, g, s_pick[w][2], node_base + irank);
	}
	__syncwarp();
	const int axis = s_pick[w][0], bin = s_pick[w][1]; const int32_t mid = s_pick[w][2];
	if (axis < 0) return;   // halved in current order: nothing moves
	int32_t nl = 0, nr = 0;
	const uint32_t lt = (1u << lane) - 1u;
	for (int32_t b0 = 0; b0 < n; b0 += 32) {
		int32_t i = b0 + lane; bool act = i < n;
		int32_t sid = act ? stage[i] : 0;
		bool left = act && bin_of(centre_of(sid, axis, pa, pb, cz), base[axis], 16.f / (o2f(bw[9 + axis]) - base[axis])) < bin;
		uint32_t bl = __ballot_sync(FB_FULL, left), br = __ballot_sync(FB_FULL, act && !left);
		if (act) idx[left ? g.lo + nl + __popc(bl & lt) : mid + nr + __popc(br & lt)] = sid;
		nl += __popc(bl); nr += __popc(br);
	}
}

// ---- numbering and the device layout ------------------------------------------------------------------------
__global__ void k_newid(int32_t nn, const int32_t* perm, int32_t* newid) { int32_t j = blockIdx.x * blockDim.x + threadIdx.x; if (j < nn) newid[perm[j]] = j; }

__device__ __forceinline__ void store_fast_node(fast_node* out, const float f[12], int32_t c0, int32_t c1, int32_t n0, int32_t n1) {
	// builder order c0.min c0.max c1.min c1.max -> (child 0, child 1) pairs per plane (scene_dev.cuh)
	float p[12];
	for (int ax = 0; ax < 3; ++ax) { p[2 * ax] = f[ax]; p[2 * ax + 1] = f[6 + ax]; p[6 + 2 * ax] = f[3 + ax]; p[6 + 2 * ax + 1] = f[9 + ax]; }
	out->a = make_float4(p[0], p[1], p[2], p[3]); out->b = make_float4(p[4], p[5], p[6], p[7]); out->c = make_float4(p[8], p[9], p[10], p[11]);
	out->d = make_int4(c0, c1, n0, n1);
}
__device__ __forceinline__ int32_t pack_ref(int32_t child, int32_t count, const int32_t* newid) {
	return child >= 0 ? newid[child] : ~(((~child) << 3) | (count & 7));
}
__global__ void k_fast_emit(int32_t nn, const int32_t* perm, const int32_t* newid, const bnode* nodes, fast_node* out) {
	int32_t j = blockIdx.x * blockDim.x + threadIdx.x;
	if (j >= nn) return;
	bnode b = nodes[perm[j]];
	store_fast_node(out + j, b.f, pack_ref(b.child[0], b.count[0], newid), pack_ref(b.child[1], b.count[1], newid), b.count[0], b.count[1]);
}
// <= 2 surfaces: a root whose child 0 is the leaf and child 1 is empty
__global__ void k_root_leaf(int32_t n, const float* root_box, fast_node* out) {
	float f[12];
	for (int k = 0; k < 6; ++k) f[k] = root_box[k];
	f[6] = f[7] = f[8] = INFINITY; f[9] = f[10] = f[11] = -INFINITY;
	store_fast_node(out, f, ~(n & 7), ~0, n, 0);
}
// leaf slot -> the surface itself (scene_dev.cuh: fast_leaf)
__global__ void k_fast_leaf(int32_t n, const int32_t* idx, const float* prim, const int32_t* meta, float* leaf) {
	int32_t k = blockIdx.x * blockDim.x + threadIdx.x;
	if (k >= n) return;
	int32_t s = idx[k], kind = meta[4 * s], rank = meta[4 * s + 3];
	const float4* src = (const float4*)(prim + (size_t)s * 12);
	float4 a = src[0], b = src[1], c = src[2];
	b.w = __int_as_float((int32_t)(((uint32_t)kind << 30) | (uint32_t)rank));
	if (kind == PLU_PRIM_SPHERE) b.x = __int_as_float(s); else a.w = __int_as_float(s);
	float4* o = (float4*)(leaf + (size_t)k * 12);
	o[0] = a; o[1] = b; o[2] = c;
}

// Part 3: the fast BVH
bool fast_bvh_on_device(int32_t n, device_tables& t, cudaStream_t st, plu_build_stats* stats, std::string& err) {
	dev_mem mem(st);
	float4* pa = mem.get<float4>(n); float4* pb = mem.get<float4>(n); float* cz = mem.get<float>(n);
	int32_t* seg_of = mem.get<int32_t>(n); int32_t* idx_tmp = mem.get<int32_t>(n); int32_t* seg_tmp = mem.get<int32_t>(n);
	int32_t* flag = mem.get<int32_t>(n); int32_t* pre = mem.get<int32_t>(n);
	const size_t seg_cap = (size_t)n + 2, big_cap = (size_t)n / FB_SMALL + 2;
	fseg* segs = mem.get<fseg>(seg_cap); fseg* next = mem.get<fseg>(seg_cap);
	unsigned long long* packed = mem.get<unsigned long long>(seg_cap); unsigned long long* ranks = mem.get<unsigned long long>(seg_cap);
	uint32_t* seg_box = mem.get<uint32_t>(seg_cap * 12);
	uint32_t* bins = mem.get<uint32_t>(big_cap * FB_BINW); fsplit* splits = mem.get<fsplit>(big_cap);
	bnode* nodes = mem.get<bnode>((size_t)n + 1); unsigned long long* node_key = mem.get<unsigned long long>((size_t)n + 1);
	uint32_t* totals = mem.get<uint32_t>(2); float* root_box = mem.get<float>(6);
	int32_t* idx = nullptr;
	if (cudaMallocAsync(&idx, (size_t)n * sizeof(int32_t), st) != cudaSuccess) { err = "cudaMalloc (fast index)"; return false; }
	t.d_fast_index = idx;
	if (!pa || !pb || !cz || !seg_of || !idx_tmp || !seg_tmp || !flag || !pre || !segs || !next || !packed || !ranks || !seg_box || !bins || !splits || !nodes || !node_key || !totals || !root_box) {
		err = "cudaMalloc (fast BVH scratch)"; return false;
	}
	size_t scan_seg = 0, scan_el = 0;
	CK(cub::DeviceScan::ExclusiveSum(nullptr, scan_seg, packed, ranks, (int)seg_cap, st));
	CK(cub::DeviceScan::ExclusiveSum(nullptr, scan_el, flag, pre, n, st));
	unsigned char* scan_tmp = mem.get<unsigned char>(std::max(scan_seg, scan_el));
	if (!scan_tmp) { err = "cudaMalloc (scan scratch)"; return false; }
	size_t scan_bytes = std::max(scan_seg, scan_el);

	k_fast_prims<<<blocks(n), 256, 0, st>>>(n, t.d_prim, t.d_prim_meta, pa, pb, cz);
	k_iota<<<blocks(n), 256, 0, st>>>(n, idx);
	k_fill_i32<<<blocks(n), 256, 0, st>>>(n, seg_of, n > FB_SMALL ? 0 : -1);
	fseg root{0, n, -1, 0};
	CK(cudaMemcpyAsync(segs, &root, sizeof(root), cudaMemcpyHostToDevice, st));
	int32_t m = 1, node_base = 0, levels = 0;
	while (m > 0) {
		k_lvl_prep<<<blocks(m), 256, 0, st>>>(m, segs, packed, seg_box);
		size_t sb = scan_bytes;
		CK(cub::DeviceScan::ExclusiveSum(scan_tmp, sb, packed, ranks, m, st));
		k_lvl_totals<<<1, 1, 0, st>>>(m, packed, ranks, totals);
		uint32_t tot[2];
		CK(cudaMemcpyAsync(tot, totals, sizeof(tot), cudaMemcpyDeviceToHost, st));
		CK(cudaStreamSynchronize(st));
		const int32_t n_int = (int32_t)tot[0], n_big = (int32_t)tot[1];
		if ((size_t)n_big > big_cap || (size_t)node_base + (size_t)n_int > (size_t)n || 2 * (size_t)n_int > seg_cap) { err = "fast BVH build: level bookkeeping out of range"; return false; }
		if (n_big) k_big_bounds<<<blocks(n), 256, 0, st>>>(n, idx, seg_of, pa, pb, cz, seg_box);
		k_small_bounds<<<blocks((size_t)m * 32), 256, 0, st>>>(m, segs, packed, idx, pa, pb, cz, seg_box);
		k_lvl_parent<<<blocks(m), 256, 0, st>>>(m, segs, packed, ranks, node_base, seg_box, nodes, node_key, root_box);
		if (n_big) {
			k_bins_init<<<blocks((size_t)n_big * FB_BINW), 256, 0, st>>>(n_big * FB_BINW, bins);
			k_big_bin<<<blocks(n), 256, 0, st>>>(n, idx, seg_of, segs, ranks, pa, pb, cz, seg_box, bins);
			k_big_sah<<<blocks(m), 256, 0, st>>>(m, segs, packed, ranks, node_base, seg_box, bins, splits, next);
			k_big_flags<<<blocks(n), 256, 0, st>>>(n, idx, seg_of, segs, ranks, splits, pa, pb, cz, seg_box, flag);
			sb = scan_bytes;
			CK(cub::DeviceScan::ExclusiveSum(scan_tmp, sb, flag, pre, n, st));
			k_big_scatter<<<blocks(n), 256, 0, st>>>(n, idx, seg_of, segs, ranks, splits, flag, pre, idx_tmp, seg_tmp);
			k_big_copyback<<<blocks(n), 256, 0, st>>>(n, idx, seg_of, idx_tmp, seg_tmp);
		}
		if (n_int > n_big) k_small_split<<<blocks((size_t)m, FS_WARPS), FS_WARPS * 32, 0, st>>>(m, segs, packed, ranks, node_base, seg_box, idx, pa, pb, cz, next);
		node_base += n_int; m = 2 * n_int; std::swap(segs, next); levels++;
		if (levels > 100) { err = "fast BVH build: depth guard"; return false; }
	}
	const int32_t nn = node_base;
	t.n_fast_nodes = std::max(nn, 1);
	CK(cudaMallocAsync(&t.d_fast, (size_t)t.n_fast_nodes * sizeof(fast_node), st));
	if (nn == 0) {
		k_root_leaf<<<1, 1, 0, st>>>(n, root_box, (fast_node*)t.d_fast);
	} else {   // breadth-first ids -> the recursive builder's pre-order: sort by (first leaf slot, depth)
		int32_t* ids = mem.get<int32_t>(nn); int32_t* perm = mem.get<int32_t>(nn); int32_t* newid = mem.get<int32_t>(nn);
		unsigned long long* key2 = mem.get<unsigned long long>(nn);
		if (!ids || !perm || !newid || !key2) { err = "cudaMalloc (node order)"; return false; }
		size_t sort_bytes = 0;
		CK(cub::DeviceRadixSort::SortPairs(nullptr, sort_bytes, node_key, key2, ids, perm, nn, 0, 40, st));
		unsigned char* sort_tmp = mem.get<unsigned char>(sort_bytes);
		if (!sort_tmp) { err = "cudaMalloc (node sort scratch)"; return false; }
		k_iota<<<blocks(nn), 256, 0, st>>>(nn, ids);
		CK(cub::DeviceRadixSort::SortPairs(sort_tmp, sort_bytes, node_key, key2, ids, perm, nn, 0, 40, st));
		k_newid<<<blocks(nn), 256, 0, st>>>(nn, perm, newid);
		k_fast_emit<<<blocks(nn), 256, 0, st>>>(nn, perm, newid, nodes, (fast_node*)t.d_fast);
		CK(cudaStreamSynchronize(st));   // scratch of this scope is freed on return
	}
	CK(cudaMallocAsync(&t.d_fast_leaf, (size_t)n * 12 * sizeof(float), st));
	k_fast_leaf<<<blocks(n), 256, 0, st>>>(n, idx, t.d_prim, t.d_prim_meta, t.d_fast_leaf);
	CK(cudaStreamSynchronize(st));
	CK(cudaGetLastError());
	if (stats) { stats->levels_fast = levels; stats->n_fast_nodes = t.n_fast_nodes; }
	return true;
}

}  // namespace

bool build_on_device(const plu_scene_desc* d, device_tables& t, cudaStream_t st, plu_build_stats* stats, std::string& err) {
	const int32_t n = d->n_prims;
	if (n <= 0) return true;
	dev_mem mem(st);
	int32_t* order = mem.get<int32_t>(n);
	if (!order) { err = "cudaMalloc (leaf order)"; return false; }
	double t0 = now_ms();
	if (d->prim_tie_rank) {   // ranks given by the caller's live tree (already in meta): only invert them
		k_order_from_rank<<<blocks(n), 256, 0, st>>>(n, t.d_prim_meta, order);
	} else if (!ref_order_on_device(d, t, order, st, stats, err)) return false;
	CK(cudaStreamSynchronize(st));
	double t1 = now_ms();
	if (!ref_nodes_on_device(n, t, order, st, err)) return false;
	CK(cudaStreamSynchronize(st));
	double t2 = now_ms();
	if (!fast_bvh_on_device(n, t, st, stats, err)) return false;
	double t3 = now_ms();
	if (stats) { stats->ms_ref_order = t1 - t0; stats->ms_ref_nodes = t2 - t1; stats->ms_fast_bvh = t3 - t2; stats->built_on_device = 1; }
	return true;
}

}  // namespace plubuild
