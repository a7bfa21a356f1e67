// bvh_build.cpp — see bvh_build.h.  Host plumbing (one-off per scene upload), no CUDA.
#include "bvh_build.h"
#include <algorithm>
#include <cmath>
#include <cstring>
#include <thread>

namespace pluhost {
namespace {

struct box3 { float mn[3], mx[3]; };

// aabb default ctor + add_point (cmmn.h:70-72,88-97): boxes START at the origin
box3 origin_box() { box3 b; for (int i = 0; i < 3; ++i) b.mn[i] = b.mx[i] = 0.f; return b; }
void add_point(box3& b, const float* p) {
	for (int i = 0; i < 3; ++i) { if (p[i] > b.mx[i]) b.mx[i] = p[i]; if (p[i] < b.mn[i]) b.mn[i] = p[i]; }
}
box3 unite_origin(const box3& a, const box3& c) {   // aabb(const aabb&, const aabb&), cmmn.h:78-85
	box3 r = origin_box(); add_point(r, a.mn); add_point(r, a.mx); add_point(r, c.mn); add_point(r, c.mx); return r;
}
// surface::bounds(): sphere.h:11-13 (center -+ radius), box.h:11-13, triangle.h:20-24 (origin-inflated)
box3 surface_bounds(const plu_scene_desc* d, int32_t s) {
	const float* f = d->prim_data + (size_t)s * 12;
	box3 b;
	if (d->prim_tag[s] == PLU_PRIM_SPHERE) { for (int i = 0; i < 3; ++i) { b.mn[i] = f[i] - f[3]; b.mx[i] = f[i] + f[3]; } }
	else if (d->prim_tag[s] == PLU_PRIM_BOX) { for (int i = 0; i < 3; ++i) { b.mn[i] = f[i]; b.mx[i] = f[4 + i]; } }
	else { b = origin_box(); add_point(b, f); add_point(b, f + 4); add_point(b, f + 8); }
	return b;
}
// exact bounds (no origin) for the fast BVH
box3 tight_bounds(const plu_scene_desc* d, int32_t s) {
	if (d->prim_tag[s] != PLU_PRIM_TRIANGLE) return surface_bounds(d, s);
	const float* f = d->prim_data + (size_t)s * 12;
	box3 b; for (int i = 0; i < 3; ++i) { b.mn[i] = std::min(f[i], std::min(f[4 + i], f[8 + i])); b.mx[i] = std::max(f[i], std::max(f[4 + i], f[8 + i])); }
	return b;
}
float centroid(const box3& b, int axis) { return (b.mn[axis] + b.mx[axis]) * 1.f / 2.f; }   // aabb::center, cmmn.h:204-207

// (centre, id) pairs: the comparator reads the element itself instead of chasing two tables per comparison
struct keyed { float k; int32_t v; };

// dfs: where this subtree's leaves go (their DFS positions are a function of the counts alone: the left half comes first).
// The two halves work on disjoint surfaces and disjoint output ranges, so while `threads_left` allows, the left one runs on a
// thread of its own — the permutation is the sequential one, std::sort by std::sort.
void order_rec(const std::vector<float>* cen, std::vector<int32_t>& s, int axis, int32_t* dfs, int threads_left) {
	if (s.size() == 1) { dfs[0] = s[0]; return; }
	if (s.size() == 2) { dfs[0] = s[0]; dfs[1] = s[1]; return; }
	// same std::sort, same comparator outcomes on the same input order as bvh_tree.cpp:23-27 => same permutation
	std::vector<keyed> kv(s.size());
	for (size_t i = 0; i < s.size(); ++i) kv[i] = keyed{cen[axis][(size_t)s[i]], s[i]};
	std::sort(kv.begin(), kv.end(), [](const keyed& a, const keyed& b) { return a.k < b.k; });
	size_t mid = s.size() / 2;
	std::vector<int32_t> lv(mid), rv(s.size() - mid);
	for (size_t i = 0; i < mid; ++i) lv[i] = kv[i].v;
	for (size_t i = mid; i < s.size(); ++i) rv[i - mid] = kv[i].v;
	std::vector<keyed>().swap(kv);
	std::vector<int32_t>().swap(s);
	const int next_axis = (axis + 1) % 3;
	if (threads_left > 1 && lv.size() >= 4096) {
		const int lt = threads_left / 2;
		std::thread left([&]() { order_rec(cen, lv, next_axis, dfs, lt); });
		order_rec(cen, rv, next_axis, dfs + mid, threads_left - lt);
		left.join();
	} else {
		order_rec(cen, lv, next_axis, dfs, 1);
		order_rec(cen, rv, next_axis, dfs + mid, 1);
	}
}

box3 emit_ref(const plu_scene_desc* d, const std::vector<int32_t>& dfs, size_t lo, size_t hi, std::vector<ref_node_h>& out) {
	size_t me = out.size();
	out.emplace_back();
	box3 b;
	if (hi - lo == 1) {
		b = surface_bounds(d, dfs[lo]);
		out[me].prim = dfs[lo];
	} else {
		size_t mid = lo + (hi - lo) / 2;   // 2 -> 1|1, n -> n/2 | n - n/2 (bvh_tree.cpp:19-32)
		box3 l = emit_ref(d, dfs, lo, mid, out), r = emit_ref(d, dfs, mid, hi, out);
		b = unite_origin(l, r);
		out[me].prim = -1;
	}
	for (int i = 0; i < 3; ++i) { out[me].lo[i] = b.mn[i]; out[me].hi[i] = b.mx[i]; }
	// leaf bounds are never tested (bvh_tree.cpp:40-42), so a leaf node carries what its surface test reads:
	// a box leaf's bounds already ARE the box (min, max); a sphere leaf holds (centre | radius)
	if (out[me].prim >= 0 && d->prim_tag[out[me].prim] == PLU_PRIM_SPHERE) {
		const float* f = d->prim_data + (size_t)out[me].prim * 12;
		for (int i = 0; i < 3; ++i) { out[me].lo[i] = f[i]; out[me].hi[i] = f[3]; }
	}
	// interior: index of the next node outside this subtree; leaf (whose successor is always me + 1):
	// the surface kind, so the walk needs no second lookup
	out[me].skip = out[me].prim >= 0 ? d->prim_tag[out[me].prim] : (int32_t)out.size();
	return b;
}

// ---- binned SAH --------------------------------------------------------------------------------------
struct build_ctx {
	const plu_scene_desc* d;
	std::vector<box3> pb;        // padded per-surface boxes
	std::vector<float> cen;      // 3 per surface
	std::vector<int32_t> idx;    // permutation being partitioned
	std::vector<fast_node_h>* nodes;
};
float half_area(const box3& b) { float x = b.mx[0] - b.mn[0], y = b.mx[1] - b.mn[1], z = b.mx[2] - b.mn[2]; return x * y + y * z + z * x; }
box3 empty_box() { box3 b; for (int i = 0; i < 3; ++i) { b.mn[i] = INFINITY; b.mx[i] = -INFINITY; } return b; }
void grow(box3& a, const box3& c) { for (int i = 0; i < 3; ++i) { a.mn[i] = std::min(a.mn[i], c.mn[i]); a.mx[i] = std::max(a.mx[i], c.mx[i]); } }

// returns child reference (>= 0 node, < 0 leaf ~first) and its count/bounds
int32_t build_rec(build_ctx& c, size_t lo, size_t hi, int depth, box3& bounds, int32_t& count) {
	bounds = empty_box();
	box3 cb = empty_box();
	for (size_t i = lo; i < hi; ++i) {
		int32_t s = c.idx[i]; grow(bounds, c.pb[(size_t)s]);
		for (int a = 0; a < 3; ++a) { float v = c.cen[(size_t)s * 3 + a]; cb.mn[a] = std::min(cb.mn[a], v); cb.mx[a] = std::max(cb.mx[a], v); }
	}
	size_t n = hi - lo;
	if (n <= PLU_LEAF_SMALL || (n <= 4 && depth > 8)) { count = (int32_t)n; return ~(int32_t)lo; }
	size_t mid = lo;
	bool split_ok = false;
	if (depth < 40) {
		const int NB = 16;
		float best_cost = INFINITY; int best_axis = -1, best_bin = -1;
		for (int a = 0; a < 3; ++a) {
			float ext = cb.mx[a] - cb.mn[a];
			if (!(ext > 0.f)) continue;
			box3 bb[NB]; int cnt[NB];
			for (int k = 0; k < NB; ++k) { bb[k] = empty_box(); cnt[k] = 0; }
			float scale = (float)NB / ext;
			for (size_t i = lo; i < hi; ++i) {
				int32_t s = c.idx[i];
				int k = std::min(NB - 1, std::max(0, (int)((c.cen[(size_t)s * 3 + a] - cb.mn[a]) * scale)));
				grow(bb[k], c.pb[(size_t)s]); cnt[k]++;
			}
			float la[NB]; int lc[NB]; box3 acc = empty_box(); int ac = 0;
			for (int k = 0; k < NB; ++k) { grow(acc, bb[k]); ac += cnt[k]; la[k] = ac ? half_area(acc) : 0.f; lc[k] = ac; }
			acc = empty_box(); ac = 0;
			for (int k = NB - 1; k > 0; --k) {
				grow(acc, bb[k]); ac += cnt[k];
				if (ac == 0 || lc[k - 1] == 0) continue;
				float cost = la[k - 1] * (float)lc[k - 1] + half_area(acc) * (float)ac;
				if (cost < best_cost) { best_cost = cost; best_axis = a; best_bin = k; }
			}
		}
		if (best_axis >= 0) {
			float ext = cb.mx[best_axis] - cb.mn[best_axis], scale = 16.f / ext, base = cb.mn[best_axis];
			auto it = std::stable_partition(c.idx.begin() + (long)lo, c.idx.begin() + (long)hi, [&](int32_t s) {
				int k = std::min(15, std::max(0, (int)((c.cen[(size_t)s * 3 + best_axis] - base) * scale)));
				return k < best_bin;
			});
			mid = (size_t)(it - c.idx.begin());
			split_ok = mid > lo && mid < hi;
		}
	}
	// equal centroids or depth guard: halve in the current order (the partition above is stable, so the order — and with
	// it every table — is the one bvh_gpu.cu's device build produces)
	if (!split_ok) mid = lo + n / 2;
	int32_t me = (int32_t)c.nodes->size();
	c.nodes->emplace_back();
	box3 b0, b1; int32_t n0, n1;
	int32_t c0 = build_rec(c, lo, mid, depth + 1, b0, n0);
	int32_t c1 = build_rec(c, mid, hi, depth + 1, b1, n1);
	fast_node_h& nd = (*c.nodes)[(size_t)me];
	// a = (c0.min.xyz, c0.max.x)  b = (c0.max.yz, c1.min.xy)  c = (c1.min.z, c1.max.xyz)
	nd.f[0] = b0.mn[0]; nd.f[1] = b0.mn[1]; nd.f[2] = b0.mn[2]; nd.f[3] = b0.mx[0];
	nd.f[4] = b0.mx[1]; nd.f[5] = b0.mx[2]; nd.f[6] = b1.mn[0]; nd.f[7] = b1.mn[1];
	nd.f[8] = b1.mn[2]; nd.f[9] = b1.mx[0]; nd.f[10] = b1.mx[1]; nd.f[11] = b1.mx[2];
	nd.child[0] = c0; nd.child[1] = c1; nd.count[0] = n0; nd.count[1] = n1;
	count = 0;
	return me;
}

}  // namespace

void reference_leaf_order(const plu_scene_desc* d, std::vector<int32_t>& rank) {
	size_t n = (size_t)d->n_prims;
	rank.assign(n, 0);
	if (n == 0) return;
	std::vector<int32_t> s(n), dfs(n);
	for (size_t i = 0; i < n; ++i) s[i] = (int32_t)i;   // load order (scene.h:268-297)
	std::vector<float> cen[3];
	for (int a = 0; a < 3; ++a) { cen[a].resize(n); for (size_t i = 0; i < n; ++i) cen[a][i] = centroid(surface_bounds(d, (int32_t)i), a); }
	order_rec(cen, s, 0, dfs.data(), (int)std::max(1u, std::min(16u, std::thread::hardware_concurrency())));
	for (size_t k = 0; k < n; ++k) rank[(size_t)dfs[k]] = (int32_t)k;
}

void build_ref_nodes(const plu_scene_desc* d, const std::vector<int32_t>& rank, std::vector<ref_node_h>& out) {
	out.clear();
	size_t n = (size_t)d->n_prims;
	if (n == 0) return;
	std::vector<int32_t> dfs(n);
	for (size_t s = 0; s < n; ++s) dfs[(size_t)rank[s]] = (int32_t)s;
	out.reserve(2 * n);
	emit_ref(d, dfs, 0, n, out);
}

void build_fast_bvh(const plu_scene_desc* d, std::vector<fast_node_h>& nodes, std::vector<int32_t>& prim_index) {
	nodes.clear(); prim_index.clear();
	size_t n = (size_t)d->n_prims;
	if (n == 0) return;
	build_ctx c; c.d = d; c.nodes = &nodes;
	c.pb.resize(n); c.cen.resize(3 * n); c.idx.resize(n);
	for (size_t s = 0; s < n; ++s) {
		box3 b = tight_bounds(d, (int32_t)s);
		float big = 0.f, diag = 0.f;
		for (int i = 0; i < 3; ++i) { big = std::max(big, std::max(std::fabs(b.mn[i]), std::fabs(b.mx[i]))); diag += b.mx[i] - b.mn[i]; }
		float pad = 1e-5f * big + 1e-5f * diag + 1e-30f;   // guard band >> rounding of the slab / hit arithmetic
		for (int i = 0; i < 3; ++i) { c.cen[3 * s + (size_t)i] = 0.5f * (b.mn[i] + b.mx[i]); b.mn[i] -= pad; b.mx[i] += pad; }
		c.pb[s] = b; c.idx[s] = (int32_t)s;
	}
	nodes.reserve(n);
	box3 rb; int32_t rc;
	int32_t root = build_rec(c, 0, n, 0, rb, rc);
	if (root < 0) {   // <= 2 surfaces: a root whose child 0 is the leaf and child 1 is empty
		fast_node_h nd; std::memset(&nd, 0, sizeof(nd));
		nd.f[0] = rb.mn[0]; nd.f[1] = rb.mn[1]; nd.f[2] = rb.mn[2]; nd.f[3] = rb.mx[0]; nd.f[4] = rb.mx[1]; nd.f[5] = rb.mx[2];
		nd.f[6] = nd.f[7] = nd.f[8] = INFINITY; nd.f[9] = nd.f[10] = nd.f[11] = -INFINITY;
		nd.child[0] = root; nd.count[0] = rc; nd.child[1] = ~0; nd.count[1] = 0;
		nodes.push_back(nd);
	}
	prim_index = c.idx;
}

}  // namespace pluhost
