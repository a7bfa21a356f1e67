// isect.cuh — primitive intersection and BVH traversal (replaces scene->hit: bvh_tree.cpp:39-76,
// aabb::hit cmmn.h:151-173, sphere.cpp:16-47, box.cpp:6-62, triangle.cpp:5-33).
//
// Traversal only decides (t, surface); the rest of hit_record (norm, uv, dpdu) is a pure function of
// (ray, t, surface) and is rebuilt by the consumer with fill_record(), so a hit travels between
// kernels as 8 bytes instead of the reference's 56-byte record.
#pragma once
#include "scene_dev.cuh"

namespace plud {

struct ray { f3 e, d; };
// ray::operator()(t) = e + d * t (cmmn.h:55): a multiply and an add, each rounded.  The path has no ray epsilon, so whether
// the next ray re-hits its origin surface at t ~ +0 is decided by the last bit of this point; it is therefore written with
// explicit round-to-nearest operations that no compiler flag can fuse into an FMA (csrc/shade_tu.cu may be built with
// -fmad=true: PLU_SHADE_FAST).
PLU_DEV f3 ray_at(const ray& r, float t) {
	return mk3(__fadd_rn(r.e.x, __fmul_rn(r.d.x, t)), __fadd_rn(r.e.y, __fmul_rn(r.d.y, t)), __fadd_rn(r.e.z, __fmul_rn(r.d.z, t)));
}

#define PLU_UNIFORM_WALK_MAX_NODES 64
#ifndef PLU_UNIFORM_UNROLL
#define PLU_UNIFORM_UNROLL 1   // (1 / 2 / 4 measured: 1 is the fastest and the smallest) unroll factor of the uniform walk's node loop (build option)
#endif
#define PLU_PRAGMA_(x) _Pragma(#x)
#define PLU_UNROLL(n) PLU_PRAGMA_(unroll n)
#define PLU_T_MAX 100000.f   // hit_record::t initial value, cmmn.h:229

struct hit_rec {   // what shading needs of hit_record (dpdv is never read by the path)
	f3 norm; f2 uv; f3 dpdu;
};

// aabb::hit / box::hit slab arithmetic, cmmn.h:155-169 == box.cpp:8-23
PLU_DEV void slabs(f3 mn, f3 mx, const ray& r, float& tmin, float& tmax) {
	f3 rrd = 1.f / r.d;
	f3 t1 = (mn - r.e) * rrd;
	f3 t2 = (mx - r.e) * rrd;
	f3 m12 = mk3(gmin(t1.x, t2.x), gmin(t1.y, t2.y), gmin(t1.z, t2.z));
	f3 x12 = mk3(gmax(t1.x, t2.x), gmax(t1.y, t2.y), gmax(t1.z, t2.z));
	tmin = m12.x; tmin = gmax(tmin, m12.y); tmin = gmax(tmin, m12.z);
	tmax = x12.x; tmax = gmin(tmax, x12.y); tmax = gmin(tmax, x12.z);
}

// same arithmetic with the reciprocal direction hoisted (1.f / r.d is the same value at every node)
PLU_DEV void slabs_rrd(f3 mn, f3 mx, f3 e, f3 rrd, float& tmin, float& tmax) {
	f3 t1 = (mn - e) * rrd;
	f3 t2 = (mx - e) * rrd;
	f3 m12 = mk3(gmin(t1.x, t2.x), gmin(t1.y, t2.y), gmin(t1.z, t2.z));
	f3 x12 = mk3(gmax(t1.x, t2.x), gmax(t1.y, t2.y), gmax(t1.z, t2.z));
	tmin = m12.x; tmin = gmax(tmin, m12.y); tmin = gmax(tmin, m12.z);
	tmax = x12.x; tmax = gmin(tmax, x12.y); tmax = gmin(tmax, x12.z);
}

// aabb::hit / box::hit slab arithmetic once more, for the hot uniform walk: the 18 compare+select pairs
// of glm::min/max become 6 FMNMX.NAN + 2 FMNMX3.NAN.  IEEE min/max agree with glm's `x<y?x:y` except
// (1) when an operand is NaN and (2) in the sign of a zero result.  The .NaN forms make any NaN among
// the six plane distances surface in BOTH results, so `tmin != tmin` is an exact hazard test; the
// caller re-does the node with slabs_rrd() on a hazard (and on tmin == 0 where the value is returned
// as a hit distance), so the outcome is the reference's bit for bit.
PLU_DEV float fmin_nan(float a, float b) { float r; asm("min.NaN.f32 %0, %1, %2;" : "=f"(r) : "f"(a), "f"(b)); return r; }
PLU_DEV float fmax_nan(float a, float b) { float r; asm("max.NaN.f32 %0, %1, %2;" : "=f"(r) : "f"(a), "f"(b)); return r; }
PLU_DEV float fmin3_nan(float a, float b, float c) { float r; asm("min.NaN.f32 %0, %1, %2, %3;" : "=f"(r) : "f"(a), "f"(b), "f"(c)); return r; }
PLU_DEV float fmax3_nan(float a, float b, float c) { float r; asm("max.NaN.f32 %0, %1, %2, %3;" : "=f"(r) : "f"(a), "f"(b), "f"(c)); return r; }
PLU_DEV void slabs_fast(f3 mn, f3 mx, f3 e, f3 rrd, float& tmin, float& tmax) {
	f3 t1 = (mn - e) * rrd;
	f3 t2 = (mx - e) * rrd;
	tmin = fmax3_nan(fmin_nan(t1.x, t2.x), fmin_nan(t1.y, t2.y), fmin_nan(t1.z, t2.z));
	tmax = fmin3_nan(fmax_nan(t1.x, t2.x), fmax_nan(t1.y, t2.y), fmax_nan(t1.z, t2.z));
}

// Packed fp32 pairs (sm_100 FADD2 / FMUL2: two IEEE round-to-nearest results per issue slot): the plane
// distances of BOTH children of a fast node along one axis side are one subtract + one multiply.
PLU_DEV unsigned long long pk2(float lo, float hi) { unsigned long long r; asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi)); return r; }
PLU_DEV void upk2(unsigned long long v, float& lo, float& hi) { asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v)); }
PLU_DEV unsigned long long submul2(unsigned long long m, unsigned long long e, unsigned long long i) {
	unsigned long long t;
	asm("{\n\t.reg .b64 d;\n\tsub.rn.f32x2 d, %1, %2;\n\tmul.rn.f32x2 %0, d, %3;\n\t}" : "=l"(t) : "l"(m), "l"(e), "l"(i));
	return t;
}
PLU_DEV float fmin3f(float a, float b, float c) { float r; asm("min.f32 %0, %1, %2, %3;" : "=f"(r) : "f"(a), "f"(b), "f"(c)); return r; }
PLU_DEV float fmax3f(float a, float b, float c) { float r; asm("max.f32 %0, %1, %2, %3;" : "=f"(r) : "f"(a), "f"(b), "f"(c)); return r; }

// t-only versions of surface::hit(r, hr) with hr->t == bound on entry.
PLU_DEV bool sphere_t(float4 cr, const ray& r, float bound, float& t) {   // sphere.cpp:17-26
	f3 v = r.e - mk3(cr);
	float b = -dot(v, r.d);
	float det = (b * b) - dot(v, v) + cr.w * cr.w;
	if (det < 0) return false;
	det = sqrtf(det);
	float i1 = b - det, i2 = b + det;
	if (i2 > 0 && i1 > 0) {
		if (bound < i1) return false;
		t = i1; return true;
	}
	return false;
}
PLU_DEV bool box_t(float4 mn, float4 mx, const ray& r, float bound, float& t) {   // box.cpp:8-27
	const f3 rrd = 1.f / r.d;
	float tmin, tmax; slabs_fast(mk3(mn), mk3(mx), r.e, rrd, tmin, tmax);
	if ((tmin != tmin) || tmin == 0.f) slabs_rrd(mk3(mn), mk3(mx), r.e, rrd, tmin, tmax);   // hazard: exact glm order
	if (tmax < tmin || tmin < 0) return false;
	if (tmin > bound) return false;
	t = tmin; return true;
}
PLU_DEV bool triangle_t(float4 a, float4 b, float4 c, const ray& r, float bound, float& t) {   // triangle.cpp:6-25
	f3 v0 = mk3(a);
	f3 e1 = mk3(b) - v0, e2 = mk3(c) - v0;
	f3 pv = cross(r.d, e2);
	float det = dot(e1, pv);
	if (det == 0) return false;
	float idet = 1.f / det;
	f3 tv = r.e - v0;
	float u = dot(tv, pv) * idet;
	if (u < 0 || u > 1.f) return false;
	f3 qv = cross(tv, e1);
	float v = dot(r.d, qv) * idet;
	if (v < 0 || u + v > 1) return false;
	float nt = dot(e2, qv) * idet;
	if (nt > 0 && nt < bound) { t = nt; return true; }
	return false;
}

PLU_DEV bool prim_t(const scene_ref& sc, int32_t s, int32_t tag, const ray& r, float bound, float& t) {
	const float4* p = sc.prim(s);
	if (tag == PLU_PRIM_BOX) return box_t(p[0], p[1], r, bound, t);
	if (tag == PLU_PRIM_SPHERE) return sphere_t(p[0], r, bound, t);
	return triangle_t(p[0], p[1], p[2], r, bound, t);
}

// box::get_normal, box.cpp:37-62
PLU_DEV f3 box_normal(f3 mn, f3 mx, f3 p) {
	f3 c = (mn + mx) * 1.f / 2.f;
	f3 ext = mx - c;
	f3 n = mk3(0.f);
	float m = 3.402823466e+38f;
	f3 np = p - c;
#pragma unroll
	for (int i = 0; i < 3; ++i) {
		float dist = fabsf(comp(ext, i) - fabsf(comp(np, i)));
		if (dist < m) {
			m = dist;
			float sg = gsign(comp(np, i));
			n = mk3(sg * (i == 0 ? 1.f : 0.f), sg * (i == 1 ? 1.f : 0.f), sg * (i == 2 ? 1.f : 0.f));
		}
	}
	return n;
}

// The remaining fields of hit_record for a hit of surface s at distance t along r.
// want_uv: the material is textured (sphere uv costs two acos + a sin).
PLU_DEV void fill_record(const scene_ref& sc, int32_t s, int32_t tag, const ray& r, float t, bool want_uv, hit_rec& hr) {
	const float4* p = sc.prim(s);
	hr.uv = mk2(0.f, 0.f);
	if (tag == PLU_PRIM_BOX) {   // box.cpp:28-34
		f3 q = ray_at(r, t);
		f3 n = box_normal(mk3(p[0]), mk3(p[1]), q);
		uint32_t mci = 0;
		if (n.x != 0.f) mci = 0; if (n.y != 0.f) mci = 1; if (n.z != 0.f) mci = 2;
		uint32_t ia = (mci - 1u) % 3u, ib = (mci + 1u) % 3u;   // unsigned wraparound, box.cpp:31-33
		hr.norm = n;
		hr.uv = mk2(comp(q, (int)ia), comp(q, (int)ib));
		hr.dpdu = mk3(ia == 0 ? 1.f : 0.f, ia == 1 ? 1.f : 0.f, ia == 2 ? 1.f : 0.f);
	} else if (tag == PLU_PRIM_SPHERE) {   // sphere.cpp:27-44
		float4 cr = p[0];
		hr.norm = normalize(ray_at(r, t) - mk3(cr));
		if (want_uv) {
			float cos_phi = -dot(hr.norm, mk3(0.f, 1.f, 0.f));
			float phi = det_acos(cos_phi);
			float sin_phi = det_sin(phi);
			hr.uv.y = phi * PLU_ONE_OVER_PI;
			float theta = det_acos(dot(mk3(0.f, 0.f, -1.f), hr.norm) / sin_phi) * PLU_TWO_OVER_PI;
			if (dot(mk3(1.f, 0.f, 0.f), hr.norm) >= 0) theta = 1.f - theta;
			hr.uv.x = theta;
		}
		f3 q = ray_at(r, t);
		hr.dpdu = mk3(-2.f * PLU_PI * q.y, 2.f * PLU_PI * q.x, 0.f);
	} else {   // triangle.cpp:26-31 (u, v recomputed exactly as in triangle_t)
		f3 v0 = mk3(p[0]);
		f3 e1 = mk3(p[1]) - v0, e2 = mk3(p[2]) - v0;
		f3 U = normalize(e1), V = normalize(e2);
		hr.norm = cross(U, V);
		hr.dpdu = U;
		if (want_uv) {
			f3 pv = cross(r.d, e2);
			float idet = 1.f / dot(e1, pv);
			f3 tv = r.e - v0;
			float u = dot(tv, pv) * idet;
			f3 qv = cross(tv, e1);
			float v = dot(r.d, qv) * idet;
			float w = 1 - (u + v);
			const float* tc = sc.tri_uv + (size_t)s * 6;
			hr.uv = mk2(tc[0] * u + tc[2] * v + tc[4] * w, tc[1] * u + tc[3] * v + tc[5] * w);
		}
	}
}

// ---- traversal of the reference's own tree ------------------------------------------------------------
// Same node set, same inflated bounds, same slab test as bvh_node::hit, therefore the same set of
// surfaces is tested; leaves are met in DFS (left first) order and a later leaf wins unless the
// earlier one is strictly nearer — `lhr.t < rhr.t ? lhr : rhr`, bvh_tree.cpp:73.  (No t-culling: the
// reference gives every child a fresh record.)
struct trace_counts { uint32_t nodes, prims; };

template <bool COUNT>
PLU_DEV bool trace_ref(const scene_ref& sc, const ray& r, float root_bound, float& t_out, int32_t& prim_out, trace_counts* tc) {
	const ref_node* nodes = sc.ref_nodes();
	int32_t n = sc.n_ref_nodes;
	float best_t = 0.f; int32_t best = -1;
	int32_t i = 0;
	while (i < n) {
		float4 lo = nodes[i].lo, hi = nodes[i].hi;
		int32_t prim = __float_as_int(hi.w);
		if (prim >= 0) {
			int32_t tag = __float_as_int(lo.w);
			float t;
			if (COUNT) tc->prims++;
			// a root that is itself a leaf sees the caller's record (renderer.cpp:60,86 reuses hr)
			if (prim_t(sc, prim, tag, r, n == 1 ? root_bound : PLU_T_MAX, t)) {
				if (best < 0 || !(best_t < t)) { best_t = t; best = prim; }
			}
			i = i + 1;
		} else {
			if (COUNT) tc->nodes++;
			float tmin, tmax; slabs(mk3(lo), mk3(hi), r, tmin, tmax);
			i = (tmax >= tmin) ? i + 1 : __float_as_int(lo.w);
		}
	}
	t_out = best_t; prim_out = best;
	return best >= 0;
}

// scene->hit: the fast BVH when the scene has one (large scenes), else the reference's own tree.
// Small scenes (all five named ones): the same walk as a WARP-UNIFORM loop over the pre-order node
// array.  Every lane steps through every node index; a lane whose ray failed an ancestor's slab test
// is merely predicated off until the loop reaches that ancestor's skip index.  Node reads are one
// shared-memory broadcast per warp, so there is no divergence left except the predicate itself (ncu
// r01: the per-lane walk ran at 18.4 of 32 active threads per instruction).  Interior nodes and box
// leaves (a box leaf's node bounds ARE the box, bvh_build.cpp) share one branch-free slab test; a
// sphere leaf carries (centre, radius) in its node; only triangles fetch primitive data.  Tested
// surfaces and the tie rule are unchanged.
// One pass of the uniform walk over the PAIRED staged nodes (stage_scene(.., pair_nodes = true): a = (min.x max.x min.y max.y),
// b = (min.z max.z aux prim)): the six plane distances are three packed fp32x2 subtract + multiply pairs (sub.rn / mul.rn per half:
// the scalar results bit for bit, in half the issue slots).  EXACT = false: min / max in the IEEE .NaN forms; a box LEAF whose tmin
// is NaN or zero (a ray that leaves a box starts on one of its planes: about once per secondary ray) is redone on the spot in glm's
// select order on the same six distances; a NaN at a live INTERIOR node (a direction component of exactly 0 with the origin on a
// plane of that node's box) is only REPORTED — the caller then repeats the whole walk with EXACT = true (select order at every
// node), which keeps a compare, a branch and a reconvergence pair out of every interior node of the common pass.
template <bool COUNT, bool EXACT>
PLU_DEV bool uniform_pass_paired(const scene_ref& sc, const ref_node* nodes, int32_t n, const ray& r, f3 rrd, float bound,
                                 float& best_t_out, int32_t& best_out, trace_counts& cnt) {
	const unsigned long long ex2 = pk2(r.e.x, r.e.x), ey2 = pk2(r.e.y, r.e.y), ez2 = pk2(r.e.z, r.e.z);
	const unsigned long long ix2 = pk2(rrd.x, rrd.x), iy2 = pk2(rrd.y, rrd.y), iz2 = pk2(rrd.z, rrd.z);
	float best_t = 0.f; int32_t best = -1;
	int32_t skip_to = 0;
	float nan_acc = 0.f;   // NaN once any interior node's tmin was (min.NaN propagates it; a node of a skipped subtree counts too: harmless)
PLU_UNROLL(PLU_UNIFORM_UNROLL)
	for (int32_t i = 0; i < n; ++i) {
		const float4 a = nodes[i].lo, b = nodes[i].hi;
		const int32_t prim = __float_as_int(b.w), aux = __float_as_int(b.z);   // aux: skip index | surface kind
		const bool live = i >= skip_to;
		const bool interior = prim < 0;
		if (interior || aux == PLU_PRIM_BOX) {
			float ax, bx, ay, by, az, bz;   // (min - e) * rrd, (max - e) * rrd per axis
			upk2(submul2(pk2(a.x, a.y), ex2, ix2), ax, bx);
			upk2(submul2(pk2(a.z, a.w), ey2, iy2), ay, by);
			upk2(submul2(pk2(b.x, b.y), ez2, iz2), az, bz);
			float tmin, tmax;
			if (EXACT) {
				tmin = gmax(gmax(gmin(ax, bx), gmin(ay, by)), gmin(az, bz));
				tmax = gmin(gmin(gmax(ax, bx), gmax(ay, by)), gmax(az, bz));
			} else {
				tmin = fmax3_nan(fmin_nan(ax, bx), fmin_nan(ay, by), fmin_nan(az, bz));
				tmax = fmin3_nan(fmax_nan(ax, bx), fmax_nan(ay, by), fmax_nan(az, bz));
			}
			if (interior) {
				if (COUNT && live) cnt.nodes++;
				if (!EXACT) nan_acc = fmin_nan(nan_acc, tmin);
				if (live && !(tmax >= tmin)) skip_to = aux;
			} else {
				if (!EXACT && ((tmin != tmin) || tmin == 0.f)) {
					tmin = gmax(gmax(gmin(ax, bx), gmin(ay, by)), gmin(az, bz));
					tmax = gmin(gmin(gmax(ax, bx), gmax(ay, by)), gmax(az, bz));
				}
				if (COUNT && live) cnt.prims++;
				bool h = !(tmax < tmin || tmin < 0) && !(tmin > bound);
				if (live && h && (best < 0 || !(best_t < tmin))) { best_t = tmin; best = prim; }
			}
		} else {
			float t; bool h;
			if (aux == PLU_PRIM_SPHERE) h = sphere_t(make_float4(a.x, a.z, b.x, a.y), r, bound, t);
			else { const float4* p = sc.prim(prim); h = triangle_t(p[0], p[1], p[2], r, bound, t); }
			if (COUNT && live) cnt.prims++;
			if (live && h && (best < 0 || !(best_t < t))) { best_t = t; best = prim; }
		}
	}
	best_t_out = best_t; best_out = best;
	return nan_acc != nan_acc;
}

// SMEM: the staged, paired nodes (uniform_pass_paired above); otherwise the table in global memory in its ref_node layout.
template <bool COUNT, bool SMEM>
PLU_DEV bool trace_ref_uniform(const scene_ref& sc, const ray& r, float root_bound, float& t_out, int32_t& prim_out, trace_counts* tc) {
	const ref_node* nodes = sc.ref_nodes();
	const int32_t n = sc.n_ref_nodes;
	const f3 rrd = 1.f / r.d;
	const float bound = n == 1 ? root_bound : PLU_T_MAX;
	float best_t = 0.f; int32_t best = -1;
	if (SMEM) {
		__builtin_assume(__isShared(nodes));
		trace_counts cnt; cnt.nodes = 0; cnt.prims = 0;
		if (uniform_pass_paired<COUNT, false>(sc, nodes, n, r, rrd, bound, best_t, best, cnt)) {
			cnt.nodes = 0; cnt.prims = 0;
			uniform_pass_paired<COUNT, true>(sc, nodes, n, r, rrd, bound, best_t, best, cnt);
		}
		if (COUNT) { tc->nodes += cnt.nodes; tc->prims += cnt.prims; }
		t_out = best_t; prim_out = best;
		return best >= 0;
	}
	int32_t skip_to = 0;
PLU_UNROLL(PLU_UNIFORM_UNROLL)
	for (int32_t i = 0; i < n; ++i) {
		const float4 lo = nodes[i].lo, hi = nodes[i].hi;
		const int32_t prim = __float_as_int(hi.w), aux = __float_as_int(lo.w);   // aux: skip index | surface kind
		const bool live = i >= skip_to;
		const bool interior = prim < 0;
		if (interior || aux == PLU_PRIM_BOX) {
			float tmin, tmax;
			slabs_fast(mk3(lo), mk3(hi), r.e, rrd, tmin, tmax);
			if ((tmin != tmin) || (!interior && tmin == 0.f)) slabs_rrd(mk3(lo), mk3(hi), r.e, rrd, tmin, tmax);   // hazard: exact glm order
			if (interior) {
				if (COUNT && live) tc->nodes++;
				if (live && !(tmax >= tmin)) skip_to = aux;
			} else {
				if (COUNT && live) tc->prims++;
				bool h = !(tmax < tmin || tmin < 0) && !(tmin > bound);
				if (live && h && (best < 0 || !(best_t < tmin))) { best_t = tmin; best = prim; }
			}
		} else {
			float t; bool h;
			if (aux == PLU_PRIM_SPHERE) h = sphere_t(make_float4(lo.x, lo.y, lo.z, hi.x), r, bound, t);
			else { const float4* p = sc.prim(prim); h = triangle_t(p[0], p[1], p[2], r, bound, t); }
			if (COUNT && live) tc->prims++;
			if (live && h && (best < 0 || !(best_t < t))) { best_t = t; best = prim; }
		}
	}
	t_out = best_t; prim_out = best;
	return best >= 0;
}

// Tight binned-SAH BVH2 (host/bvh_build.cpp), nearest-child-first with a short stack and distance
// culling.  It must return what the reference's exhaustive walk returns, i.e. the minimum-t hit and,
// among equal t, the surface that comes LAST in the reference tree's DFS leaf order (tie rank, kept
// in prim_meta.w).  Exactness argument (DESIGN.md "fast BVH"): primitives are tested with the same
// arithmetic as the reference (prim_t, bound 1e5); a subtree is skipped only if its PADDED box is
// missed, lies behind the origin, or starts strictly beyond the best t so far (+ a relative guard
// band), none of which can hide a primitive hit with t <= best; NaN slabs (origin on a slab plane,
// d == 0 on that axis) drop out of fminf/fmaxf, i.e. count as "inside".  What it does NOT reproduce is
// the reference losing a true hit because one of its own origin-inflated ancestor boxes rejects the
// ray by rounding; tests compare both walks on every named scene and on the synthetic ones.
// Traversal stack of the fast walk: the lowest S levels live in shared memory (column threadIdx.x of an
// [S][256] array: lanes at different depths still hit 32 different banks), deeper levels in the caller's local
// array.  A local-only stack (S = 0) sends every push/pop through L1, where it competes with the node loads.
#ifndef PLU_SSTK
#define PLU_SSTK 8
#endif
#define PLU_WALK_CTA 256   // == PLU_CTA and k_trace_rays' block size
#ifndef PLU_PREFETCH_FAR
#define PLU_PREFETCH_FAR 0
#endif
// Capacity: both builders split by SAH down to depth 40 and halve the surface count from there on (host/bvh_build.cpp,
// bvh_gpu.cu), so a tree over n surfaces is at most 41 + ceil(log2 n) <= PLU_WALK_DEPTH_MAX levels deep for every int32
// n, and the walks push at most one entry per level; plu_gpu_upload_scene re-checks the bound against the built tree.
#define PLU_WALK_DEPTH_MAX 72
template <int S> struct walk_stack {
	int32_t* sm;        // &shared[0][threadIdx.x] (unused when S == 0)
	int32_t loc[PLU_WALK_DEPTH_MAX - S];
	PLU_DEV void push(int sp, int32_t v) { if (S > 0 && sp < S) sm[sp * PLU_WALK_CTA] = v; else loc[sp - S] = v; }
	PLU_DEV int32_t pop(int sp) const { return (S > 0 && sp < S) ? sm[sp * PLU_WALK_CTA] : loc[sp - S]; }
};

// The walk as a resumable state machine: k_extend_pt (wavefront.cuh) keeps one per lane across refills, trace_fast
// runs one to completion.  cur: >= 0 inner node, < 0 leaf reference, DONE when the stack is exhausted.
struct fast_walker {
	static constexpr int32_t DONE = (int32_t)0x80000000;
	unsigned long long ex2, ey2, ez2, ix2, iy2, iz2;
	float best_t; int32_t best, best_rank, cur; int sp;
	// the stack itself (walk_stack) is a separate object of the caller, so that these scalars stay in registers

	template <bool COUNT>
	PLU_DEV void begin_walk(const scene_ref& sc, const ray& r, trace_counts*) { begin(sc, r); }
	PLU_DEV void begin(const scene_ref& sc, const ray& r) {
		const float ix = 1.f / r.d.x, iy = 1.f / r.d.y, iz = 1.f / r.d.z;
		ex2 = pk2(r.e.x, r.e.x); ey2 = pk2(r.e.y, r.e.y); ez2 = pk2(r.e.z, r.e.z);
		ix2 = pk2(ix, ix); iy2 = pk2(iy, iy); iz2 = pk2(iz, iz);
		best_t = 3.0e38f; best = -1; best_rank = -1; sp = 0;
		cur = sc.n_fast_nodes > 0 ? 0 : DONE;
	}
	// one inner node (cur >= 0): both children in one pass, nearer child first, the other on the stack
	template <bool COUNT, class STK>
	PLU_DEV void node_step(const fast_node* nodes, STK& stack, trace_counts* tc) {
		const fast_node* nd = nodes + cur;
		float4 a = __ldg(&nd->a), b = __ldg(&nd->b), c = __ldg(&nd->c); int4 d = __ldg(&nd->d);
		if (COUNT) tc->nodes++;
		float cull = best_t + fabsf(best_t) * 1e-5f;
		// (child 0, child 1) pairs: a = min.x min.y, b = min.z max.x, c = max.y max.z  (plugpu.cu: upload)
		float lx0, lx1, ly0, ly1, lz0, lz1, hx0, hx1, hy0, hy1, hz0, hz1;
		upk2(submul2(pk2(a.x, a.y), ex2, ix2), lx0, lx1); upk2(submul2(pk2(a.z, a.w), ey2, iy2), ly0, ly1);
		upk2(submul2(pk2(b.x, b.y), ez2, iz2), lz0, lz1); upk2(submul2(pk2(b.z, b.w), ex2, ix2), hx0, hx1);
		upk2(submul2(pk2(c.x, c.y), ey2, iy2), hy0, hy1); upk2(submul2(pk2(c.z, c.w), ez2, iz2), hz0, hz1);
		float n0 = fmax3f(fminf(lx0, hx0), fminf(ly0, hy0), fminf(lz0, hz0)), f0 = fmin3f(fmaxf(lx0, hx0), fmaxf(ly0, hy0), fmaxf(lz0, hz0));
		float n1 = fmax3f(fminf(lx1, hx1), fminf(ly1, hy1), fminf(lz1, hz1)), f1 = fmin3f(fmaxf(lx1, hx1), fmaxf(ly1, hy1), fmaxf(lz1, hz1));
		bool h0 = !(f0 < n0) && !(f0 < 0.f) && !(n0 > cull);
		bool h1 = !(f1 < n1) && !(f1 < 0.f) && !(n1 > cull);
		if (h0 && h1) {
			bool first1 = n1 < n0;
			const int32_t far = first1 ? d.x : d.y;
			stack.push(sp++, far);
#if PLU_PREFETCH_FAR
			// experiment (profiles/r02_summary.md 6): request the stacked child now, so that it is in L1 / L2 when the walk returns to it
			if (far >= 0) { const char* q = (const char*)(nodes + far); asm volatile("prefetch.global.L2 [%0];" ::"l"(q)); }
#endif
			cur = first1 ? d.y : d.x;
		} else if (h0) cur = d.x;
		else if (h1) cur = d.y;
		else { if (sp > 0) cur = stack.pop(--sp); else cur = DONE; }
	}
	// one leaf: its <= 4 surface records
	template <bool COUNT>
	PLU_DEV void leaf_test(const float4* leaves, const ray& r, int32_t ref, trace_counts* tc) {
		const int32_t first = (~ref) >> 3, count = (~ref) & 7;
		for (int32_t k = 0; k < count; ++k) {
			const float4* rec = leaves + 3 * (size_t)(first + k);
			float4 a = __ldg(rec), b = __ldg(rec + 1);
			const int32_t kr = __float_as_int(b.w), kind = (int32_t)((uint32_t)kr >> 30), rank = kr & 0x3fffffff;
			float t; bool h; int32_t id;
			if (COUNT) tc->prims++;
			if (kind == PLU_PRIM_BOX) { id = __float_as_int(a.w); h = box_t(a, b, r, PLU_T_MAX, t); }
			else if (kind == PLU_PRIM_SPHERE) { id = __float_as_int(b.x); h = sphere_t(a, r, PLU_T_MAX, t); }
			else { id = __float_as_int(a.w); h = triangle_t(a, b, __ldg(rec + 2), r, PLU_T_MAX, t); }
			if (h && (best < 0 || t < best_t || (t == best_t && rank > best_rank))) { best_t = t; best = id; best_rank = rank; }
		}
	}
	// One round of the "while-while" walk: descend to the next leaf, then test it.  (Postponing the leaf while the
	// lane keeps descending to a second one — "speculative traversal" — was measured: 6 % more nodes, 14 % more
	// surface tests through the stale best t, 3 % slower; profiles/r01_summary.md.)
	// any_hit: the caller only asks WHETHER a surface is hit (point-light visibility, renderer.cpp:17 with
	// is_delta()): the first hit ends the walk.  Whether a hit exists does not depend on the visiting order, and
	// distance culling only starts after a first hit, so the answer is the closest-hit walk's.
	// keep >= 0: the walk was SEEDED with the hit of surface `keep` (seed(): an area light's own surface, whose hit distance the
	// caller computed with the same prim_t) and the caller only asks whether that surface is the closest hit (renderer.cpp:17,
	// :42): culling starts at the seed's distance and the walk ends as soon as anything replaces it under the same
	// (t, tie rank) rule — what replaced it need not be the closest hit, the answer "not the light" is already final.
	template <bool COUNT, class STK>
	PLU_DEV void round(const scene_ref& sc, const ray& r, STK& stack, trace_counts* tc, bool any_hit = false, int32_t keep = -1) {
		const fast_node* nodes = sc.fast_nodes;
		while (cur >= 0) node_step<COUNT>(nodes, stack, tc);
		if (cur == DONE) return;
		leaf_test<COUNT>(sc.fast_leaf, r, cur, tc);
		const bool stop = (any_hit && best >= 0) || (keep >= 0 && best != keep);
		if (sp > 0 && !stop) cur = stack.pop(--sp); else cur = DONE;
	}
	static constexpr bool CAN_SEED = true;
	PLU_DEV void seed(float t, int32_t prim, int32_t rank) { best_t = t; best = prim; best_rank = rank; }
};

template <bool COUNT, int S>
PLU_DEV bool trace_fast(const scene_ref& sc, const ray& r, float& t_out, int32_t& prim_out, trace_counts* tc, bool any_hit, int32_t light_prim = -1) {
	fast_walker w; w.begin(sc, r);
	int32_t keep = -1;
	if (light_prim >= 0) {   // "is the light's surface the closest hit": seeded walk, see fast_walker::round
		const int4 lm = sc.prim_meta(light_prim);
		float tl;
		if (prim_t(sc, light_prim, lm.x, r, PLU_T_MAX, tl)) { w.seed(tl, light_prim, lm.w); keep = light_prim; }
	}
	__shared__ int32_t s_stack[S > 0 ? S : 1][S > 0 ? PLU_WALK_CTA : 1];
	walk_stack<S> stack; stack.sm = &s_stack[0][S > 0 ? threadIdx.x : 0];
	// "while-while" walk: every lane first descends to its next leaf, then the warp tests leaves together;
	// with one if/else per step (r01f: 8.6 of 32 threads active per instruction) node and leaf work of
	// different lanes alternated serially
	while (w.cur != fast_walker::DONE) w.round<COUNT>(sc, r, stack, tc, any_hit, keep);
	t_out = w.best_t; prim_out = w.best;
	return w.best >= 0;
}

// ---- scenes with spheres: the walk that follows sphere::hit for ANY direction length ----------------------------
// sphere::hit assumes a unit direction (sphere.cpp:17-19: b = -dot(v, d), det = b*b - dot(v, v) + r*r) but the reference
// also feeds it NON-unit ones (its shading frames are not orthonormal, Q9).  With |d| = s, tau = the centre's offset along
// the ray and rho its distance from the ray's axis, the quadratic accepts the sphere iff the origin is outside it,
// tau > 0 and rho^2 <= r^2 + (s^2 - 1) tau^2: for s > 1 a CONE of half-angle atan(sqrt(s^2 - 1)) around the ray (spheres
// far off the ray "hit"), for s < 1 an ellipsoid around the origin.  No geometric BVH describes that set, and which of
// those spheres the reference finds also depends on ITS tree: a leaf is reached only through ancestors whose
// origin-inflated boxes pass `tmax >= tmin` (bvh_tree.cpp:43-47), a real filter for hits that are not on the ray.
//
// trace_ref_pruned therefore walks the REFERENCE tree (same topology, same inflated boxes, same exact slab test, hence
// the same reachable leaves) and adds to every interior node the TIGHT box of its surfaces (walk_node, built at upload)
// to skip what cannot matter: a subtree is entered iff the reference enters it AND its tight box passes
//   * the slab test against the ray (boxes, triangles: their tests are scale-invariant), or
//   * the sphere criterion: the ray hits the box grown by the cone's width at the box's far end (general_walker),
// and its lower bound of the hit distance does not exceed the best hit so far: the slab entry distance, or for spheres
// i1 = b - sqrt(b^2 - q) = q / (b + sqrt(b^2 - q)) >= q / (2 b) >= (dist^2 - rmax^2) / (2 max w.d), q = |w|^2 - r^2.
// Children are visited nearest bound first.  Candidates are tested with the unchanged sphere_t / box_t / triangle_t;
// equal distances resolve to the later leaf in DFS order (bvh_tree.cpp:73).  Same (t, surface) as trace_ref, including
// the rays the reference loses to its own inflated boxes (which a walk of the SAH tree cannot reproduce).
struct general_walker {
	f3 e, d, rrd, u;          // u = d / |d|
	float s, g2;              // |d|;  max(|d|^2 - 1, 0) + margin: growth of the accepted radius^2 per unit of tau^2
	float rmin, rmax2, best_t; int32_t best, best_rank;

	PLU_DEV void begin(const ray& r, float rmin_, float rmax) {
		e = r.e; d = r.d; rrd = 1.f / r.d;   // the value slabs() computes at every node (cmmn.h:155)
		const float l2 = d.x * d.x + d.y * d.y + d.z * d.z;
		s = sqrtf(l2);
		// the reference evaluates det = b*b - dot(v,v) + r*r in float: its rounding (~1e-7 of b*b + |v|^2) acts like a
		// direction that is up to ~2e-7 too long, so even "unit" rays get that much cone
		g2 = fmaxf(l2 - 1.f, 0.f) + 5e-7f;
		const float inv = 1.f / s;
		u = mk3(d.x * inv, d.y * inv, d.z * inv);
		rmin = rmin_; rmax2 = rmax < 0.f ? -1.f : rmax * rmax;
		best_t = 0.f; best = -1; best_rank = -1;
	}
	// One tight box (lo, hi, rb = half its diagonal): lower bound of every acceptable hit distance inside it, or +inf when
	// nothing in it can be accepted.  A sphere (c, r) is accepted only if the ray passes within R = sqrt(r^2 + g2 tau^2)
	// of c at tau > 0, i.e. hits the sphere's own box c -+ r grown by R - r <= sqrt(rmin^2 + g2 tau_max^2) - rmin: ONE
	// slab test on the grown box serves spheres (for |d| <= 1: no growth beyond the rounding margin), the plain one
	// boxes and triangles.
	PLU_DEV float child_key(float lx, float ly, float lz, float hx, float hy, float hz, float rb, float cull) const {
		const float wlx = lx - e.x, wly = ly - e.y, wlz = lz - e.z, whx = hx - e.x, why = hy - e.y, whz = hz - e.z;
		// slab test (boxes, triangles): entry / exit distance; NaN planes (0 * inf) drop out of fminf / fmaxf
		const float ax = wlx * rrd.x, bx = whx * rrd.x, ay = wly * rrd.y, by = why * rrd.y, az = wlz * rrd.z, bz = whz * rrd.z;
		const float n = fmax3f(fminf(ax, bx), fminf(ay, by), fminf(az, bz)), f = fmin3f(fmaxf(ax, bx), fmaxf(ay, by), fmaxf(az, bz));
		float key = INFINITY;
		if (!(f < n) && !(f < 0.f) && !(n > cull)) key = n;
		if (rmax2 < 0.f) return key;   // no spheres in the scene
		const float mx = 0.5f * (wlx + whx), my = 0.5f * (wly + why), mz = 0.5f * (wlz + whz);
		const float tau = mx * u.x + my * u.y + mz * u.z;
		const float t_hi = tau + rb, t_lo = tau - rb;   // range of tau over the centres in the box (via its bounding ball)
		bool sph; float bmax = s * t_hi;
		if (!(s > 0.f) || !(s < 3.0e38f)) { sph = true; bmax = 0.f; }   // zero / non-finite direction: no pruning
		else if (!(t_hi > 0.f)) sph = false;                            // every centre lies behind the origin: b <= 0
		else {
			const float pad = (sqrtf(rmin * rmin + g2 * (t_hi * t_hi)) - rmin) * 1.01f + 1e-5f * (fabsf(t_hi) + rb);
			const float px = pad * fabsf(rrd.x), py = pad * fabsf(rrd.y), pz = pad * fabsf(rrd.z);   // the grown box's planes move by pad / |d_k|
			const float n2 = fmax3f(fminf(ax, bx) - px, fminf(ay, by) - py, fminf(az, bz) - pz);
			const float f2 = fmin3f(fmaxf(ax, bx) + px, fmaxf(ay, by) + py, fmaxf(az, bz) + pz);
			sph = !(f2 < n2) && !(f2 < 0.f);
		}
		if (sph) {
			const float gx = fmaxf(fmaxf(wlx, -whx), 0.f), gy = fmaxf(fmaxf(wly, -why), 0.f), gz = fmaxf(fmaxf(wlz, -whz), 0.f);
			const float q = gx * gx + gy * gy + gz * gz - rmax2;   // <= |w|^2 - r^2 of every sphere in the box
			float lb = (q > 0.f && bmax > 0.f) ? (q / (2.f * bmax)) * 0.999f : 0.f;   // i1 = q / (b + sqrt(b^2 - q)) >= q / (2 b)
			// and, i1 growing with rho, i1 >= s tau - sqrt((s^2 - 1) tau^2 + r^2), which itself grows with tau: the value at t_lo
			if (t_lo > 0.f) lb = fmaxf(lb, (s * t_lo - sqrtf(g2 * (t_lo * t_lo) + rmax2)) * 0.999f - 1e-5f * t_lo);
			if (!(lb == lb)) lb = 0.f;
			if (!(lb > cull)) key = fminf(key, lb);
		}
		return key;
	}
};

// The walk as a resumable state machine (like fast_walker): k_extend_pt keeps one per lane across refills — the cost of
// these rays is heavy-tailed (a cone ray whose near candidates the reference's own boxes reject looks at thousands of
// surfaces, the median ray at a dozen), so one ray per lane per iteration would idle most of a warp.
struct pruned_walker : general_walker {
	static constexpr int32_t DONE = (int32_t)0x80000000;
	int32_t cur; int sp;

	// one child of an interior node: a leaf is tested at once (the reference tests leaves without a box, bvh_tree.cpp:40),
	// an interior child yields its key (+inf: the reference does not enter it, or nothing in it can improve the result)
	template <bool COUNT>
	PLU_DEV float visit(const scene_ref& sc, const ray& r, int32_t node, int32_t k, float cull, trace_counts* tc) {
		const walk_node* nd = sc.walk_nodes + node;
		const float4 a = __ldg(&nd->a), b = __ldg(&nd->b);
		if (k == 1) {
			const int32_t tag = __float_as_int(a.w), prim = __float_as_int(b.w);
			float t; bool h;
			if (COUNT) tc->prims++;
			if (tag == PLU_PRIM_BOX) h = box_t(a, b, r, PLU_T_MAX, t);
			else if (tag == PLU_PRIM_SPHERE) h = sphere_t(make_float4(a.x, a.y, a.z, b.x), r, PLU_T_MAX, t);
			else { const float4* p = sc.prim(prim); h = triangle_t(p[0], p[1], p[2], r, PLU_T_MAX, t); }   // (the blob may be staged in shared memory: no __ldg)
			if (h) {
				const int32_t rank = __float_as_int(__ldg(&nd->d).w);
				if (best < 0 || t < best_t || (t == best_t && rank > best_rank)) { best_t = t; best = prim; best_rank = rank; }
			}
			return INFINITY;
		}
		float tmin, tmax; slabs_rrd(mk3(a), mk3(b), r.e, rrd, tmin, tmax);   // the reference's own test on its own box
		if (!(tmax >= tmin)) return INFINITY;
		const float4 c = __ldg(&nd->c), d4 = __ldg(&nd->d);
		return child_key(c.x, c.y, c.z, d4.x, d4.y, d4.z, d4.w, cull);
	}
	template <bool COUNT>
	PLU_DEV void begin_walk(const scene_ref& sc, const ray& r, trace_counts* tc) {
		begin(r, sc.sphere_rmin, sc.has_spheres ? sc.sphere_rmax : -1.f);
		sp = 0; cur = 0;
		if (sc.n_prims == 1) { visit<COUNT>(sc, r, 0, 1, 3.0e38f, tc); cur = DONE; }
		else {   // the root is tested like any interior node
			float tmin, tmax; const walk_node* nd = sc.walk_nodes;
			slabs_rrd(mk3(__ldg(&nd->a)), mk3(__ldg(&nd->b)), r.e, rrd, tmin, tmax);
			if (!(tmax >= tmin)) cur = DONE;
		}
	}
	// one round: expand the interior node `cur` (both children: leaves are tested, interior ones ranked), move on
	// (never seeded: the reference reaches the light's leaf only through ancestors that pass ITS slab test, which the seed
	// would assume)
	static constexpr bool CAN_SEED = false;
	PLU_DEV void seed(float, int32_t, int32_t) {}
	template <bool COUNT, class STK>
	PLU_DEV void round(const scene_ref& sc, const ray& r, STK& stack, trace_counts* tc, bool any_hit = false, int32_t = -1) {
		if (cur == DONE) return;
		const int32_t k = __float_as_int(__ldg(&sc.walk_nodes[cur].c).w);
		const int32_t kl = k >> 1, L = cur + 1, R = cur + 2 * kl;
		if (COUNT) tc->nodes++;
		float cull = best >= 0 ? best_t + fabsf(best_t) * 1e-5f : 3.0e38f;
		const float k0 = visit<COUNT>(sc, r, L, kl, cull, tc);
		if (best >= 0) cull = best_t + fabsf(best_t) * 1e-5f;
		const float k1 = visit<COUNT>(sc, r, R, k - kl, cull, tc);
		if (any_hit && best >= 0) { cur = DONE; return; }
		const bool h0 = k0 < INFINITY, h1 = k1 < INFINITY;
		if (h0 && h1) {
			const bool first1 = k1 < k0;
			stack.push(sp++, first1 ? L : R);
			cur = first1 ? R : L;
		} else if (h0) cur = L;
		else if (h1) cur = R;
		else {
			// a stacked subtree may have fallen behind the best hit found since it was pushed: its children are ranked
			// again when it is expanded, so popping it unconditionally costs those two visits at most
			if (sp > 0) cur = stack.pop(--sp); else cur = DONE;
		}
	}
};

template <bool COUNT, int S>
PLU_DEV bool trace_ref_pruned(const scene_ref& sc, const ray& r, float& t_out, int32_t& prim_out, trace_counts* tc, bool any_hit) {
	pruned_walker w;
	__shared__ int32_t s_stack[S > 0 ? S : 1][S > 0 ? PLU_WALK_CTA : 1];
	walk_stack<S> stack; stack.sm = &s_stack[0][S > 0 ? threadIdx.x : 0];
	w.begin_walk<COUNT>(sc, r, tc);
	while (w.cur != pruned_walker::DONE) w.round<COUNT>(sc, r, stack, tc, any_hit);
	t_out = w.best_t; prim_out = w.best;
	return w.best >= 0;
}

// scene->hit.  Kernels of scenes staged in shared memory (SMEM: every named scene) only ever walk the reference's own
// small tree; the large-scene walks live in the global-memory kernels alone (PLU_RENDER_BVH_FAST runs a small scene
// through those kernels, which is how the tests exercise them on the reference's scenes) — keeping them out of the staged
// kernels keeps those at 64 registers without spills.
// WALKS: which large-scene walks the caller compiles in (a hot kernel is instantiated per walk so that it carries one).
#define PLU_WALKS_BOTH 0
#define PLU_WALKS_FAST 1     // sphere-free scene: the SAH tree
#define PLU_WALKS_PRUNED 2   // scene with spheres: the pruned reference tree
template <bool COUNT, bool SMEM, int WALKS = PLU_WALKS_BOTH>
// light_prim >= 0: the caller only asks whether that surface (an area light's) is the closest hit; walks that can answer
// without finding the closest hit do (the SAH tree), the others ignore it — the returned surface is the light's iff it is.
PLU_DEV bool trace_scene(const scene_ref& sc, const ray& r, float root_bound, float& t, int32_t& prim, trace_counts* tc, bool any_hit = false, int32_t light_prim = -1) {
	if (!SMEM && sc.fast_nodes != nullptr) {
		if (WALKS == PLU_WALKS_PRUNED || (WALKS == PLU_WALKS_BOTH && sc.has_spheres)) return trace_ref_pruned<COUNT, PLU_SSTK>(sc, r, t, prim, tc, any_hit);
		return trace_fast<COUNT, PLU_SSTK>(sc, r, t, prim, tc, any_hit, light_prim);
	}
	if (sc.n_ref_nodes <= PLU_UNIFORM_WALK_MAX_NODES) return trace_ref_uniform<COUNT, SMEM>(sc, r, root_bound, t, prim, tc);
	return trace_ref<COUNT>(sc, r, root_bound, t, prim, tc);
}

}  // namespace plud
