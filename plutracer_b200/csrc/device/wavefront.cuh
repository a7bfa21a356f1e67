// wavefront.cuh — the wavefront kernels that replace the per-pixel recursive path loop
// (reference src/renderer.cpp:59-150: render -> ray_color -> uniform_sample_one_light ->
// estimate_direct_light).
//
//   k_generate    camera::generate_ray for every path slot of the wave            (camera.h:25-37)
//   k_extend      scene->hit for the queued paths, compacts the hits              (renderer.cpp:61,86)
//   k_shade       Le, material -> bsdf, light pick + light/BSDF sampling (the NEE
//                 terms of estimate_direct_light up to the visibility test), BSDF
//                 continuation sample, throughput update; emits shadow-ray
//                 entries and the next extend queue                               (renderer.cpp:5-57,66-83)
//   k_connect     shadow / MIS connection rays, adds the NEE term to the path     (renderer.cpp:16-24,41-47)
//   k_accumulate  target->pixel(px) += L * (1/n^2) in sample order               (renderer.cpp:132)
//
// Paths live in slot-indexed structure-of-arrays buffers (128-bit rows); kernels are linked by
// queues of slot ids compacted with a warp ballot + popc prefix and ONE atomicAdd per warp.
// All kernels are persistent grid-stride loops whose trip count is read from a device counter, so a
// whole wave is enqueued without any host synchronisation.
#pragma once
#include "shade.cuh"

namespace plud {

#ifndef PLU_STAGED_WARP_SCOPE
#define PLU_STAGED_WARP_SCOPE 0   // bit 0: staged k_extend, bit 1: staged k_shade push per warp as well (build option; A/B in profiles/r02_summary.md 13)
#endif

// Exact division of a numerator below 2^31 by a wave-uniform divisor: q = umulhi(n, m) >> sh with m = floor(2^(31 + l) / d) + 1,
// l = ceil(log2 d), sh = l - 1 (the error m * d - 2^(31 + l) is at most d <= 2^l, so n * error < 2^(31 + l) for every n < 2^31);
// d == 1 is passed through.  Five instructions where the generic 32-bit division takes ~20; set by the host (plugpu.cu: set_div).
struct udiv32 { uint32_t d, m, sh; };
PLU_DEV uint32_t div_by(uint32_t n, const udiv32& v) { return v.d == 1u ? n : (__umulhi(n, v.m) >> v.sh); }

struct wave_desc {
	int32_t width, height;      // target->size
	int32_t n_strata;           // sample_count.x
	int32_t pix0, npix;         // pixel range of this wave (row-major; >= W*H: overflow-column strip, row = pix - W*H)
	int32_t k0, kcount;         // ordinals of the samples of this wave
	int32_t sample_begin, sample_stride;   // sample index s = begin + (k0 + k) * stride
	int32_t max_vertices;
	uint32_t seed_lo, seed_hi;
	uint32_t n_slots;           // npix * kcount
	int32_t use_lens;
	int32_t use_dirty;          // 1: the wave started with k_primary — rad[slot] is only valid where the slot's dirty bit is set
	udiv32 div_npix, div_width, div_strata;   // divisions by npix, width, n_strata
	float inv_strata;           // 1.f / (float)n_strata (sampler.h:60; an IEEE division, the same on the host)
};

struct wave_buffers {
	// Path state: one 64-byte record per slot = one DRAM access atom.  Queues hold slot ids, so state is
	// GATHERED by index; with one array per field every 16-byte gather dragged a 64-byte atom from HBM
	// (ncu r01e: k_shade read 340 B per vertex for 60 B of payload).
	//   rec[4*slot+0] = (ray origin, hit_record::t the next scene->hit starts from)
	//   rec[4*slot+1] = (ray direction, flags)
	//   rec[4*slot+2] = (path_throughput for even vertices, -)
	//   rec[4*slot+3] = (path_throughput for odd vertices, -)
	// path_throughput ping-pongs by vertex parity: k_connect of vertex v still needs T before the vertex.
	float4* rec;
	float4* rad;          // L
	uint32_t* dirty;      // one bit per slot (k_primary): 0 = the camera ray missed, L == 0 and rad[slot] was never written
	uint32_t* queue[2];   // extend queues, ping-pong
	// paths whose extension ray hit, by material class (CLS_DIFFUSE, CLS_SPECULAR): (slot, t, surface, -).  The hit travels IN
	// the queue entry (written and read as coalesced 16-byte rows) instead of being scattered into the path record, where
	// two 4-byte stores per ray cost a read-for-ownership of a DRAM sector (ncu r01: k_extend moved 2.7x its bytes)
	uint4* q_hit[2];
	float4 *sh_p, *sh_w1, *sh_c1, *sh_w2, *sh_c2;   // shadow entries (one per shaded vertex that needs a connection)
	uint32_t* counters;   // [bounce*CNT_STRIDE + {CNT_EXTEND, CNT_SHADOW, CNT_HIT_DIFFUSE, CNT_HIT_SPECULAR}]
	unsigned long long* stats;   // 0 extend rays, 1 shadow rays, 2 nodes visited, 3 prims tested, 4 shaded vertices, 5 shadow entries, 6 graph loop rounds
	// The wave's descriptor lives in DEVICE memory (copied there ahead of the wave), not in kernel parameters: the launch
	// sequence of a wave is then identical for every wave of a render and is replayed as ONE CUDA graph launch per wave
	// (plugpu.cu: wave_graph) instead of ~60 driver calls — short-path scenes were bound by the host's launch rate.
	const wave_desc* desc;
};

// one 128-byte line per bounce, one 32-byte sector per counter a kernel adds to; the two hit counters are
// the halves of one 64-bit word so that k_extend adds to both with a single atomic
#define CNT_STRIDE 32
#define CNT_EXTEND 0
#define CNT_SHADOW 8
#define CNT_HIT_DIFFUSE 16
#define CNT_HIT_SPECULAR 17

// Material classes: extend sorts the hits so that each shade launch runs one code path.
//   CLS_DIFFUSE   diffuse_material: NEE (light + BSDF sampling, MIS) and a Lambert continuation
//   CLS_SPECULAR  perfect_reflection / perfect_refraction / glass: every component is specular, so
//                 bsdf::F == 0 and bsdf::sampleF(all & ~specular) has no matching component: the NEE
//                 term is identically zero (renderer.cpp:11-12,31-32) and only the continuation is left
//   emission      empty bsdf: Le, then the path ends (sampleF: pdf = 0) — done inside k_extend
#define CLS_DIFFUSE 0
#define CLS_SPECULAR 1

#define FLAG_SPEC 1      // spec_bounce (renderer.cpp:79)
#define FLAG_NONFINITE 2 // path_throughput holds an inf/NaN: `L += T * 0` terms are no longer no-ops

#define SH_NEED1 1u
#define SH_NEED2 2u
#define SH_DELTA 4u


PLU_DEV void slot_to_path(const wave_desc& w, uint32_t slot, philox_key& key, int32_t& px, int32_t& py, int32_t& s) {
	const uint32_t k = div_by(slot, w.div_npix), q = slot - k * (uint32_t)w.npix;
	int32_t pix = w.pix0 + (int32_t)q;
	s = w.sample_begin + (w.k0 + (int32_t)k) * w.sample_stride;
	int32_t wh = w.width * w.height;
	if (pix >= wh) { px = w.width; py = pix - wh; } else { py = (int32_t)div_by((uint32_t)pix, w.div_width); px = pix - py * w.width; }
	key.pixel = (uint32_t)pix; key.sample = (uint32_t)s; key.k0 = w.seed_lo; key.k1 = w.seed_hi;
}

// Queue pushes are aggregated in two levels.  ncu (profiles/r01b): with one global atomicAdd per warp
// the ~1.2 G atomics/s that all land on the same counter paced k_extend (the shuffle that consumes the
// atomic's result still stalled when it was issued a whole loop iteration earlier).  Now: warp ballot +
// popc prefix -> one SHARED-memory atomic per warp -> one global atomic per CTA and loop iteration,
// issued by thread 0 after a barrier and consumed one barrier later, so its round trip overlaps the
// next block of work.  warp_tally() is the warp part; the CTA part is written out in each kernel.
struct warp_tally {
	uint32_t ofs;    // this lane's offset inside the CTA's batch (valid where pred was set)
	PLU_DEV void add(uint32_t* s_counter, bool pred) {
		unsigned mask = __ballot_sync(0xffffffffu, pred);
		unsigned lane = threadIdx.x & 31u;
		uint32_t w = 0;
		if (mask != 0) {
			if (lane == 0) w = atomicAdd(s_counter, (uint32_t)__popc(mask));
			w = __shfl_sync(0xffffffffu, w, 0);
		}
		ofs = w + (uint32_t)__popc(mask & ((1u << lane) - 1u));
	}
};

// Per-thread asynchronous gathers into shared memory (LDGSTS): the inputs of the NEXT loop iteration
// are fetched while the current one is traced / shaded, without holding registers.  Each thread only
// ever reads the rows it copied itself, so cp.async.wait_group is the only synchronisation needed.
PLU_DEV void cp_async_16(void* smem_dst, const void* gsrc) {
	unsigned s = (unsigned)__cvta_generic_to_shared(smem_dst);
	asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(s), "l"(gsrc) : "memory");
}
PLU_DEV void cp_async_8(void* smem_dst, const void* gsrc) {
	unsigned s = (unsigned)__cvta_generic_to_shared(smem_dst);
	asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(s), "l"(gsrc) : "memory");
}
PLU_DEV void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N> PLU_DEV void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

// Path vertex of a launch: a literal index, or — inside the while loop of a wave's CUDA graph, where every iteration runs
// the same kernel nodes — the loop counter in device memory (row max_vertices + 1 of the counters, cleared with them).
#define PLU_B_LOOP (-1)        // the vertex the loop is at
#define PLU_B_LOOP_NEXT (-2)   // the one after it (the extension that follows shade + connect)
PLU_DEV int wave_vertex(const wave_buffers& wb, int b) {
	if (b >= 0) return b;
	return (int)wb.counters[(wb.desc->max_vertices + 1) * CNT_STRIDE] + (-1 - b);
}

#define PLU_CTA 256
#ifndef PLU_ANYHIT
#define PLU_ANYHIT 1   // 1: point-light shadow rays on the fast BVH stop at their first hit (k_connect)
#endif
// staging bytes per kernel (two stages of per-thread rows), added to the scene blob in dynamic shared memory
#define PLU_STAGE_EXTEND (2 * 2 * PLU_CTA * 16)
#define PLU_STAGE_SHADE_IN (2 * 4 * PLU_CTA * 16)
#define PLU_STAGE_SHADE (PLU_STAGE_SHADE_IN + 2 * 5 * PLU_CTA * 16)   // + two stages of shadow-entry payload
#define PLU_STAGE_CONNECT (2 * 3 * PLU_CTA * 16)

extern __shared__ uint4 plu_smem[];

// ---- k_generate -----------------------------------------------------------------------------------------------
// stratified_sampler::generate_samples for one sample (sampler.h:58-69): stratum (s % n, s / n),
// px = pixel + ((sx,sy) + xi*0.999)/n, lens likewise; then camera::generate_ray.  Only waves that extend with k_extend_pt
// (fast BVH, persistent warps) start here; all others start with k_primary (k_extend<.., PRIMARY>).
PLU_DEV ray primary_ray(const scene_ref& sc, const wave_desc& w, uint32_t slot);
static __global__ void __launch_bounds__(256) k_generate(const __grid_constant__ scene_ref sc_in, const __grid_constant__ wave_buffers wb) {
	const wave_desc w = *wb.desc;
	for (uint32_t slot = blockIdx.x * blockDim.x + threadIdx.x; slot < w.n_slots; slot += gridDim.x * blockDim.x) {
		ray r = primary_ray(sc_in, w, slot);
		wb.rec[4 * (size_t)slot] = make_float4(r.e.x, r.e.y, r.e.z, PLU_T_MAX);
		wb.rec[4 * (size_t)slot + 1] = make_float4(r.d.x, r.d.y, r.d.z, __int_as_float(0));
		wb.rad[slot] = make_float4(0.f, 0.f, 0.f, 0.f);
	}
}

// What k_extend does with the outcome of scene->hit for one path: returns the material class whose hit queue the path
// joins with (slot, t, surface) (0: the path ends here).
//   emission_material: the vertex adds Le (first vertex / after a specular bounce, renderer.cpp:66); its
//   empty bsdf gives NEE = 0 and sampleF pdf = 0, i.e. the path ends (renderer.cpp:78)
//   miss: renderer.cpp:87-90, after a specular bounce `L += T * l->Le(r)` with Le == 0 for every light —
//   a no-op unless T is not finite
PLU_DEV int extend_finish(const scene_ref& sc, float4* recs, float4* rad, int bounce, uint32_t slot, const ray& r, int flags,
                          bool hit, float t, int32_t prim, float nl) {
	if (hit) {
		int4 meta = sc.prim_meta(prim);   // kind, material, light, tie rank
		int32_t mtag = sc.material(meta.y).tag;
		if (mtag == PLU_MAT_EMISSION) {
			bool le = bounce == 0 || (flags & FLAG_SPEC);
			if (le || (flags & FLAG_NONFINITE)) {
				f3 T = bounce == 0 ? mk3(1.f) : mk3(recs[4 * (size_t)slot + 2 + (bounce & 1)]);
				f3 L = mk3(rad[slot]);
				if (le) {
					hit_rec hr; fill_record(sc, prim, meta.x, r, t, false, hr);
					L = L + T * area_L(sc.light(meta.z), hr.norm, -r.d);
				}
				if (flags & FLAG_NONFINITE) L = L + T * (nl * mk3(0.f));
				rad[slot] = make_float4(L.x, L.y, L.z, 0.f);
			}
			return 0;
		}
		return mtag == PLU_MAT_DIFFUSE ? 1 : 2;   // the caller queues (slot, t, surface)
	}
	if ((flags & (FLAG_SPEC | FLAG_NONFINITE)) == (FLAG_SPEC | FLAG_NONFINITE)) {
		f3 T = mk3(recs[4 * (size_t)slot + 2 + (bounce & 1)]); f3 L = mk3(rad[slot]);
		for (int32_t i = 0; i < sc.n_lights; ++i) L = L + T * mk3(0.f);
		rad[slot] = make_float4(L.x, L.y, L.z, 0.f);
	}
	return 0;
}

// stratified_sampler::generate_samples for one sample (sampler.h:58-69) + camera::generate_ray: the camera ray of a slot
PLU_DEV ray primary_ray(const scene_ref& sc, const wave_desc& w, uint32_t slot) {
	philox_key key; int32_t px_x, px_y, s;
	slot_to_path(w, slot, key, px_x, px_y, s);
	u4 o = philox_call(key, 0u);
	int32_t n = w.n_strata;
	const float dp = w.inv_strata;
	const int32_t sq = (int32_t)div_by((uint32_t)s, w.div_strata);
	float sx = (float)(s - sq * n), sy = (float)sq;
	f2 px = mk2((float)(uint32_t)px_x + (sx + u01(o.x) * 0.999f) * dp, (float)(uint32_t)px_y + (sy + u01(o.y) * 0.999f) * dp);
	f2 lens = mk2((sx + u01(o.z) * 0.999f) * dp, (sy + u01(o.w) * 0.999f) * dp);
	return generate_ray(sc.cam, px, lens, w.use_lens != 0);
}

// extend_finish for the camera ray (k_primary): nothing of the path exists in memory yet, so a miss writes NOTHING
// (renderer.cpp:61: `if (!scene->hit) return 0` — the slot's dirty bit stays 0 and k_accumulate adds 0 without
// reading), an emissive hit writes L = 0 + 1 * Le, any other hit writes the whole 64-byte record and L = 0.
PLU_DEV int primary_finish(const scene_ref& sc, float4* recs, float4* rad, uint32_t slot, const ray& r, bool hit, float t, int32_t prim, bool& dirty) {
	dirty = hit;
	if (!hit) return 0;
	int4 meta = sc.prim_meta(prim);
	int32_t mtag = sc.material(meta.y).tag;
	if (mtag == PLU_MAT_EMISSION) {
		hit_rec hr; fill_record(sc, prim, meta.x, r, t, false, hr);
		f3 L = mk3(0.f) + mk3(1.f) * area_L(sc.light(meta.z), hr.norm, -r.d);
		rad[slot] = make_float4(L.x, L.y, L.z, 0.f);
		return 0;
	}
	float4* rec = recs + 4 * (size_t)slot;   // (path_throughput is 1 at the first vertex: its quarters are not read there)
	rec[0] = make_float4(r.e.x, r.e.y, r.e.z, PLU_T_MAX);
	rec[1] = make_float4(r.d.x, r.d.y, r.d.z, __int_as_float(0));
	rad[slot] = make_float4(0.f, 0.f, 0.f, 0.f);
	return mtag == PLU_MAT_DIFFUSE ? 1 : 2;
}

// ---- k_extend -------------------------------------------------------------------------------------------------
// PRIMARY (k_primary): the first kernel of a wave — camera ray generation fused into the extension of bounce 0.  The ray
// of a slot is computed in registers and traced at once; it reaches memory only if it hits (k_generate + k_extend moved
// 48 + 32 bytes per sample through HBM for rays of which, on the open scenes, three in four hit nothing).
// Loop skeleton shared by extend / shade / connect: iteration i computes on rows that were gathered
// into shared memory by cp.async during iteration i-1, issues the gathers of iteration i+1 and loads
// the queue entry of iteration i+2; queue pushes started in iteration i are completed in iteration i+1.
template <bool SMEM, bool COUNT, bool PRIMARY = false>
__global__ void __launch_bounds__(PLU_CTA, 4) k_extend(const __grid_constant__ scene_ref sc_in, const __grid_constant__ wave_buffers wb, int bounce_arg) {
	const int bounce = PRIMARY ? 0 : wave_vertex(wb, bounce_arg);
	scene_ref sc = sc_in;
	float4* const recs = wb.rec;
	unsigned char* stage_base = (unsigned char*)plu_smem;
	wave_desc w;
	if (PRIMARY) w = *wb.desc;
	const uint32_t n = bounce == 0 ? wb.desc->n_slots : wb.counters[bounce * CNT_STRIDE + CNT_EXTEND];
	if (blockIdx.x * PLU_CTA >= n) return;   // a CTA without a first entry has no work: leave before staging the scene
	if (SMEM) { stage_scene(sc, (unsigned char*)plu_smem, sc.n_ref_nodes <= PLU_UNIFORM_WALK_MAX_NODES); stage_base += (sc.blob_bytes + 127u) & ~127u; }
	float4* st_o = (float4*)stage_base;            // [2][PLU_CTA]
	float4* st_d = st_o + 2 * PLU_CTA;             // [2][PLU_CTA]
	const uint32_t* q = wb.queue[bounce & 1];
	uint32_t* cnt_diffuse = wb.counters + bounce * CNT_STRIDE + CNT_HIT_DIFFUSE;
	uint32_t* cnt_specular = wb.counters + bounce * CNT_STRIDE + CNT_HIT_SPECULAR;
	if (blockIdx.x == 0 && threadIdx.x == 0) atomicAdd(wb.stats + 0, (unsigned long long)n);
	trace_counts tc; tc.nodes = 0; tc.prims = 0;
	const uint32_t tid = threadIdx.x, lane = tid & 31u, stride = gridDim.x * PLU_CTA;
	const float nl = (float)sc.n_lights;
	const bool ident = bounce == 0;

	uint32_t j = blockIdx.x * PLU_CTA + tid;
	uint32_t slot = 0, slot_next = 0;
	if (!PRIMARY && j < n) {
		slot = ident ? j : q[j];
		cp_async_16(st_o + tid, recs + 4 * (size_t)slot); cp_async_16(st_d + tid, recs + 4 * (size_t)slot + 1);
	}
	cp_async_commit();
	if (PRIMARY) { slot = j; slot_next = j + stride; }
	else if (j + stride < n) slot_next = ident ? j + stride : q[j + stride];
	// Push state of one GROUP (see warp_tally): the whole CTA when the scene is staged in shared memory (uniform
	// walk: every warp takes the same time, one barrier per iteration is free), a single warp otherwise
	// (large scenes: r01f had 34 % of k_extend's stall samples on the CTA barrier, rays of different length).
	// counts [3] rotate so that the leader can clear one while the warps already add to the next; s_base
	// is double-buffered for the same reason.
	constexpr bool WARP_SCOPE = !SMEM || (PLU_STAGED_WARP_SCOPE & 1);
	const uint32_t grp = WARP_SCOPE ? (tid >> 5) : 0u;
	const bool leader = WARP_SCOPE ? lane == 0 : tid == 0;
	__shared__ uint32_t s_cnt[PLU_CTA / 32][3][2];
	__shared__ uint2 s_base[PLU_CTA / 32][2];
	if (tid < (PLU_CTA / 32) * 6) (&s_cnt[0][0][0])[tid] = 0;
	__syncthreads();
	unsigned long long pending = 0;   // leader: result of the group's global atomic of the previous iteration
	uint32_t pslot = 0, p_ofs_d = 0, p_ofs_s = 0; bool p_d = false, p_s = false;
	uint32_t p_t = 0, p_prim = 0, c_t = 0, c_prim = 0;   // hit of the previous / current iteration (t as bits)
	uint32_t buf = 0, it = 0, k3 = 0;
	unsigned long long* cnt_pair = (unsigned long long*)cnt_diffuse;   // (diffuse, specular) = (low, high)
	for (uint32_t base = blockIdx.x * PLU_CTA; base < n; base += stride) {
		const uint32_t jn = j + stride, jnn = jn + stride;
		const bool has_next = jn < n;
		if (!PRIMARY && has_next) {
			cp_async_16(st_o + (buf ^ 1u) * PLU_CTA + tid, recs + 4 * (size_t)slot_next);
			cp_async_16(st_d + (buf ^ 1u) * PLU_CTA + tid, recs + 4 * (size_t)slot_next + 1);
		}
		if (!PRIMARY) cp_async_commit();
		uint32_t slot_nn = 0;
		if (PRIMARY) slot_nn = jnn;
		else if (has_next && jnn < n) slot_nn = ident ? jnn : q[jnn];
		if (!PRIMARY) cp_async_wait<1>();
		bool active = j < n, to_diffuse = false, to_specular = false;
		if (PRIMARY) {
			bool dirty = false;
			if (active) {
				ray r = primary_ray(sc, w, slot);
				float t; int32_t prim;
				bool hit = trace_scene<COUNT, SMEM>(sc, r, PLU_T_MAX, t, prim, &tc);
				int cls = primary_finish(sc, recs, wb.rad, slot, r, hit, t, prim, dirty);
				to_diffuse = cls == 1; to_specular = cls == 2; c_t = __float_as_uint(t); c_prim = (uint32_t)prim;
			}
			// slots of a warp are 32 consecutive, 32-aligned indices: one mask word per warp and iteration
			const uint32_t mask = __ballot_sync(0xffffffffu, dirty);
			if (lane == 0 && j < n) wb.dirty[j >> 5] = mask;
		} else if (active) {
			float4 o = st_o[buf * PLU_CTA + tid], d = st_d[buf * PLU_CTA + tid];
			ray r; r.e = mk3(o); r.d = mk3(d);
			int flags = __float_as_int(d.w);
			float t; int32_t prim;
			bool hit = trace_scene<COUNT, SMEM>(sc, r, o.w, t, prim, &tc);
			int cls = extend_finish(sc, recs, wb.rad, bounce, slot, r, flags, hit, t, prim, nl);
			to_diffuse = cls == 1; to_specular = cls == 2; c_t = __float_as_uint(t); c_prim = (uint32_t)prim;
		}
		warp_tally td, ts; td.add(&s_cnt[grp][k3][0], to_diffuse); ts.add(&s_cnt[grp][k3][1], to_specular);
		if (leader && it > 0) s_base[grp][(it - 1u) & 1u] = make_uint2((uint32_t)pending, (uint32_t)(pending >> 32));
		if (WARP_SCOPE) __syncwarp(); else __syncthreads();
		if (it > 0) {   // the previous iteration's entries: its global base has arrived meanwhile
			uint2 gb = s_base[grp][(it - 1u) & 1u];
			if (p_d) wb.q_hit[CLS_DIFFUSE][gb.x + p_ofs_d] = make_uint4(pslot, p_t, p_prim, 0u);
			if (p_s) wb.q_hit[CLS_SPECULAR][gb.y + p_ofs_s] = make_uint4(pslot, p_t, p_prim, 0u);
		}
		if (leader) {
			uint32_t cd = s_cnt[grp][k3][0], cs = s_cnt[grp][k3][1];
			pending = (cd | cs) ? atomicAdd(cnt_pair, ((unsigned long long)cs << 32) | cd) : 0ull;
			uint32_t kz = k3 == 0 ? 2u : k3 - 1u;   // (it + 2) % 3: idle until the iteration after next
			s_cnt[grp][kz][0] = 0; s_cnt[grp][kz][1] = 0;
		}
		p_d = to_diffuse; p_s = to_specular; p_ofs_d = td.ofs; p_ofs_s = ts.ofs; pslot = slot; p_t = c_t; p_prim = c_prim;
		slot = slot_next; slot_next = slot_nn; j = jn; buf ^= 1u;
		it++; k3 = k3 == 2 ? 0u : k3 + 1u;
	}
	if (it > 0) {
		if (leader) s_base[grp][(it - 1u) & 1u] = make_uint2((uint32_t)pending, (uint32_t)(pending >> 32));
		if (WARP_SCOPE) __syncwarp(); else __syncthreads();
		uint2 gb = s_base[grp][(it - 1u) & 1u];
		if (p_d) wb.q_hit[CLS_DIFFUSE][gb.x + p_ofs_d] = make_uint4(pslot, p_t, p_prim, 0u);
		if (p_s) wb.q_hit[CLS_SPECULAR][gb.y + p_ofs_s] = make_uint4(pslot, p_t, p_prim, 0u);
	}
	if (COUNT) {
		unsigned long long a = tc.nodes, b = tc.prims;
		for (int o = 16; o > 0; o >>= 1) { a += __shfl_down_sync(0xffffffffu, a, o); b += __shfl_down_sync(0xffffffffu, b, o); }
		if (lane == 0) { atomicAdd(wb.stats + 2, a); atomicAdd(wb.stats + 3, b); }
	}
}

// ---- k_extend_pt ----------------------------------------------------------------------------------------------
// (__launch_bounds__(.., 4) on the traversal kernels: 64 registers = 32 warps per SM; the walks are latency-bound and at 68
// registers / 24 warps the 1M-box scene ran 7 % slower, profiles/r02_summary.md)
// k_extend for scenes that walk the fast BVH: persistent warps that REFILL finished lanes.  With one ray per lane
// per iteration a warp runs as long as its longest ray (profiles/r01_summary.md: 10-13 of 32 threads active per
// instruction on the 1M-surface scene, coherent camera rays included); here a lane whose walk ends waits only
// until PT_REFILL lanes of its warp are idle, then those lanes retire their rays (extend_finish + queue push,
// one 64-bit atomic per round) and take the next entries of the warp's chunk of the queue (PT_CHUNK entries per
// atomic on the launch's work counter).  The walk state (fast_walker) lives across rounds.  Same per-ray
// arithmetic as k_extend; only the order of the queue entries differs.
#define CNT_WORK 24
#define PT_CHUNK 128u
#ifndef PT_REFILL
#define PT_REFILL 16u
#endif

#ifndef PLU_PT_BLOCKS
#define PLU_PT_BLOCKS 4   // resident CTAs per SM the persistent-warp kernels are compiled for (4: 64 registers)
#endif
template <bool COUNT, class WALKER = fast_walker>
__global__ void __launch_bounds__(PLU_CTA, PLU_PT_BLOCKS) k_extend_pt(const __grid_constant__ scene_ref sc, const __grid_constant__ wave_buffers wb, int bounce_arg, uint32_t refill) {
	constexpr uint32_t FULL = 0xffffffffu;
	const int bounce = wave_vertex(wb, bounce_arg);
	float4* const recs = wb.rec;
	const bool ident = bounce == 0;
	const uint32_t n = ident ? wb.desc->n_slots : wb.counters[bounce * CNT_STRIDE + CNT_EXTEND];
	const uint32_t* q = wb.queue[bounce & 1];
	unsigned long long* cnt_pair = (unsigned long long*)(wb.counters + bounce * CNT_STRIDE + CNT_HIT_DIFFUSE);   // (diffuse, specular)
	uint32_t* work = wb.counters + bounce * CNT_STRIDE + CNT_WORK;
	if (blockIdx.x == 0 && threadIdx.x == 0) atomicAdd(wb.stats + 0, (unsigned long long)n);
	trace_counts tc; tc.nodes = 0; tc.prims = 0;
	const uint32_t lane = threadIdx.x & 31u, lt = (1u << lane) - 1u;
	const float nl = (float)sc.n_lights;
	WALKER w; w.cur = WALKER::DONE; w.sp = 0; w.best = -1; w.best_t = 0.f; w.best_rank = -1;
	__shared__ int32_t s_stack[PLU_SSTK > 0 ? PLU_SSTK : 1][PLU_SSTK > 0 ? PLU_CTA : 1];
	walk_stack<PLU_SSTK> stack; stack.sm = &s_stack[0][PLU_SSTK > 0 ? threadIdx.x : 0];
	bool have = false, exhausted = false;
	uint32_t slot = 0; int flags = 0;
	ray r; r.e = mk3(0.f); r.d = mk3(0.f);
	uint32_t c_next = 0, c_end = 0;   // the warp's chunk of the queue (warp-uniform)
	for (;;) {
		// retire the lanes whose walk has ended
		int cls = 0;
		if (have && w.cur == WALKER::DONE) {
			cls = extend_finish(sc, recs, wb.rad, bounce, slot, r, flags, w.best >= 0, w.best_t, w.best, nl);
			have = false;
		}
		const uint32_t md = __ballot_sync(FULL, cls == 1), ms = __ballot_sync(FULL, cls == 2);
		if (md | ms) {
			unsigned long long base = 0;
			if (lane == 0) base = atomicAdd(cnt_pair, ((unsigned long long)__popc(ms) << 32) | (unsigned long long)__popc(md));
			base = __shfl_sync(FULL, base, 0);
			const uint4 entry = make_uint4(slot, __float_as_uint(w.best_t), (uint32_t)w.best, 0u);
			if (cls == 1) wb.q_hit[CLS_DIFFUSE][(uint32_t)base + __popc(md & lt)] = entry;
			if (cls == 2) wb.q_hit[CLS_SPECULAR][(uint32_t)(base >> 32) + __popc(ms & lt)] = entry;
		}
		// refill: idle lanes take consecutive entries of the chunk; a second pass opens the next chunk at a chunk's end
		for (int pass = 0; pass < 2 && !exhausted; ++pass) {
			const uint32_t idle = __ballot_sync(FULL, !have);
			if (!idle) break;
			if (c_next == c_end) {
				uint32_t b = 0;
				if (lane == 0) b = atomicAdd(work, PT_CHUNK);
				b = __shfl_sync(FULL, b, 0);
				if (b >= n) { exhausted = true; break; }
				c_next = b; c_end = min(b + PT_CHUNK, n);
			}
			const uint32_t idx = c_next + __popc(idle & lt);
			if (!have && idx < c_end) {
				slot = ident ? idx : q[idx];
				float4 o = recs[4 * (size_t)slot], d = recs[4 * (size_t)slot + 1];
				r.e = mk3(o); r.d = mk3(d); flags = __float_as_int(d.w);
				w.template begin_walk<COUNT>(sc, r, &tc);
				have = true;
			}
			c_next = min(c_next + (uint32_t)__popc(idle), c_end);
		}
		if (!__any_sync(FULL, have)) break;
		// walk until fewer than min_active lanes are still walking (to the end once the queue is exhausted)
		const uint32_t min_active = exhausted ? 1u : 33u - refill;
		for (;;) {
			w.template round<COUNT>(sc, r, stack, &tc);
			if ((uint32_t)__popc(__ballot_sync(FULL, w.cur != WALKER::DONE)) < min_active) break;
		}
	}
	if (COUNT) {
		unsigned long long a = tc.nodes, b = tc.prims;
		for (int o = 16; o > 0; o >>= 1) { a += __shfl_down_sync(FULL, a, o); b += __shfl_down_sync(FULL, b, o); }
		if (lane == 0) { atomicAdd(wb.stats + 2, a); atomicAdd(wb.stats + 3, b); }
	}
}

// ---- k_shade --------------------------------------------------------------------------------------------------
// One path vertex: everything renderer.cpp:66-83 does between two scene->hit calls, except the
// visibility tests of estimate_direct_light, which become shadow entries for k_connect.  One
// instantiation per material class; the queue it reads holds only vertices of that class.
#ifndef PLU_SHADE_BLOCKS
#define PLU_SHADE_BLOCKS 3   // resident CTAs per SM k_shade is compiled for (3: 80 registers; 4: 64 registers and spills, a build option)
#endif
template <bool SMEM, int CLS>
__global__ void __launch_bounds__(PLU_CTA, PLU_SHADE_BLOCKS) k_shade(const __grid_constant__ scene_ref sc_in, const __grid_constant__ wave_buffers wb, int bounce_arg) {
	const int bounce = wave_vertex(wb, bounce_arg);
	scene_ref sc = sc_in;
	const wave_desc w = *wb.desc;
	float4* const recs = wb.rec;
	unsigned char* stage_base = (unsigned char*)plu_smem;
	const uint32_t n = wb.counters[bounce * CNT_STRIDE + (CLS == CLS_DIFFUSE ? CNT_HIT_DIFFUSE : CNT_HIT_SPECULAR)];
	if (blockIdx.x * PLU_CTA >= n) return;   // no work for this CTA: leave before staging the scene
	if (SMEM) { stage_scene(sc, (unsigned char*)plu_smem); stage_base += (sc.blob_bytes + 127u) & ~127u; }
	float4* st_rec = (float4*)stage_base;          // [2][4][PLU_CTA]: record quarters 0, 1, the incoming throughput, the hit entry
	if (blockIdx.x == 0 && threadIdx.x == 0) atomicAdd(wb.stats + 4, (unsigned long long)n);
	const uint4* q_in = wb.q_hit[CLS];
	uint32_t* cnt_shadow = wb.counters + bounce * CNT_STRIDE + CNT_SHADOW;
	uint32_t* cnt_next = wb.counters + (bounce + 1) * CNT_STRIDE + CNT_EXTEND;
	uint32_t* q_next = wb.queue[(bounce + 1) & 1];
	const uint32_t tid = threadIdx.x, stride = gridDim.x * PLU_CTA;
	const float nl = (float)sc.n_lights;
	const uint32_t q_T_in = 2u + ((uint32_t)bounce & 1u), q_T_out = 2u + (((uint32_t)bounce + 1u) & 1u);

	auto gather = [&](uint32_t b, uint32_t sl, uint32_t jj) {   // ray + incoming throughput of the slot (one 64-byte record), and the hit entry
		float4* dst = st_rec + (size_t)b * 4 * PLU_CTA + tid;
		const float4* src = recs + 4 * (size_t)sl;
		cp_async_16(dst, src); cp_async_16(dst + PLU_CTA, src + 1); cp_async_16(dst + 2 * PLU_CTA, src + q_T_in);
		cp_async_16(dst + 3 * PLU_CTA, q_in + jj);
	};
	uint32_t j = blockIdx.x * PLU_CTA + tid;
	uint32_t slot = 0, slot_next = 0;
	if (j < n) { slot = q_in[j].x; gather(0, slot, j); }
	cp_async_commit();
	if (j + stride < n) slot_next = q_in[j + stride].x;
	// CTA-level push state.  Both queues of an iteration (shadow entries, next extend queue) are counted
	// into shared memory, reserved by ONE thread after the barrier and written in the FOLLOWING iteration;
	// the shadow payload waits in shared memory meanwhile (st_pay), not in registers.
	float4* st_pay = (float4*)(stage_base + PLU_STAGE_SHADE_IN);   // [2][5][PLU_CTA]: p|slot, w1|flags, c1, w2, c2
	constexpr bool WARP_SCOPE = !SMEM || (PLU_STAGED_WARP_SCOPE & 2);   // group = CTA for staged scenes, warp otherwise (see k_extend)
	const uint32_t lane = tid & 31u, grp = WARP_SCOPE ? (tid >> 5) : 0u;
	const bool leader = WARP_SCOPE ? lane == 0 : tid == 0;
	__shared__ uint32_t s_sh_cnt[PLU_CTA / 32][3], s_n_cnt[PLU_CTA / 32][3];
	__shared__ uint2 s_base[PLU_CTA / 32][2];   // (next-queue base, shadow-queue base)
	if (tid < (PLU_CTA / 32) * 3) { (&s_sh_cnt[0][0])[tid] = 0; (&s_n_cnt[0][0])[tid] = 0; }
	__syncthreads();
	uint32_t pending_sh = 0, pending_n = 0;   // leader
	uint32_t pslot = 0, p_ofs_n = 0, p_ofs_sh = 0, buf = 0, it = 0, k3 = 0; bool p_go = false, p_need = false;
	auto flush = [&](uint32_t prev) {   // entries of iteration `prev`
		uint2 gb = s_base[grp][prev & 1u];
		if (p_go) q_next[gb.x + p_ofs_n] = pslot;
		if (CLS == CLS_DIFFUSE && p_need) {
			const float4* row = st_pay + (size_t)(prev & 1u) * 5 * PLU_CTA + tid;
			uint32_t at = gb.y + p_ofs_sh;
			float4 w1 = row[PLU_CTA];
			wb.sh_p[at] = row[0]; wb.sh_w1[at] = w1; wb.sh_c1[at] = row[2 * PLU_CTA];
			if (__float_as_uint(w1.w) & SH_NEED2) { wb.sh_w2[at] = row[3 * PLU_CTA]; wb.sh_c2[at] = row[4 * PLU_CTA]; }
		}
	};
	for (uint32_t base = blockIdx.x * PLU_CTA; base < n; base += stride) {
		const uint32_t jn = j + stride, jnn = jn + stride;
		const bool has_next = jn < n;
		if (has_next) gather(buf ^ 1u, slot_next, jn);
		cp_async_commit();
		uint32_t slot_nn = 0;
		if (has_next && jnn < n) slot_nn = q_in[jnn].x;
		cp_async_wait<1>();
		bool active = j < n, need_entry = false, go_on = false;
		uint32_t sh_flags = 0;
		f3 p = mk3(0.f), wi1 = mk3(0.f), c1 = mk3(0.f), wi2 = mk3(0.f), c2 = mk3(0.f);
		f3 T = mk3(1.f), nrm = mk3(0.f), wo = mk3(0.f);
		float t = 0.f; int flags = 0;
		bsdf b; vertex_rng rng;
		b.tag = PLU_MAT_DIFFUSE;
		if (active) {
			const float4* row = st_rec + (size_t)buf * 4 * PLU_CTA + tid;
			float4 o4 = row[0], d4 = row[PLU_CTA], qT = row[2 * PLU_CTA], he = row[3 * PLU_CTA];
			ray r; r.e = mk3(o4); r.d = mk3(d4);
			flags = __float_as_int(d4.w);
			t = he.y; int32_t prim = __float_as_int(he.z);   // (slot, t, surface, -)
			int4 meta = sc.prim_meta(prim);   // kind, material, light (-1 here: emissive hits end in k_extend), tie rank
			const plu_material_desc& m = sc.material(meta.y);
			hit_rec hr; fill_record(sc, prim, meta.x, r, t, m.texture >= 0, hr);
			if (bounce != 0) T = mk3(qT);
			philox_key key; int32_t px_x, px_y, s;
			slot_to_path(w, slot, key, px_x, px_y, s);
			if (CLS == CLS_DIFFUSE) rng.begin(key, bounce);
			else rng.c = philox_call(key, 3u + 3u * (uint32_t)bounce);   // only the continuation draws are read
			p = ray_at(r, t);
			nrm = hr.norm; wo = -r.d;
			// renderer.cpp:66: `L += T * Le` with Le == 0 on a non-emissive surface: only observable when T is not finite
			if ((bounce == 0 || (flags & FLAG_SPEC)) && (flags & FLAG_NONFINITE)) {
				f3 L = mk3(wb.rad[slot]);
				L = L + T * mk3(0.f);
				wb.rad[slot] = make_float4(L.x, L.y, L.z, 0.f);
			}
			make_bsdf(sc, m, hr, b, meta.x == PLU_PRIM_BOX);
			if (CLS == CLS_DIFFUSE) b.tag = PLU_MAT_DIFFUSE;   // compile-time constant for everything below
			// renderer.cpp:53-57 + :5-51 up to the visibility tests
			if (CLS == CLS_DIFFUSE && sc.n_lights > 0) {
				int32_t li = (int32_t)gmin(floorf(rng.light_pick() * nl), nl - 1.f);
				const plu_light_desc& l = sc.light(li);
				bool is_delta = l.tag == PLU_LIGHT_POINT;
				int32_t ltag = is_delta ? 0 : sc.prim_meta(l.prim).x;
				f3 Li, wi; float light_pdf;
				if (is_delta) {   // point_light::sampleL, light.h:20-27
					f3 l2p = mk3(l.a[0], l.a[1], l.a[2]) - p;
					float len2 = dot(l2p, l2p);
					wi = l2p / sqrtf(len2);
					light_pdf = 1.f;
					Li = mk3(l.b[0], l.b[1], l.b[2]) / len2;
				} else {   // diffuse_area_light::sampleL, area_light.h:25-31
					f3 ln, ps = surface_sample(sc, l.prim, ltag, rng.light_u(), ln, rng);
					wi = normalize(ps - p);
					light_pdf = surface_pdf(sc, l.prim, ltag, l.area, p, wi);
					Li = area_L(l, ln, -wi);
				}
				if (light_pdf > 0.f && dot(Li, Li) > 0.f) {
					f3 f = b.F(nrm, wo, wi, BX_NEE_TYPES);
					if (dot(f, f) > 0.f) {
						sh_flags |= SH_NEED1; wi1 = wi;
						if (is_delta) c1 = f * Li * (fabsf(dot(wi, nrm)) / light_pdf);
						else {
							float bsdf_pdf = b.pdf(wo, wi, BX_NEE_TYPES);
							float wgt = (bsdf_pdf * bsdf_pdf) / (bsdf_pdf * bsdf_pdf + light_pdf * light_pdf);
							c1 = f * Li * (fabsf(dot(wi, nrm)) * wgt / light_pdf);
						}
					}
				}
				if (!is_delta) {   // BSDF-sampled MIS term, renderer.cpp:29-49
					uint32_t sampled = 0; float bsdf_pdf;
					f3 f = b.sampleF(rng, 0, nrm, wo, wi, bsdf_pdf, BX_NEE_TYPES, sampled);
					if (dot(f, f) > 0.f && bsdf_pdf > 0.f) {
						float wgt = 1.f; bool live = true;
						if (!(sampled & BX_SPECULAR)) {
							float lp = surface_pdf(sc, l.prim, ltag, l.area, p, wi);
							if (lp == 0.f) live = false;
							else wgt = (bsdf_pdf * bsdf_pdf) / (bsdf_pdf * bsdf_pdf + lp * lp);
						}
						if (live) {
							// Li = lihr.surf->mat->Le(p, n, -wi) uses the SHADING point's normal (renderer.cpp:43), so it
							// is known before the ray is cast; a zero Li makes the cast unobservable and it is skipped
							f3 Li2 = area_L(l, nrm, -wi);
							if (dot(Li2, Li2) > 0.f) {
								sh_flags |= SH_NEED2; wi2 = wi;
								c2 = f * Li2 * fabsf(dot(wi, nrm)) * wgt / bsdf_pdf;
							}
						}
					}
				}
				if (is_delta) sh_flags |= SH_DELTA;
				sh_flags |= (uint32_t)li << 8;
				need_entry = (sh_flags & (SH_NEED1 | SH_NEED2)) != 0;
			}
			if (!need_entry && (flags & FLAG_NONFINITE)) {   // L += T * (nl * 0): only observable when T is not finite
				f3 L = mk3(wb.rad[slot]);
				L = L + T * (nl * mk3(0.f));
				wb.rad[slot] = make_float4(L.x, L.y, L.z, 0.f);
			}
			if (CLS == CLS_DIFFUSE && need_entry) {   // park the shadow entry in shared memory until its queue position is known
				float4* row = st_pay + (size_t)(it & 1u) * 5 * PLU_CTA + tid;
				row[0] = make_float4(p.x, p.y, p.z, __uint_as_float(slot));
				row[PLU_CTA] = make_float4(wi1.x, wi1.y, wi1.z, __uint_as_float(sh_flags));
				row[2 * PLU_CTA] = make_float4(c1.x, c1.y, c1.z, 0.f);
				if (sh_flags & SH_NEED2) {
					row[3 * PLU_CTA] = make_float4(wi2.x, wi2.y, wi2.z, 0.f);
					row[4 * PLU_CTA] = make_float4(c2.x, c2.y, c2.z, 0.f);
				}
			}
			// renderer.cpp:76-83: continuation
			if (bounce < w.max_vertices - 1) {
				uint32_t types = 0; float pdf; f3 wi;
				f3 f = b.sampleF(rng, 1, nrm, wo, wi, pdf, BX_ALL, types);
				if (!(dot(f, f) == 0.f || pdf == 0.f)) {
					T = T * (f * fabsf(dot(wi, nrm)) / pdf);
					int nf = ((types & BX_SPECULAR) ? FLAG_SPEC : 0) | (finite3(T) ? 0 : FLAG_NONFINITE);
					float4* rec = recs + 4 * (size_t)slot;
					rec[0] = make_float4(p.x, p.y, p.z, t);
					rec[1] = make_float4(wi.x, wi.y, wi.z, __int_as_float(nf));
					rec[q_T_out] = make_float4(T.x, T.y, T.z, 0.f);
					go_on = true;
				}
			}
		}
		// count this iteration's batches; publish the bases of the previous one; ONE barrier
		warp_tally tsh, tn; tsh.ofs = 0;
		if (CLS == CLS_DIFFUSE) tsh.add(&s_sh_cnt[grp][k3], need_entry);
		tn.add(&s_n_cnt[grp][k3], go_on);
		if (leader && it > 0) s_base[grp][(it - 1u) & 1u] = make_uint2(pending_n, pending_sh);
		if (WARP_SCOPE) __syncwarp(); else __syncthreads();
		if (it > 0) flush(it - 1u);
		if (leader) {   // the global atomics travel during the next iteration
			uint32_t cn = s_n_cnt[grp][k3], cs = CLS == CLS_DIFFUSE ? s_sh_cnt[grp][k3] : 0u;
			pending_n = cn ? atomicAdd(cnt_next, cn) : 0u;
			pending_sh = cs ? atomicAdd(cnt_shadow, cs) : 0u;
			uint32_t kz = k3 == 0 ? 2u : k3 - 1u;   // (it + 2) % 3
			s_n_cnt[grp][kz] = 0; s_sh_cnt[grp][kz] = 0;
		}
		p_go = go_on; p_ofs_n = tn.ofs; p_need = need_entry; p_ofs_sh = tsh.ofs; pslot = slot;
		slot = slot_next; slot_next = slot_nn; j = jn; buf ^= 1u;
		it++; k3 = k3 == 2 ? 0u : k3 + 1u;
	}
	if (it > 0) {
		if (leader) s_base[grp][(it - 1u) & 1u] = make_uint2(pending_n, pending_sh);
		if (WARP_SCOPE) __syncwarp(); else __syncthreads();
		flush(it - 1u);
	}
}

#ifndef PLU_SEED_AREA
#define PLU_SEED_AREA 1   // 1: area-light shadow rays on the SAH tree end at the first surface that beats the light's own hit
#endif
// ---- k_connect ------------------------------------------------------------------------------------------------
// The scene->hit calls of estimate_direct_light (renderer.cpp:16,41) and the radiance update
// `L += path_throughput * (nlights * Ld)` (renderer.cpp:56,74).
template <bool SMEM, bool COUNT, int WALKS = PLU_WALKS_BOTH>
__global__ void __launch_bounds__(PLU_CTA, 4) k_connect(const __grid_constant__ scene_ref sc_in, const __grid_constant__ wave_buffers wb, int bounce_arg) {
	const int bounce = wave_vertex(wb, bounce_arg);
	scene_ref sc = sc_in;
	float4* const recs = wb.rec;
	unsigned char* stage_base = (unsigned char*)plu_smem;
	const uint32_t n = wb.counters[bounce * CNT_STRIDE + CNT_SHADOW];
	if (blockIdx.x * PLU_CTA >= n) return;   // no work for this CTA: leave before staging the scene
	if (SMEM) { stage_scene(sc, (unsigned char*)plu_smem, sc.n_ref_nodes <= PLU_UNIFORM_WALK_MAX_NODES); stage_base += (sc.blob_bytes + 127u) & ~127u; }
	float4* st = (float4*)stage_base;   // [2][3][PLU_CTA]: p, w1, c1 (the rare BSDF-sampled MIS ray's w2, c2 are fetched on demand)
	if (blockIdx.x == 0 && threadIdx.x == 0) atomicAdd(wb.stats + 5, (unsigned long long)n);
	const float nl = (float)sc.n_lights;
	trace_counts tc; tc.nodes = 0; tc.prims = 0;
	uint32_t rays = 0;
	const uint32_t tid = threadIdx.x, lane = tid & 31u, stride = gridDim.x * PLU_CTA;
	auto seed_prim = [&](int32_t light) -> int32_t {   // surface of an area light for the seeded walk (k_connect_pt does the same)
#if PLU_SEED_AREA
		return (!SMEM && light >= 0) ? sc.light(light).prim : -1;
#else
		return -1;
#endif
	};
	auto gather = [&](uint32_t b, uint32_t jj) {
		float4* row = st + (size_t)b * 3 * PLU_CTA + tid;
		cp_async_16(row, wb.sh_p + jj); cp_async_16(row + PLU_CTA, wb.sh_w1 + jj); cp_async_16(row + 2 * PLU_CTA, wb.sh_c1 + jj);
	};
	uint32_t j = blockIdx.x * PLU_CTA + tid, buf = 0;
	if (j < n) gather(0, j);
	cp_async_commit();
	for (uint32_t j0 = j - lane; j0 < n; j0 += stride) {
		const uint32_t jn = j + stride;
		if (jn < n) gather(buf ^ 1u, jn);
		cp_async_commit();
		cp_async_wait<1>();
		if (j < n) {
			const float4* row = st + (size_t)buf * 3 * PLU_CTA + tid;
			float4 p4 = row[0], w1 = row[PLU_CTA];
			uint32_t slot = __float_as_uint(p4.w), fl = __float_as_uint(w1.w);
			// path state is only needed after the casts: the loads travel meanwhile
			float4 T4 = bounce == 0 ? make_float4(1.f, 1.f, 1.f, 0.f) : recs[4 * (size_t)slot + 2 + (bounce & 1)];
			float4 L4 = wb.rad[slot];
			int32_t li = (int32_t)(fl >> 8);
			f3 Ld = mk3(0.f);
			bool any = false;
			ray r; r.e = mk3(p4);
			if (fl & SH_NEED1) {
				r.d = mk3(w1);
				float t; int32_t prim;
				rays++;
#if PLU_ANYHIT
				const bool any_hit = (fl & SH_DELTA) != 0;   // a point light is visible iff NOTHING is hit: the first hit decides
#else
				const bool any_hit = false;
#endif
				bool hit = trace_scene<COUNT, SMEM, WALKS>(sc, r, PLU_T_MAX, t, prim, &tc, any_hit, seed_prim(fl & SH_DELTA ? -1 : li));
				// visible: nothing hit, or (area light) the surface hit is the light's own (renderer.cpp:17)
				if (!hit || (!(fl & SH_DELTA) && sc.prim_meta(prim).z == li)) { Ld = Ld + mk3(row[2 * PLU_CTA]); any = true; }
			}
			if (fl & SH_NEED2) {
				// Li of this term uses the SHADING normal against -wi (renderer.cpp:43), which vanishes for every wi an orthonormal
				// frame can sample: the entries that carry it are few, so its two rows are not prefetched for every entry
				const float4 c2 = wb.sh_c2[j];
				r.d = mk3(wb.sh_w2[j]);
				float t; int32_t prim;
				rays++;
				bool hit = trace_scene<COUNT, SMEM, WALKS>(sc, r, PLU_T_MAX, t, prim, &tc, false, seed_prim(li));
				if (hit && sc.prim_meta(prim).z == li) { Ld = Ld + mk3(c2); any = true; }
			}
			f3 T = mk3(T4);
			if (any || !finite3(T)) {
				f3 L = mk3(L4);
				L = L + T * (nl * Ld);
				wb.rad[slot] = make_float4(L.x, L.y, L.z, 0.f);
			}
		}
		j = jn; buf ^= 1u;
	}
	unsigned long long a = rays, b = tc.nodes, c = tc.prims;
	for (int o = 16; o > 0; o >>= 1) {
		a += __shfl_down_sync(0xffffffffu, a, o);
		if (COUNT) { b += __shfl_down_sync(0xffffffffu, b, o); c += __shfl_down_sync(0xffffffffu, c, o); }
	}
	if (lane == 0) {
		if (a) atomicAdd(wb.stats + 1, a);
		if (COUNT) { atomicAdd(wb.stats + 2, b); atomicAdd(wb.stats + 3, c); }
	}
}

// ---- k_connect_pt ---------------------------------------------------------------------------------------------
// k_connect for scenes on the large-scene walks: persistent warps that refill finished lanes, as k_extend_pt (ncu r02a,
// 1M boxes: k_connect ran at 9.2 of 32 threads per instruction with one entry per lane per iteration — a shadow ray's walk
// is as uneven as an extension ray's).  A lane holds one shadow entry; its one or two rays (light sample, then the rare
// BSDF-sampled MIS ray) are walked one after the other, then L += T * (nlights * Ld) as in k_connect.  Same per-entry
// arithmetic; only the order in which entries are taken differs.
#define CNT_WORK_CONNECT 25
template <bool COUNT, class WALKER>
__global__ void __launch_bounds__(PLU_CTA, PLU_PT_BLOCKS) k_connect_pt(const __grid_constant__ scene_ref sc, const __grid_constant__ wave_buffers wb, int bounce_arg, uint32_t refill) {
	constexpr uint32_t FULL = 0xffffffffu;
	const int bounce = wave_vertex(wb, bounce_arg);
	float4* const recs = wb.rec;
	const uint32_t n = wb.counters[bounce * CNT_STRIDE + CNT_SHADOW];
	uint32_t* work = wb.counters + bounce * CNT_STRIDE + CNT_WORK_CONNECT;
	if (blockIdx.x == 0 && threadIdx.x == 0) atomicAdd(wb.stats + 5, (unsigned long long)n);
	trace_counts tc; tc.nodes = 0; tc.prims = 0;
	uint32_t rays = 0;
	const uint32_t lane = threadIdx.x & 31u, lt = (1u << lane) - 1u;
	const float nl = (float)sc.n_lights;
	WALKER w; w.cur = WALKER::DONE; w.sp = 0; w.best = -1; w.best_t = 0.f; w.best_rank = -1;
	__shared__ int32_t s_stack[PLU_SSTK > 0 ? PLU_SSTK : 1][PLU_SSTK > 0 ? PLU_CTA : 1];
	walk_stack<PLU_SSTK> stack; stack.sm = &s_stack[0][PLU_SSTK > 0 ? threadIdx.x : 0];
	bool have = false, exhausted = false, any = false;
	uint32_t j = 0, fl = 0; int phase = 0;
	f3 Ld = mk3(0.f);
	ray r; r.e = mk3(0.f); r.d = mk3(0.f);
	int32_t keep = -1;   // surface the current walk was seeded with (an area light's own surface), -1: closest-hit walk
	// start the walk of the lane's current ray; area-light rays on the SAH tree are seeded with the light surface's own hit
	auto start_ray = [&]() {
		w.template begin_walk<COUNT>(sc, r, &tc);
		keep = -1;
#if PLU_SEED_AREA
		if (WALKER::CAN_SEED && !(phase == 0 && (fl & SH_DELTA))) {
			const int32_t lp = sc.light((int32_t)(fl >> 8)).prim;
			const int4 lm = sc.prim_meta(lp);
			float tl;
			if (prim_t(sc, lp, lm.x, r, PLU_T_MAX, tl)) { w.seed(tl, lp, lm.w); keep = lp; }
		}
#endif
	};
	uint32_t c_next = 0, c_end = 0;
	for (;;) {
		// lanes whose walk has ended: visibility rule of the ray, next ray of the entry or the radiance update
		if (have && w.cur == WALKER::DONE) {
			const bool hit = w.best >= 0;
			const int32_t li = (int32_t)(fl >> 8);
			bool more = false;
			if (phase == 0) {
				// visible: nothing hit, or (area light) the surface hit is the light's own (renderer.cpp:17)
				if (!hit || (!(fl & SH_DELTA) && sc.prim_meta(w.best).z == li)) { Ld = Ld + mk3(wb.sh_c1[j]); any = true; }
				more = (fl & SH_NEED2) != 0;
			} else if (hit && sc.prim_meta(w.best).z == li) { Ld = Ld + mk3(wb.sh_c2[j]); any = true; }
			if (more) {
				phase = 1; r.d = mk3(wb.sh_w2[j]); rays++;
				start_ray();
			} else {
				const uint32_t slot = __float_as_uint(wb.sh_p[j].w);
				const f3 T = bounce == 0 ? mk3(1.f) : mk3(recs[4 * (size_t)slot + 2 + (bounce & 1)]);
				if (any || !finite3(T)) {
					f3 L = mk3(wb.rad[slot]);
					L = L + T * (nl * Ld);
					wb.rad[slot] = make_float4(L.x, L.y, L.z, 0.f);
				}
				have = false;
			}
		}
		for (int pass = 0; pass < 2 && !exhausted; ++pass) {
			const uint32_t idle = __ballot_sync(FULL, !have);
			if (!idle) break;
			if (c_next == c_end) {
				uint32_t b = 0;
				if (lane == 0) b = atomicAdd(work, PT_CHUNK);
				b = __shfl_sync(FULL, b, 0);
				if (b >= n) { exhausted = true; break; }
				c_next = b; c_end = min(b + PT_CHUNK, n);
			}
			const uint32_t idx = c_next + __popc(idle & lt);
			if (!have && idx < c_end) {
				j = idx;
				const float4 p4 = wb.sh_p[j], w1 = wb.sh_w1[j];
				fl = __float_as_uint(w1.w);
				r.e = mk3(p4); Ld = mk3(0.f); any = false;
				if (fl & SH_NEED1) { phase = 0; r.d = mk3(w1); }
				else { phase = 1; r.d = mk3(wb.sh_w2[j]); }
				rays++;
				start_ray();
				have = true;
			}
			c_next = min(c_next + (uint32_t)__popc(idle), c_end);
		}
		if (!__any_sync(FULL, have)) break;
		const uint32_t min_active = exhausted ? 1u : 33u - refill;
#if PLU_ANYHIT
		const bool any_hit = phase == 0 && (fl & SH_DELTA) != 0;   // a point light is visible iff NOTHING is hit: the first hit decides
#else
		const bool any_hit = false;
#endif
		for (;;) {
			w.template round<COUNT>(sc, r, stack, &tc, any_hit, keep);
			if ((uint32_t)__popc(__ballot_sync(FULL, have && w.cur != WALKER::DONE)) < min_active) break;
		}
	}
	unsigned long long a = rays, b = tc.nodes, c = tc.prims;
	for (int o = 16; o > 0; o >>= 1) {
		a += __shfl_down_sync(FULL, a, o);
		if (COUNT) { b += __shfl_down_sync(FULL, b, o); c += __shfl_down_sync(FULL, c, o); }
	}
	if (lane == 0) {
		if (a) atomicAdd(wb.stats + 1, a);
		if (COUNT) { atomicAdd(wb.stats + 2, b); atomicAdd(wb.stats + 3, c); }
	}
}

// ---- k_loop_next ----------------------------------------------------------------------------------------------
// Last node of the graph's while body: the extension of the next vertex has run; step the loop counter and decide whether
// another round is needed.  No path is alive when that extension found no surface (its hit queues are empty): every later
// launch of the wave would find empty queues, so the loop ends there (renderer.cpp:86-91: a miss ends the path).
static __global__ void k_loop_next(const __grid_constant__ wave_buffers wb, cudaGraphConditionalHandle handle) {
	if (threadIdx.x != 0 || blockIdx.x != 0) return;
	const int mv = wb.desc->max_vertices;
	uint32_t* loop = wb.counters + (mv + 1) * CNT_STRIDE;
	const uint32_t v = *loop + 1u;
	*loop = v;
	atomicAdd(wb.stats + 6, 1ull);   // rounds run: the host derives the launch count from it
	const uint32_t hits = wb.counters[v * CNT_STRIDE + CNT_HIT_DIFFUSE] + wb.counters[v * CNT_STRIDE + CNT_HIT_SPECULAR];
	cudaGraphSetConditional(handle, ((int)v < mv && hits > 0u) ? 1u : 0u);
}

// ---- k_accumulate ---------------------------------------------------------------------------------------------
// target->pixel(floor(px)) += ray_color * smp_wgt, samples in ascending order (renderer.cpp:132).
// The overflow-column strip (pixel index >= W*H, sampler.h:74-78) lands in pixel (0, row+1).
static __global__ void __launch_bounds__(256) k_accumulate(const __grid_constant__ wave_desc w, const __grid_constant__ wave_buffers wb, float* fb, float smp_wgt) {
	for (uint32_t q = blockIdx.x * blockDim.x + threadIdx.x; q < (uint32_t)w.npix; q += gridDim.x * blockDim.x) {
		int32_t pix = w.pix0 + (int32_t)q, wh = w.width * w.height;
		int32_t dst = pix < wh ? pix : (pix - wh + 1) * w.width;
		if (dst >= wh) continue;   // the reference writes one vec3 past the buffer here
		float* o = fb + (size_t)dst * 3;
		f3 acc = mk3(o[0], o[1], o[2]);
		for (int32_t k = 0; k < w.kcount; ++k) {
			const size_t slot = (size_t)k * (size_t)w.npix + q;
			// a clean slot (camera ray missed, k_primary) holds L == 0: the add is done, the 16-byte read is not
			const bool have = !w.use_dirty || ((wb.dirty[slot >> 5] >> (slot & 31u)) & 1u);
			f3 c = have ? mk3(wb.rad[slot]) : mk3(0.f);
			acc = acc + c * smp_wgt;
		}
		o[0] = acc.x; o[1] = acc.y; o[2] = acc.z;
	}
}

// ---- parity hooks ---------------------------------------------------------------------------------------------
// scene->hit for explicit rays: surface id, t, hit_record::norm.
template <bool SMEM, bool COUNT = false>
__global__ void __launch_bounds__(256) k_trace_rays(const __grid_constant__ scene_ref sc_in, int64_t n, const float* o3, const float* d3,
                                                    int primary_w, int primary_h, int32_t* id, float* t_out, float* n3) {
	scene_ref sc = sc_in;
	if (SMEM) stage_scene(sc, (unsigned char*)plu_smem, sc.n_ref_nodes <= PLU_UNIFORM_WALK_MAX_NODES);
	for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
		ray r;
		if (primary_w > 0) {   // pixel-centre pinhole rays
			int32_t x = (int32_t)(i % primary_w), y = (int32_t)(i / primary_w);
			r = generate_ray(sc.cam, mk2((float)x + .5f, (float)y + .5f), mk2(.5f, .5f), false);
		} else {
			r.e = mk3(o3[3 * i], o3[3 * i + 1], o3[3 * i + 2]); r.d = mk3(d3[3 * i], d3[3 * i + 1], d3[3 * i + 2]);
		}
		float t; int32_t prim; trace_counts tc; tc.nodes = 0; tc.prims = 0;
		if (trace_scene<COUNT, SMEM>(sc, r, PLU_T_MAX, t, prim, &tc)) {
			hit_rec hr; fill_record(sc, prim, sc.prim_meta(prim).x, r, t, false, hr);
			id[i] = prim; t_out[i] = t; n3[3 * i] = hr.norm.x; n3[3 * i + 1] = hr.norm.y; n3[3 * i + 2] = hr.norm.z;
		} else { id[i] = -1; t_out[i] = 0.f; n3[3 * i] = 0.f; n3[3 * i + 1] = 0.f; n3[3 * i + 2] = 0.f; }
		if (COUNT) { n3[3 * i] = (float)tc.nodes; n3[3 * i + 1] = (float)tc.prims; n3[3 * i + 2] = 0.f; }   // the walk's cost instead of the normal
	}
}

}  // namespace plud
