// shade.cuh — materials, BSDF, textures, lights and sampling warps as tag-indexed device functions.
// Restates (operation for operation, quirks included — SURVEY.md 8a Q9-Q15):
//   inc/material.h:8-191, src/material.cpp:5-70, inc/lights/area_light.h, inc/light.h,
//   inc/props.h:24-30, inc/textures/textures.h, inc/texture.h:53-60, inc/cmmn.h:257-298,
//   inc/surface.h:26-33, inc/surfaces/box.h:23-34, sphere.h:18-22, src/surfaces/triangle.cpp:35-39.
#pragma once
#include "isect.cuh"

namespace plud {

// bxdf::type bits, material.h:50-60
#define BX_REFLECTION 1u
#define BX_TRANSMISSION 2u
#define BX_DIFFUSE 4u
#define BX_GLOSSY 8u
#define BX_SPECULAR 16u
#define BX_ALL 31u
#define BX_NEE_TYPES (BX_ALL & ~BX_SPECULAR)   // renderer.cpp:56

enum { K_LAMBERT = 0, K_SPEC_REFL = 1, K_SPEC_TRANS = 2 };

// ---- random draws of one path vertex (layout: rng comment in pmath.cuh / DESIGN.md) ----------------
struct vertex_rng {
	u4 a, b, c;
	PLU_DEV void begin(const philox_key& k, int v) {
		a = philox_call(k, 1u + 3u * (uint32_t)v);
		b = philox_call(k, 2u + 3u * (uint32_t)v);
		c = philox_call(k, 3u + 3u * (uint32_t)v);
	}
	PLU_DEV float light_pick() const { return u01(a.x); }
	PLU_DEV f2 light_u() const { return mk2(u01(a.y), u01(a.z)); }
	PLU_DEV float box_randf() const { return u01(a.w); }
	PLU_DEV int box_randi() const { return (int)__umulhi(b.x, 3u); }
	PLU_DEV float which(int purpose) const { return purpose == 0 ? u01(b.y) : u01(c.x); }
	PLU_DEV f2 bsdf_u(int purpose) const { return purpose == 0 ? mk2(u01(b.z), u01(b.w)) : mk2(u01(c.y), u01(c.z)); }
};

// ---- sampling warps, cmmn.h:257-298 --------------------------------------------------------------------
PLU_DEV f2 concentric_disk_sample(f2 u) {
	u = mk2(2.f * u.x - 1.f, 2.f * u.y - 1.f);
	if (u.x == 0.f && u.y == 0.f) return mk2(0.f, 0.f);
	// cmmn.h:262-279 has one division per quadrant of the square (u.y / u.x or u.x / u.y, then q, 8 + q, 2 - q, 4 - q or 6 - q).  The
	// lanes of a warp fall into all four, so the quadrant only SELECTS the operands here and the division is issued once: the same
	// operands meet the same operations, the result is the reference's bit for bit.
	const bool A = u.x >= -u.y;
	const bool B = A ? (u.x > u.y) : (u.x <= u.y);   // B: u.y / u.x, radius +-u.x; else u.x / u.y, radius +-u.y
	const float num = B ? u.y : u.x, den = B ? u.x : u.y;
	const float q = num / den;
	f2 rt;
	rt.x = A ? den : -den;
	if (A && B) rt.y = u.y > 0.f ? q : 8.f + q;
	else rt.y = (A ? 2.f : (B ? 4.f : 6.f)) - q;
	rt.y *= PLU_PI * .25f;
	float sn, cs; det_sincos(rt.y, sn, cs);
	return mk2(cs * rt.x, sn * rt.x);
}
PLU_DEV f3 uniform_sphere_sample(f2 u) {
	float z = 1.f - 2.f * u.x;
	float r = sqrtf(gmax(0.f, 1.f - z * z));
	float phi = 2.f * PLU_PI * u.y;
	float sn, cs; det_sincos(phi, sn, cs);
	return mk3(r * cs, r * sn, z);
}
PLU_DEV f3 cosine_hemisphere_sample(f2 u) {
	f2 d = concentric_disk_sample(u);
	return mk3(d.x, d.y, sqrtf(gmax(0.f, 1.f - dot(d, d))));
}

// ---- camera::generate_ray, camera.h:25-37 ---------------------------------------------------------------
PLU_DEV ray generate_ray(const plu_camera_desc& c, f2 px, f2 lens, bool use_lens) {
	f3 pos = mk3(c.pos[0], c.pos[1], c.pos[2]), look = mk3(c.look[0], c.look[1], c.look[2]);
	f3 up = mk3(c.up[0], c.up[1], c.up[2]), right = mk3(c.right[0], c.right[1], c.right[2]);
	f2 uv = mk2((px.x * c.inv_image_size[0]) * 2.f - 1.f, (px.y * c.inv_image_size[1]) * 2.f - 1.f);
	uv.y *= -1.f;
	ray r;
	r.e = pos;
	r.d = normalize(c.w * look + uv.x * right + uv.y * up);
	if (use_lens && c.lens_radius > 0.f) {
		f2 l = concentric_disk_sample(lens);
		l = mk2(l.x * c.lens_radius, l.y * c.lens_radius);
		f3 pof = ray_at(r, c.focal_distance / r.d.z);
		r.e.x += l.x; r.e.y += l.y;
		r.d = normalize(pof - r.e);
	}
	return r;
}

// ---- props::color / textures ---------------------------------------------------------------------------------
PLU_DEV f3 texel(const scene_ref& sc, const plu_texture_desc& t, f2 uv) {
	f3 c0 = mk3(t.c0[0], t.c0[1], t.c0[2]), c1 = mk3(t.c1[0], t.c1[1], t.c1[2]);
	if (t.tag == PLU_TEX_CHECKERBOARD) {   // textures.h:15-18
		float fx = floorf(uv.x * t.scale), fy = floorf(uv.y * t.scale);
		float m = gmod(fx + fy, 2.f);
		return ((m >= 1.f) && (m < 2.f)) ? c1 : c0;
	}
	if (t.tag == PLU_TEX_GRID) {   // textures.h:27-30; c0 = fg, c1 = bg
		float fx = gfract(uv.x * t.scale), fy = gfract(uv.y * t.scale);
		float gx = t.line < fx ? 0.f : 1.f, gy = t.line < fy ? 0.f : 1.f;
		float a = gmax(gx, gy);
		return c1 + a * (c0 - c1);
	}
	// texture2d::texel, texture.h:53-60
	float cx = gmod(uv.x, 1.f), cy = gmod(uv.y, 1.f);
	uint32_t ix = (uint32_t)floorf(cx * (float)(uint32_t)t.img_w), iy = (uint32_t)floorf(cy * (float)(uint32_t)t.img_h);
	if (ix == (uint32_t)t.img_w) ix--;
	if (iy == (uint32_t)t.img_h) iy--;
	if (ix >= (uint32_t)t.img_w || iy >= (uint32_t)t.img_h) return mk3(0.f);
	const float* px = sc.texels + ((size_t)t.img_offset + ix + (size_t)iy * (size_t)t.img_w) * 3;
	return mk3(__ldg(px), __ldg(px + 1), __ldg(px + 2));
}

// ---- Fresnel terms, material.h:26-45,84-107 --------------------------------------------------------------------
PLU_DEV f3 fresnel_dielectric(float eta_i, float eta_t, float cos_i) {
	cos_i = gclamp(cos_i, -1.f, 1.f);
	bool entering = cos_i > 0.f;
	float ei = eta_i, et = eta_t;
	if (entering) { float s = ei; ei = et; et = s; }   // swapped when ENTERING (reverse of PBRT) — kept
	float sin_t = ei / et * sqrtf(gmax(0.f, 1.f - cos_i * cos_i));
	if (sin_t >= 1.f) return mk3(1.f);
	float cos_t = sqrtf(gmax(0.f, 1.f - sin_t * sin_t));
	float ci = fabsf(cos_i);
	float Rparl = ((et * ci) - (ei * cos_t)) / ((et * ci) + (ei * cos_t));
	float Rperp = ((ei * ci) - (et * cos_t)) / ((ei * ci) + (et * cos_t));
	return mk3((Rparl * Rparl + Rperp * Rperp) * 0.5f);
}
PLU_DEV f3 fresnel_conductor(f3 eta, f3 k, float cos_i) {
	cos_i = fabsf(cos_i);
	f3 tmp0 = (eta * eta + k * k);
	f3 tmp1 = tmp0 * cos_i * cos_i;
	f3 X = (2.f * eta * cos_i);
	f3 Rparl2 = (tmp1 - X + 1.f) / (tmp1 + X + 1.f);
	f3 A = tmp0 - X + cos_i * cos_i;
	f3 Rperp2 = A / A;   // numerator == denominator in the reference (material.h:41-43)
	return (Rparl2 + Rperp2) / 2.f;
}

// ---- BSDF ---------------------------------------------------------------------------------------------------------
// material::operator() (material.h:213-254, area_light.h:52-54) as a tag: the component list is
// implied by the material tag, the colour is evaluated once, the frame is bsdf::complete().
struct bsdf {
	int32_t tag;            // PLU_MAT_*
	f3 R;                   // props::color at the hit
	f3 S, T, N;             // material.h:166-171, NOT orthonormalised
	f3 eta, k;              // conductor Fresnel (perfect-reflection)
	float fr_eta_i, fr_eta_t;   // dielectric Fresnel members as constructed (glass / perfect-refraction)

	PLU_DEV int num_components() const { return tag == PLU_MAT_GLASS ? 2 : (tag == PLU_MAT_EMISSION ? 0 : 1); }
	PLU_DEV int kind(int i) const {
		if (tag == PLU_MAT_DIFFUSE) return K_LAMBERT;
		if (tag == PLU_MAT_PERFECT_REFLECTION) return K_SPEC_REFL;
		if (tag == PLU_MAT_PERFECT_REFRACTION) return K_SPEC_TRANS;
		return i == 0 ? K_SPEC_REFL : K_SPEC_TRANS;
	}
	PLU_DEV static uint32_t type_of(int kind) {
		return kind == K_LAMBERT ? (BX_REFLECTION | BX_DIFFUSE) : (kind == K_SPEC_REFL ? (BX_REFLECTION | BX_SPECULAR) : (BX_TRANSMISSION | BX_SPECULAR));
	}
	PLU_DEV static bool matches(int kind, uint32_t T) { uint32_t ty = type_of(kind); return (ty & T) == ty; }
	PLU_DEV int count_matching(uint32_t t) const { int n = 0; for (int i = 0; i < num_components(); ++i) if (matches(kind(i), t)) n++; return n; }

	PLU_DEV f3 w2l(f3 v) const { return mk3(dot(v, S), dot(v, T), dot(v, N)); }
	PLU_DEV f3 l2w(f3 v) const {
		return mk3(S.x * v.x + T.x * v.y + N.x * v.z, S.y * v.x + T.y * v.y + N.y * v.z, S.z * v.x + T.z * v.y + N.z * v.z);
	}
	PLU_DEV f3 comp_F(int kind) const { return kind == K_LAMBERT ? R * PLU_ONE_OVER_PI : mk3(0.f); }
	PLU_DEV static float comp_pdf(int kind, f3 wo, f3 wi) {   // material.h:67-69; specular: 0
		if (kind != K_LAMBERT) return 0.f;
		return (wo.z * wi.z > 0.f) ? fabsf(wi.z) * PLU_ONE_OVER_PI : 0.f;
	}
	PLU_DEV f3 Fr_refl(float cos_i) const {
		return tag == PLU_MAT_PERFECT_REFLECTION ? fresnel_conductor(eta, k, cos_i) : fresnel_dielectric(fr_eta_i, fr_eta_t, cos_i);
	}
	// bxdf::sampleF: material.h:71-76 (default), :124-128 (reflection), :136-150 (transmission)
	PLU_DEV f3 comp_sampleF(int kind, f3 wo, f3& wi, f2 u, float& pdf_) const {
		if (kind == K_LAMBERT) {
			wi = cosine_hemisphere_sample(u);
			if (wo.z < 0.f) wi.z *= -1.f;
			pdf_ = comp_pdf(kind, wo, wi);
			return comp_F(kind);
		}
		if (kind == K_SPEC_REFL) {
			wi = mk3(-wo.x, -wo.y, wo.z);
			pdf_ = 1.f;
			return Fr_refl(wo.z) * R / fabsf(wi.z);
		}
		bool entering = wo.z > 0.f;
		float ei = fr_eta_i, et = fr_eta_t;
		if (!entering) { float s = ei; ei = et; et = s; }
		float sin2_i = gmax(0.f, 1.f - wo.z * wo.z);
		float eta_ = ei / et;
		float sin2_t = eta_ * eta_ * sin2_i;
		if (sin2_t >= 1.f) return mk3(0.f);   // total internal reflection: pdf_ stays 0
		float cos_t = sqrtf(gmax(0.f, 1.f - sin2_t));
		if (entering) cos_t *= -1.f;
		wi = mk3(eta_ * -wo.x, eta_ * -wo.y, cos_t);
		pdf_ = 1.f;
		return (et * et) / (ei * ei) * (1.f - fresnel_dielectric(fr_eta_i, fr_eta_t, wo.z)) * R / fabsf(wi.z);
	}

	// bsdf::F, material.cpp:5-16
	PLU_DEV f3 F(f3 n, f3 wwo, f3 wwi, uint32_t types) const {
		if (dot(wwi, n) * dot(wwo, n) > 0) types = types & ~BX_TRANSMISSION;
		else types = types & ~BX_REFLECTION;
		f3 f = mk3(0.f);
		for (int i = 0; i < num_components(); ++i)
			if (matches(kind(i), types)) f = f + comp_F(kind(i));
		return f;
	}
	// bsdf::pdf, material.cpp:56-65
	PLU_DEV float pdf(f3 wo, f3 wi, uint32_t t) const {
		float p = 0.f, w = 0.f;
		wo = w2l(wo); wi = w2l(wi);
		for (int i = 0; i < num_components(); ++i)
			if (matches(kind(i), t)) { p += comp_pdf(kind(i), wo, wi); w++; }
		return w > 0.f ? p / w : 0.f;
	}
	// bsdf::sampleF, material.cpp:17-55
	PLU_DEV f3 sampleF(const vertex_rng& smp, int purpose, f3 n, f3 wwo, f3& wwi, float& pdf_, uint32_t types, uint32_t& sampled_type) const {
		int matching = count_matching(types);
		if (matching == 0) { pdf_ = 0.f; return mk3(0.f); }
		int which = (int)gmin(floorf(smp.which(purpose) * (float)matching), (float)(matching - 1));
		int bi = 0, c = which;
		for (int i = 0; i < num_components(); ++i)
			if (matches(kind(i), types) && c-- == 0) { bi = i; break; }
		int bk = kind(bi);
		uint32_t btype = type_of(bk);
		f3 wo = w2l(wwo);
		f3 wi = mk3(0.f);
		pdf_ = 0.f;
		f3 f = comp_sampleF(bk, wo, wi, smp.bsdf_u(purpose), pdf_);
		if (pdf_ == 0.f) return mk3(0.f);
		sampled_type = btype;
		wwi = l2w(wi);
		if (!(btype & BX_SPECULAR) && matching > 1)
			for (int i = 0; i < num_components(); ++i)
				if (i != bi && matches(kind(i), types)) pdf_ += comp_pdf(kind(i), wo, wi);
		if (matching > 1) pdf_ /= (float)matching;
		if (!(btype & BX_SPECULAR)) {
			f = mk3(0.f);
			if (dot(wwi, n) * dot(wwo, n) > 0) types = types & ~BX_TRANSMISSION;
			else types = types & ~BX_REFLECTION;
			for (int i = 0; i < num_components(); ++i)
				if (matches(kind(i), types)) f = f + comp_F(kind(i));
		}
		return f;
	}
};

// unit_dpdu: the caller knows that hr.dpdu is an axis unit vector (box::hit, box.cpp:31-36): normalize() of it multiplies by
// 1 / sqrt(1) = 1 and is skipped
PLU_DEV void make_bsdf(const scene_ref& sc, const plu_material_desc& m, const hit_rec& hr, bsdf& b, bool unit_dpdu = false) {
	b.tag = m.tag;
	b.R = (m.texture >= 0) ? texel(sc, sc.texture(m.texture), hr.uv) : mk3(m.color[0], m.color[1], m.color[2]);
	b.eta = mk3(m.p0[0], m.p0[1], m.p0[2]); b.k = mk3(m.p1[0], m.p1[1], m.p1[2]);
	// specular_transmission(t, et, ei) : Fr(et, ei) and fresnel_dielectric(i, t) (material.h:90,133):
	// perfect-refraction(color, eta_t, eta_i) -> Fr.eta_i = eta_t, Fr.eta_t = eta_i; glass -> (1, ior)
	if (m.tag == PLU_MAT_PERFECT_REFRACTION) { b.fr_eta_i = m.p0[0]; b.fr_eta_t = m.p0[1]; }
	else { b.fr_eta_i = 1.f; b.fr_eta_t = m.p0[0]; }
	b.N = hr.norm; b.S = unit_dpdu ? hr.dpdu : normalize(hr.dpdu); b.T = cross(b.N, b.S);
}

// ---- area-light surfaces: surface::sample / surface::pdf ---------------------------------------------------
PLU_DEV f3 surface_sample(const scene_ref& sc, int32_t s, int32_t tag, f2 u, f3& n, const vertex_rng& rng) {
	const float4* p = sc.prim(s);
	if (tag == PLU_PRIM_SPHERE) {   // sphere.h:18-22
		f3 np = uniform_sphere_sample(u);
		n = np;
		return mk3(p[0]) + np * p[0].w;
	}
	if (tag == PLU_PRIM_BOX) {   // box.h:23-34 (two draws outside the sample stream)
		f3 mn = mk3(p[0]), mx = mk3(p[1]);
		f3 U = mk3(u.x, rng.box_randf(), u.y);
		int mi = rng.box_randi();
		setc(U, mi, comp(U, mi) > 0.5f ? 1.f : 0.f);
		f3 q = mn + U * (mx - mn);
		n = mk3(0.f);
		setc(n, mi, comp(U, mi) > 0.5f ? 1.f : -1.f);
		return q;
	}
	const float* nr = sc.tri_normal + (size_t)s * 9;   // triangle.cpp:35-39
	float w = 1 - (u.x + u.y);
	n = mk3(nr[0], nr[1], nr[2]) * u.x + mk3(nr[3], nr[4], nr[5]) * u.y + mk3(nr[6], nr[7], nr[8]) * w;
	return mk3(p[0]) * u.x + mk3(p[1]) * u.y + mk3(p[2]) * w;
}
// surface::pdf(p, wi), surface.h:26-33: one ray cast on the light's surface alone; D is measured from
// the WORLD ORIGIN (dot(D,D) quirk, kept)
PLU_DEV float surface_pdf(const scene_ref& sc, int32_t s, int32_t tag, float area, f3 p, f3 wi) {
	ray r; r.e = p; r.d = wi;
	float t;
	if (!prim_t(sc, s, tag, r, PLU_T_MAX, t)) return 0.f;
	hit_rec hr; fill_record(sc, s, tag, r, t, false, hr);
	f3 D = p + wi * t;
	return dot(D, D) / (fabsf(dot(hr.norm, -wi)) * area);
}
PLU_DEV f3 area_L(const plu_light_desc& l, f3 n, f3 w) {   // diffuse_area_light::L, area_light.h:21-23
	return dot(n, w) > 0.f ? mk3(l.b[0], l.b[1], l.b[2]) : mk3(0.f);
}

}  // namespace plud
