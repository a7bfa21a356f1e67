// pmath.cuh — float32 vector math, deterministic trig and the Philox stream of the device path.
//
// The reference (andrew-pa/plutracer) is scalar GLM code compiled for x86-64 without FMA; its
// converged images depend on rounding (no ray epsilon anywhere: self-intersection at t~+0 is decided
// by the last bit, SURVEY.md 7 "hard parts").  Every helper here therefore keeps GLM's operation
// ORDER (depd/glm/glm/detail/func_geometric.inl:58-94, func_common.inl:18-27,187,256) and this
// translation unit is compiled with -fmad=false -prec-div=true -prec-sqrt=true -ftz=false so that
// + - * / sqrt are IEEE-754 round-to-nearest exactly as on the host.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

#define PLU_DEV __device__ __forceinline__

namespace plud {

struct f2 { float x, y; };
struct f3 { float x, y, z; };

PLU_DEV f3 mk3(float x, float y, float z) { f3 r; r.x = x; r.y = y; r.z = z; return r; }
PLU_DEV f3 mk3(float s) { return mk3(s, s, s); }
PLU_DEV f3 mk3(const float4& v) { return mk3(v.x, v.y, v.z); }
PLU_DEV f2 mk2(float x, float y) { f2 r; r.x = x; r.y = y; return r; }
PLU_DEV float comp(const f3& v, int i) { return i == 0 ? v.x : (i == 1 ? v.y : v.z); }
PLU_DEV void setc(f3& v, int i, float s) { if (i == 0) v.x = s; else if (i == 1) v.y = s; else v.z = s; }

PLU_DEV f3 operator+(f3 a, f3 b) { return mk3(a.x + b.x, a.y + b.y, a.z + b.z); }
PLU_DEV f3 operator-(f3 a, f3 b) { return mk3(a.x - b.x, a.y - b.y, a.z - b.z); }
PLU_DEV f3 operator*(f3 a, f3 b) { return mk3(a.x * b.x, a.y * b.y, a.z * b.z); }
PLU_DEV f3 operator/(f3 a, f3 b) { return mk3(a.x / b.x, a.y / b.y, a.z / b.z); }
PLU_DEV f3 operator*(f3 a, float s) { return mk3(a.x * s, a.y * s, a.z * s); }
PLU_DEV f3 operator*(float s, f3 a) { return mk3(s * a.x, s * a.y, s * a.z); }
PLU_DEV f3 operator/(f3 a, float s) { return mk3(a.x / s, a.y / s, a.z / s); }
PLU_DEV f3 operator/(float s, f3 a) { return mk3(s / a.x, s / a.y, s / a.z); }
PLU_DEV f3 operator+(f3 a, float s) { return mk3(a.x + s, a.y + s, a.z + s); }
PLU_DEV f3 operator-(f3 a, float s) { return mk3(a.x - s, a.y - s, a.z - s); }
PLU_DEV f3 operator-(float s, f3 a) { return mk3(s - a.x, s - a.y, s - a.z); }
PLU_DEV f3 operator-(f3 a) { return mk3(-a.x, -a.y, -a.z); }

// glm::min / max are `x<y?x:y` / `x>y?x:y` (a NaN in x falls through to y) — NOT fminf/fmaxf
PLU_DEV float gmin(float x, float y) { return x < y ? x : y; }
PLU_DEV float gmax(float x, float y) { return x > y ? x : y; }
PLU_DEV float gclamp(float x, float lo, float hi) { return gmin(gmax(x, lo), hi); }
PLU_DEV float gsign(float x) { return (float)(0.f < x) - (float)(x < 0.f); }
PLU_DEV float gmod(float x, float y) { return x - y * floorf(x / y); }
PLU_DEV float gfract(float x) { return x - floorf(x); }
PLU_DEV float dot(f3 a, f3 b) { f3 t = a * b; return t.x + t.y + t.z; }
PLU_DEV float dot(f2 a, f2 b) { return a.x * b.x + a.y * b.y; }
PLU_DEV f3 cross(f3 x, f3 y) { return mk3(x.y * y.z - y.y * x.z, x.z * y.x - y.z * x.x, x.x * y.y - y.x * x.y); }
PLU_DEV f3 normalize(f3 v) { return v * (1.f / sqrtf(dot(v, v))); }
PLU_DEV bool finite3(f3 v) { return isfinite(v.x) && isfinite(v.y) && isfinite(v.z); }

#define PLU_PI 3.14159265358979323846264338327950288f
#define PLU_ONE_OVER_PI 0.318309886183790671537767526745028724f
#define PLU_TWO_OVER_PI 0.636619772367581343075535053490057448f

// ---- deterministic sin / cos / acos ---------------------------------------------------------------
// Single-precision Cephes algorithms (S. Moshier, Cephes 2.2 sinf.c / asinf.c): 3-term Cody-Waite
// pi/4 reduction and minimax polynomials, written with IEEE + - * / sqrt only, so the device and
// the CPU oracle (which carries an independent copy) agree to the bit; ~1 ulp from libm on the
// ranges the path uses (angles in [0, 2*pi], acos on [-1, 1]).
PLU_DEV int trig_reduce(float ax, float& r) {
	const float FOPI = 1.27323954473516f;
	const float DP1 = 0.78515625f, DP2 = 2.4187564849853515625e-4f, DP3 = 3.77489497744594108e-8f;
	int j = (int)(FOPI * ax);
	float y = (float)j;
	if (j & 1) { j += 1; y += 1.f; }
	r = ((ax - y * DP1) - y * DP2) - y * DP3;
	return j;
}
PLU_DEV float trig_poly_sin(float r) {
	float z = r * r;
	return ((-1.9515295891e-4f * z + 8.3321608736e-3f) * z - 1.6666654611e-1f) * z * r + r;
}
PLU_DEV float trig_poly_cos(float r) {
	float z = r * r;
	return ((2.443315711809948e-5f * z - 1.388731625493765e-3f) * z + 4.166664568298827e-2f) * z * z - 0.5f * z + 1.f;
}
PLU_DEV float det_sin(float x) {
	if (!(fabsf(x) <= 8192.f)) return x - x;
	float sign = 1.f, ax = x;
	if (x < 0.f) { sign = -1.f; ax = -x; }
	float r; int j = trig_reduce(ax, r) & 7;
	if (j > 3) { sign = -sign; j -= 4; }
	float y = (j == 1 || j == 2) ? trig_poly_cos(r) : trig_poly_sin(r);
	return sign * y;
}
PLU_DEV float det_cos(float x) {
	if (!(fabsf(x) <= 8192.f)) return x - x;
	float sign = 1.f, ax = x < 0.f ? -x : x;
	float r; int j = trig_reduce(ax, r) & 7;
	if (j > 3) { sign = -sign; j -= 4; }
	if (j > 1) sign = -sign;
	float y = (j == 1 || j == 2) ? trig_poly_sin(r) : trig_poly_cos(r);
	return sign * y;
}
// det_cos(x) and det_sin(x) of the same angle with ONE reduction and one evaluation of each polynomial (the two calls side by
// side evaluate them up to four times: the compiler does not merge the selects).  Same operations on the same values, so the
// results are det_cos / det_sin bit for bit.
PLU_DEV void det_sincos(float x, float& s_out, float& c_out) {
	if (!(fabsf(x) <= 8192.f)) { s_out = c_out = x - x; return; }
	float sign_s = 1.f, sign_c = 1.f, ax = x;
	if (x < 0.f) { sign_s = -1.f; ax = -x; }
	float r; int j = trig_reduce(ax, r) & 7;
	if (j > 3) { sign_s = -sign_s; sign_c = -sign_c; j -= 4; }
	if (j > 1) sign_c = -sign_c;
	const float ps = trig_poly_sin(r), pc = trig_poly_cos(r);
	const bool swap = (j == 1 || j == 2);
	s_out = sign_s * (swap ? pc : ps);
	c_out = sign_c * (swap ? ps : pc);
}
PLU_DEV float det_asin(float x) {
	float a = x < 0.f ? -x : x;
	if (!(a <= 1.f)) return (x - x) / (x - x);
	if (a < 1.0e-4f) return x;
	float z, s; bool flag = false;
	if (a > 0.5f) { z = 0.5f * (1.f - a); s = sqrtf(z); flag = true; }
	else { s = a; z = s * s; }
	z = ((((4.2163199048e-2f * z + 2.4181311049e-2f) * z + 4.5470025998e-2f) * z + 7.4953002686e-2f) * z + 1.6666752422e-1f) * z * s + s;
	if (flag) { z = z + z; z = 1.5707963267948966192f - z; }
	return x < 0.f ? -z : z;
}
PLU_DEV float det_acos(float x) {
	if (!(x >= -1.f && x <= 1.f)) return (x - x) / (x - x);
	if (x < -0.5f) return 3.14159265358979323846f - 2.f * det_asin(sqrtf(0.5f * (1.f + x)));
	if (x > 0.5f) return 2.f * det_asin(sqrtf(0.5f * (1.f - x)));
	return 1.5707963267948966192f - det_asin(x);
}

// ---- Philox4x32-10 (Salmon et al., SC'11) — the counter-based sampler stream -------------------------
// Replaces sample::next() / rnd::randf() / rnd::randi() (reference inc/sampler.h:12-22,
// inc/cmmn.h:239-255: one racy global mt19937).  counter = (pixel, sample, call, 'PLU1'), key = seed.
struct u4 { uint32_t x, y, z, w; };
PLU_DEV u4 philox4x32_10(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t k0, uint32_t k1) {
#pragma unroll
	for (int r = 0; r < 10; ++r) {
		uint32_t h0 = __umulhi(0xD2511F53u, c0), l0 = 0xD2511F53u * c0;
		uint32_t h1 = __umulhi(0xCD9E8D57u, c2), l1 = 0xCD9E8D57u * c2;
		uint32_t n0 = h1 ^ c1 ^ k0, n2 = h0 ^ c3 ^ k1;
		c0 = n0; c1 = l1; c2 = n2; c3 = l0;
		k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
	}
	u4 o; o.x = c0; o.y = c1; o.z = c2; o.w = c3; return o;
}
// one 32-bit word -> [0,1) the way libstdc++ maps an mt19937 word for the reference (cmmn.h:245-248:
// uniform_real_distribution<float> = float(x) / 2^32, round to nearest, 1.0 stepped down)
PLU_DEV float u01(uint32_t x) { float u = __uint2float_rn(x) * 2.3283064365386963e-10f; return u >= 1.f ? 0.99999994f : u; }

struct philox_key { uint32_t pixel, sample, k0, k1; };
PLU_DEV u4 philox_call(const philox_key& k, uint32_t call) { return philox4x32_10(k.pixel, k.sample, call, 0x504c5531u, k.k0, k.k1); }

}  // namespace plud
