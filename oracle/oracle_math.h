// oracle_math.h — scalar float32 restatement of the handful of GLM / libm operations the reference
// render path uses.  TEST INFRASTRUCTURE ONLY (see oracle_pt.cpp); nothing in the product path
// includes, links or runs anything under oracle/.
//
// Every function restates the arithmetic ORDER of the GLM 0.9.9.0-dev code vendored by the
// reference (depd/glm @85bd368), so that IEEE-754 results are bit-identical:
//   dot        glm/detail/func_geometric.inl:58-59   tmp = a*b; tmp.x + tmp.y + tmp.z
//   cross      glm/detail/func_geometric.inl:80-83
//   normalize  glm/detail/func_geometric.inl:94      v * inversesqrt(dot(v,v)), inversesqrt = 1/sqrt
//   min / max  glm/detail/func_common.inl:18,27      x<y?x:y  /  x>y?x:y   (NaN falls to y)
//   sign       glm/detail/func_common.inl:187        (0<x) - (x<0)
//   mod        glm/detail/func_common.inl:~256       x - y*floor(x/y);   fract = x - floor(x)
// Build with -ffp-contract=off (no FMA), as the reference's x86-64 -O2 build has none.
#pragma once
#include <cmath>
#include <cstdint>

namespace orc {

struct v2 { float x, y; };
struct v3 {
	float x, y, z;
	float operator[](int i) const { return i == 0 ? x : (i == 1 ? y : z); }
	float& at(int i) { return i == 0 ? x : (i == 1 ? y : z); }
};

inline v3 V3(float x, float y, float z) { return v3{x, y, z}; }
inline v3 V3(float s) { return v3{s, s, s}; }
inline v3 operator+(v3 a, v3 b) { return V3(a.x + b.x, a.y + b.y, a.z + b.z); }
inline v3 operator-(v3 a, v3 b) { return V3(a.x - b.x, a.y - b.y, a.z - b.z); }
inline v3 operator*(v3 a, v3 b) { return V3(a.x * b.x, a.y * b.y, a.z * b.z); }
inline v3 operator/(v3 a, v3 b) { return V3(a.x / b.x, a.y / b.y, a.z / b.z); }
inline v3 operator*(v3 a, float s) { return V3(a.x * s, a.y * s, a.z * s); }
inline v3 operator*(float s, v3 a) { return V3(s * a.x, s * a.y, s * a.z); }
inline v3 operator/(v3 a, float s) { return V3(a.x / s, a.y / s, a.z / s); }
inline v3 operator/(float s, v3 a) { return V3(s / a.x, s / a.y, s / a.z); }
inline v3 operator+(v3 a, float s) { return V3(a.x + s, a.y + s, a.z + s); }
inline v3 operator-(v3 a, float s) { return V3(a.x - s, a.y - s, a.z - s); }
inline v3 operator-(float s, v3 a) { return V3(s - a.x, s - a.y, s - a.z); }
inline v3 operator-(v3 a) { return V3(-a.x, -a.y, -a.z); }

inline float gmin(float x, float y) { return x < y ? x : y; }
inline float gmax(float x, float y) { return x > y ? x : y; }
inline float gclamp(float x, float lo, float hi) { return gmin(gmax(x, lo), hi); }
inline float gsign(float x) { return (float)(0.f < x) - (float)(x < 0.f); }
inline float gmod(float x, float y) { return x - y * std::floor(x / y); }
inline float gfract(float x) { return x - std::floor(x); }
inline float dot(v3 a, v3 b) { v3 t = a * b; return t.x + t.y + t.z; }
inline float dot(v2 a, v2 b) { return a.x * b.x + a.y * b.y; }
inline v3 cross(v3 x, v3 y) {
	return V3(x.y * y.z - y.y * x.z, x.z * y.x - y.z * x.x, x.x * y.y - y.x * x.y);
}
inline v3 normalize(v3 v) { return v * (1.f / std::sqrt(dot(v, v))); }
inline float length(v3 v) { return std::sqrt(dot(v, v)); }
inline float distance(v3 a, v3 b) { return length(b - a); }

const float PI = 3.14159265358979323846264338327950288f;       // glm::pi<float>()
const float ONE_OVER_PI = 0.318309886183790671537767526745028724f;
const float TWO_OVER_PI = 0.636619772367581343075535053490057448f;

// ---- transcendental functions ---------------------------------------------------------------
// libm mode (glibc sinf/cosf/acosf) is what the compiled reference executes; it is used when the
// oracle is pinned against oracle/_ref/plu_ref.  det mode is a portable single-precision
// restatement of the Cephes sinf/cosf/asinf/acosf algorithms (S. Moshier, Cephes Math Library
// 2.2, single-precision "sinf.c"/"asinf.c": Cody-Waite 3-term pi/4 reduction + degree-7/6
// minimax polynomials; max error ~1 ulp on the ranges used here) made of IEEE + - * / sqrt only,
// so the CUDA path (which carries its own copy, plutracer_b200/csrc/device/pmath.cuh) and this
// oracle produce the same bits.
struct libm_trig {
	static float sin(float x) { return ::sinf(x); }
	static float cos(float x) { return ::cosf(x); }
	static float acos(float x) { return ::acosf(x); }
};

struct det_trig {
	// shared reduction: returns octant j (even), leaves reduced argument in r (|r| <= pi/4)
	static inline int reduce(float ax, float& r) {
		const float FOPI = 1.27323954473516f;
		const float DP1 = 0.78515625f, DP2 = 2.4187564849853515625e-4f, DP3 = 3.77489497744594108e-8f;
		int j = (int)(FOPI * ax);
		float y = (float)j;
		if (j & 1) { j += 1; y += 1.f; }
		r = ((ax - y * DP1) - y * DP2) - y * DP3;
		return j;
	}
	static inline float poly_sin(float r) {
		float z = r * r;
		return ((-1.9515295891e-4f * z + 8.3321608736e-3f) * z - 1.6666654611e-1f) * z * r + r;
	}
	static inline float poly_cos(float r) {
		float z = r * r;
		return ((2.443315711809948e-5f * z - 1.388731625493765e-3f) * z + 4.166664568298827e-2f) * z * z - 0.5f * z + 1.f;
	}
	static float sin(float x) {
		if (!(std::fabs(x) <= 8192.f)) return x - x;  // NaN for NaN/inf; (huge arguments never occur)
		float sign = 1.f, ax = x;
		if (x < 0.f) { sign = -1.f; ax = -x; }
		float r; int j = reduce(ax, r) & 7;
		if (j > 3) { sign = -sign; j -= 4; }
		float y = (j == 1 || j == 2) ? poly_cos(r) : poly_sin(r);
		return sign * y;
	}
	static float cos(float x) {
		if (!(std::fabs(x) <= 8192.f)) return x - x;
		float sign = 1.f, ax = x < 0.f ? -x : x;
		float r; int j = reduce(ax, r) & 7;
		if (j > 3) { sign = -sign; j -= 4; }
		if (j > 1) sign = -sign;
		float y = (j == 1 || j == 2) ? poly_sin(r) : poly_cos(r);
		return sign * y;
	}
	static float asin(float x) {
		float a = x < 0.f ? -x : x;
		if (!(a <= 1.f)) return (x - x) / (x - x);  // NaN: outside the domain or NaN input (glibc: NaN)
		if (a < 1.0e-4f) return x;
		float z, s; bool flag = false;
		if (a > 0.5f) { z = 0.5f * (1.f - a); s = std::sqrt(z); flag = true; }
		else { s = a; z = s * s; }
		z = ((((4.2163199048e-2f * z + 2.4181311049e-2f) * z + 4.5470025998e-2f) * z + 7.4953002686e-2f) * z + 1.6666752422e-1f) * z * s + s;
		if (flag) { z = z + z; z = 1.5707963267948966192f - z; }
		return x < 0.f ? -z : z;
	}
	static float acos(float x) {
		if (!(x >= -1.f && x <= 1.f)) return (x - x) / (x - x);
		if (x < -0.5f) return 3.14159265358979323846f - 2.f * asin(std::sqrt(0.5f * (1.f + x)));
		if (x > 0.5f) return 2.f * asin(std::sqrt(0.5f * (1.f - x)));
		return 1.5707963267948966192f - asin(x);
	}
};

// ---- Philox4x32-10 (Salmon, Moraes, Dror, Shaw: "Parallel random numbers: as easy as 1, 2, 3",
// SC'11; Random123 v1.x philox.h).  Published algorithm, pinned by the Random123 known-answer
// vectors in tests/test_oracle.py.
inline void philox4x32_10(const uint32_t ctr[4], const uint32_t key[2], uint32_t out[4]) {
	uint32_t c0 = ctr[0], c1 = ctr[1], c2 = ctr[2], c3 = ctr[3], k0 = key[0], k1 = key[1];
	for (int r = 0; r < 10; ++r) {
		uint64_t p0 = (uint64_t)0xD2511F53u * c0, p1 = (uint64_t)0xCD9E8D57u * c2;
		uint32_t n0 = (uint32_t)(p1 >> 32) ^ c1 ^ k0, n1 = (uint32_t)p1;
		uint32_t n2 = (uint32_t)(p0 >> 32) ^ c3 ^ k1, n3 = (uint32_t)p0;
		c0 = n0; c1 = n1; c2 = n2; c3 = n3;
		k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
	}
	out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}
// 32-bit word -> [0,1) exactly as libstdc++'s uniform_real_distribution<float> maps one mt19937 word
// (generate_canonical<float,24>: float(x) / 2^32 rounded to nearest, a result of 1.0 stepped down to
// nextafter(1, 0)), so the Philox stream has the reference's resolution near 0 — the estimator has 1/cos
// singularities at the rim of the cosine-hemisphere warp and its tail depends on that resolution
inline float u01(uint32_t x) { float u = (float)x * 2.3283064365386963e-10f; return u >= 1.f ? 0.99999994f : u; }

}  // namespace orc
