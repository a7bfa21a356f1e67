// post.cuh — the step right after the path: tone map and image encode.
// Restates postprocesser::whitePreservingLumaBasedReinhardToneMapping (reference src/main.cpp:77-86)
// and texture2d::write_bmp + stbi_write_bmp (src/texture.cpp:104-194) as two small HBM-bound kernels:
// 12 B read (+12 B written, or 3 B) per pixel, one thread per pixel, rows of the BMP written coalesced.
#pragma once
#include "pmath.cuh"

namespace plud {

// glm::pow(color, vec3(1/2.2)) is std::pow(float, float) per component on the host.  The device evaluates
// pow in double and rounds once: the correctly rounded float except for double-rounding ties, which is
// also what glibc's powf delivers for all but a vanishing fraction of inputs (tests bound the difference).
PLU_DEV float pow_gamma(float v) { return (float)pow((double)v, (double)(1.f / 2.2f)); }

PLU_DEV f3 tone_map(f3 c) {   // main.cpp:77-86
	const float white = 2.f;
	float luma = dot(c, mk3(0.2126f, 0.7152f, 0.0722f));
	float toneMappedLuma = luma * (1.f + luma / (white * white)) / (1.f + luma);
	float s = toneMappedLuma / luma;   // a black pixel: 0/0 = NaN, kept (write_bmp turns it into 0)
	c = c * s;
	return mk3(pow_gamma(c.x), pow_gamma(c.y), pow_gamma(c.z));
}

// texture.cpp:183-186: clamp(d, 0, 1) = min(max(x, 0), 1) with glm's selects (max(NaN, 0) = 0), then
// (unsigned char)(d * 255.f), i.e. truncation
PLU_DEV unsigned char quantise(float x) {
	float d = gmin(gmax(x, 0.f), 1.f);
	return (unsigned char)(d * 255.f);
}

__global__ void __launch_bounds__(256) k_tonemap(int64_t n_pixels, float* rgb) {
	for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n_pixels; i += (int64_t)gridDim.x * blockDim.x) {
		f3 c = tone_map(mk3(rgb[3 * i], rgb[3 * i + 1], rgb[3 * i + 2]));
		rgb[3 * i] = c.x; rgb[3 * i + 1] = c.y; rgb[3 * i + 2] = c.z;
	}
}

// One thread per output pixel of the BMP body: row j of the file is image row H-1-j (vdir = -1,
// texture.cpp:113-116), bytes B,G,R (rgb_dir = -1, :137), rows padded to a multiple of 4 (:169).
__global__ void __launch_bounds__(256) k_encode_bmp(int32_t W, int32_t H, const float* rgb, int tonemap, unsigned char* body) {
	const int64_t n = (int64_t)W * H;
	const int64_t row_bytes = (int64_t)W * 3 + ((-W * 3) & 3);
	for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
		int64_t frow = i / W, x = i - frow * W;
		const float* p = rgb + 3 * ((int64_t)(H - 1 - frow) * W + x);
		f3 c = mk3(p[0], p[1], p[2]);
		if (tonemap) c = tone_map(c);
		unsigned char* o = body + frow * row_bytes + 3 * x;
		o[0] = quantise(c.z); o[1] = quantise(c.y); o[2] = quantise(c.x);
		if (x == W - 1) for (int64_t k = 3 * (int64_t)W; k < row_bytes; ++k) body[frow * row_bytes + k] = 0;
	}
}

}  // namespace plud
