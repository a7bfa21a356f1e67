// shade_tu.cu — k_shade in its own translation unit, so that it can be built with other floating-point flags than the
// rest of the path (plutracer_b200/_build.py).
//
// Default build: the same flags as everything else (-fmad=false -prec-div=true -prec-sqrt=true): the whole path is the
// reference's arithmetic, operation for operation, and the bit-exact parity tests run on this build.
//
// PLU_SHADE_FAST=1 build (measured option, profiles/r02_summary.md): this unit alone gets -fmad=true -prec-div=false
// -prec-sqrt=false.  Intersection, ray set-up, the camera and the framebuffer adds keep the exact flags (north_star needs
// bit-exact primary hits, and the self-intersection statistics of a path without ray epsilon depend on rounding — the
// hit point e + d * t is written with explicit rounding and stays exact here too); what relaxes is the BSDF / light /
// texture / MIS arithmetic between two scene->hit calls, within the converged-image tolerance only.
#include "device/wavefront.cuh"
#include "shade_tu.h"

using namespace plud;

namespace plushade {

int max_blocks_per_sm(bool smem, int cls, size_t dyn_smem) {
	int occ = 0;
	auto q = [&](auto kernel) {
		cudaFuncAttributes fa; size_t fixed = 0;
		if (cudaFuncGetAttributes(&fa, kernel) == cudaSuccess) fixed = fa.sharedSizeBytes;
		if (dyn_smem + fixed > 48 * 1024) cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)dyn_smem);
		if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kernel, PLU_CTA, dyn_smem) != cudaSuccess || occ < 1) occ = 1;
	};
	if (smem) { if (cls == CLS_DIFFUSE) q(k_shade<true, CLS_DIFFUSE>); else q(k_shade<true, CLS_SPECULAR>); }
	else { if (cls == CLS_DIFFUSE) q(k_shade<false, CLS_DIFFUSE>); else q(k_shade<false, CLS_SPECULAR>); }
	return occ;
}

void launch(bool smem, int cls, int grid, size_t dyn_smem, cudaStream_t st, const scene_ref& sc, const wave_buffers& wb, int vertex) {
	if (smem) {
		if (cls == CLS_DIFFUSE) k_shade<true, CLS_DIFFUSE><<<grid, PLU_CTA, dyn_smem, st>>>(sc, wb, vertex);
		else k_shade<true, CLS_SPECULAR><<<grid, PLU_CTA, dyn_smem, st>>>(sc, wb, vertex);
	} else {
		if (cls == CLS_DIFFUSE) k_shade<false, CLS_DIFFUSE><<<grid, PLU_CTA, dyn_smem, st>>>(sc, wb, vertex);
		else k_shade<false, CLS_SPECULAR><<<grid, PLU_CTA, dyn_smem, st>>>(sc, wb, vertex);
	}
}

int fast_math() {
#ifdef PLU_SHADE_FAST
	return 1;
#else
	return 0;
#endif
}

}  // namespace plushade
