// shade_tu.h — host entry points of the translation unit that holds k_shade (shade_tu.cu).
#pragma once
#include "device/wavefront.cuh"

namespace plushade {
int max_blocks_per_sm(bool smem, int cls, size_t dyn_smem);   // occupancy of k_shade<smem, cls> at PLU_CTA threads
void launch(bool smem, int cls, int grid, size_t dyn_smem, cudaStream_t st, const plud::scene_ref& sc, const plud::wave_buffers& wb, int vertex);
int fast_math();   // 1: this unit was built with PLU_SHADE_FAST (relaxed floating-point flags)
}  // namespace plushade
