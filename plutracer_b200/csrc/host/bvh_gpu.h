// bvh_gpu.h — construction of both traversal structures ON THE DEVICE, see ../bvh_gpu.cu.
#pragma once
#include "plugpu.h"
#include <cstdint>
#include <cuda_runtime.h>
#include <string>

namespace plubuild {

struct device_tables {
	// in: device copies the caller already owns
	const float* d_prim;        // [n*12] surface data (the blob's prim table)
	int32_t* d_prim_meta;       // [n*4] tag, material, light, tie rank (rank written here)
	void* d_ref_nodes;          // [2n-1] plud::ref_node, written here
	// out: allocated here (cudaMalloc), owned by the caller afterwards
	void* d_fast = nullptr;             // plud::fast_node[n_fast_nodes], device layout (paired planes, packed leaf refs)
	int32_t* d_fast_index = nullptr;    // [n] leaf slot -> surface id
	float* d_fast_leaf = nullptr;       // [n*12] leaf-ordered surface records
	int32_t n_fast_nodes = 0;
};

// Builds, for the scene `d` (host tables, used for sizes and for the rare host tie fix-up):
//   * the reference tree's leaf order = tie ranks (bvh_tree.cpp:7-36), unless d->prim_tie_rank gives them,
//   * the reference tree in pre-order with skip links,
//   * the binned-SAH fast BVH with its leaf-ordered surface records.
// Same tables, byte for byte, as host/bvh_build.cpp produces.  Returns false with `err` set on a CUDA error.
bool build_on_device(const plu_scene_desc* d, device_tables& t, cudaStream_t st, plu_build_stats* stats, std::string& err);

}  // namespace plubuild
