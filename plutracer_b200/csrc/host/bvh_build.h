// bvh_build.h — host-side construction of the two traversal structures of libplugpu.
#pragma once
#include "plugpu.h"
#include <cstdint>
#include <vector>

// leaf rule of the fast BVH, shared by the host builder and bvh_gpu.cu: a segment of up to PLU_LEAF_SMALL surfaces is a
// leaf anywhere, one of up to 4 below depth 8
#ifndef PLU_LEAF_SMALL
#define PLU_LEAF_SMALL 2
#endif

namespace pluhost {

struct ref_node_h { float lo[3]; int32_t skip; float hi[3]; int32_t prim; };            // == plud::ref_node
struct fast_node_h { float f[12]; int32_t child[2]; int32_t count[2]; };                 // == plud::fast_node

// bvh_tree.cpp:7-36 restated on the flat tables: rank[surface] = position of the surface among the
// leaves of the reference tree in DFS order.
void reference_leaf_order(const plu_scene_desc* d, std::vector<int32_t>& rank);

// The reference tree itself (same topology, same origin-inflated bounds, cmmn.h:78-85) in pre-order
// with skip links.  `rank` as above (given by the caller's live tree, or derived).
void build_ref_nodes(const plu_scene_desc* d, const std::vector<int32_t>& rank, std::vector<ref_node_h>& out);

// Tight binned-SAH BVH2 with padded leaf boxes, <= 4 surfaces per leaf, depth <= 60.
void build_fast_bvh(const plu_scene_desc* d, std::vector<fast_node_h>& nodes, std::vector<int32_t>& prim_index);

}  // namespace pluhost
