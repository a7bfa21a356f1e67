// scene_dev.cuh — device-side scene tables (tag-indexed; replaces the reference's virtual
// surface / material / light / texture object graph).
//
// The host packs every table into ONE 16-byte-aligned blob (host/scene_pack.cpp).  Kernels get a
// `scene_ref` by value (kernel parameter = constant bank, uniform reads) whose pointers address the
// blob in global memory; when the blob is small (all five named scenes: < 2 KB) each CTA first
// stages it into shared memory with 128-bit copies and re-bases the pointers, so BVH nodes,
// primitives, materials and lights are then read from on-chip memory by divergent lanes.
#pragma once
#include "pmath.cuh"
#include "plugpu.h"

namespace plud {

// Reference-topology BVH node (bvh_tree.cpp:7-36), pre-order, stackless:
//   lo = (min.xyz, skip)   hi = (max.xyz, prim)      prim >= 0: leaf (no bounds test, bvh_tree.cpp:40)
// `skip` = index of the next node in pre-order that is NOT in this node's subtree; in a leaf (whose
// successor is the next index anyway) the field holds the surface kind (PLU_PRIM_*).
struct ref_node { float4 lo, hi; };

// Tight two-child node of the fast BVH (host/bvh_build.cpp), 64 B = four 128-bit loads:
//   (child 0, child 1) pairs, so that one packed fp32x2 instruction serves both children:
//   a = (min.x0 min.x1 min.y0 min.y1)  b = (min.z0 min.z1 max.x0 max.x1)  c = (max.y0 max.y1 max.z0 max.z1)
//   d = (child0, child1, -, -) as int: >= 0 inner node index, < 0 leaf: ~((first_leaf_slot << 3) | count)
// Leaf slots hold the surfaces themselves in leaf order (fast_leaf, 3 x float4 each), so a leaf test is one
// dependent load level instead of index -> meta -> data:
//   box      a = (min | surface id)      b = (max | kind<<30 | tie rank)
//   sphere   a = (centre | radius)       b = (surface id, -, - | kind<<30 | tie rank)
//   triangle a = (v0 | surface id)       b = (v1 | kind<<30 | tie rank)       c = (v2 | -)
struct fast_node { float4 a, b, c; int4 d; };

// Node of the pruned reference-tree walk (scenes with spheres, isect.cuh: trace_ref_pruned), 64 B, same pre-order index
// space as ref_node:
//   interior  a = ref lo | skip     b = ref hi | -1      c = tight min | leaf count k (int)   d = tight max | half diagonal
//   leaf      a = ref lo | kind     b = ref hi | surface c = - | 1                            d = - | DFS rank (tie rule)
// (tight = padded union of the subtree's surface bounds; ref lo/hi = the reference's origin-inflated box, for a sphere
// leaf centre / radius as in ref_node).  Children of interior node i with k leaves: i + 1 (k / 2 leaves) and
// i + 2 * (k / 2) (the rest) — the tree's shape is a function of the surface count alone (bvh_tree.cpp:27-31).
struct walk_node { float4 a, b, c, d; };

struct scene_ref {
	const unsigned char* blob;      // base (global, or shared after staging)
	uint32_t blob_bytes;
	uint32_t off_ref_nodes, off_prim, off_prim_meta, off_material, off_texture, off_light;
	int32_t n_ref_nodes, n_prims, n_materials, n_textures, n_lights;
	// large / optional tables stay in global memory
	const float* tri_uv;            // [n_prims*6] or null
	const float* tri_normal;        // [n_prims*9] or null
	const float* texels;            // image textures, rgb float
	const fast_node* fast_nodes;    // null: traverse the reference tree
	const int32_t* fast_prim_index; // leaf slots -> prim ids
	const float4* fast_leaf;        // leaf slots -> surface records (see fast_node)
	const walk_node* walk_nodes;    // non-null: scenes with spheres walk the reference tree with tight-box pruning
	int32_t n_fast_nodes;
	int32_t has_spheres;
	float sphere_rmax, sphere_rmin; // largest / smallest sphere radius of the scene (trace_ref_pruned's node criterion)
	plu_camera_desc cam;

	PLU_DEV const ref_node* ref_nodes() const { return (const ref_node*)(blob + off_ref_nodes); }
	PLU_DEV const float4* prim(int32_t s) const { return (const float4*)(blob + off_prim) + 3 * s; }
	PLU_DEV int4 prim_meta(int32_t s) const { return ((const int4*)(blob + off_prim_meta))[s]; }   // tag, material, light, tie rank
	PLU_DEV const plu_material_desc& material(int32_t m) const { return ((const plu_material_desc*)(blob + off_material))[m]; }
	PLU_DEV const plu_texture_desc& texture(int32_t t) const { return ((const plu_texture_desc*)(blob + off_texture))[t]; }
	PLU_DEV const plu_light_desc& light(int32_t l) const { return ((const plu_light_desc*)(blob + off_light))[l]; }
};

// Blobs up to this size are staged to shared memory by every CTA.
#define PLU_SMEM_BLOB_MAX 16384

// Stage the blob (call from all threads of the CTA; contains __syncthreads()).
// pair_nodes: the staged copy of the reference-tree nodes is rewritten in place from (min.xyz, skip | max.xyz, prim) to
// (min.x max.x min.y max.y | min.z max.z skip prim), so that the uniform walk (isect.cuh: trace_ref_uniform<.., SMEM>) gets
// the two planes of an axis as one 64-bit register pair for the packed fp32x2 subtract / multiply.  Only the CTA's own
// shared-memory copy changes; the table in global memory (plu_gpu_download_table) keeps the layout above.
PLU_DEV void stage_scene(scene_ref& sc, unsigned char* smem, bool pair_nodes = false) {
	const uint4* src = (const uint4*)sc.blob;
	uint4* dst = (uint4*)smem;
	uint32_t n = sc.blob_bytes >> 4;
	for (uint32_t i = threadIdx.x; i < n; i += blockDim.x) dst[i] = __ldg(src + i);
	__syncthreads();
	sc.blob = smem;
	if (pair_nodes) {
		float4* nd = (float4*)(smem + sc.off_ref_nodes);
		for (uint32_t i = threadIdx.x; i < (uint32_t)sc.n_ref_nodes; i += blockDim.x) {
			float4 lo = nd[2 * i], hi = nd[2 * i + 1];
			nd[2 * i] = make_float4(lo.x, hi.x, lo.y, hi.y);
			nd[2 * i + 1] = make_float4(lo.z, hi.z, lo.w, hi.w);
		}
		__syncthreads();
	}
}

}  // namespace plud
