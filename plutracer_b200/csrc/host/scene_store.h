// scene_store.h — host-side owner of the arrays a plu_scene_desc points at (the `plu_scene`
// handle of include/plugpu.h).  Host plumbing only; no CUDA here.
#pragma once
#include "plugpu.h"
#include <string>
#include <vector>

struct plu_scene {
	plu_camera_desc cam{};
	std::vector<int32_t> prim_tag, prim_material, prim_tie_rank;
	std::vector<float> prim_data;        // 12 per prim
	std::vector<float> prim_tri_uv;      // 6 per prim (empty when the scene has no triangles)
	std::vector<float> prim_tri_normal;  // 9 per prim (ditto)
	std::vector<plu_material_desc> materials;
	std::vector<plu_texture_desc> textures;
	std::vector<plu_light_desc> lights;
	std::vector<float> image_texels;
	int32_t width = 0, height = 0, n_strata = 0;  // scene-file defaults (resolution, antialiasing-samples)
	plu_scene_desc desc{};

	// (re)point desc at the vectors; call after any mutation
	void seal();
	int32_t add_prim(int32_t tag, int32_t material, const float data[12]);
};

namespace pluhost {
// validate a desc coming over the ABI; returns "" or a message
std::string validate_desc(const plu_scene_desc* d);
void copy_desc(const plu_scene_desc* d, plu_scene& out);
void set_err(char* err, int err_len, const std::string& msg);
}
