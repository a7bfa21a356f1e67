// plu_gpu_renderer.h — the reference-side binding for libplugpu.
//
// A plutracer maintainer drops this header into inc/ and changes ONE type name in
// src/main.cpp:171 (`plu::renderer r(...)` -> `plu::gpu_renderer r(...)`); `r.render(tx)` at
// :177 then runs on the B200 through the C ABI of include/plugpu.h.  This file compiles ONLY
// against the reference's own headers (surface.h, material.h, light.h, ...); it is not part of
// libplugpu and nothing on the GPU box needs it.
//
// It walks the reference's object graph with dynamic_cast (all plugin types are polymorphic)
// and writes the flat tables of plu_scene_desc.  Two of the reference's types hide what the
// flattener must read: props::color::{col,tex} (props.h:10-11) and bvh_tree::{root,bvh_node}
// (bvh_tree.h:11,35).  Either add `friend struct plu::gpu::flattener;` to those two classes, or
// include the reference headers for this one TU under
//     #define private public / #define protected public / #define class struct
// (after all std + GLM headers) as oracle/ref_driver.cpp does.
#pragma once
#include "plugpu.h"
#include <cstdio>
#include <cstdlib>
#include <map>
#include <stdexcept>
#include <string>
#include <vector>

namespace plu {
namespace gpu {

struct flat_scene {
	plu_camera_desc cam;
	std::vector<int32_t> prim_tag, prim_material, prim_tie_rank;
	std::vector<float> prim_data, tri_uv, tri_normal;
	std::vector<plu_material_desc> materials;
	std::vector<plu_texture_desc> textures;
	std::vector<plu_light_desc> lights;
	std::vector<float> image_texels;
	bool has_triangles = false;

	plu_scene_desc desc() const {
		plu_scene_desc d = {};
		d.abi_version = PLUGPU_ABI_VERSION;
		d.cam = cam;
		d.n_prims = (int32_t)prim_tag.size();
		d.n_materials = (int32_t)materials.size();
		d.n_textures = (int32_t)textures.size();
		d.n_lights = (int32_t)lights.size();
		d.prim_tag = prim_tag.data();
		d.prim_material = prim_material.data();
		d.prim_data = prim_data.data();
		d.prim_tri_uv = has_triangles ? tri_uv.data() : nullptr;
		d.prim_tri_normal = has_triangles ? tri_normal.data() : nullptr;
		d.materials = materials.data();
		d.textures = textures.data();
		d.lights = lights.data();
		d.image_texels = image_texels.empty() ? nullptr : image_texels.data();
		d.n_image_texels = (int64_t)(image_texels.size() / 3);
		d.prim_tie_rank = prim_tie_rank.empty() ? nullptr : prim_tie_rank.data();
		return d;
	}
};

struct flattener {
	flat_scene& out;
	std::map<const material*, int32_t> mat_ids;
	std::map<const light*, int32_t> light_ids;
	std::map<const surface*, int32_t> surf_ids;

	explicit flattener(flat_scene& o) : out(o) {}

	static void v3(float* dst, const vec3& v) { dst[0] = v.x; dst[1] = v.y; dst[2] = v.z; }

	int32_t texture_of(const props::color& c, float* col) {
		v3(col, c.col);
		if (c.tex == nullptr) return -1;
		plu_texture_desc t = {};
		const texture<vec3, vec2>* tp = c.tex.get();
		if (auto cb = dynamic_cast<const textures::checkerboard_texture*>(tp)) {
			t.tag = PLU_TEX_CHECKERBOARD;
			v3(t.c0, cb->colors[0]); v3(t.c1, cb->colors[1]); t.scale = cb->scale;
		} else if (auto g = dynamic_cast<const textures::grid_texture*>(tp)) {
			t.tag = PLU_TEX_GRID;
			v3(t.c0, g->fg_color); v3(t.c1, g->bg_color); t.scale = g->scale; t.line = g->line_size;
		} else if (auto im = dynamic_cast<const texture2d*>(tp)) {
			t.tag = PLU_TEX_IMAGE;
			t.img_w = (int32_t)im->size.x; t.img_h = (int32_t)im->size.y;
			t.img_offset = (int32_t)(out.image_texels.size() / 3);
			for (uint32_t y = 0; y < im->size.y; ++y)
				for (uint32_t x = 0; x < im->size.x; ++x) {
					const vec3& p = im->pixel(uvec2(x, y));
					out.image_texels.push_back(p.x); out.image_texels.push_back(p.y); out.image_texels.push_back(p.z);
				}
		} else {
			throw std::runtime_error("plu::gpu: unknown texture type");
		}
		out.textures.push_back(t);
		return (int32_t)out.textures.size() - 1;
	}

	int32_t material_of(const material* m) {
		while (auto fw = dynamic_cast<const forward_material*>(m)) m = fw->mat->get();
		if (m == nullptr) throw std::runtime_error("plu::gpu: surface without a material");
		auto it = mat_ids.find(m);
		if (it != mat_ids.end()) return it->second;
		plu_material_desc d = {};
		d.texture = -1; d.light = -1;
		if (auto df = dynamic_cast<const materials::diffuse_material*>(m)) {
			d.tag = PLU_MAT_DIFFUSE; d.texture = texture_of(df->diffuse, d.color);
		} else if (auto rf = dynamic_cast<const materials::perfect_reflection_material*>(m)) {
			d.tag = PLU_MAT_PERFECT_REFLECTION; d.texture = texture_of(rf->color, d.color);
			v3(d.p0, rf->eta); v3(d.p1, rf->k);
		} else if (auto rr = dynamic_cast<const materials::perfect_refraction_material*>(m)) {
			d.tag = PLU_MAT_PERFECT_REFRACTION; d.texture = texture_of(rr->color, d.color);
			d.p0[0] = rr->eta_t; d.p0[1] = rr->eta_i;
		} else if (auto gl = dynamic_cast<const materials::glass_material*>(m)) {
			d.tag = PLU_MAT_GLASS; d.texture = texture_of(gl->color, d.color);
			d.p0[0] = gl->ior;
		} else if (auto em = dynamic_cast<const materials::emission_material*>(m)) {
			d.tag = PLU_MAT_EMISSION;
			auto li = light_ids.find(em->L.get());
			if (li == light_ids.end()) throw std::runtime_error("plu::gpu: emission material whose light is not in the light list");
			d.light = li->second;
		} else {
			throw std::runtime_error("plu::gpu: unknown material type");
		}
		out.materials.push_back(d);
		int32_t id = (int32_t)out.materials.size() - 1;
		mat_ids[m] = id;
		return id;
	}

	// surfs: the surfaces in id order; lights: renderer::lights order (renderer.h:16)
	void run(const std::vector<const surface*>& surfs, const std::vector<shared_ptr<light>>& lights, const camera& cam) {
		v3(out.cam.pos, cam.pos); v3(out.cam.look, cam.look); v3(out.cam.up, cam.up); v3(out.cam.right, cam.right);
		out.cam.inv_image_size[0] = cam.inv_image_size.x; out.cam.inv_image_size[1] = cam.inv_image_size.y;
		out.cam.w = cam.w; out.cam.lens_radius = cam.lens_radius; out.cam.focal_distance = cam.focal_distance;

		for (size_t i = 0; i < surfs.size(); ++i) surf_ids[surfs[i]] = (int32_t)i;
		for (size_t i = 0; i < lights.size(); ++i) light_ids[lights[i].get()] = (int32_t)i;

		size_t n = surfs.size();
		out.prim_tag.resize(n); out.prim_material.resize(n);
		out.prim_data.assign(n * 12, 0.f); out.tri_uv.assign(n * 6, 0.f); out.tri_normal.assign(n * 9, 0.f);
		for (size_t i = 0; i < n; ++i) {
			const surface* s = surfs[i];
			float* d = &out.prim_data[i * 12];
			if (auto sp = dynamic_cast<const surfaces::sphere*>(s)) {
				out.prim_tag[i] = PLU_PRIM_SPHERE;
				v3(d, sp->center); d[3] = sp->radius;
			} else if (auto bx = dynamic_cast<const surfaces::box*>(s)) {
				out.prim_tag[i] = PLU_PRIM_BOX;
				v3(d, bx->box_bounds._min); v3(d + 4, bx->box_bounds._max);
			} else if (auto tr = dynamic_cast<const surfaces::triangle*>(s)) {
				out.prim_tag[i] = PLU_PRIM_TRIANGLE;
				out.has_triangles = true;
				for (int k = 0; k < 3; ++k) {
					v3(d + 4 * k, *tr->vertices[k]);
					v3(&out.tri_normal[i * 9 + 3 * k], *tr->normals[k]);
					out.tri_uv[i * 6 + 2 * k] = tr->texture_coords[k]->x;
					out.tri_uv[i * 6 + 2 * k + 1] = tr->texture_coords[k]->y;
				}
			} else {
				throw std::runtime_error("plu::gpu: unknown surface type");
			}
			out.prim_material[i] = material_of(s->mat.get());
		}

		out.lights.resize(lights.size());
		for (size_t i = 0; i < lights.size(); ++i) {
			plu_light_desc l = {};
			l.prim = -1;
			if (auto pl = dynamic_cast<const lights::point_light*>(lights[i].get())) {
				l.tag = PLU_LIGHT_POINT; v3(l.a, pl->pos); v3(l.b, pl->intensity);
			} else if (auto al = dynamic_cast<const lights::diffuse_area_light*>(lights[i].get())) {
				l.tag = PLU_LIGHT_DIFFUSE_AREA; v3(l.b, al->Lemit); l.area = al->area;
				auto si = surf_ids.find(al->surf.get());
				if (si == surf_ids.end()) throw std::runtime_error("plu::gpu: area light whose surface is not in the scene");
				l.prim = si->second;
			} else {
				throw std::runtime_error("plu::gpu: unknown light type");
			}
			out.lights[i] = l;
		}
	}

	// leaves of the reference tree in DFS (left first) order — bvh_tree.cpp:39-76 visits them so
	static void collect_leaves(const surfaces::bvh_tree::bvh_node* n, std::vector<const surface*>& dst) {
		if (n == nullptr) return;
		if (n->object != nullptr) { dst.push_back(n->object.get()); return; }
		collect_leaves(n->left_child.get(), dst);
		collect_leaves(n->right_child.get(), dst);
	}
};

}  // namespace gpu

namespace gpu {
// ONE libplugpu context per process, created on first use and kept alive across render / postprocess / write_bmp
// (CUDA set-up, the wave buffers and the device framebuffer are paid once, not per image).  Devices:
//   PLU_DEVICES unset      every visible GPU of the box (plu_gpu_create_multi: samples split k = g (mod n), scene
//                          replicated, ONE ncclReduce of the float framebuffers to the first device)
//   PLU_DEVICES="0,2,5"    exactly these;  PLU_DEVICES="0": a single GPU (no NCCL is loaded)
inline plu_gpu_ctx* shared_ctx() {
	struct holder {
		plu_gpu_ctx* c = nullptr;
		~holder() { plu_gpu_destroy(c); }
	};
	static holder h;
	if (h.c) return h.c;
	std::vector<int> devs;
	if (const char* e = std::getenv("PLU_DEVICES")) {
		for (const char* p = e; *p;) {
			char* end = nullptr; long v = std::strtol(p, &end, 10);
			if (end == p) break;
			devs.push_back((int)v);
			p = (*end == ',') ? end + 1 : end;
		}
	}
	int rc = devs.empty() ? plu_gpu_create_multi(nullptr, 0, &h.c) : plu_gpu_create_multi(devs.data(), (int)devs.size(), &h.c);
	if (rc != PLU_OK) throw std::runtime_error(plu_gpu_last_error(nullptr));
	return h.c;
}

// Page-locks the pixels of an existing texture2d (`new vec3[w*h]`, texture.h:29,36) for the guard's lifetime, so that the two
// PCIe copies of gpu_renderer::render run at the link rate; worth it when the same target is rendered into more than once or
// the render itself takes milliseconds.  An optimisation only: if the driver refuses, the target simply stays pageable.
struct pinned_target {
	void* p = nullptr;
	explicit pinned_target(const shared_ptr<texture2d>& t) {
		void* q = &t->pixel(uvec2(0, 0)).x;
		if (plu_host_register(q, (size_t)t->size.x * t->size.y * sizeof(vec3)) == PLU_OK) p = q;
	}
	~pinned_target() { if (p) plu_host_unregister(p); }
	pinned_target(const pinned_target&) = delete;
	pinned_target& operator=(const pinned_target&) = delete;
};
}  // namespace gpu

// Same constructor and render() signature as plu::renderer (renderer.h:20-25).
struct gpu_renderer {
	const uvec2 tile_size;  // kept for signature compatibility; the GPU path has no tiles
	surface* scene;
	vector<shared_ptr<light>> lights;
	camera cam;
	sampler* main_sampler;
	uint64_t seed = 20261104ull;
	uint32_t flags = 0;
	mutable plu_render_stats last_stats = {};

	gpu_renderer(surface* S, const camera& C, uvec2 tile_size, sampler* smp, const vector<shared_ptr<light>>& l)
		: tile_size(tile_size), scene(S), lights(l), cam(C), main_sampler(smp) {}

	void render(shared_ptr<texture2d> target) const {
		gpu::flat_scene fs;
		std::vector<const surface*> surfs;
		if (auto tree = dynamic_cast<const surfaces::bvh_tree*>(scene)) {
			gpu::flattener::collect_leaves(tree->root.get(), surfs);
			// ids are DFS positions, so the tie rank (bvh_tree.cpp:73) is the identity
			fs.prim_tie_rank.resize(surfs.size());
			for (size_t i = 0; i < surfs.size(); ++i) fs.prim_tie_rank[i] = (int32_t)i;
		} else {
			surfs.push_back(scene);
		}
		gpu::flattener(fs).run(surfs, lights, cam);
		plu_scene_desc d = fs.desc();

		plu_gpu_ctx* ctx = gpu::shared_ctx();
		if (plu_gpu_upload_scene(ctx, &d) != PLU_OK) throw std::runtime_error(plu_gpu_last_error(ctx));
		plu_render_params p = {};
		p.width = (int32_t)target->size.x; p.height = (int32_t)target->size.y;
		p.n_strata = (int32_t)main_sampler->sample_count.x;
		p.max_vertices = 8; p.seed = seed; p.flags = flags;
		// texture2d::_pixels is a packed vec3 array (texture.h:29); pixel(0,0) is its first float
		float* rgb = &target->pixel(uvec2(0, 0)).x;
		if (plu_gpu_render(ctx, &p, rgb, &last_stats) != PLU_OK) throw std::runtime_error(plu_gpu_last_error(ctx));
	}
};

// Drop-in for plu::postprocesser (main.cpp:76-112): same postprocess(shared_ptr<texture2d>) call.
struct gpu_postprocesser {
	void postprocess(shared_ptr<texture2d> rt) {
		plu_gpu_ctx* ctx = gpu::shared_ctx();
		if (plu_gpu_postprocess(ctx, (int32_t)rt->size.x, (int32_t)rt->size.y, &rt->pixel(uvec2(0, 0)).x) != PLU_OK)
			throw std::runtime_error(plu_gpu_last_error(ctx));
	}
};

inline void gpu_write_file(const std::string& filename, const std::vector<unsigned char>& bytes) {
	FILE* f = fopen(filename.c_str(), "wb");
	if (!f) throw std::runtime_error("cannot open " + filename);
	fwrite(bytes.data(), 1, bytes.size(), f);
	fclose(f);
}

// texture2d::write_bmp(filename) (texture.cpp:178-194) with clamp / quantise / BMP layout on the device.
inline void gpu_write_bmp(shared_ptr<texture2d> rt, const std::string& filename) {
	plu_gpu_ctx* ctx = gpu::shared_ctx();
	int64_t n = plu_bmp_size((int32_t)rt->size.x, (int32_t)rt->size.y);
	std::vector<unsigned char> bytes((size_t)n);
	if (plu_gpu_encode_bmp(ctx, (int32_t)rt->size.x, (int32_t)rt->size.y, &rt->pixel(uvec2(0, 0)).x, 0, bytes.data(), n) != PLU_OK)
		throw std::runtime_error(plu_gpu_last_error(ctx));
	gpu_write_file(filename, bytes);
}

// main.cpp:177-209 without the watermark, with the pixels crossing PCIe once: r.render(tx) leaves the image resident on
// the device; the tone map (main.cpp:183) and the .bmp bytes (main.cpp:209) are produced from that resident copy.  On
// return `tx` holds the tone-mapped image exactly as after plu::postprocesser, and `filename` the reference's .bmp.
inline void gpu_render_post_write(const gpu_renderer& r, shared_ptr<texture2d> tx, const std::string& filename) {
	r.render(tx);
	plu_gpu_ctx* ctx = gpu::shared_ctx();
	if (plu_gpu_postprocess_resident(ctx, &tx->pixel(uvec2(0, 0)).x) != PLU_OK) throw std::runtime_error(plu_gpu_last_error(ctx));
	int64_t n = plu_bmp_size((int32_t)tx->size.x, (int32_t)tx->size.y);
	std::vector<unsigned char> bytes((size_t)n);
	if (plu_gpu_encode_bmp_resident(ctx, 0, bytes.data(), n) != PLU_OK) throw std::runtime_error(plu_gpu_last_error(ctx));
	gpu_write_file(filename, bytes);
}
}  // namespace plu
