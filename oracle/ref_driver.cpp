// ref_driver.cpp — driver around the UNMODIFIED reference render path (oracle/_ref/plu_ref).
//
// TEST INFRASTRUCTURE ONLY: used to pin oracle/oracle_pt.cpp, to generate tests/golden/* and as
// the timed CPU baseline (bench.py --impl reference / cpu_baseline.kind == "reference").
// Nothing in the product path links, loads or runs this.
//
// It is a unity TU: the reference's own .cpp files are #included from where they lie (the -I
// paths come from oracle/build_ref.sh), replacing only src/main.cpp (which post-processes,
// watermarks and blocks on getchar(), main.cpp:183-210).  One TU => ONE `static mt19937 RNG`
// (cmmn.h:240), so PLU_SEED makes single-threaded runs reproducible.
//
//   plu_ref primary    <scene.urn> <W> <H> <out.bin>            pixel-centre pinhole rays -> id,t,n
//   plu_ref rays       <scene.urn> <rays.bin> <out.bin>         arbitrary rays (n x 6 f32) -> id,t,n
//   plu_ref render     <scene.urn> <W> <H> <smp> <out.bin>      renderer::render (its own threads)
//   plu_ref render-st  <scene.urn> <W> <H> <smp> <seed> <out.bin> [x0 y0 x1 y1 [samples.bin]]
//                                                               1 thread, RNG reseeded per pixel
//   plu_ref render-flat <scene.pluscene> <W> <H> <smp> <out.bin|->  renderer::render on a scene rebuilt from flat tables
//   plu_ref dump-scene <scene.urn> <W> <H> <smp> <out.pluscene> flat tables via the integration adapter
//   plu_ref leaf-order <scene.urn>                              load-order ids of the bvh leaves in DFS order
#include "cmmn.h"
#include <array>
#include <set>
#include <tuple>
#include <list>
#include <fstream>
#include <sstream>
#include <cassert>
#include <functional>
#include <map>
#include <chrono>

// open up props::color and bvh_tree for the flattener (see integration/plu_gpu_renderer.h)
#define private public
#define protected public
#define class struct
#include "surfaces/triangle.h"  // main.cpp:5 includes it ahead of scene.h, which needs it
#include "scene.h"
#include "renderer.h"
#include "renderer.cpp"
#include "material.cpp"
#include "texture.cpp"
#include "urn.cpp"
#include "surfaces/box.cpp"
#include "surfaces/sphere.cpp"
#include "surfaces/triangle.cpp"
#include "surfaces/bvh_tree.cpp"
#include "../integration/plu_gpu_renderer.h"
#undef class
#undef protected
#undef private

// plu_ref (checker): the flat-scene store is compiled in.  plu_ref_gpu (-DPLU_WITH_GPU, the drop-in proof): the same
// symbols and the plu_gpu_* entry points come from libplugpu.so instead.
#ifndef PLU_WITH_GPU
#include "../plutracer_b200/csrc/host/scene_store.cpp"
#endif

namespace {

struct loaded {
	std::unique_ptr<plu::scenes::scene> sc;
	std::vector<const plu::surface*> load_order;  // ids
	std::map<const plu::surface*, int32_t> id_of;
	std::unique_ptr<plu::surfaces::bvh_tree> tree;
};

void load(loaded& L, const std::string& path, int W, int H, int smp) {
	auto ts = urn::token_stream(std::ifstream(path));
	urn::value tlv = urn::value(ts);
	std::list<std::string> args;
	if (W > 0) { args.push_back("/res"); args.push_back(std::to_string(W) + "x" + std::to_string(H)); }
	if (smp > 0) { args.push_back("/smp"); args.push_back(std::to_string(smp)); }
	L.sc.reset(new plu::scenes::scene(tlv, args));
	for (auto& s : L.sc->surfs) { L.id_of[s.get()] = (int32_t)L.load_order.size(); L.load_order.push_back(s.get()); }
	L.tree.reset(new plu::surfaces::bvh_tree(L.sc->surfs));  // sorts sc->surfs in place (bvh_tree.cpp:23)
}

// Rebuild the reference's OWN object graph (its sphere / box / material / light classes) from a flat
// ".pluscene" fixture, so the unmodified renderer can run where the .urn files are not available (the
// GPU box has no /root/reference).  Inverse of integration/plu_gpu_renderer.h's flattener.
struct flat_loaded {
	plu::camera cam;
	std::vector<std::shared_ptr<plu::surface>> surfs;
	std::vector<std::shared_ptr<plu::light>> lights;
	int samples = 0; int W = 0, H = 0;
	std::vector<glm::vec3> tri_v, tri_n; std::vector<glm::vec2> tri_t;   // storage the triangles point into (scene.h:145-148 keeps it per mesh)
	std::unique_ptr<plu::surfaces::bvh_tree> tree;
};

glm::vec3 v3of(const float* f) { return glm::vec3(f[0], f[1], f[2]); }

void load_flat(flat_loaded& L, const std::string& path, int W, int H, int smp) {
	plu_scene* ps = nullptr; char err[256];
	if (plu_scene_load_flat(path.c_str(), &ps, err, sizeof(err)) != PLU_OK) throw std::runtime_error(err);
	if (W > 0 || smp > 0) plu_scene_override(ps, W, H, smp);
	const plu_scene_desc* d = plu_scene_get_desc(ps);
	int32_t w, h, n; plu_scene_get_defaults(ps, &w, &h, &n);
	L.W = w; L.H = h; L.samples = n;
	L.cam.pos = v3of(d->cam.pos); L.cam.look = v3of(d->cam.look); L.cam.up = v3of(d->cam.up); L.cam.right = v3of(d->cam.right);
	L.cam.inv_image_size = glm::vec2(d->cam.inv_image_size[0], d->cam.inv_image_size[1]);
	L.cam.w = d->cam.w; L.cam.lens_radius = d->cam.lens_radius; L.cam.focal_distance = d->cam.focal_distance;
	auto color_of = [&](const plu_material_desc& m) -> plu::props::color {
		if (m.texture < 0) return plu::props::color(v3of(m.color));
		const plu_texture_desc& t = d->textures[m.texture];
		if (t.tag == PLU_TEX_CHECKERBOARD) return plu::props::color(std::make_shared<plu::textures::checkerboard_texture>(v3of(t.c0), v3of(t.c1), t.scale));
		if (t.tag == PLU_TEX_GRID) return plu::props::color(std::make_shared<plu::textures::grid_texture>(v3of(t.c0), v3of(t.c1), t.scale, t.line));
		throw std::runtime_error("render-flat: image textures are not rebuilt");
	};
	std::vector<std::shared_ptr<plu::material>> mats((size_t)d->n_materials);
	std::vector<std::shared_ptr<plu::materials::emission_material>> emis((size_t)d->n_materials);
	for (int32_t i = 0; i < d->n_materials; ++i) {
		const plu_material_desc& m = d->materials[i];
		switch (m.tag) {
			case PLU_MAT_DIFFUSE: mats[i] = std::make_shared<plu::materials::diffuse_material>(color_of(m)); break;
			case PLU_MAT_PERFECT_REFLECTION: mats[i] = std::make_shared<plu::materials::perfect_reflection_material>(color_of(m), v3of(m.p0), v3of(m.p1)); break;
			case PLU_MAT_PERFECT_REFRACTION: mats[i] = std::make_shared<plu::materials::perfect_refraction_material>(color_of(m), m.p0[0], m.p0[1]); break;
			case PLU_MAT_GLASS: mats[i] = std::make_shared<plu::materials::glass_material>(color_of(m), m.p0[0]); break;
			default: emis[i] = std::make_shared<plu::materials::emission_material>(nullptr); mats[i] = emis[i]; break;
		}
	}
	L.tri_v.resize((size_t)d->n_prims * 3); L.tri_n.resize((size_t)d->n_prims * 3); L.tri_t.resize((size_t)d->n_prims * 3);
	for (int32_t s = 0; s < d->n_prims; ++s) {
		const float* f = d->prim_data + (size_t)s * 12;
		std::shared_ptr<plu::surface> surf;
		if (d->prim_tag[s] == PLU_PRIM_SPHERE) surf = std::make_shared<plu::surfaces::sphere>(v3of(f), f[3], nullptr);
		else if (d->prim_tag[s] == PLU_PRIM_BOX) {
			auto b = std::make_shared<plu::surfaces::box>(glm::vec3(0.f), glm::vec3(0.f), nullptr);
			b->box_bounds._min = v3of(f); b->box_bounds._max = v3of(f + 4);   // exact bits, not c -+ e again
			surf = b;
		} else {
			glm::vec3* v = &L.tri_v[(size_t)s * 3]; glm::vec3* nn = &L.tri_n[(size_t)s * 3]; glm::vec2* tt = &L.tri_t[(size_t)s * 3];
			for (int k = 0; k < 3; ++k) {
				v[k] = v3of(f + 4 * k);
				nn[k] = v3of(d->prim_tri_normal + (size_t)s * 9 + 3 * k);
				tt[k] = glm::vec2(d->prim_tri_uv[(size_t)s * 6 + 2 * k], d->prim_tri_uv[(size_t)s * 6 + 2 * k + 1]);
			}
			surf = std::make_shared<plu::surfaces::triangle>(v, v + 1, v + 2, nn, nn + 1, nn + 2, tt, tt + 1, tt + 2, nullptr);
		}
		surf->mat = mats[(size_t)d->prim_material[s]];
		L.surfs.push_back(surf);
	}
	L.lights.resize((size_t)d->n_lights);
	for (int32_t i = 0; i < d->n_lights; ++i) {
		const plu_light_desc& l = d->lights[i];
		if (l.tag == PLU_LIGHT_POINT) L.lights[i] = std::make_shared<plu::lights::point_light>(v3of(l.a), v3of(l.b));
		else {
			auto al = std::make_shared<plu::lights::diffuse_area_light>(v3of(l.b), L.surfs[(size_t)l.prim]);   // area = surf->area()
			for (int32_t m = 0; m < d->n_materials; ++m) if (d->materials[m].tag == PLU_MAT_EMISSION && d->materials[m].light == i) emis[m]->L = al;
			L.lights[i] = al;
		}
	}
	plu_scene_free(ps);
	std::vector<std::shared_ptr<plu::surface>> sorted = L.surfs;   // bvh_tree sorts its argument in place (bvh_tree.cpp:23)
	L.tree.reset(new plu::surfaces::bvh_tree(sorted));
}

void write_all(const std::string& path, const void* p, size_t bytes, bool append = false) {
	FILE* f = fopen(path.c_str(), append ? "ab" : "wb");
	if (!f || fwrite(p, 1, bytes, f) != bytes) { fprintf(stderr, "plu_ref: cannot write %s\n", path.c_str()); exit(2); }
	fclose(f);
}

void dump_hits(const loaded& L, const std::vector<plu::ray>& rays, const std::string& out) {
	size_t n = rays.size();
	std::vector<int32_t> id(n); std::vector<float> t(n), nrm(n * 3);
	for (size_t i = 0; i < n; ++i) {
		plu::hit_record hr;
		if (L.tree->hit(rays[i], &hr)) {
			id[i] = L.id_of.at(hr.surf); t[i] = hr.t;
			nrm[3 * i] = hr.norm.x; nrm[3 * i + 1] = hr.norm.y; nrm[3 * i + 2] = hr.norm.z;
		} else { id[i] = -1; t[i] = 0.f; nrm[3 * i] = nrm[3 * i + 1] = nrm[3 * i + 2] = 0.f; }
	}
	write_all(out, id.data(), n * 4);
	write_all(out, t.data(), n * 4, true);
	write_all(out, nrm.data(), n * 12, true);
}

// texture2d whose buffer has one spare vec3: renderer::render writes pixel(W, H-1), one element
// past the end (sampler.h:74-78 overflow column, texture.h:44-47)
struct padded_target : public plu::texture2d {
	padded_target(glm::uvec2 s) : plu::texture2d(s) {
		delete[] _pixels;
		_pixels = new glm::vec3[(size_t)s.x * s.y + 1];
	}
};

}  // namespace

int main(int argc, char** argv) {
	if (argc < 3) { fprintf(stderr, "usage: plu_ref <mode> <scene.urn> ...\n"); return 1; }
	std::string mode = argv[1], scene_path = argv[2];
	try {
		if (mode == "primary" && argc == 6) {
			int W = atoi(argv[3]), H = atoi(argv[4]);
			loaded L; load(L, scene_path, W, H, 0);
			plu::camera cam = L.sc->cam; cam.lens_radius = 0.f;
			std::vector<plu::ray> rays((size_t)W * H);
			for (int y = 0; y < H; ++y) for (int x = 0; x < W; ++x) {
				plu::sample s(glm::vec2(x + .5f, y + .5f)); s.lens = glm::vec2(.5f);
				cam.generate_ray(rays[(size_t)y * W + x], s);
			}
			dump_hits(L, rays, argv[5]);
		} else if (mode == "rays" && argc == 5) {
			loaded L; load(L, scene_path, 0, 0, 0);
			std::ifstream in(argv[3], std::ios::binary);
			std::vector<char> buf((std::istreambuf_iterator<char>(in)), std::istreambuf_iterator<char>());
			size_t n = buf.size() / 24; const float* f = (const float*)buf.data();
			std::vector<plu::ray> rays(n);
			for (size_t i = 0; i < n; ++i) rays[i] = plu::ray(glm::vec3(f[6*i], f[6*i+1], f[6*i+2]), glm::vec3(f[6*i+3], f[6*i+4], f[6*i+5]));
			dump_hits(L, rays, argv[4]);
		} else if (mode == "render" && argc == 7) {
			int W = atoi(argv[3]), H = atoi(argv[4]), smp = atoi(argv[5]);
			loaded L; load(L, scene_path, W, H, smp);
			auto tx = std::make_shared<padded_target>(glm::uvec2(W, H));
			plu::samplers::stratified_sampler sampler(tx->size, glm::uvec2((unsigned)L.sc->samples), true);
			plu::renderer r(L.tree.get(), L.sc->cam, glm::uvec2(32), &sampler, L.sc->lights);
			std::streambuf* old = std::cout.rdbuf(); std::ostringstream sink; std::cout.rdbuf(sink.rdbuf());
			auto t0 = std::chrono::high_resolution_clock::now();
			r.render(tx);  // same span as main.cpp:176-178
			auto t1 = std::chrono::high_resolution_clock::now();
			std::cout.rdbuf(old);
			write_all(argv[6], &tx->pixel(glm::uvec2(0, 0)), (size_t)W * H * 12);
			double ms = std::chrono::duration<double, std::milli>(t1 - t0).count();
			printf("{\"mode\": \"render\", \"width\": %d, \"height\": %d, \"smp\": %d, \"samples\": %llu, \"ms\": %.3f, \"threads\": %u}\n",
			       W, H, smp, (unsigned long long)W * H * smp * smp, ms, std::thread::hardware_concurrency());
		} else if (mode == "render-flat" && argc == 7) {
			// renderer::render of the unmodified reference on a scene rebuilt from a .pluscene fixture
			int W = atoi(argv[3]), H = atoi(argv[4]), smp = atoi(argv[5]);
			flat_loaded L; load_flat(L, scene_path, W, H, smp);
			auto tx = std::make_shared<padded_target>(glm::uvec2(L.W, L.H));
			plu::samplers::stratified_sampler sampler(tx->size, glm::uvec2((unsigned)L.samples), true);
			plu::renderer r(L.tree.get(), L.cam, glm::uvec2(32), &sampler, L.lights);
			std::streambuf* old = std::cout.rdbuf(); std::ostringstream sink; std::cout.rdbuf(sink.rdbuf());
			auto t0 = std::chrono::high_resolution_clock::now();
			r.render(tx);
			auto t1 = std::chrono::high_resolution_clock::now();
			std::cout.rdbuf(old);
			if (std::string(argv[6]) != "-") write_all(argv[6], &tx->pixel(glm::uvec2(0, 0)), (size_t)L.W * L.H * 12);
			double ms = std::chrono::duration<double, std::milli>(t1 - t0).count();
			printf("{\"mode\": \"render-flat\", \"width\": %d, \"height\": %d, \"smp\": %d, \"samples\": %llu, \"ms\": %.3f, \"threads\": %u}\n",
			       L.W, L.H, L.samples, (unsigned long long)L.W * L.H * L.samples * L.samples, ms, std::thread::hardware_concurrency());
#ifdef PLU_WITH_GPU
		} else if ((mode == "render-gpu" || mode == "render-gpu-flat") && (argc == 7 || argc == 8)) {
			// THE DROP-IN: the reference's own process (its loader or its object graph, its live bvh_tree, its
			// sampler and texture2d) with plu::gpu_renderer in place of plu::renderer (main.cpp:171,177), then
			// optionally its post-process + BMP writer through the device path (main.cpp:183,209)
			int W = atoi(argv[3]), H = atoi(argv[4]), smp = atoi(argv[5]);
			loaded L; flat_loaded F;
			const bool flat = mode == "render-gpu-flat";
			if (flat) load_flat(F, scene_path, W, H, smp); else load(L, scene_path, W, H, smp);
			const int w = flat ? F.W : (int)L.sc->resolution.x, h = flat ? F.H : (int)L.sc->resolution.y;
			const int n = flat ? F.samples : (int)L.sc->samples;
			auto tx = std::make_shared<plu::texture2d>(glm::uvec2(w, h));
			plu::samplers::stratified_sampler sampler(tx->size, glm::uvec2((unsigned)n), true);
			plu::gpu_renderer r(flat ? (plu::surface*)F.tree.get() : (plu::surface*)L.tree.get(), flat ? F.cam : L.sc->cam, glm::uvec2(32), &sampler,
			                    flat ? F.lights : L.sc->lights);
			// PLU_REF_REPEAT=k: render k times into fresh targets (the context, wave buffers and NCCL communicators of
			// the drop-in live across calls), report the LAST call's time; PLU_REF_FUSED=1 (with a .bmp argument): the
			// resident-framebuffer flow gpu_render_post_write, <out.bin> then holds the tone-mapped image; PLU_REF_PIN=1: the
			// target's pixels are page-locked first (plu::gpu::pinned_target, outside the timed call)
			const int repeat = std::getenv("PLU_REF_REPEAT") ? std::max(1, atoi(std::getenv("PLU_REF_REPEAT"))) : 1;
			const bool fused = argc == 8 && std::getenv("PLU_REF_FUSED") != nullptr;
			double ms = 0.0, ms_first = 0.0;
			for (int it = 0; it < repeat; ++it) {
				if (it > 0) tx = std::make_shared<plu::texture2d>(glm::uvec2(w, h));
				std::unique_ptr<plu::gpu::pinned_target> pin;
				if (std::getenv("PLU_REF_PIN")) pin.reset(new plu::gpu::pinned_target(tx));
				auto t0 = std::chrono::high_resolution_clock::now();
				if (fused) plu::gpu_render_post_write(r, tx, argv[7]); else r.render(tx);
				auto t1 = std::chrono::high_resolution_clock::now();
				ms = std::chrono::duration<double, std::milli>(t1 - t0).count();
				if (it == 0) ms_first = ms;
			}
			write_all(argv[6], &tx->pixel(glm::uvec2(0, 0)), (size_t)w * h * 12);
			if (argc == 8 && !fused) { plu::gpu_postprocesser pp; pp.postprocess(tx); plu::gpu_write_bmp(tx, argv[7]); }
			printf("{\"mode\": \"%s\", \"width\": %d, \"height\": %d, \"smp\": %d, \"samples\": %llu, \"ms\": %.3f, \"ms_first_call\": %.3f, "
			       "\"devices\": %d, \"ms_device\": %.3f, \"ms_reduce\": %.3f}\n", mode.c_str(),
			       w, h, n, (unsigned long long)w * h * n * n, ms, ms_first, r.last_stats.n_devices, r.last_stats.ms_device, r.last_stats.ms_reduce);
#endif
		} else if (mode == "render-st" && (argc == 8 || argc == 12 || argc == 13)) {
			int W = atoi(argv[3]), H = atoi(argv[4]), smp = atoi(argv[5]);
			unsigned long seed = strtoul(argv[6], nullptr, 10);
			int x0 = 0, y0 = 0, x1 = W, y1 = H;
			if (argc >= 12) { x0 = atoi(argv[8]); y0 = atoi(argv[9]); x1 = atoi(argv[10]); y1 = atoi(argv[11]); }
			loaded L; load(L, scene_path, W, H, smp);
			plu::renderer r(L.tree.get(), L.sc->cam, glm::uvec2(32), nullptr, L.sc->lights);
			std::vector<float> img((size_t)(x1 - x0) * (y1 - y0) * 3, 0.f), per_sample;
			float smp_wgt = 1.f / (float)(smp * smp);
			plu::memory::arena arena; std::vector<plu::sample> smps; plu::ray ray;
			auto t0 = std::chrono::high_resolution_clock::now();
			for (int y = y0; y < y1; ++y) for (int x = x0; x < x1; ++x) {
				plu::rnd::RNG.seed(seed + (unsigned long)y * W + x);
				plu::samplers::stratified_sampler px(glm::uvec2(0, 1), glm::uvec2(smp), true, glm::uvec2(x, y));
				px.generate_samples(smps);
				glm::vec3 acc(0.f);
				for (auto& s : smps) {
					L.sc->cam.generate_ray(ray, s);
					glm::vec3 c = r.ray_color(ray, s, arena);
					acc += c * smp_wgt;  // renderer.cpp:132
					if (argc == 13) { per_sample.push_back(c.x); per_sample.push_back(c.y); per_sample.push_back(c.z); }
				}
				arena.free_all();
				size_t o = ((size_t)(y - y0) * (x1 - x0) + (x - x0)) * 3;
				img[o] = acc.x; img[o + 1] = acc.y; img[o + 2] = acc.z;
			}
			auto t1 = std::chrono::high_resolution_clock::now();
			write_all(argv[7], img.data(), img.size() * 4);
			if (argc == 13) write_all(argv[12], per_sample.data(), per_sample.size() * 4);
			printf("{\"mode\": \"render-st\", \"samples\": %llu, \"ms\": %.3f, \"threads\": 1}\n",
			       (unsigned long long)(x1 - x0) * (y1 - y0) * smp * smp, std::chrono::duration<double, std::milli>(t1 - t0).count());
		} else if (mode == "dump-scene" && argc == 7) {
			int W = atoi(argv[3]), H = atoi(argv[4]), smp = atoi(argv[5]);
			loaded L; load(L, scene_path, W, H, smp);
			plu::gpu::flat_scene fs;
			plu::gpu::flattener(fs).run(L.load_order, L.sc->lights, L.sc->cam);
			// real DFS order of the reference tree, in load-order ids
			std::vector<const plu::surface*> dfs; plu::gpu::flattener::collect_leaves(L.tree->root.get(), dfs);
			fs.prim_tie_rank.resize(dfs.size());
			for (size_t k = 0; k < dfs.size(); ++k) fs.prim_tie_rank[L.id_of.at(dfs[k])] = (int32_t)k;
			plu_scene_desc d = fs.desc();
			int rc = plu_scene_save_flat(&d, (int32_t)L.sc->resolution.x, (int32_t)L.sc->resolution.y, (int32_t)L.sc->samples, argv[6]);
			if (rc != PLU_OK) { fprintf(stderr, "plu_ref: save_flat failed (%d)\n", rc); return 2; }
		} else if (mode == "leaf-order" && argc == 3) {
			loaded L; load(L, scene_path, 0, 0, 0);
			std::vector<const plu::surface*> dfs; plu::gpu::flattener::collect_leaves(L.tree->root.get(), dfs);
			for (auto s : dfs) printf("%d ", L.id_of.at(s));
			printf("\n");
		} else {
			fprintf(stderr, "plu_ref: bad mode/arguments\n"); return 1;
		}
	} catch (const std::exception& e) {
		fprintf(stderr, "plu_ref: %s\n", e.what()); return 3;
	}
	return 0;
}
