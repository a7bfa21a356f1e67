// scene_store.cpp — plu_scene handle: deep copies, validation and the flat ".pluscene" file
// form of a plu_scene_desc.  Host plumbing only; no CUDA here.
#include "scene_store.h"
#include <cstdio>
#include <cstring>
#include <memory>
#include <new>
#include <stdexcept>
#include <sys/stat.h>

void plu_scene::seal() {
	desc = plu_scene_desc{};
	desc.abi_version = PLUGPU_ABI_VERSION;
	desc.cam = cam;
	desc.n_prims = (int32_t)prim_tag.size();
	desc.n_materials = (int32_t)materials.size();
	desc.n_textures = (int32_t)textures.size();
	desc.n_lights = (int32_t)lights.size();
	desc.prim_tag = prim_tag.data();
	desc.prim_material = prim_material.data();
	desc.prim_data = prim_data.data();
	desc.prim_tri_uv = prim_tri_uv.empty() ? nullptr : prim_tri_uv.data();
	desc.prim_tri_normal = prim_tri_normal.empty() ? nullptr : prim_tri_normal.data();
	desc.materials = materials.data();
	desc.textures = textures.data();
	desc.lights = lights.data();
	desc.image_texels = image_texels.empty() ? nullptr : image_texels.data();
	desc.n_image_texels = (int64_t)(image_texels.size() / 3);
	desc.prim_tie_rank = prim_tie_rank.empty() ? nullptr : prim_tie_rank.data();
}

int32_t plu_scene::add_prim(int32_t tag, int32_t material, const float data[12]) {
	prim_tag.push_back(tag);
	prim_material.push_back(material);
	prim_data.insert(prim_data.end(), data, data + 12);
	return (int32_t)prim_tag.size() - 1;
}

namespace pluhost {

void set_err(char* err, int err_len, const std::string& msg) {
	if (err && err_len > 0) {
		std::snprintf(err, (size_t)err_len, "%s", msg.c_str());
	}
}

std::string validate_desc(const plu_scene_desc* d) {
	if (!d) return "scene desc is NULL";
	if (d->abi_version != PLUGPU_ABI_VERSION) return "scene desc abi_version mismatch";
	if (d->n_prims < 0 || d->n_materials < 0 || d->n_textures < 0 || d->n_lights < 0)
		return "negative table size";
	if (d->n_prims > 0 && (!d->prim_tag || !d->prim_material || !d->prim_data))
		return "prim tables missing";
	if (d->n_materials > 0 && !d->materials) return "material table missing";
	if (d->n_textures > 0 && !d->textures) return "texture table missing";
	if (d->n_lights > 0 && !d->lights) return "light table missing";
	for (int32_t i = 0; i < d->n_prims; ++i) {
		int32_t t = d->prim_tag[i];
		if (t < PLU_PRIM_SPHERE || t > PLU_PRIM_TRIANGLE) return "unknown prim tag";
		int32_t m = d->prim_material[i];
		if (m < 0 || m >= d->n_materials) return "prim material index out of range";
		if (t == PLU_PRIM_TRIANGLE && (!d->prim_tri_uv || !d->prim_tri_normal))
			return "triangle prim without uv/normal tables";
	}
	for (int32_t i = 0; i < d->n_materials; ++i) {
		const plu_material_desc& m = d->materials[i];
		if (m.tag < PLU_MAT_DIFFUSE || m.tag > PLU_MAT_EMISSION) return "unknown material tag";
		if (m.texture >= d->n_textures) return "material texture index out of range";
		if (m.tag == PLU_MAT_EMISSION && (m.light < 0 || m.light >= d->n_lights))
			return "emission material without a light";
	}
	for (int32_t i = 0; i < d->n_textures; ++i) {
		const plu_texture_desc& t = d->textures[i];
		if (t.tag < PLU_TEX_CHECKERBOARD || t.tag > PLU_TEX_IMAGE) return "unknown texture tag";
		if (t.tag == PLU_TEX_IMAGE) {
			if (t.img_w <= 0 || t.img_h <= 0 || t.img_offset < 0 ||
			    (int64_t)t.img_offset + (int64_t)t.img_w * t.img_h > d->n_image_texels || !d->image_texels)
				return "image texture outside image_texels";
		}
	}
	for (int32_t i = 0; i < d->n_lights; ++i) {
		const plu_light_desc& l = d->lights[i];
		if (l.tag < PLU_LIGHT_POINT || l.tag > PLU_LIGHT_DIFFUSE_AREA) return "unknown light tag";
		if (l.tag == PLU_LIGHT_DIFFUSE_AREA && (l.prim < 0 || l.prim >= d->n_prims))
			return "area light surface index out of range";
	}
	return "";
}

void copy_desc(const plu_scene_desc* d, plu_scene& s) {
	s.cam = d->cam;
	size_t n = (size_t)d->n_prims;
	s.prim_tag.assign(d->prim_tag, d->prim_tag + n);
	s.prim_material.assign(d->prim_material, d->prim_material + n);
	s.prim_data.assign(d->prim_data, d->prim_data + n * 12);
	if (d->prim_tri_uv) s.prim_tri_uv.assign(d->prim_tri_uv, d->prim_tri_uv + n * 6); else s.prim_tri_uv.clear();
	if (d->prim_tri_normal) s.prim_tri_normal.assign(d->prim_tri_normal, d->prim_tri_normal + n * 9); else s.prim_tri_normal.clear();
	if (d->prim_tie_rank) s.prim_tie_rank.assign(d->prim_tie_rank, d->prim_tie_rank + n); else s.prim_tie_rank.clear();
	s.materials.assign(d->materials, d->materials + d->n_materials);
	s.textures.assign(d->textures, d->textures + d->n_textures);
	s.lights.assign(d->lights, d->lights + d->n_lights);
	if (d->image_texels) s.image_texels.assign(d->image_texels, d->image_texels + d->n_image_texels * 3); else s.image_texels.clear();
	s.seal();
}

}  // namespace pluhost

// ---- ".pluscene" flat file ------------------------------------------------------------------
// magic[8] "PLUSCN1\0" | i32 width height n_strata flags | plu_camera_desc | i32 n_prims
// n_materials n_textures n_lights | i64 n_image_texels | arrays in desc order.  flags bit0 =
// tri tables present, bit1 = tie ranks present.  Little-endian, no padding between sections.
namespace {
const char kMagic[8] = {'P', 'L', 'U', 'S', 'C', 'N', '1', '\0'};

template <class T> bool wr(FILE* f, const T* p, size_t n) { return n == 0 || std::fwrite(p, sizeof(T), n, f) == n; }
template <class T> bool rd(FILE* f, T* p, size_t n) { return n == 0 || std::fread(p, sizeof(T), n, f) == n; }
}  // namespace

extern "C" int plu_scene_save_flat(const plu_scene_desc* d, int32_t width, int32_t height,
                                   int32_t n_strata, const char* path) {
	if (!path || !pluhost::validate_desc(d).empty()) return PLU_E_ARG;
	FILE* f = std::fopen(path, "wb");
	if (!f) return PLU_E_IO;
	size_t n = (size_t)d->n_prims;
	bool tri = d->prim_tri_uv && d->prim_tri_normal;
	int32_t head[4] = {width, height, n_strata, (tri ? 1 : 0) | (d->prim_tie_rank ? 2 : 0)};
	int32_t counts[4] = {d->n_prims, d->n_materials, d->n_textures, d->n_lights};
	int64_t ntex = d->image_texels ? d->n_image_texels : 0;
	bool ok = wr(f, kMagic, 8) && wr(f, head, 4) && wr(f, &d->cam, 1) && wr(f, counts, 4) && wr(f, &ntex, 1) &&
	          wr(f, d->prim_tag, n) && wr(f, d->prim_material, n) && wr(f, d->prim_data, n * 12) &&
	          (!tri || (wr(f, d->prim_tri_uv, n * 6) && wr(f, d->prim_tri_normal, n * 9))) &&
	          (!d->prim_tie_rank || wr(f, d->prim_tie_rank, n)) &&
	          wr(f, d->materials, (size_t)d->n_materials) && wr(f, d->textures, (size_t)d->n_textures) &&
	          wr(f, d->lights, (size_t)d->n_lights) && wr(f, d->image_texels, (size_t)ntex * 3);
	ok = (std::fclose(f) == 0) && ok;
	return ok ? PLU_OK : PLU_E_IO;
}

// No exception crosses the C ABI: allocation failures and length errors of the host plumbing become error codes.
#define PLU_GUARD_BEGIN try {
#define PLU_GUARD_END(err, err_len)                                                                         \
	} catch (const std::bad_alloc&) { pluhost::set_err(err, err_len, "out of host memory"); return PLU_E_NOMEM; } \
	catch (const std::exception& e) { pluhost::set_err(err, err_len, e.what()); return PLU_E_PARSE; }

static int load_flat_impl(const char* path, plu_scene** out, char* err, int err_len) {
	FILE* f = std::fopen(path, "rb");
	if (!f) { pluhost::set_err(err, err_len, std::string("cannot open ") + path); return PLU_E_IO; }
	struct stat sb; int64_t file_bytes = (fstat(fileno(f), &sb) == 0) ? (int64_t)sb.st_size : -1;
	std::unique_ptr<plu_scene> s(new plu_scene);
	char magic[8]; int32_t head[4], counts[4]; int64_t ntex = 0;
	bool ok = rd(f, magic, 8) && std::memcmp(magic, kMagic, 8) == 0 && rd(f, head, 4) && rd(f, &s->cam, 1) &&
	          rd(f, counts, 4) && rd(f, &ntex, 1) && counts[0] >= 0 && counts[1] >= 0 && counts[2] >= 0 &&
	          counts[3] >= 0 && ntex >= 0;
	if (ok && file_bytes >= 0) {   // the counts of a truncated or corrupt header must not size the vectors
		const int64_t n = counts[0];
		__int128 need = (__int128)n * (4 + 4 + 48) + ((head[3] & 1) ? (__int128)n * 60 : 0) + ((head[3] & 2) ? (__int128)n * 4 : 0) +
		                (__int128)counts[1] * (int64_t)sizeof(plu_material_desc) + (__int128)counts[2] * (int64_t)sizeof(plu_texture_desc) +
		                (__int128)counts[3] * (int64_t)sizeof(plu_light_desc) + (__int128)ntex * 12;
		ok = need <= (__int128)file_bytes;
	}
	if (ok) {
		size_t n = (size_t)counts[0];
		s->width = head[0]; s->height = head[1]; s->n_strata = head[2];
		s->prim_tag.resize(n); s->prim_material.resize(n); s->prim_data.resize(n * 12);
		ok = rd(f, s->prim_tag.data(), n) && rd(f, s->prim_material.data(), n) && rd(f, s->prim_data.data(), n * 12);
		if (ok && (head[3] & 1)) {
			s->prim_tri_uv.resize(n * 6); s->prim_tri_normal.resize(n * 9);
			ok = rd(f, s->prim_tri_uv.data(), n * 6) && rd(f, s->prim_tri_normal.data(), n * 9);
		}
		if (ok && (head[3] & 2)) { s->prim_tie_rank.resize(n); ok = rd(f, s->prim_tie_rank.data(), n); }
		s->materials.resize((size_t)counts[1]); s->textures.resize((size_t)counts[2]); s->lights.resize((size_t)counts[3]);
		s->image_texels.resize((size_t)ntex * 3);
		ok = ok && rd(f, s->materials.data(), s->materials.size()) && rd(f, s->textures.data(), s->textures.size()) &&
		     rd(f, s->lights.data(), s->lights.size()) && rd(f, s->image_texels.data(), s->image_texels.size());
	}
	std::fclose(f);
	if (!ok) { pluhost::set_err(err, err_len, std::string("malformed .pluscene file ") + path); return PLU_E_PARSE; }
	s->seal();
	std::string v = pluhost::validate_desc(&s->desc);
	if (!v.empty()) { pluhost::set_err(err, err_len, v); return PLU_E_PARSE; }
	*out = s.release();
	return PLU_OK;
}

extern "C" int plu_scene_load_flat(const char* path, plu_scene** out, char* err, int err_len) {
	if (!path || !out) return PLU_E_ARG;
	*out = nullptr;
	PLU_GUARD_BEGIN
	return load_flat_impl(path, out, err, err_len);
	PLU_GUARD_END(err, err_len)
}

extern "C" int plu_scene_from_desc(const plu_scene_desc* desc, int32_t width, int32_t height,
                                   int32_t n_strata, plu_scene** out) {
	if (!out || !pluhost::validate_desc(desc).empty()) return PLU_E_ARG;
	*out = nullptr;
	PLU_GUARD_BEGIN
	std::unique_ptr<plu_scene> s(new plu_scene);
	pluhost::copy_desc(desc, *s);
	s->width = width; s->height = height; s->n_strata = n_strata;
	*out = s.release();
	return PLU_OK;
	PLU_GUARD_END(nullptr, 0)
}

extern "C" const plu_scene_desc* plu_scene_get_desc(const plu_scene* s) { return s ? &s->desc : nullptr; }

extern "C" void plu_scene_get_defaults(const plu_scene* s, int32_t* width, int32_t* height, int32_t* n_strata) {
	if (!s) return;
	if (width) *width = s->width;
	if (height) *height = s->height;
	if (n_strata) *n_strata = s->n_strata;
}

extern "C" int plu_scene_override(plu_scene* s, int32_t width, int32_t height, int32_t n_strata) {
	if (!s) return PLU_E_ARG;
	if (width > 0 && height > 0) {
		s->width = width; s->height = height;
		// camera ctor: inv_image_size(1.f / (vec2)target_size), camera.h:18
		s->cam.inv_image_size[0] = 1.f / (float)(uint32_t)width;
		s->cam.inv_image_size[1] = 1.f / (float)(uint32_t)height;
	}
	if (n_strata > 0) s->n_strata = n_strata;
	s->seal();
	return PLU_OK;
}

extern "C" void plu_scene_free(plu_scene* s) { delete s; }

extern "C" uint32_t plu_abi_version(void) { return PLUGPU_ABI_VERSION; }
