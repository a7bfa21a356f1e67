// ref_post_driver.cpp — the reference's own post-process and BMP writer (oracle/_ref/plu_ref_post).
//
// TEST INFRASTRUCTURE ONLY (pins oracle_pt.cpp's orc_postprocess / orc_encode_bmp and generates
// tests/golden/post.npz).  plu::postprocesser lives inside src/main.cpp (:76-112), so this unity TU
// #includes main.cpp itself, from where it lies, with its main() renamed; nothing is copied.
//
//   plu_ref_post post <W> <H> <in.f32> <out.bmp> <out_tonemapped.f32>
//       in.f32: W*H*3 linear float RGB -> postprocesser::postprocess (main.cpp:88-111) -> floats out
//       -> texture2d::write_bmp (texture.cpp:178-194)
//   plu_ref_post bmp  <W> <H> <in.f32> <out.bmp>        write_bmp only (no tone map)
#define main plu_reference_main
#include "main.cpp"
#undef main
#include "renderer.cpp"
#include "material.cpp"
#include "texture.cpp"
#include "urn.cpp"
#include "surfaces/box.cpp"
#include "surfaces/sphere.cpp"
#include "surfaces/triangle.cpp"
#include "surfaces/bvh_tree.cpp"

int main(int argc, char** argv) {
	if (argc < 6) { fprintf(stderr, "usage: plu_ref_post post|bmp <W> <H> <in.f32> <out.bmp> [out.f32]\n"); return 1; }
	std::string mode = argv[1];
	int W = atoi(argv[2]), H = atoi(argv[3]);
	auto tx = std::make_shared<plu::texture2d>(glm::uvec2(W, H));
	FILE* f = fopen(argv[4], "rb");
	if (!f || fread(&tx->pixel(glm::uvec2(0, 0)), 12, (size_t)W * H, f) != (size_t)W * H) { fprintf(stderr, "plu_ref_post: cannot read %s\n", argv[4]); return 1; }
	fclose(f);
	if (mode == "post") {
		plu::postprocesser pp;
		pp.postprocess(tx);
		if (argc > 6) { FILE* o = fopen(argv[6], "wb"); fwrite(&tx->pixel(glm::uvec2(0, 0)), 12, (size_t)W * H, o); fclose(o); }
	}
	tx->write_bmp(argv[5]);
	return 0;
}
