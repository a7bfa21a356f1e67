/* oracle_pt.h — C entry points of the CPU oracle (oracle/_build/liborc.so).
 * TEST INFRASTRUCTURE ONLY: see the header of oracle_pt.cpp. */
#ifndef ORACLE_PT_H
#define ORACLE_PT_H
#include "../include/plugpu.h"
#ifdef __cplusplus
extern "C" {
#endif

/* trig: 0 = det (bit-identical with the CUDA path), 1 = libm (what the compiled reference runs) */

/* scene->hit for n rays (o, d: n*3 float): surface id (-1 miss), t, hit_record::norm. */
int orc_trace_rays(const plu_scene_desc* d, int64_t n, const float* origin3, const float* dir3,
                   int trig, int32_t* surf_id, float* t, float* normal3);
/* pixel-centre pinhole rays through camera::generate_ray (lens_radius forced to 0). */
int orc_trace_primary(const plu_scene_desc* d, int32_t width, int32_t height, int trig,
                      int32_t* surf_id, float* t, float* normal3);
/* Leaves of the restated bvh_tree in DFS order: rank[surface] = position. */
int orc_leaf_order(const plu_scene_desc* d, int32_t* rank);
/* node table of the restated tree: per node left,right,object + min.xyz,max.xyz; returns count */
int orc_bvh_nodes(const plu_scene_desc* d, int32_t max_nodes, int32_t* lro3, float* bounds6);

/* Philox stream (the CUDA path's): per-sample radiance of samples [p->sample_begin ..) of pixels
 * [pixel_begin, +pixel_count): out[(pix*count + k)*3].  Same contract as plu_gpu_sample_radiance. */
int orc_sample_radiance(const plu_scene_desc* d, const plu_render_params* p, int64_t pixel_begin,
                        int64_t pixel_count, int trig, float* out_rgb);
/* Philox stream: whole image, accumulated into out_rgb (+=) like renderer::render, n_threads
 * host threads over rows; stats->samples / rays_extend / rays_shadow filled. */
int orc_render(const plu_scene_desc* d, const plu_render_params* p, int trig, int n_threads,
               float* out_rgb, plu_render_stats* stats);

/* The reference's own stream: std::mt19937 reseeded with seed + y*W + x at every pixel, samples
 * generated and consumed in the reference's order (sampler.h:58-80) — reproduces
 * `plu_ref render-st` bit for bit.  img: (x1-x0)*(y1-y0)*3; per_sample (may be NULL):
 * pixels*n^2*3.  x_first: evaluation order of vec2(next(), next()) in the compiled reference. */
int orc_render_mt(const plu_scene_desc* d, int32_t width, int32_t height, int32_t n_strata,
                  uint32_t seed, int32_t x0, int32_t y0, int32_t x1, int32_t y1, int trig,
                  int x_first, float* img, float* per_sample);

/* postprocesser::postprocess (main.cpp:76-111): white-preserving luma Reinhard (white = 2) + gamma 1/2.2,
 * in place over n pixels.  pow_mode: 0 = powf of the host libm (what the compiled reference runs),
 * 1 = double pow rounded to float (what the CUDA path computes). */
void orc_postprocess(int64_t n_pixels, float* rgb, int pow_mode);
/* texture2d::write_bmp (texture.cpp:178-194) + stbi_write_bmp (:167-173): the bytes of the .bmp file
 * (54-byte header, 24-bit BGR, bottom-up, rows padded to 4 bytes).  Returns the byte count; writes at
 * most cap bytes. */
int64_t orc_encode_bmp(int32_t width, int32_t height, const float* rgb, unsigned char* out, int64_t cap);

void orc_philox4x32_10(const uint32_t ctr[4], const uint32_t key[2], uint32_t out[4]);
/* det trig for tests: which = 0 sin, 1 cos, 2 acos */
void orc_det_trig(int which, int64_t n, const float* x, float* y);
#ifdef __cplusplus
}
#endif
#endif
