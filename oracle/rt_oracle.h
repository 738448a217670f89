/* rt_oracle.h — CPU restatement of the reference render path.  TEST INFRASTRUCTURE ONLY.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may load
 * liboracle.so; the product library (librt_b200.so) never links, loads or calls anything here.
 *
 * What it restates (reference paths relative to the reference repo root):
 *   render loop + ray_color        gpu-version/main.cu:17-105, cmake-cpu-version/main.cpp:13-55
 *   hittable_list (linear closest)  gpu-version/object.cuh:23-37
 *   sphere / rects / cylinder       gpu-version/object.cuh:47-75, 105-192, 199-290
 *   materials                       gpu-version/material.cuh:37-178, cmake-cpu-version/material.h:25-92
 *   textures                        gpu-version/texture.cuh:24-52
 *   camera                          gpu-version/camera.cuh:9-39, cmake-cpu-version/camera.h:7-39
 *   transforms                      gpu-version/vec3.cuh:303-427, parser.hpp:423-440
 *   tonemap                         gpu-version/color.cuh:15-35,70-95
 * The RNG is NOT the reference's (cuRAND XORWOW / mt19937): the north star replaces it by Philox4x32-10
 * keyed on (pixel, sample, bounce); the oracle implements that keying so the CUDA wavefront renderer can be
 * compared with it sample for sample.  The scene is intersected by BRUTE FORCE in object order (the
 * reference's semantics), never through a BVH, so the comparison also checks LBVH build + traversal.
 *
 * Parity status: pinned against the real reference compiled in oracle/_ref (tests/golden, made by
 * tests/golden/make_golden.py): per-function known-answer vectors, first-hit AOV images on every shipped
 * scene, and converged renders.  box / triangle / image-texture code has no C++ counterpart in the
 * reference: "parity unpinned by reference" for those three.
 */
#ifndef RT_ORACLE_H
#define RT_ORACLE_H
#include "rt.h"
#ifdef __cplusplus
extern "C" {
#endif

/* W*H*3 float sums (row 0 = bottom) of samples [spp_begin, spp_begin+spp_count); returns 0 on success.
 * n_rays (may be NULL) receives the number of path segments traced. */
int orc_render(const rt_scene_desc *scene, uint32_t seed, int spp_begin, int spp_count, float *rgb_sum,
               int n_threads, int64_t *n_rays);
/* as orc_render, with rt_render flags (RT_RENDER_NEE: next-event estimation, reference_cpu sampling only) */
int orc_render_ex(const rt_scene_desc *scene, uint32_t seed, int spp_begin, int spp_count, float *rgb_sum,
                  int n_threads, int flags, int64_t *n_rays, int64_t *n_shadow_rays);
/* samples [spp_begin, spp_begin + spp_count) of an image of total_spp samples (chunk boundaries of the fixed-order accumulation
 * from total_spp, rt.h: rt_sample_chunk) */
int orc_render_range(const rt_scene_desc *scene, uint32_t seed, int spp_begin, int spp_count, int total_spp, float *rgb_sum,
                     int n_threads, int64_t *n_rays);
/* first-hit record, 24 floats per pixel (same layout as oracle/ref_build/ref_gpu_harness.cpp `aov`) */
int orc_aov(const rt_scene_desc *scene, float *aov);
/* color.cuh:15-35 (gamma=0) / color.cuh:70-95 (gamma=1): rgb8 rows flipped so row 0 = top */
void orc_tonemap(const float *rgb_sum, int width, int height, int spp, int gamma, unsigned char *rgb8);

/* object-to-world = translate(t) * rotate(axis, angle_deg) as parser.hpp:423-440 builds it for cylinders */
void orc_make_xform(const float axis[3], double angle_deg, int has_rotate, const float tr[3], int has_translate,
                    rt_xform_desc *out);

/* Philox4x32-10 block function (key 2 words, counter 4 words) */
void orc_philox4x32_10(const uint32_t ctr[4], const uint32_t key[2], uint32_t out[4]);
/* the n-th..(n+count-1)-th uniforms of stream (seed, pixel, sample, bounce) in the given sampling mode */
void orc_uniforms(uint32_t seed, uint32_t pixel, uint32_t sample, uint32_t bounce, int sampling, int count, float *out);
/* n points of random_in_unit_sphere, stream (seed, first_pixel + i, sample, bounce) each: out[3n] */
void orc_ball_samples(uint32_t seed, uint32_t first_pixel, uint32_t sample, uint32_t bounce, int sampling, int n, float *out);

/* known-answer entry points; record formats identical to ref_gpu_harness.cpp kat_* modes */
void orc_kat_sphere(const float *in, float *out, int n);
void orc_kat_rect(const float *in, float *out, int n);
void orc_kat_cyl(const float *in, float *out, int n);
void orc_kat_scatter(const float *in, float *out, int n, int sampling);
void orc_kat_camera(const float *in, float *out, int n);
void orc_kat_xform(const float *in, float *out, int n);
/* shared-algorithm transcendental routines (exposed for accuracy tests against libm) */
float orc_sinf(float x);
float orc_acosf(float x);
float orc_atan2f(float y, float x);

/* Host twin of the GPU LBVH build (lbvh_host.c).  Array sizes as rt_bvh_download in rt.h;
 * parent has 2n-1 entries (n-1 internal nodes then n leaves). */
int orc_lbvh_build(const rt_scene_desc *scene, uint64_t *morton_sorted, int32_t *prim_sorted, int32_t *left,
                   int32_t *right, int32_t *parent, float *node_aabb, float *leaf_aabb, float *scene_aabb);
/* as orc_lbvh_build; info[4] = { leaves in the tree, giants kept out of it (they occupy the last leaf slots), tree depth, 0 } */
int orc_lbvh_build_ex(const rt_scene_desc *scene, uint64_t *morton_sorted, int32_t *prim_sorted, int32_t *left,
                      int32_t *right, int32_t *parent, float *node_aabb, float *leaf_aabb, float *scene_aabb, int32_t *info);
/* orc_render_ex with the closest hit found by the host twin of the device traversal (orc_trav.c) instead of the linear walk;
 * counters[10] (may be NULL) = box tests, inner-node visits, sphere reject / accept, rect, cylinder reject / accept, triangle,
 * box-primitive tests, deepest stack — the serial statement of the device's event counters (rt_stats). */
int orc_render_bvh(const rt_scene_desc *scene, uint32_t seed, int spp_begin, int spp_count, float *rgb_sum, int n_threads,
                   int flags, int64_t *n_rays, int64_t *counters);
/* orc_render_bvh restricted to the pixels whose index is a multiple of `stride` (others untouched): full-size parity checks */
int orc_render_bvh_strided(const rt_scene_desc *scene, uint32_t seed, float *rgb_sum, int n_threads, int stride, int64_t *n_rays);
/* only the listed pixel indices; use_bvh = 0: brute-force linear walk, 1: host twin of the traversal */
int orc_render_pixels(const rt_scene_desc *scene, uint32_t seed, float *rgb_sum, int n_threads, const int32_t *pixels, int n_pixels, int use_bvh);
/* stable LSD radix sort of (key, value) pairs — the host statement of the device sort, exposed for tests */
void orc_radix_sort_pairs(uint64_t *keys, int32_t *vals, int n);

#ifdef __cplusplus
}
#endif
#endif
