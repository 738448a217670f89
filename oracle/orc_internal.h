/* orc_internal.h — shared helpers of the CPU oracle (TEST INFRASTRUCTURE ONLY, see rt_oracle.h).
 *
 * Numeric contract, identical to the CUDA kernels':
 *   IEEE-754 binary32, round-to-nearest; a multiply-add is fused ONLY where fmaf() is written
 *   (gcc -ffp-contract=off here, nvcc -fmad=false there); sqrtf and / are correctly rounded;
 *   sin/acos/atan2 are the polynomial routines below, never libm.
 */
#ifndef ORC_INTERNAL_H
#define ORC_INTERNAL_H
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include "rt_oracle.h"

typedef struct { float x, y, z; } v3;

static inline v3 V(float x, float y, float z) { v3 r = {x, y, z}; return r; }
static inline v3 vadd(v3 a, v3 b) { return V(a.x + b.x, a.y + b.y, a.z + b.z); }
static inline v3 vsub(v3 a, v3 b) { return V(a.x - b.x, a.y - b.y, a.z - b.z); }
static inline v3 vmul(v3 a, v3 b) { return V(a.x * b.x, a.y * b.y, a.z * b.z); }
static inline v3 vscale(float t, v3 a) { return V(t * a.x, t * a.y, t * a.z); }
static inline v3 vneg(v3 a) { return V(-a.x, -a.y, -a.z); }
/* device-path dot product: two fused steps (vec3.cuh:110-112 leaves contraction to the compiler; we pin it) */
static inline float dot3(v3 a, v3 b) { return fmaf(a.z, b.z, fmaf(a.y, b.y, a.x * b.x)); }
/* a + t*b, one fma per component (ray::at, ray.cuh:19) */
static inline v3 vfma(float t, v3 b, v3 a) { return V(fmaf(t, b.x, a.x), fmaf(t, b.y, a.y), fmaf(t, b.z, a.z)); }
/* unit_vector = v / v.length() = (1/len) * v   (vec3.cuh:108,120) */
static inline v3 vunit(v3 a) { float r = 1.0f / sqrtf(dot3(a, a)); return vscale(r, a); }
static inline v3 vcross(v3 u, v3 v) {
    return V(u.y * v.z - u.z * v.y, u.z * v.x - u.x * v.z, u.x * v.y - u.y * v.x);
}
/* host-side (scene set-up) dot: plain left-to-right, no fusion — camera ctor / matrices run on the host in the reference */
static inline float hdot(v3 a, v3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
static inline v3 hunit(v3 a) { float r = 1.0f / sqrtf(hdot(a, a)); return vscale(r, a); }

#define ORC_PI_F 3.14159274101257324f
#define ORC_2PI_F 6.28318548202514648f
#define ORC_INV_PI_F 0.318309873342514038f
#define ORC_INV_2PI_F 0.159154936671257019f
#define ORC_INV_4PI_F 0.0795774683356285095f

/* ---- Philox4x32-10 stream keyed (seed, pixel | sample, bounce) ---- */
typedef struct {
    uint32_t key[2];
    uint32_t ctr[4];
    uint32_t out[4];
    int idx;
    int sampling;          /* family, RT_SAMPLING_REFERENCE_GPU / _CPU : [0,1) for CPU semantics, (0,1] for GPU semantics */
    int direct;            /* RT_SAMPLING_DIRECT: direct draws instead of the reference's rejection loops */
    const float *replay;   /* known-answer tests: replayed uniforms instead of Philox */
    int replay_n, replay_pos;
    int draws;
} orc_rng;

void orc_rng_init(orc_rng *r, uint32_t seed, uint32_t pixel, uint32_t sample, uint32_t bounce, int sampling);
float orc_rng_next(orc_rng *r);
void orc_rng_block(orc_rng *r, uint32_t out[4]);
/* direct draws (RT_SAMPLING_DIRECT; csrc/rt_shade.cuh:ball_direct): uniform unit vector + radius of a uniform ball point */
void orc_sincos_turn24(uint32_t k, float *c, float *s);
void orc_ball_direct(orc_rng *rng, v3 *unit, float *rad);

/* ---- prepared scene ---- */
typedef struct {
    v3 origin, llc, horizontal, vertical, u, v, w;
    float lens_radius;
} orc_camera;

typedef struct {
    const rt_scene_desc *d;
    orc_camera cam;
} orc_scene;

typedef struct {
    int hit;
    float t;
    v3 p, n;          /* n: shading normal (face-flipped) */
    int front;
    float u, v;
    int prim, mat;
} orc_hit;

void orc_camera_setup(const rt_camera_desc *c, orc_camera *out);
void orc_get_ray(const orc_camera *cam, float s, float t, int defocus, orc_rng *rng, v3 *o, v3 *d);

int orc_hit_object(const rt_scene_desc *sc, int i, v3 o, v3 d, float inv_a, float tmin, float tmax, orc_hit *rec);
int orc_hit_world(const rt_scene_desc *sc, v3 o, v3 d, float tmin, float tmax, orc_hit *rec);

v3 orc_texture_value(const rt_scene_desc *sc, int tex, float u, float v, v3 p);
v3 orc_emitted(const rt_scene_desc *sc, int mat, float u, float v, v3 p);
int orc_scatter(const rt_scene_desc *sc, int mat, v3 rd, const orc_hit *rec, orc_rng *rng, v3 *att, v3 *so, v3 *sd);
v3 orc_background(const rt_scene_desc *sc, v3 d);

/* next-event estimation (orc_nee.c) */
int orc_is_nee_light(const rt_scene_desc *sc, int i);
float orc_light_area(const rt_scene_desc *sc, int i);
int orc_light_list(const rt_scene_desc *sc, int *lights);
int orc_nee_sample(const rt_scene_desc *sc, const int *lights, int n_lights, v3 x, v3 n, uint32_t seed, uint32_t pixel,
                   uint32_t sample, uint32_t bounce, int mis, v3 *f, int64_t *shadow_rays);
float orc_mis_hit_weight(const rt_scene_desc *sc, int obj, int n_lights, v3 d, float t, v3 ny);

v3 orc_xform_point(const float m[12], v3 p);
v3 orc_xform_vec(const float m[12], v3 p);
v3 orc_xform_normal(const float minv[12], v3 p);

/* host twin of the device traversal (orc_trav.c): the LBVH of lbvh_host.c walked with the device's box arithmetic and order */
typedef struct orc_bvh orc_bvh;
typedef struct { int64_t box, nodes, sph_rej, sph_acc, rect, cyl_rej, cyl_acc, tri, boxprim, overflow; int max_stack; } orc_trav_counters;
orc_bvh *orc_bvh_create(const rt_scene_desc *sc);
void orc_bvh_destroy(orc_bvh *b);
int orc_bvh_hit(const orc_bvh *b, v3 o, v3 d, float tmin, float tmax, int any_hit, orc_hit *rec, orc_trav_counters *cnt);

/* object AABB exactly as the GPU build computes it (lbvh_host.c) */
void orc_object_aabb(const rt_scene_desc *sc, int i, float bmin[3], float bmax[3]);

#endif
