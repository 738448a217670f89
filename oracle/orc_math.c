/* orc_math.c — transcendental routines, Philox RNG, transforms and camera of the CPU oracle.
 * TEST INFRASTRUCTURE ONLY (see rt_oracle.h). */
#include "orc_internal.h"

/* ---------------- shared-algorithm transcendentals ----------------
 * The reference calls sin / acos / atan2 (texture.cuh:46, object.cuh:88-89,283).  CUDA's and glibc's
 * versions differ in the last ulp, which would break sample-for-sample comparison, so both sides
 * evaluate the SAME polynomials with the SAME fma sequence.  Accuracy (~1e-7 abs) is tested against libm. */
float orc_sinf(float x) {
    /* Cody-Waite reduction to [-pi/4, pi/4] by multiples of pi/2 (3-term, fma) */
    float k = rintf(x * 0.636619746685028076f);
    float r = fmaf(-k, 1.57079637050628662f, x);
    r = fmaf(-k, -4.37113882867379289e-08f, r);
    r = fmaf(-k, -1.77635683940025046e-15f, r);
    int q = (int)k & 3;
    float s = r * r;
    float ps = fmaf(s, 2.72425904768169858e-06f, -1.98400259250774980e-04f);
    ps = fmaf(s, ps, 8.33333190530538559e-03f);
    ps = fmaf(s, ps, -1.66666671633720398e-01f);
    float sn = fmaf(r * s, ps, r);
    float pc = fmaf(s, -2.72958857294724998e-07f, 2.48005617322633043e-05f);
    pc = fmaf(s, pc, -1.38888880610466003e-03f);
    pc = fmaf(s, pc, 4.16666679084300995e-02f);
    float cs = fmaf(s * s, pc, fmaf(-0.5f, s, 1.0f));
    float res = (q & 1) ? cs : sn;
    return (q & 2) ? -res : res;
}

static inline float asin_core(float z) {   /* |z| <= 0.5 */
    float s = z * z;
    float p = fmaf(s, 3.38299162685871124e-02f, 1.70643702149391174e-02f);
    p = fmaf(s, p, 3.11187468469142914e-02f);
    p = fmaf(s, p, 4.45978343486785889e-02f);
    p = fmaf(s, p, 7.50009939074516296e-02f);
    p = fmaf(s, p, 1.66666656732559204e-01f);
    return fmaf(z * s, p, z);
}

float orc_acosf(float x) {
    float ax = fabsf(x);
    if (ax <= 0.5f) return 1.57079637050628662f - asin_core(x);
    if (ax > 1.0f) ax = 1.0f;                       /* clamp: normals are unit up to rounding */
    float z = sqrtf(fmaf(-0.5f, ax, 0.5f));         /* sqrt((1-|x|)/2) */
    float a = 2.0f * asin_core(z);
    return x > 0.0f ? a : ORC_PI_F - a;
}

float orc_atan2f(float y, float x) {
    float ax = fabsf(x), ay = fabsf(y);
    float mx = ax > ay ? ax : ay, mn = ax > ay ? ay : ax;
    float a = mx > 0.0f ? mn / mx : 0.0f;
    float s = a * a;
    float p = fmaf(s, -2.44705588556826115e-03f, 1.37503892183303833e-02f);
    p = fmaf(s, p, -3.62703576683998108e-02f);
    p = fmaf(s, p, 6.28437772393226624e-02f);
    p = fmaf(s, p, -8.67317989468574524e-02f);
    p = fmaf(s, p, 1.10379971563816071e-01f);
    p = fmaf(s, p, -1.42791107296943665e-01f);
    p = fmaf(s, p, 1.99997663497924805e-01f);
    p = fmaf(s, p, -3.33333313465118408e-01f);
    float r = fmaf(a * s, p, a);
    if (ay > ax) r = 1.57079637050628662f - r;
    if (x < 0.0f) r = ORC_PI_F - r;
    return y < 0.0f ? -r : r;
}

/* ---------------- Philox4x32-10 (Salmon et al., SC'11) ---------------- */
void orc_philox4x32_10(const uint32_t ctr[4], const uint32_t key[2], uint32_t out[4]) {
    uint32_t c0 = ctr[0], c1 = ctr[1], c2 = ctr[2], c3 = ctr[3], k0 = key[0], k1 = key[1];
    for (int i = 0; i < 10; ++i) {
        uint64_t p0 = (uint64_t)0xD2511F53u * c0, p1 = (uint64_t)0xCD9E8D57u * c2;
        uint32_t n0 = (uint32_t)(p1 >> 32) ^ c1 ^ k0, n1 = (uint32_t)p1;
        uint32_t n2 = (uint32_t)(p0 >> 32) ^ c3 ^ k1, n3 = (uint32_t)p0;
        c0 = n0; c1 = n1; c2 = n2; c3 = n3;
        k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
    }
    out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}

/* key = (seed, pixel); counter = (sample, bounce, block, 'RT'): bounce 0 is the camera sample, k>=1 the k-th shade */
void orc_rng_init(orc_rng *r, uint32_t seed, uint32_t pixel, uint32_t sample, uint32_t bounce, int sampling) {
    memset(r, 0, sizeof *r);
    r->key[0] = seed; r->key[1] = pixel;
    r->ctr[0] = sample; r->ctr[1] = bounce; r->ctr[2] = 0; r->ctr[3] = 0x5254u;
    r->idx = 4; r->sampling = sampling & 1; r->direct = (sampling & RT_SAMPLING_DIRECT) ? 1 : 0;
}

float orc_rng_next(orc_rng *r) {
    r->draws++;
    if (r->replay) { float v = r->replay[r->replay_pos % r->replay_n]; r->replay_pos++; return v; }
    if (r->idx == 4) { orc_philox4x32_10(r->ctr, r->key, r->out); r->ctr[2]++; r->idx = 0; }
    uint32_t x = r->out[r->idx++] >> 8;
    /* cmake-cpu-version/rtweekend.h:18-21: [0,1)   |   curand_uniform, gpu-version/rtweekend.cuh:23-25: (0,1] */
    if (r->sampling == RT_SAMPLING_REFERENCE_GPU) x += 1u;
    return (float)x * 5.9604644775390625e-08f;
}

/* the next whole block of the stream (unread words of the current block are dropped; the next draw starts a new block) */
void orc_rng_block(orc_rng *r, uint32_t out[4]) {
    orc_philox4x32_10(r->ctr, r->key, out);
    r->ctr[2]++; r->idx = 4;
}

void orc_uniforms(uint32_t seed, uint32_t pixel, uint32_t sample, uint32_t bounce, int sampling, int count, float *out) {
    orc_rng r; orc_rng_init(&r, seed, pixel, sample, bounce, sampling);
    for (int i = 0; i < count; ++i) out[i] = orc_rng_next(&r);
}

/* ---------------- transforms (vec3.cuh:331-370, 382-416; parser.hpp:423-440) ---------------- */
v3 orc_xform_point(const float m[12], v3 p) {
    return V(fmaf(m[2], p.z, fmaf(m[1], p.y, m[0] * p.x)) + m[3],
             fmaf(m[6], p.z, fmaf(m[5], p.y, m[4] * p.x)) + m[7],
             fmaf(m[10], p.z, fmaf(m[9], p.y, m[8] * p.x)) + m[11]);
}
v3 orc_xform_vec(const float m[12], v3 p) {
    return V(fmaf(m[2], p.z, fmaf(m[1], p.y, m[0] * p.x)),
             fmaf(m[6], p.z, fmaf(m[5], p.y, m[4] * p.x)),
             fmaf(m[10], p.z, fmaf(m[9], p.y, m[8] * p.x)));
}
/* apply_normal: transpose of the inverse (vec3.cuh:364-370) */
v3 orc_xform_normal(const float mi[12], v3 p) {
    return V(fmaf(mi[8], p.z, fmaf(mi[4], p.y, mi[0] * p.x)),
             fmaf(mi[9], p.z, fmaf(mi[5], p.y, mi[1] * p.x)),
             fmaf(mi[10], p.z, fmaf(mi[6], p.y, mi[2] * p.x)));
}

static void mat4_mul(const float a[16], const float b[16], float r[16]) {   /* vec3.cuh:219-226 */
    for (int i = 0; i < 4; ++i)
        for (int j = 0; j < 4; ++j)
            r[i * 4 + j] = a[i * 4 + 0] * b[0 * 4 + j] + a[i * 4 + 1] * b[1 * 4 + j] + a[i * 4 + 2] * b[2 * 4 + j] +
                           a[i * 4 + 3] * b[3 * 4 + j];
}

void orc_make_xform(const float axis[3], double angle_deg, int has_rotate, const float tr[3], int has_translate,
                    rt_xform_desc *out) {
    /* o2w starts as identity() = translate(0,0,0) (object.cuh:296, vec3.cuh:426) */
    float m[16] = {1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1}, mi[16];
    memcpy(mi, m, sizeof m);
    if (has_rotate) {   /* o2w = rotate(axis, rad) * o2w   (object.cuh:225-227) */
        float theta = (float)(angle_deg / 180.0 * M_PI);           /* parser.hpp:427-429 */
        float len = sqrtf(axis[0] * axis[0] + axis[1] * axis[1] + axis[2] * axis[2]);   /* vec3::normalize, vec3.cuh:70-73 */
        float ax = axis[0] / len, ay = axis[1] / len, az = axis[2] / len;
        float s = sinf(theta), c = cosf(theta);
        float r[16] = {ax * ax + (1 - ax * ax) * c, ax * ay * (1 - c) - az * s, ax * az * (1 - c) + ay * s, 0,
                       ax * ay * (1 - c) + az * s, ay * ay + (1 - ay * ay) * c, ay * az * (1 - c) - ax * s, 0,
                       ax * az * (1 - c) - ay * s, ay * az * (1 - c) + ax * s, az * az + (1 - az * az) * c, 0,
                       0, 0, 0, 1};
        float rt[16];
        for (int i = 0; i < 4; ++i) for (int j = 0; j < 4; ++j) rt[i * 4 + j] = r[j * 4 + i];
        float nm[16], nmi[16];
        mat4_mul(r, m, nm); mat4_mul(mi, rt, nmi);                           /* (A*B).m_inv = B.m_inv * A.m_inv, vec3.cuh:327-329 */
        memcpy(m, nm, sizeof m); memcpy(mi, nmi, sizeof mi);
    }
    if (has_translate) {   /* o2w = translate(offset) * o2w   (object.cuh:229-231) */
        float t[16] = {1, 0, 0, tr[0], 0, 1, 0, tr[1], 0, 0, 1, tr[2], 0, 0, 0, 1};
        float ti[16] = {1, 0, 0, -tr[0], 0, 1, 0, -tr[1], 0, 0, 1, -tr[2], 0, 0, 0, 1};
        float nm[16], nmi[16];
        mat4_mul(t, m, nm); mat4_mul(mi, ti, nmi);
        memcpy(m, nm, sizeof m); memcpy(mi, nmi, sizeof mi);
    }
    memcpy(out->m, m, 12 * sizeof(float));
    memcpy(out->m_inv, mi, 12 * sizeof(float));
}

/* ---------------- camera (camera.cuh:9-39; camera.h:7-39) ---------------- */
void orc_camera_setup(const rt_camera_desc *c, orc_camera *cam) {
    float theta = (float)((double)c->vfov_deg * M_PI / (double)180.0f);     /* rtweekend.cuh:17-19 */
    float h = tanf(theta / 2.0f);
    float viewport_height = 2.0f * h;
    float viewport_width = c->aspect * viewport_height;
    v3 from = V(c->lookfrom[0], c->lookfrom[1], c->lookfrom[2]);
    v3 at = V(c->lookat[0], c->lookat[1], c->lookat[2]);
    v3 vup = V(c->vup[0], c->vup[1], c->vup[2]);
    cam->w = hunit(vsub(from, at));
    cam->u = hunit(vcross(vup, cam->w));
    cam->v = vcross(cam->w, cam->u);
    cam->origin = from;
    cam->horizontal = vscale(c->focus_dist * viewport_width, cam->u);
    cam->vertical = vscale(c->focus_dist * viewport_height, cam->v);
    /* origin - horizontal/2 - vertical/2 - focus_dist*w ; operator/ is (1/t)*v (vec3.cuh:108) */
    v3 hh = vscale(1.0f / 2.0f, cam->horizontal), vh = vscale(1.0f / 2.0f, cam->vertical);
    cam->llc = vsub(vsub(vsub(from, hh), vh), vscale(c->focus_dist, cam->w));
    cam->lens_radius = c->aperture / 2.0f;
}

/* get_ray: camera.h:32-39 with defocus, camera.cuh:31-39 without */
/* ---- direct draws: the distributions of the reference's rejection samplers without the loops (our definition of how the
 * Philox words are cut; the reference draws from XORWOW / mt19937 and is matched statistically only) ---- */
void orc_sincos_turn24(uint32_t k, float *c, float *s) {   /* angle 2 pi k / 2^24 - pi / 4 */
    const float th = (float)((int)(k & 0x3fffffu) - 0x200000) * 3.74507398e-07f;
    const float s2 = th * th;
    float ps = fmaf(s2, 2.72425904768169858e-06f, -1.98400259250774980e-04f);
    ps = fmaf(s2, ps, 8.33333190530538559e-03f);
    ps = fmaf(s2, ps, -1.66666671633720398e-01f);
    const float sn = fmaf(th * s2, ps, th);
    float pc = fmaf(s2, -2.72958857294724998e-07f, 2.48005617322633043e-05f);
    pc = fmaf(s2, pc, -1.38888880610466003e-03f);
    pc = fmaf(s2, pc, 4.16666679084300995e-02f);
    const float cs = fmaf(s2 * s2, pc, fmaf(-0.5f, s2, 1.0f));
    const uint32_t q = k >> 22;
    *c = (q & 1u) ? sn : cs; *s = (q & 1u) ? cs : sn;
    if (q == 1u || q == 2u) *c = -*c;
    if (q >= 2u) *s = -*s;
}
static uint32_t k24_of(float u) { uint32_t k = (uint32_t)(u * 16777216.0f); return k > 0xffffffu ? 0xffffffu : k; }
static void ball_direct_k(uint32_t k1, uint32_t k2, uint32_t k3, uint32_t k4, uint32_t k5, v3 *unit, float *rad) {
    const float z = fmaf(-2.0f, (float)k1 * 5.9604644775390625e-08f, 1.0f);
    const float r = sqrtf(fmaxf(fmaf(-z, z, 1.0f), 0.0f));
    float c, s;
    orc_sincos_turn24(k2, &c, &s);
    *unit = V(r * c, r * s, z);
    uint32_t m = k3 > k4 ? k3 : k4; if (k5 > m) m = k5;
    *rad = (float)(m + 1u) * 5.9604644775390625e-08f;
}
void orc_ball_direct(orc_rng *rng, v3 *unit, float *rad) {
    if (rng->replay) {          /* replayed uniforms: five draws as they come */
        float a = orc_rng_next(rng), b = orc_rng_next(rng), c = orc_rng_next(rng), d = orc_rng_next(rng), e = orc_rng_next(rng);
        ball_direct_k(k24_of(a), k24_of(b), k24_of(c), k24_of(d), k24_of(e), unit, rad);
        return;
    }
    uint32_t w[4];
    orc_rng_block(rng, w);      /* one whole block; the next draw starts a new one */
    rng->draws += 5;
    ball_direct_k(w[0] >> 8, w[1] >> 8, w[2] >> 8, w[3] >> 8, (w[0] & 0xffu) | ((w[1] & 0xffu) << 8) | ((w[2] & 0xffu) << 16), unit, rad);
}

void orc_get_ray(const orc_camera *cam, float s, float t, int defocus, orc_rng *rng, v3 *o, v3 *d) {
    float rx = 0.0f, ry = 0.0f;
    if (defocus && rng->direct) {   /* uniform disk: radius sqrt(u1), angle from u2 */
        float u1 = orc_rng_next(rng), u2 = orc_rng_next(rng), c, sn;
        orc_sincos_turn24(k24_of(u2), &c, &sn);
        float rr = sqrtf(u1);
        rx = cam->lens_radius * (rr * c); ry = cam->lens_radius * (rr * sn);
    } else if (defocus) {
        for (;;) {   /* random_in_unit_disk, vec3.h:157-165 */
            float px = fmaf(2.0f, orc_rng_next(rng), -1.0f), py = fmaf(2.0f, orc_rng_next(rng), -1.0f);
            if (fmaf(py, py, px * px) >= 1.0f) continue;
            rx = cam->lens_radius * px; ry = cam->lens_radius * py;
            break;
        }
    }
    v3 off = V(fmaf(cam->v.x, ry, cam->u.x * rx), fmaf(cam->v.y, ry, cam->u.y * rx), fmaf(cam->v.z, ry, cam->u.z * rx));
    *o = vadd(cam->origin, off);
    v3 a = vfma(t, cam->vertical, vfma(s, cam->horizontal, cam->llc));
    *d = vsub(vsub(a, cam->origin), off);
}
