/* orc_render.c — textures, materials, path loop, AOV pass, tonemap of the CPU oracle.
 * TEST INFRASTRUCTURE ONLY (see rt_oracle.h). */
#include "orc_internal.h"
#include <float.h>
#include <pthread.h>

/* ---------------- textures (texture.cuh:24-52; image: taichi-version/material.py:137-144 semantics, nearest) ---------------- */
v3 orc_texture_value(const rt_scene_desc *sc, int tex, float u, float v, v3 p) {
    for (int guard = 0; guard < 8; ++guard) {
        const rt_texture_desc *t = &sc->textures[tex];
        if (t->type == RT_TEX_SOLID) return V(t->color[0], t->color[1], t->color[2]);
        if (t->type == RT_TEX_CHECKER) {
            float sines = orc_sinf(10.0f * p.x) * orc_sinf(10.0f * p.y) * orc_sinf(10.0f * p.z);
            tex = sines < 0.0f ? t->odd : t->even;
            continue;
        }
        if (t->type == RT_TEX_IMAGE) {
            const rt_image_desc *im = &sc->images[t->image];
            /* clamp u,v to [0,1], flip v (row 0 = top), nearest texel */
            float uu = u < 0.0f ? 0.0f : (u > 1.0f ? 1.0f : u);
            float vv = 1.0f - (v < 0.0f ? 0.0f : (v > 1.0f ? 1.0f : v));
            int i = (int)(uu * (float)im->width), j = (int)(vv * (float)im->height);
            if (i >= im->width) i = im->width - 1;
            if (j >= im->height) j = im->height - 1;
            const uint8_t *px = im->rgb8 + 3 * ((size_t)j * im->width + i);
            return V((float)px[0] / 255.0f, (float)px[1] / 255.0f, (float)px[2] / 255.0f);
        }
        break;
    }
    return V(0, 0, 0);
}

v3 orc_emitted(const rt_scene_desc *sc, int mat, float u, float v, v3 p) {   /* material.cuh:23-25,175-178 */
    const rt_material_desc *m = &sc->materials[mat];
    if (m->type == RT_MAT_DIFFUSE_LIGHT) return orc_texture_value(sc, m->texture, u, v, p);
    return V(0, 0, 0);
}

/* random_in_unit_sphere: vec3.cuh:133-141 (positive octant, GPU semantics) | vec3.h:121-129 (full cube, CPU semantics).
 * The rejection loop is the reference's; how uniforms are cut out of the Philox stream is ours (the reference draws from
 * XORWOW / mt19937): the sampler consumes whole 128-bit blocks, two trials of three 21-bit uniforms per block — trial 1 =
 * the low 21 bits of words 0..2, trial 2 = their high 11 bits extended by bits 0-9 / 10-19 / 20-29 of word 3 — and the
 * next draw after it starts a new block.  Replayed uniforms (known-answer tests) are taken three per trial as they come. */
static int ball_trial(const orc_rng *rng, float a, float b, float c, v3 *p) {
    *p = rng->sampling == RT_SAMPLING_REFERENCE_GPU ? V(a, b, c)
                                                    : V(fmaf(2.0f, a, -1.0f), fmaf(2.0f, b, -1.0f), fmaf(2.0f, c, -1.0f));
    return !(dot3(*p, *p) >= 1.0f);
}
static float u21(const orc_rng *rng, uint32_t x) {
    if (rng->sampling == RT_SAMPLING_REFERENCE_GPU) x += 1u;     /* (0,1] | [0,1), as orc_rng_next */
    return (float)x * 4.76837158203125e-07f;                     /* 2^-21 */
}
static v3 random_in_unit_sphere(orc_rng *rng) {
    v3 p;
    if (rng->replay) {
        for (;;) {
            float a = orc_rng_next(rng), b = orc_rng_next(rng), c = orc_rng_next(rng);
            if (ball_trial(rng, a, b, c, &p)) return p;
        }
    }
    for (;;) {
        uint32_t w[4];
        orc_rng_block(rng, w);
        rng->draws += 3;
        if (ball_trial(rng, u21(rng, w[0] & 0x1fffffu), u21(rng, w[1] & 0x1fffffu), u21(rng, w[2] & 0x1fffffu), &p)) return p;
        rng->draws += 3;
        if (ball_trial(rng, u21(rng, (w[0] >> 21) | ((w[3] & 0x3ffu) << 11)), u21(rng, (w[1] >> 21) | (((w[3] >> 10) & 0x3ffu) << 11)),
                       u21(rng, (w[2] >> 21) | (((w[3] >> 20) & 0x3ffu) << 11)), &p)) return p;
    }
}

/* test hook: n points of the sampler, one stream per point (pixel = first_pixel + i) */
void orc_ball_samples(uint32_t seed, uint32_t first_pixel, uint32_t sample, uint32_t bounce, int sampling, int n, float *out) {
    for (int i = 0; i < n; ++i) {
        orc_rng r; orc_rng_init(&r, seed, first_pixel + (uint32_t)i, sample, bounce, sampling);
        v3 p;
        if (r.direct) {       /* the ball point of the direct sampler: radius * unit vector (componentwise |.| in the GPU family) */
            float rad; orc_ball_direct(&r, &p, &rad);
            p = r.sampling == RT_SAMPLING_REFERENCE_GPU ? vscale(rad, V(fabsf(p.x), fabsf(p.y), fabsf(p.z))) : vscale(rad, p);
        } else p = random_in_unit_sphere(&r);
        out[3 * i] = p.x; out[3 * i + 1] = p.y; out[3 * i + 2] = p.z;
    }
}

static v3 reflect(v3 v, v3 n) {   /* vec3.cuh:121-123 */
    float k = 2.0f * dot3(v, n);
    return V(fmaf(-k, n.x, v.x), fmaf(-k, n.y, v.y), fmaf(-k, n.z, v.z));
}

int orc_scatter(const rt_scene_desc *sc, int mat, v3 rd, const orc_hit *rec, orc_rng *rng, v3 *att, v3 *so, v3 *sd) {
    const rt_material_desc *m = &sc->materials[mat];
    *so = rec->p;
    switch (m->type) {
    case RT_MAT_LAMBERTIAN: {   /* material.cuh:37-52 (GPU) | material.h:25-35 (CPU) */
        v3 r;
        if (rng->direct) {        /* same distributions, drawn directly: octant ball (GPU family) | unit vector (CPU family) */
            float rad; orc_ball_direct(rng, &r, &rad);
            if (rng->sampling == RT_SAMPLING_REFERENCE_GPU) r = vscale(rad, V(fabsf(r.x), fabsf(r.y), fabsf(r.z)));
        } else {
            r = random_in_unit_sphere(rng);
            if (rng->sampling == RT_SAMPLING_REFERENCE_CPU) r = vunit(r);     /* random_unit_vector, vec3.h:131-134 */
        }
        v3 dir = vadd(rec->n, r);
        if (fabsf(dir.x) < 1e-8f && fabsf(dir.y) < 1e-8f && fabsf(dir.z) < 1e-8f) dir = rec->n;   /* near_zero, vec3.cuh:66-68 */
        *sd = dir;
        *att = orc_texture_value(sc, m->texture, rec->u, rec->v, rec->p);
        return 1;
    }
    case RT_MAT_METAL: {        /* material.cuh:63-70 */
        v3 refl = reflect(vunit(rd), rec->n);
        float fuzz = m->fuzz < 1.0f ? m->fuzz : 1.0f;
        v3 r;
        if (rng->direct) {
            float rad; orc_ball_direct(rng, &r, &rad);
            r = rng->sampling == RT_SAMPLING_REFERENCE_GPU ? vscale(rad, V(fabsf(r.x), fabsf(r.y), fabsf(r.z))) : vscale(rad, r);
        } else r = random_in_unit_sphere(rng);
        *sd = vfma(fuzz, r, refl);
        *att = V(m->albedo[0], m->albedo[1], m->albedo[2]);
        return dot3(*sd, rec->n) > 0.0f;
    }
    case RT_MAT_DIELECTRIC: {   /* material.cuh:130-147,152-158; refract: vec3.cuh:125-131 */
        *att = V(1.0f, 1.0f, 1.0f);
        float ratio = rec->front ? (1.0f / m->ior) : m->ior;
        v3 ud = vunit(rd);
        float cos_theta = fminf(-dot3(ud, rec->n), 1.0f);
        float sin_theta = sqrtf(fmaf(-cos_theta, cos_theta, 1.0f));
        int cannot_refract = ratio * sin_theta > 1.0f;
        int do_reflect = cannot_refract;
        if (!do_reflect) {
            float r0 = (1.0f - ratio) / (1.0f + ratio);
            r0 = r0 * r0;
            float x = 1.0f - cos_theta, x2 = x * x;
            float refl = fmaf(1.0f - r0, x2 * x2 * x, r0);    /* pow(1-cos,5) as exact products */
            do_reflect = refl > orc_rng_next(rng);
        }
        if (do_reflect) *sd = reflect(ud, rec->n);
        else {
            v3 perp = vscale(ratio, vfma(cos_theta, rec->n, ud));
            float k = -sqrtf(fabsf(1.0f - dot3(perp, perp)));
            *sd = vfma(k, rec->n, perp);
        }
        return 1;
    }
    default:                    /* diffuse_light, material.cuh:169-173 */
        *att = V(0, 0, 0); *sd = V(0, 0, 0);
        return 0;
    }
}

v3 orc_background(const rt_scene_desc *sc, v3 d) {
    if (sc->background_mode == RT_BG_SKY_GRADIENT) {   /* main.cpp:36-38 */
        v3 ud = vunit(d);
        float t = 0.5f * (ud.y + 1.0f), w = 1.0f - t;
        return V(fmaf(t, 0.5f, w), fmaf(t, 0.7f, w), fmaf(t, 1.0f, w));
    }
    return V(sc->background[0], sc->background[1], sc->background[2]);   /* main.cu:61-63 */
}

/* ray_color, main.cu:17-70 / main.cpp:13-43.  Emission is non-zero only on diffuse_light, which never
 * scatters, so the radiance of a path is throughput * (E | background | 0) at its last vertex. */
typedef struct { int on, mis, n_lights; const int *lights; const unsigned char *is_light; } nee_t;

static v3 ray_color(const rt_scene_desc *sc, v3 o, v3 d, uint32_t seed, uint32_t pixel, uint32_t sample, int64_t *nrays,
                    const nee_t *nee, int64_t *nshadow, const orc_bvh *bvh, orc_trav_counters *cnt) {
    v3 thr = V(1, 1, 1), L = V(0, 0, 0);
    int prev_nee = 0;       /* the previous vertex sampled the lights: do not count a sampled light again when hit */
    for (int bounce = 0; bounce < sc->max_depth; ++bounce) {
        orc_hit rec;
        ++*nrays;
        const int hit = bvh ? orc_bvh_hit(bvh, o, d, 0.001f, FLT_MAX, 0, &rec, cnt) : orc_hit_world(sc, o, d, 0.001f, FLT_MAX, &rec);
        if (!hit) { L = vadd(L, vmul(thr, orc_background(sc, d))); break; }
        v3 E = orc_emitted(sc, rec.mat, rec.u, rec.v, rec.p);
        if (prev_nee && nee->is_light[rec.prim])       /* a sampled light found by the scattered ray: dropped, or MIS-weighted */
            E = nee->mis ? vscale(orc_mis_hit_weight(sc, rec.prim, nee->n_lights, d, rec.t, rec.n), E) : V(0, 0, 0);
        orc_rng rng; orc_rng_init(&rng, seed, pixel, sample, (uint32_t)bounce + 1u, sc->sampling);
        v3 att, so, sd;
        int ok = orc_scatter(sc, rec.mat, d, &rec, &rng, &att, &so, &sd);
        L = vadd(L, vmul(thr, E));
        prev_nee = 0;
        if (nee->on && nee->n_lights > 0 && sc->materials[rec.mat].type == RT_MAT_LAMBERTIAN) {
            v3 f;
            if (orc_nee_sample(sc, nee->lights, nee->n_lights, rec.p, rec.n, seed, pixel, sample, (uint32_t)bounce, nee->mis, &f, nshadow))
                L = vadd(L, vmul(vmul(thr, att), f));
            prev_nee = 1;
        }
        if (!ok) break;
        thr = vmul(thr, att);
        o = so; d = sd;
    }
    return L;
}

typedef struct {
    const rt_scene_desc *sc; orc_camera cam; uint32_t seed; int s0, sn, total; float *out;
    int row_next; pthread_mutex_t mu; int64_t rays, shadow; nee_t nee;
    const orc_bvh *bvh; orc_trav_counters cnt;
    int stride;            /* > 1: only pixels whose index is a multiple of stride are rendered (parity checks at full image sizes) */
    const int32_t *only; int n_only;   /* when set: only these pixel indices (ascending) are rendered */
} job_t;

static void *worker(void *arg) {
    job_t *j = (job_t *)arg;
    const rt_scene_desc *sc = j->sc;
    const int W = sc->width, H = sc->height;
    int64_t rays = 0, shadow = 0;
    orc_trav_counters cnt; memset(&cnt, 0, sizeof cnt);
    for (;;) {
        pthread_mutex_lock(&j->mu);
        int y = j->row_next++;
        pthread_mutex_unlock(&j->mu);
        if (y >= H) break;
        for (int x = 0; x < W; ++x) {
            uint32_t pixel = (uint32_t)(y * W + x);
            if (j->stride > 1 && pixel % (uint32_t)j->stride != 0) continue;
            if (j->only) {
                int found = 0;
                for (int q = 0; q < j->n_only && !found; ++q) found = j->only[q] == (int32_t)pixel;
                if (!found) continue;
            }
            v3 sum = V(0, 0, 0);
            int first_chunk = 1;
            /* fixed-order accumulation (rt.h: rt_sample_chunk): the samples fall into 8 chunks by the TOTAL sample count; a chunk is
             * summed sample by sample (main.cu:95-103), the chunk sums are combined in chunk order */
            for (int k = 0; k < RT_SAMPLE_CHUNKS; ++k) {
            int cb = (int)((long long)k * j->total / RT_SAMPLE_CHUNKS), ce = (int)((long long)(k + 1) * j->total / RT_SAMPLE_CHUNKS);
            if (k == RT_SAMPLE_CHUNKS - 1 && ce < j->s0 + j->sn) ce = j->s0 + j->sn;
            const int pb = cb > j->s0 ? cb : j->s0, pe = ce < j->s0 + j->sn ? ce : j->s0 + j->sn;
            if (pb >= pe) continue;
            v3 csum = V(0, 0, 0);
            for (int s = pb; s < pe; ++s) {
                orc_rng rng; orc_rng_init(&rng, j->seed, pixel, (uint32_t)s, 0u, sc->sampling);
                float u = ((float)x + orc_rng_next(&rng)) / (float)(W - 1);
                float v = ((float)y + orc_rng_next(&rng)) / (float)(H - 1);
                v3 o, d;
                int defocus = (sc->sampling & 1) == RT_SAMPLING_REFERENCE_CPU && sc->camera.aperture > 0.0f;
                orc_get_ray(&j->cam, u, v, defocus, &rng, &o, &d);
                v3 c = ray_color(sc, o, d, j->seed, pixel, (uint32_t)s, &rays, &j->nee, &shadow, j->bvh, &cnt);
                csum = vadd(csum, c);                                 /* fixed order: ascending sample index */
            }
            sum = first_chunk ? csum : vadd(sum, csum);
            first_chunk = 0;
            }
            float *q = j->out + 3 * (size_t)pixel;
            q[0] = sum.x; q[1] = sum.y; q[2] = sum.z;
        }
    }
    pthread_mutex_lock(&j->mu);
    j->rays += rays; j->shadow += shadow;
    j->cnt.box += cnt.box; j->cnt.nodes += cnt.nodes; j->cnt.sph_rej += cnt.sph_rej; j->cnt.sph_acc += cnt.sph_acc; j->cnt.rect += cnt.rect;
    j->cnt.cyl_rej += cnt.cyl_rej; j->cnt.cyl_acc += cnt.cyl_acc; j->cnt.tri += cnt.tri; j->cnt.boxprim += cnt.boxprim;
    j->cnt.overflow += cnt.overflow; if (cnt.max_stack > j->cnt.max_stack) j->cnt.max_stack = cnt.max_stack;
    pthread_mutex_unlock(&j->mu);
    return NULL;
}

static const int32_t *g_only = NULL; static int g_n_only = 0;     /* orc_render_pixels (same caveat as g_stride) */
static int g_stride = 1;      /* set by orc_render_bvh_strided around its render_impl call (test infrastructure: not re-entrant) */
static int render_impl(const rt_scene_desc *sc, uint32_t seed, int spp_begin, int spp_count, float *rgb_sum, int n_threads,
                       int flags, int64_t *n_rays, int64_t *n_shadow_rays, int use_bvh, int64_t *counters, int total_spp) {
    if (!sc || !rgb_sum || sc->width < 1 || sc->height < 1) return RT_ERR_INVALID;
    if ((flags & (RT_RENDER_NEE | RT_RENDER_NEE_MIS)) && (sc->sampling & 1) != RT_SAMPLING_REFERENCE_CPU) return RT_ERR_INVALID;
    job_t j; memset(&j, 0, sizeof j);
    j.sc = sc; j.seed = seed; j.s0 = spp_begin; j.sn = spp_count > 0 ? spp_count : sc->samples_per_pixel; j.out = rgb_sum;
    j.total = total_spp > 0 ? total_spp : (sc->samples_per_pixel > j.s0 + j.sn ? sc->samples_per_pixel : j.s0 + j.sn);
    orc_camera_setup(&sc->camera, &j.cam);
    int *lights = (int *)malloc(sizeof(int) * (size_t)(sc->n_objects > 0 ? sc->n_objects : 1));
    unsigned char *is_light = (unsigned char *)calloc((size_t)(sc->n_objects > 0 ? sc->n_objects : 1), 1);
    j.nee.on = (flags & (RT_RENDER_NEE | RT_RENDER_NEE_MIS)) != 0; j.nee.mis = (flags & RT_RENDER_NEE_MIS) != 0;
    if (j.nee.on) {
        j.nee.n_lights = orc_light_list(sc, lights);
        for (int k = 0; k < j.nee.n_lights; ++k) is_light[lights[k]] = 1;
    }
    j.nee.lights = lights; j.nee.is_light = is_light;
    orc_bvh *bvh = use_bvh ? orc_bvh_create(sc) : NULL;
    j.bvh = bvh; j.stride = g_stride; j.only = g_only; j.n_only = g_n_only;
    pthread_mutex_init(&j.mu, NULL);
    if (n_threads < 1) n_threads = 1;
    if (n_threads > 256) n_threads = 256;
    pthread_t th[256];
    for (int i = 0; i < n_threads; ++i) pthread_create(&th[i], NULL, worker, &j);
    for (int i = 0; i < n_threads; ++i) pthread_join(th[i], NULL);
    pthread_mutex_destroy(&j.mu);
    free(lights); free(is_light);
    orc_bvh_destroy(bvh);
    if (n_rays) *n_rays = j.rays;
    if (n_shadow_rays) *n_shadow_rays = j.shadow;
    if (counters) {
        counters[0] = j.cnt.box; counters[1] = j.cnt.nodes; counters[2] = j.cnt.sph_rej; counters[3] = j.cnt.sph_acc; counters[4] = j.cnt.rect;
        counters[5] = j.cnt.cyl_rej; counters[6] = j.cnt.cyl_acc; counters[7] = j.cnt.tri; counters[8] = j.cnt.boxprim; counters[9] = j.cnt.max_stack;
    }
    return 0;
}

int orc_render_ex(const rt_scene_desc *sc, uint32_t seed, int spp_begin, int spp_count, float *rgb_sum, int n_threads,
                  int flags, int64_t *n_rays, int64_t *n_shadow_rays) {
    return render_impl(sc, seed, spp_begin, spp_count, rgb_sum, n_threads, flags, n_rays, n_shadow_rays, 0, NULL, 0);
}

/* the samples [spp_begin, spp_begin + spp_count) of an image of total_spp samples: one rank's share of a multi-GPU render */
int orc_render_range(const rt_scene_desc *sc, uint32_t seed, int spp_begin, int spp_count, int total_spp, float *rgb_sum, int n_threads,
                     int64_t *n_rays) {
    return render_impl(sc, seed, spp_begin, spp_count, rgb_sum, n_threads, 0, n_rays, NULL, 0, NULL, total_spp);
}

/* as orc_render_bvh for the pixels whose index is a multiple of `stride` only (the others stay untouched in rgb_sum) */
int orc_render_bvh_strided(const rt_scene_desc *sc, uint32_t seed, float *rgb_sum, int n_threads, int stride, int64_t *n_rays) {
    g_stride = stride > 1 ? stride : 1;
    const int rc = render_impl(sc, seed, 0, 0, rgb_sum, n_threads, 0, n_rays, NULL, 1, NULL, 0);
    g_stride = 1;
    return rc;
}

/* only the listed pixels (the others stay untouched in rgb_sum); use_bvh = 0: the reference's linear walk (brute force) */
int orc_render_pixels(const rt_scene_desc *sc, uint32_t seed, float *rgb_sum, int n_threads, const int32_t *pixels, int n_pixels, int use_bvh) {
    g_only = pixels; g_n_only = n_pixels;
    const int rc = render_impl(sc, seed, 0, 0, rgb_sum, n_threads, 0, NULL, NULL, use_bvh, NULL, 0);
    g_only = NULL; g_n_only = 0;
    return rc;
}

int orc_render_bvh(const rt_scene_desc *sc, uint32_t seed, int spp_begin, int spp_count, float *rgb_sum, int n_threads,
                   int flags, int64_t *n_rays, int64_t *counters) {
    return render_impl(sc, seed, spp_begin, spp_count, rgb_sum, n_threads, flags, n_rays, NULL, 1, counters, 0);
}

int orc_render(const rt_scene_desc *sc, uint32_t seed, int spp_begin, int spp_count, float *rgb_sum, int n_threads,
               int64_t *n_rays) {
    return orc_render_ex(sc, seed, spp_begin, spp_count, rgb_sum, n_threads, 0, n_rays, NULL);
}

/* first-hit record per pixel: the "deterministic path" (pixel-centre ray, no lens offset, no scattering) */
int orc_aov(const rt_scene_desc *sc, float *aov) {
    orc_camera cam; orc_camera_setup(&sc->camera, &cam);
    const int W = sc->width, H = sc->height;
    static const float fixed_u[8] = {0.5f, 0.5f, 0.5f, 0.25f, 0.25f, 0.25f, 0.1f, 0.1f};
    for (int y = 0; y < H; ++y) for (int x = 0; x < W; ++x) {
        float *q = aov + 24 * ((size_t)y * W + x);
        memset(q, 0, 24 * sizeof(float));
        float u = ((float)x + 0.5f) / (float)(W - 1), v = ((float)y + 0.5f) / (float)(H - 1);
        v3 o, d; orc_get_ray(&cam, u, v, 0, NULL, &o, &d);
        orc_hit rec;
        if (orc_hit_world(sc, o, d, 0.001f, FLT_MAX, &rec)) {
            q[0] = 1.0f; q[1] = rec.t; q[2] = rec.p.x; q[3] = rec.p.y; q[4] = rec.p.z;
            q[5] = rec.n.x; q[6] = rec.n.y; q[7] = rec.n.z; q[8] = (float)rec.front; q[9] = rec.u; q[10] = rec.v;
            q[11] = (float)rec.prim; q[12] = (float)rec.mat;
            orc_rng rng; memset(&rng, 0, sizeof rng); rng.replay = fixed_u; rng.replay_n = 8; rng.sampling = sc->sampling & 1; rng.direct = (sc->sampling & RT_SAMPLING_DIRECT) ? 1 : 0;
            v3 att, so, sd;
            int ok = orc_scatter(sc, rec.mat, d, &rec, &rng, &att, &so, &sd);
            q[13] = att.x; q[14] = att.y; q[15] = att.z;
            v3 E = orc_emitted(sc, rec.mat, rec.u, rec.v, rec.p);
            q[16] = E.x; q[17] = E.y; q[18] = E.z; q[22] = (float)ok;
        } else {
            v3 b = orc_background(sc, d);
            q[19] = b.x; q[20] = b.y; q[21] = b.z;
        }
    }
    return 0;
}

/* write_image (PNG, no gamma) color.cuh:15-35 ; write_color (PPM, sqrt gamma) color.cuh:70-95 */
void orc_tonemap(const float *rgb_sum, int W, int H, int spp, int gamma, unsigned char *rgb8) {
    float scale = 1.0f / (float)spp;   /* PPM path multiplies by 1/spp (color.cuh:82-85); PNG path divides (color.cuh:24) */
    for (int j = H - 1, row = 0; j >= 0; --j, ++row)
        for (int i = 0; i < W; ++i)
            for (int c = 0; c < 3; ++c) {
                float x = rgb_sum[3 * ((size_t)j * W + i) + c];
                x = gamma ? sqrtf(x * scale) : x / (float)spp;
                x = x < 0.0f ? 0.0f : (x > 0.999f ? 0.999f : x);   /* clamp(x,0,0.999); NaN falls through as in the reference */
                rgb8[3 * ((size_t)row * W + i) + c] = (unsigned char)(int)(256.0f * x);
            }
}

/* debug aid (tests / tools only): follows the paths of one pixel with the linear walk and reports the first ray on which the host
 * twin of the traversal disagrees with it: out[0..2] = origin, [3..5] = direction, [6] = t linear, [7] = prim linear, [8] = t twin,
 * [9] = prim twin, [10] = sample, [11] = bounce.  Returns 1 when a disagreement was found. */
int orc_debug_first_mismatch(const rt_scene_desc *sc, uint32_t seed, int pixel, int n_samples, float *out) {
    orc_camera cam; orc_camera_setup(&sc->camera, &cam);
    orc_bvh *bvh = orc_bvh_create(sc);
    const int W = sc->width, H = sc->height, x = pixel % W, y = pixel / W;
    int found = 0;
    for (int s = 0; s < n_samples && !found; ++s) {
        orc_rng rng; orc_rng_init(&rng, seed, (uint32_t)pixel, (uint32_t)s, 0u, sc->sampling);
        float u = ((float)x + orc_rng_next(&rng)) / (float)(W - 1), v = ((float)y + orc_rng_next(&rng)) / (float)(H - 1);
        v3 o, d;
        int defocus = (sc->sampling & 1) == RT_SAMPLING_REFERENCE_CPU && sc->camera.aperture > 0.0f;
        orc_get_ray(&cam, u, v, defocus, &rng, &o, &d);
        for (int bounce = 0; bounce < sc->max_depth; ++bounce) {
            orc_hit a, b;
            int ha = orc_hit_world(sc, o, d, 0.001f, FLT_MAX, &a), hb = orc_bvh_hit(bvh, o, d, 0.001f, FLT_MAX, 0, &b, NULL);
            if (ha != hb || (ha && (a.prim != b.prim || a.t != b.t))) {
                out[0] = o.x; out[1] = o.y; out[2] = o.z; out[3] = d.x; out[4] = d.y; out[5] = d.z;
                out[6] = ha ? a.t : -1.0f; out[7] = ha ? (float)a.prim : -1.0f; out[8] = hb ? b.t : -1.0f; out[9] = hb ? (float)b.prim : -1.0f;
                out[10] = (float)s; out[11] = (float)bounce;
                found = 1; break;
            }
            if (!ha) break;
            orc_rng r2; orc_rng_init(&r2, seed, (uint32_t)pixel, (uint32_t)s, (uint32_t)bounce + 1u, sc->sampling);
            v3 att, so, sd;
            if (!orc_scatter(sc, a.mat, d, &a, &r2, &att, &so, &sd)) break;
            o = so; d = sd;
        }
    }
    orc_bvh_destroy(bvh);
    (void)H;
    return found;
}
