/* orc_nee.c — next-event estimation (the "shadow" stage of the wavefront renderer), CPU statement.
 * TEST INFRASTRUCTURE ONLY (see rt_oracle.h).
 *
 * The reference has no light sampling (ray_color only follows scattered rays, gpu-version/main.cu:17-70); its unused
 * hit_shadow (朴素光线追踪/ray_tracing_models.py:124-143) and the README's "加入更多的发光源" show the intent.  This is OUR
 * definition (parity unpinned by reference), valid for the cosine-weighted lambertian of the CPU sampling convention
 * (material.h:25-35: dir = n + random_unit_vector, i.e. BRDF albedo/pi with pdf cos/pi), where it has the same
 * expectation as the plain estimator:
 *   at every lambertian hit x: pick one light uniformly, one point y uniformly on its surface, and if the segment
 *   x->y is unoccluded add  thr * albedo * Le(y) * cos_x * cos_y / |y-x|^2 * (area * n_lights / pi);
 *   the scattered ray that leaves x then does NOT collect the emission of a sampled light it happens to hit.
 * Lights = objects with a diffuse_light material of type sphere / rect / cylinder / box / triangle, in scene-file order;
 * sphere, rect and box lights may carry a rotate / translate instance (sampled in object space, taken to world space).  Emitters emit on both faces (material.cuh:175-178).
 * Draws: a separate Philox stream per vertex (bounce word | 0x10000): light index, then two surface coordinates.
 */
#include "orc_internal.h"
#include <float.h>

int orc_is_nee_light(const rt_scene_desc *sc, int i) {
    const rt_object_desc *o = &sc->objects[i];
    if (sc->materials[o->material].type != RT_MAT_DIFFUSE_LIGHT) return 0;
    if (!(o->type == RT_OBJ_SPHERE || o->type == RT_OBJ_XY_RECT || o->type == RT_OBJ_XZ_RECT || o->type == RT_OBJ_YZ_RECT ||
          o->type == RT_OBJ_CYLINDER || o->type == RT_OBJ_TRIANGLE || o->type == RT_OBJ_BOX)) return 0;
    return orc_light_area(sc, i) > 0.0f;       /* a degenerate emitter has no surface to sample */
}

float orc_light_area(const rt_scene_desc *sc, int i) {
    const rt_object_desc *o = &sc->objects[i];
    const float *p = o->p;
    switch (o->type) {
    case RT_OBJ_SPHERE: return 4.0f * ORC_PI_F * (p[3] * p[3]);
    case RT_OBJ_XY_RECT: case RT_OBJ_XZ_RECT: case RT_OBJ_YZ_RECT: return fabsf((p[1] - p[0]) * (p[3] - p[2]));
    case RT_OBJ_CYLINDER: return ORC_2PI_F * fabsf(p[0]) * fabsf(p[2] - p[1]);
    case RT_OBJ_TRIANGLE: {
        v3 e1 = V(p[3] - p[0], p[4] - p[1], p[5] - p[2]), e2 = V(p[6] - p[0], p[7] - p[1], p[8] - p[2]);
        v3 c = V(fmaf(e1.y, e2.z, -(e1.z * e2.y)), fmaf(e1.z, e2.x, -(e1.x * e2.z)), fmaf(e1.x, e2.y, -(e1.y * e2.x)));
        return 0.5f * sqrtf(dot3(c, c));
    }
    case RT_OBJ_BOX: {
        float ex = fabsf(p[3] - p[0]), ey = fabsf(p[4] - p[1]), ez = fabsf(p[5] - p[2]);
        return 2.0f * ((ey * ez + ez * ex) + ex * ey);
    }
    }
    return 0.0f;
}

int orc_light_list(const rt_scene_desc *sc, int *lights) {
    int n = 0;
    for (int i = 0; i < sc->n_objects; ++i)
        if (orc_is_nee_light(sc, i)) { if (lights) lights[n] = i; ++n; }
    return n;
}

static float orc_cosf(float x) { return orc_sinf(x + 1.57079637050628662f); }

/* point y, unit normal ny and texture coordinates on light object i from two uniforms */
static void light_point(const rt_scene_desc *sc, int i, float u1, float u2, v3 *y, v3 *ny, float *tu, float *tv) {
    const rt_object_desc *o = &sc->objects[i];
    const float *p = o->p;
    *tu = 0.0f; *tv = 0.0f;
    switch (o->type) {
    case RT_OBJ_SPHERE: {
        float z = fmaf(-2.0f, u1, 1.0f);
        float r = sqrtf(fmaxf(0.0f, fmaf(-z, z, 1.0f)));
        float phi = ORC_2PI_F * u2;
        v3 dir = V(r * orc_cosf(phi), r * orc_sinf(phi), z);
        *ny = dir;
        *y = vfma(fabsf(p[3]), dir, V(p[0], p[1], p[2]));
        float theta = orc_acosf(-dir.y);                       /* get_sphere_uv, object.cuh:87-93 */
        float ph = orc_atan2f(-dir.z, dir.x) + ORC_PI_F;
        *tu = ph * ORC_INV_2PI_F; *tv = theta * ORC_INV_PI_F;
        if (o->xform >= 0) {       /* rotate / translate instance: the point and the normal go to world space (rigid: the area is unchanged) */
            const rt_xform_desc *xf = &sc->xforms[o->xform];
            *y = orc_xform_point(xf->m, *y); *ny = vunit(orc_xform_normal(xf->m_inv, *ny));
        }
        return;
    }
    case RT_OBJ_XY_RECT: case RT_OBJ_XZ_RECT: case RT_OBJ_YZ_RECT: {
        float a = fmaf(u1, p[1] - p[0], p[0]), b = fmaf(u2, p[3] - p[2], p[2]), k = p[4];
        if (o->type == RT_OBJ_XY_RECT) { *y = V(a, b, k); *ny = V(0, 0, 1); }
        else if (o->type == RT_OBJ_XZ_RECT) { *y = V(a, k, b); *ny = V(0, 1, 0); }
        else { *y = V(k, a, b); *ny = V(1, 0, 0); }
        *tu = (a - p[0]) / (p[1] - p[0]); *tv = (b - p[2]) / (p[3] - p[2]);    /* object.cuh:113-114 */
        if (o->xform >= 0) {
            const rt_xform_desc *xf = &sc->xforms[o->xform];
            *y = orc_xform_point(xf->m, *y); *ny = vunit(orc_xform_normal(xf->m_inv, *ny));
        }
        return;
    }
    case RT_OBJ_CYLINDER: {
        const rt_xform_desc *xf = &sc->xforms[o->xform];
        float phi = ORC_2PI_F * u1, z = fmaf(u2, p[2] - p[1], p[1]);
        float c = orc_cosf(phi), s = orc_sinf(phi);
        v3 op = V(p[0] * c, p[0] * s, z);
        *y = orc_xform_point(xf->m, op);
        *ny = vunit(orc_xform_normal(xf->m_inv, V(c, s, 0.0f)));
        *tu = (orc_atan2f(op.y, op.x) + ORC_2PI_F) * ORC_INV_4PI_F; *tv = (z - p[1]) / (p[2] - p[1]);   /* object.cuh:286-287 */
        return;
    }
    case RT_OBJ_BOX: {      /* u1 picks the face in proportion to its area and is re-used inside it; uv as orc_hit.c gives them for a box */
        const float *L = p, *H = p + 3;
        float e[3] = {fabsf(p[3] - p[0]), fabsf(p[4] - p[1]), fabsf(p[5] - p[2])};
        float wf[3] = {e[1] * e[2], e[2] * e[0], e[0] * e[1]};
        float v = u1 * (2.0f * ((wf[0] + wf[1]) + wf[2]));
        int fc = 0; float c = 0.0f;
        while (fc < 5 && !(v < c + wf[fc >> 1])) { c += wf[fc >> 1]; ++fc; }
        float s1 = fminf(fmaxf((v - c) / wf[fc >> 1], 0.0f), 0.99999994f);
        int ax = fc >> 1, ua = (ax + 1) % 3, va = (ax + 2) % 3;
        float P[3], nn[3] = {0.0f, 0.0f, 0.0f};
        P[ax] = (fc & 1) ? H[ax] : L[ax]; nn[ax] = (fc & 1) ? 1.0f : -1.0f;
        P[ua] = fmaf(s1, H[ua] - L[ua], L[ua]); P[va] = fmaf(u2, H[va] - L[va], L[va]);
        *tu = (P[ua] - L[ua]) / (H[ua] - L[ua]); *tv = (P[va] - L[va]) / (H[va] - L[va]);
        *y = V(P[0], P[1], P[2]); *ny = V(nn[0], nn[1], nn[2]);
        if (o->xform >= 0) {
            const rt_xform_desc *xf = &sc->xforms[o->xform];
            *y = orc_xform_point(xf->m, *y); *ny = vunit(orc_xform_normal(xf->m_inv, *ny));
        }
        return;
    }
    default: {   /* RT_OBJ_TRIANGLE */
        v3 v0 = V(p[0], p[1], p[2]);
        v3 e1 = V(p[3] - p[0], p[4] - p[1], p[5] - p[2]), e2 = V(p[6] - p[0], p[7] - p[1], p[8] - p[2]);
        float su = sqrtf(u1);
        float b1 = u2 * su, b2 = 1.0f - su;
        *y = vfma(b2, e2, vfma(b1, e1, v0));
        v3 c = V(fmaf(e1.y, e2.z, -(e1.z * e2.y)), fmaf(e1.z, e2.x, -(e1.x * e2.z)), fmaf(e1.x, e2.y, -(e1.y * e2.x)));
        *ny = vunit(c);
        *tu = b1; *tv = b2;
        return;
    }
    }
}

/* occlusion: does any object accept a root in [tmin, tmax] (each object's own hit() range test, brute force) */
static int occluded(const rt_scene_desc *sc, v3 o, v3 d, float tmin, float tmax) {
    orc_hit tmp;
    float inv_a = 1.0f / dot3(d, d);
    for (int i = 0; i < sc->n_objects; ++i)
        if (orc_hit_object(sc, i, o, d, inv_a, tmin, tmax, &tmp)) return 1;
    return 0;
}

/* Returns 1 and the radiance factor f (to be multiplied by thr * albedo) when the sampled light is visible from x.
 * *shadow_rays is incremented when an occlusion ray was traced. */
/* mis: the light sample is weighted with the balance heuristic against the cosine-weighted scatter direction that could have found
 * the same point, w = p_l / (p_l + p_b) in solid-angle measure: p_l = |w|^2 / (cos_y * area * n_lights), p_b = cos_x / pi; the
 * emission a scattered ray finds on a sampled light then carries the complementary weight (orc_mis_hit_weight). */
int orc_nee_sample(const rt_scene_desc *sc, const int *lights, int n_lights, v3 x, v3 n, uint32_t seed, uint32_t pixel,
                   uint32_t sample, uint32_t bounce, int mis, v3 *f, int64_t *shadow_rays) {
    orc_rng rng; orc_rng_init(&rng, seed, pixel, sample, (bounce + 1u) | 0x10000u, sc->sampling);
    float u0 = orc_rng_next(&rng), u1 = orc_rng_next(&rng), u2 = orc_rng_next(&rng);
    int k = (int)(u0 * (float)n_lights);
    if (k > n_lights - 1) k = n_lights - 1;
    const int obj = lights[k];
    v3 y, ny; float tu, tv;
    light_point(sc, obj, u1, u2, &y, &ny, &tu, &tv);
    v3 w = vsub(y, x);
    float d2 = dot3(w, w);
    if (!(d2 > 0.0f)) return 0;
    float inv_d = 1.0f / sqrtf(d2);
    float cs = dot3(n, w) * inv_d, cl = fabsf(dot3(ny, w)) * inv_d;
    if (!(cs > 0.0f) || !(cl > 0.0f)) return 0;
    ++*shadow_rays;
    if (occluded(sc, x, w, 0.001f, 0.999f)) return 0;
    float g = ((cs * cl) / d2) * (orc_light_area(sc, obj) * ((float)n_lights * ORC_INV_PI_F));
    if (mis) {
        float pl = d2 / (cl * (orc_light_area(sc, obj) * (float)n_lights)), pb = cs * ORC_INV_PI_F;
        g = g * (pl / (pl + pb));
    }
    v3 Le = orc_emitted(sc, sc->objects[obj].material, tu, tv, y);
    *f = vscale(g, Le);
    return 1;
}

/* Balance-heuristic weight of the emission found by a ray scattered at a lambertian vertex (direction d = n + unit vector, NOT
 * normalised: cos_x = |d| / 2, so p_b = |d| / (2 pi)) that hits sampled light `obj` at parameter t with unit normal ny. */
float orc_mis_hit_weight(const rt_scene_desc *sc, int obj, int n_lights, v3 d, float t, v3 ny) {
    float dd = dot3(d, d), len = sqrtf(dd);
    float dist2 = (t * t) * dd;
    float cy = fabsf(dot3(ny, d)) / len;
    float pl = dist2 / (cy * (orc_light_area(sc, obj) * (float)n_lights)), pb = (len * 0.5f) * ORC_INV_PI_F;
    return pb / (pb + pl);
}
