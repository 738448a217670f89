/* orc_hit.c — primitive intersection and the linear closest-hit walk of the CPU oracle.
 * TEST INFRASTRUCTURE ONLY (see rt_oracle.h). */
#include "orc_internal.h"
#include <float.h>

static void set_face_normal(orc_hit *rec, v3 d, v3 outward) {   /* hittable.cuh:20-23 */
    rec->front = dot3(d, outward) < 0.0f;
    rec->n = rec->front ? outward : vneg(outward);
}

/* sphere::hit, object.cuh:47-75.  inv_a = 1/dot(d,d) is hoisted per ray (one rounding differs from the
 * reference's division; covered by the 1e-5 tolerance against oracle/_ref). */
static int hit_sphere(const float *p, const rt_xform_desc *xf, v3 o, v3 d, float inv_a, float tmin, float tmax, orc_hit *rec) {
    v3 c = V(p[0], p[1], p[2]);
    float r = p[3];
    const v3 wd = d;
    if (xf) {      /* rotate / translate instance (vec3.cuh:303-427, applied as object.cuh:225-231 does for the cylinder) */
        o = orc_xform_point(xf->m_inv, o); d = orc_xform_vec(xf->m_inv, d);
        inv_a = 1.0f / dot3(d, d);
    }
    v3 oc = vsub(o, c);
    float a = dot3(d, d);
    float hb = dot3(oc, d);
    float cc = fmaf(-r, r, dot3(oc, oc));
    float disc = fmaf(hb, hb, -(a * cc));
    if (disc < 0.0f) return 0;
    float sq = sqrtf(disc);
    float root = (-hb - sq) * inv_a;
    if (root < tmin || tmax < root) {
        root = (-hb + sq) * inv_a;
        if (root < tmin || tmax < root) return 0;
    }
    rec->t = root;
    rec->p = vfma(root, d, o);
    v3 n = vscale(1.0f / r, vsub(rec->p, c));
    if (xf) {      /* the uv below follow the object-space normal: the texture turns with the sphere */
        rec->p = orc_xform_point(xf->m, rec->p);
        set_face_normal(rec, wd, orc_xform_normal(xf->m_inv, n));
    } else set_face_normal(rec, d, n);
    /* get_sphere_uv(outward normal), object.cuh:87-93 */
    float theta = orc_acosf(-n.y);
    float phi = orc_atan2f(-n.z, n.x) + ORC_PI_F;
    rec->u = phi * ORC_INV_2PI_F;
    rec->v = theta * ORC_INV_PI_F;
    return 1;
}

/* xy_rect / xz_rect / yz_rect ::hit, object.cuh:105-122,142-159,175-192 */
static int hit_rect(int type, const float *p, const rt_xform_desc *xf, v3 o, v3 d, float tmin, float tmax, orc_hit *rec) {
    const v3 wd = d;
    if (xf) { o = orc_xform_point(xf->m_inv, o); d = orc_xform_vec(xf->m_inv, d); }
    float oa, ob, ok, da, db, dk;
    if (type == RT_OBJ_XY_RECT) { oa = o.x; ob = o.y; ok = o.z; da = d.x; db = d.y; dk = d.z; }
    else if (type == RT_OBJ_XZ_RECT) { oa = o.x; ob = o.z; ok = o.y; da = d.x; db = d.z; dk = d.y; }
    else { oa = o.y; ob = o.z; ok = o.x; da = d.y; db = d.z; dk = d.x; }
    float a0 = p[0], a1 = p[1], b0 = p[2], b1 = p[3], k = p[4];
    float t = (k - ok) / dk;
    if (t < tmin || t > tmax) return 0;
    float a = fmaf(t, da, oa), b = fmaf(t, db, ob);
    if (a < a0 || a > a1 || b < b0 || b > b1) return 0;
    rec->u = (a - a0) / (a1 - a0);
    rec->v = (b - b0) / (b1 - b0);
    rec->t = t;
    v3 n = type == RT_OBJ_XY_RECT ? V(0, 0, 1) : type == RT_OBJ_XZ_RECT ? V(0, 1, 0) : V(1, 0, 0);
    if (xf) {
        rec->p = orc_xform_point(xf->m, vfma(t, d, o));
        set_face_normal(rec, wd, orc_xform_normal(xf->m_inv, n));
        return 1;
    }
    rec->front = dk < 0.0f;                 /* dot(d, axis) == d.axis */
    rec->n = rec->front ? n : vneg(n);
    rec->p = vfma(t, d, o);
    return 1;
}

/* cylinder::hit + quadratic, object.cuh:199-214,233-290 */
static int hit_cylinder(const float *p, const rt_xform_desc *xf, v3 o, v3 d, float tmin, float tmax, orc_hit *rec) {
    float radius = p[0], zmin = p[1], zmax = p[2];
    v3 oo = orc_xform_point(xf->m_inv, o), od = orc_xform_vec(xf->m_inv, d);
    float a = fmaf(od.y, od.y, od.x * od.x);
    float b = 2.0f * fmaf(od.y, oo.y, od.x * oo.x);
    float c = fmaf(-radius, radius, fmaf(oo.y, oo.y, oo.x * oo.x));
    float delta = fmaf(b, b, -(4.0f * a * c));
    if (delta < 0.0f) return 0;
    float sd = sqrtf(delta);
    float t0 = -0.5f * (b - sd) / a, t1 = -0.5f * (b + sd) / a;
    if (t0 > t1) { float tmp = t0; t0 = t1; t1 = tmp; }
    if (t0 > tmax || t1 < tmin) return 0;
    float t = t0;
    if (t0 < tmin) { t = t1; if (t > tmax) return 0; }
    v3 op = vfma(t, od, oo);
    if (op.z < zmin || op.z > zmax) {
        if (t == t1) return 0;
        t = t1;
        if (t > tmax || t < tmin) return 0;
        op = vfma(t, od, oo);
        if (op.z < zmin || op.z > zmax) return 0;
    }
    float len = sqrtf(fmaf(op.y, op.y, op.x * op.x));
    v3 on = V(op.x / len, op.y / len, 0.0f / len);
    rec->p = orc_xform_point(xf->m, op);
    set_face_normal(rec, d, orc_xform_normal(xf->m_inv, on));
    rec->t = t;
    float phi = orc_atan2f(op.y, op.x) + ORC_2PI_F;
    rec->u = phi * ORC_INV_4PI_F;
    rec->v = (op.z - zmin) / (zmax - zmin);
    return 1;
}

/* BOX (extension, RTNW semantics: closed axis-aligned box, optional object-to-world transform).
 * Not in the reference's C++: parity unpinned by reference. */
static int hit_box(const float *p, const rt_xform_desc *xf, v3 o, v3 d, float tmin, float tmax, orc_hit *rec) {
    v3 oo = o, od = d;
    if (xf) { oo = orc_xform_point(xf->m_inv, o); od = orc_xform_vec(xf->m_inv, d); }
    float lo[3] = {p[0], p[1], p[2]}, hi[3] = {p[3], p[4], p[5]};
    float O[3] = {oo.x, oo.y, oo.z}, D[3] = {od.x, od.y, od.z};
    float tn = -INFINITY, tf = INFINITY; int an = 0, af = 0;
    for (int k = 0; k < 3; ++k) {
        float t0 = (lo[k] - O[k]) / D[k], t1 = (hi[k] - O[k]) / D[k];
        if (D[k] == 0.0f) { if (O[k] < lo[k] || O[k] > hi[k]) return 0; continue; }
        if (t0 > t1) { float tmp = t0; t0 = t1; t1 = tmp; }
        if (t0 > tn) { tn = t0; an = k; }
        if (t1 < tf) { tf = t1; af = k; }
    }
    if (tn > tf) return 0;
    float t = tn; int ax = an;
    if (t < tmin || tmax < t) { t = tf; ax = af; if (t < tmin || tmax < t) return 0; }
    v3 op = vfma(t, od, oo);
    float P[3] = {op.x, op.y, op.z};
    float nn[3] = {0, 0, 0};
    nn[ax] = (P[ax] - lo[ax]) < (hi[ax] - P[ax]) ? -1.0f : 1.0f;
    int ua = (ax + 1) % 3, va = (ax + 2) % 3;
    rec->u = (P[ua] - lo[ua]) / (hi[ua] - lo[ua]);
    rec->v = (P[va] - lo[va]) / (hi[va] - lo[va]);
    rec->t = t;
    v3 on = V(nn[0], nn[1], nn[2]);
    if (xf) { rec->p = orc_xform_point(xf->m, op); on = orc_xform_normal(xf->m_inv, on); }
    else rec->p = op;
    set_face_normal(rec, d, on);
    return 1;
}

/* TRIANGLE (extension): Moeller-Trumbore, two-sided, uv = barycentrics.  Parity unpinned by reference. */
static int hit_triangle(const float *p, v3 o, v3 d, float tmin, float tmax, orc_hit *rec) {
    v3 v0 = V(p[0], p[1], p[2]);
    v3 e1 = vsub(V(p[3], p[4], p[5]), v0), e2 = vsub(V(p[6], p[7], p[8]), v0);
    v3 pv = V(fmaf(d.y, e2.z, -(d.z * e2.y)), fmaf(d.z, e2.x, -(d.x * e2.z)), fmaf(d.x, e2.y, -(d.y * e2.x)));
    float det = dot3(e1, pv);
    if (det == 0.0f) return 0;
    float inv = 1.0f / det;
    v3 tv = vsub(o, v0);
    float u = dot3(tv, pv) * inv;
    if (u < 0.0f || u > 1.0f) return 0;
    v3 qv = V(fmaf(tv.y, e1.z, -(tv.z * e1.y)), fmaf(tv.z, e1.x, -(tv.x * e1.z)), fmaf(tv.x, e1.y, -(tv.y * e1.x)));
    float v = dot3(d, qv) * inv;
    if (v < 0.0f || u + v > 1.0f) return 0;
    float t = dot3(e2, qv) * inv;
    if (t < tmin || tmax < t) return 0;
    /* A grazing ray (|det| tiny) carries a barycentric error of ~1e-7 |o - v0| |d x e2| / |det|, so the float test above can accept a
     * point well outside the triangle (found on the 1 M-triangle terrain: 6e-5 outside a 0.03-wide triangle, once per ~1e9 rays).
     * A hit counts only if its point lies in the triangle's own box widened by 2^-11 of its largest extent: that keeps the
     * acceptance region inside the leaf box of the LBVH (2^-10 of the extent + 2^-20 of the scene), whatever the rounding. */
    {
        const float lx = v0.x + fminf(0.0f, fminf(e1.x, e2.x)), hx = v0.x + fmaxf(0.0f, fmaxf(e1.x, e2.x));
        const float ly = v0.y + fminf(0.0f, fminf(e1.y, e2.y)), hy = v0.y + fmaxf(0.0f, fmaxf(e1.y, e2.y));
        const float lz = v0.z + fminf(0.0f, fminf(e1.z, e2.z)), hz = v0.z + fmaxf(0.0f, fmaxf(e1.z, e2.z));
        const float s = 4.8828125e-04f * fmaxf(hx - lx, fmaxf(hy - ly, hz - lz));
        const v3 q = vfma(t, d, o);
        if (!(q.x >= lx - s && q.x <= hx + s && q.y >= ly - s && q.y <= hy + s && q.z >= lz - s && q.z <= hz + s)) return 0;
    }
    v3 n = V(fmaf(e1.y, e2.z, -(e1.z * e2.y)), fmaf(e1.z, e2.x, -(e1.x * e2.z)), fmaf(e1.x, e2.y, -(e1.y * e2.x)));
    /* per-vertex texture coordinates (OBJ `vt`): 16-bit unorm pairs in p[9..11]; none: (0,0), (1,0), (0,1) = the barycentrics */
    uint32_t q[3];
    memcpy(q, p + 9, sizeof q);
    if ((q[0] | q[1] | q[2]) == 0u) { q[1] = 0x0000ffffu; q[2] = 0xffff0000u; }
    const float w0 = (1.0f - u) - v;
    rec->u = fmaf(v, (float)(q[2] & 0xffffu) / 65535.0f, fmaf(u, (float)(q[1] & 0xffffu) / 65535.0f, w0 * ((float)(q[0] & 0xffffu) / 65535.0f)));
    rec->v = fmaf(v, (float)(q[2] >> 16) / 65535.0f, fmaf(u, (float)(q[1] >> 16) / 65535.0f, w0 * ((float)(q[0] >> 16) / 65535.0f)));
    rec->t = t;
    rec->p = vfma(t, d, o);
    set_face_normal(rec, d, vunit(n));
    return 1;
}

int orc_hit_object(const rt_scene_desc *sc, int i, v3 o, v3 d, float inv_a, float tmin, float tmax, orc_hit *rec) {
    const rt_object_desc *ob = &sc->objects[i];
    const rt_xform_desc *xf = ob->xform >= 0 ? &sc->xforms[ob->xform] : NULL;
    static const rt_xform_desc ident = {{1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0}, {1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0}};
    int h = 0;
    switch (ob->type) {
    case RT_OBJ_SPHERE: h = hit_sphere(ob->p, xf, o, d, inv_a, tmin, tmax, rec); break;
    case RT_OBJ_XY_RECT: case RT_OBJ_XZ_RECT: case RT_OBJ_YZ_RECT: h = hit_rect(ob->type, ob->p, xf, o, d, tmin, tmax, rec); break;
    case RT_OBJ_CYLINDER: h = hit_cylinder(ob->p, xf ? xf : &ident, o, d, tmin, tmax, rec); break;
    case RT_OBJ_BOX: h = hit_box(ob->p, xf, o, d, tmin, tmax, rec); break;
    case RT_OBJ_TRIANGLE: h = hit_triangle(ob->p, o, d, tmin, tmax, rec); break;
    default: break;
    }
    if (h) { rec->hit = 1; rec->prim = i; rec->mat = ob->material; }
    return h;
}

/* hittable_list::hit, object.cuh:23-37: in object order, shrinking t_max; a later object with EQUAL t wins */
int orc_hit_world(const rt_scene_desc *sc, v3 o, v3 d, float tmin, float tmax, orc_hit *rec) {
    orc_hit tmp;
    int any = 0;
    float closest = tmax;
    float inv_a = 1.0f / dot3(d, d);
    rec->hit = 0;
    for (int i = 0; i < sc->n_objects; ++i) {
        if (orc_hit_object(sc, i, o, d, inv_a, tmin, closest, &tmp)) { any = 1; closest = tmp.t; *rec = tmp; }
    }
    return any;
}
