/* orc_kat.c — known-answer entry points of the CPU oracle; record formats are those of
 * oracle/ref_build/ref_gpu_harness.cpp (kat_* modes).  TEST INFRASTRUCTURE ONLY (see rt_oracle.h). */
#include "orc_internal.h"

static void put_rec(float *q, int hit, const orc_hit *rec) {
    memset(q, 0, 11 * sizeof(float));
    q[0] = (float)hit;
    if (!hit) return;
    q[1] = rec->t; q[2] = rec->p.x; q[3] = rec->p.y; q[4] = rec->p.z;
    q[5] = rec->n.x; q[6] = rec->n.y; q[7] = rec->n.z; q[8] = (float)rec->front; q[9] = rec->u; q[10] = rec->v;
}

static void one_object(rt_scene_desc *sc, rt_object_desc *ob, rt_xform_desc *xf) {
    memset(sc, 0, sizeof *sc);
    sc->n_objects = 1; sc->objects = ob; sc->xforms = xf; sc->n_xforms = xf ? 1 : 0;
}

void orc_kat_sphere(const float *in, float *out, int n) {
    for (int i = 0; i < n; ++i) {
        const float *p = in + 12 * i;
        rt_object_desc ob; memset(&ob, 0, sizeof ob); ob.type = RT_OBJ_SPHERE; ob.xform = -1;
        ob.p[0] = p[6]; ob.p[1] = p[7]; ob.p[2] = p[8]; ob.p[3] = p[9];
        rt_scene_desc sc; one_object(&sc, &ob, NULL);
        v3 o = V(p[0], p[1], p[2]), d = V(p[3], p[4], p[5]);
        orc_hit rec;
        int h = orc_hit_object(&sc, 0, o, d, 1.0f / dot3(d, d), p[10], p[11], &rec);
        put_rec(out + 11 * i, h, &rec);
    }
}

void orc_kat_rect(const float *in, float *out, int n) {
    for (int i = 0; i < n; ++i) {
        const float *p = in + 14 * i;
        rt_object_desc ob; memset(&ob, 0, sizeof ob); ob.xform = -1;
        ob.type = p[0] == 0.f ? RT_OBJ_XY_RECT : p[0] == 1.f ? RT_OBJ_XZ_RECT : RT_OBJ_YZ_RECT;
        for (int k = 0; k < 5; ++k) ob.p[k] = p[7 + k];
        rt_scene_desc sc; one_object(&sc, &ob, NULL);
        v3 o = V(p[1], p[2], p[3]), d = V(p[4], p[5], p[6]);
        orc_hit rec;
        int h = orc_hit_object(&sc, 0, o, d, 1.0f / dot3(d, d), p[12], p[13], &rec);
        put_rec(out + 11 * i, h, &rec);
    }
}

void orc_kat_cyl(const float *in, float *out, int n) {
    for (int i = 0; i < n; ++i) {
        const float *p = in + 18 * i;
        rt_object_desc ob; memset(&ob, 0, sizeof ob); ob.type = RT_OBJ_CYLINDER; ob.xform = 0;
        ob.p[0] = p[6]; ob.p[1] = p[7]; ob.p[2] = p[8];
        rt_xform_desc xf; orc_make_xform(p + 9, p[12], 1, p + 13, 1, &xf);
        rt_scene_desc sc; one_object(&sc, &ob, &xf);
        v3 o = V(p[0], p[1], p[2]), d = V(p[3], p[4], p[5]);
        orc_hit rec;
        int h = orc_hit_object(&sc, 0, o, d, 1.0f / dot3(d, d), p[16], p[17], &rec);
        put_rec(out + 11 * i, h, &rec);
    }
}

void orc_kat_scatter(const float *in, float *out, int n, int sampling) {
    for (int i = 0; i < n; ++i) {
        const float *p = in + 32 * i;
        rt_texture_desc tx; memset(&tx, 0, sizeof tx); tx.type = RT_TEX_SOLID;
        tx.color[0] = p[11]; tx.color[1] = p[12]; tx.color[2] = p[13];
        rt_material_desc m; memset(&m, 0, sizeof m);
        int t = (int)p[0];
        m.type = t == 0 ? RT_MAT_LAMBERTIAN : t == 1 ? RT_MAT_METAL : t == 2 ? RT_MAT_DIELECTRIC : RT_MAT_DIFFUSE_LIGHT;
        m.texture = 0; m.albedo[0] = p[11]; m.albedo[1] = p[12]; m.albedo[2] = p[13]; m.fuzz = p[14]; m.ior = p[15];
        rt_scene_desc sc; memset(&sc, 0, sizeof sc);
        sc.materials = &m; sc.n_materials = 1; sc.textures = &tx; sc.n_textures = 1; sc.sampling = sampling;
        orc_hit rec; memset(&rec, 0, sizeof rec);
        rec.p = V(p[4], p[5], p[6]); rec.n = V(p[7], p[8], p[9]); rec.front = p[10] != 0.f; rec.t = 1.f;
        orc_rng rng; memset(&rng, 0, sizeof rng); rng.replay = p + 16; rng.replay_n = 16; rng.sampling = sampling;
        v3 att = V(0, 0, 0), so = V(0, 0, 0), sd = V(0, 0, 0);
        int ok = orc_scatter(&sc, 0, V(p[1], p[2], p[3]), &rec, &rng, &att, &so, &sd);
        if (t == 3) { att = orc_emitted(&sc, 0, 0.f, 0.f, rec.p); so = V(0, 0, 0); }
        float *q = out + 11 * i;
        q[0] = (float)ok; q[1] = att.x; q[2] = att.y; q[3] = att.z; q[4] = so.x; q[5] = so.y; q[6] = so.z;
        q[7] = sd.x; q[8] = sd.y; q[9] = sd.z; q[10] = (float)rng.draws;
    }
}

void orc_kat_camera(const float *in, float *out, int n) {
    for (int i = 0; i < n; ++i) {
        const float *p = in + 14 * i;
        rt_camera_desc c; memset(&c, 0, sizeof c);
        for (int k = 0; k < 3; ++k) { c.lookfrom[k] = p[k]; c.lookat[k] = p[3 + k]; c.vup[k] = p[6 + k]; }
        c.vfov_deg = p[9]; c.aspect = p[10]; c.aperture = p[11];
        v3 df = vsub(V(p[0], p[1], p[2]), V(p[3], p[4], p[5]));
        c.focus_dist = sqrtf(hdot(df, df));                   /* parser.hpp:124 */
        orc_camera cam; orc_camera_setup(&c, &cam);
        v3 o, d; orc_get_ray(&cam, p[12], p[13], 0, NULL, &o, &d);
        float *q = out + 6 * i;
        q[0] = o.x; q[1] = o.y; q[2] = o.z; q[3] = d.x; q[4] = d.y; q[5] = d.z;
    }
}

void orc_kat_xform(const float *in, float *out, int n) {
    for (int i = 0; i < n; ++i) {
        const float *p = in + 10 * i;
        rt_xform_desc xf; orc_make_xform(p, p[3], 1, p + 4, 1, &xf);
        v3 v = V(p[7], p[8], p[9]);
        v3 r0 = orc_xform_point(xf.m, v), r1 = orc_xform_vec(xf.m, v), r2 = orc_xform_normal(xf.m_inv, v),
           r3 = orc_xform_point(xf.m_inv, v);
        float *q = out + 12 * i;
        q[0] = r0.x; q[1] = r0.y; q[2] = r0.z; q[3] = r1.x; q[4] = r1.y; q[5] = r1.z;
        q[6] = r2.x; q[7] = r2.y; q[8] = r2.z; q[9] = r3.x; q[10] = r3.y; q[11] = r3.z;
    }
}
