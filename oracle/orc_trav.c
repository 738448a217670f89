/* orc_trav.c — host twin of the device closest-hit traversal (csrc/rt_traverse.cuh) over the host-built LBVH.
 * TEST INFRASTRUCTURE ONLY (see rt_oracle.h).
 *
 * The reference has no tree (hittable_list::hit is a linear walk, gpu-version/object.cuh:23-37); this file exists so that
 *   (1) the conservative box arithmetic and the visiting order of the device traversal can be checked on the CPU against the
 *       brute-force walk of orc_hit.c (same hits on every ray, `-m "not gpu"` tests), and
 *   (2) the device's event counters (inner-node visits, box tests, leaf tests) can be compared with a serial count of the same
 *       walk: integer work, expected bit-exact, and the basis of the node-visits-per-ray figures in DESIGN.md.
 * Same float operation sequence as rt_traverse.cuh: reciprocal directions with the 1e18 clamp, near planes evaluated with
 * (ood (1-2^-20), oo (1-2^-20) - c), far planes with (ood (1+2^-20), oo (1+2^-20) + c), c = 2^-21 |oo| (+ 1e10 on a clamped axis);
 * giants first, then the stack walk: leaf step, ordered descent, far child pushed with its entry distance.
 */
#include "orc_internal.h"
#include <float.h>

struct orc_bvh {
    const rt_scene_desc *sc;
    int n, nt, ng, depth;
    int32_t *prim, *left, *right;
    float *node_aabb, *leaf_aabb;
};

/* ---- experiment only (ORC_BVH_SAH=1): top-down full-sweep SAH tree over the same leaf boxes, to size the headroom of tree quality ---- */
static float box_area(const float *b) { float x = b[3] - b[0], y = b[4] - b[1], z = b[5] - b[2]; return 2.0f * (x * y + y * z + z * x); }
static int sah_axis;
static const float *sah_lb;
static int sah_cmp(const void *a, const void *b) {
    const int i = *(const int *)a, j = *(const int *)b;
    const float ci = sah_lb[6 * i + sah_axis] + sah_lb[6 * i + 3 + sah_axis], cj = sah_lb[6 * j + sah_axis] + sah_lb[6 * j + 3 + sah_axis];
    return ci < cj ? -1 : ci > cj ? 1 : (i - j);
}
static int sah_build(orc_bvh *b, int *idx, int n, int *next_node) {      /* returns child code */
    if (n == 1) return ~idx[0];
    const int me = (*next_node)++;
    float best = INFINITY; int best_axis = 0, best_k = n / 2;
    int *tmp = (int *)malloc(sizeof(int) * (size_t)n);
    float *ra = (float *)malloc(sizeof(float) * (size_t)n);
    for (int ax = 0; ax < 3; ++ax) {
        memcpy(tmp, idx, sizeof(int) * (size_t)n);
        sah_axis = ax; sah_lb = b->leaf_aabb;
        qsort(tmp, (size_t)n, sizeof(int), sah_cmp);
        float bb[6] = {INFINITY, INFINITY, INFINITY, -INFINITY, -INFINITY, -INFINITY};
        for (int k = n - 1; k >= 1; --k) {
            const float *l = b->leaf_aabb + 6 * tmp[k];
            for (int c = 0; c < 3; ++c) { bb[c] = fminf(bb[c], l[c]); bb[3 + c] = fmaxf(bb[3 + c], l[3 + c]); }
            ra[k] = box_area(bb);
        }
        float lbx[6] = {INFINITY, INFINITY, INFINITY, -INFINITY, -INFINITY, -INFINITY};
        for (int k = 1; k < n; ++k) {
            const float *l = b->leaf_aabb + 6 * tmp[k - 1];
            for (int c = 0; c < 3; ++c) { lbx[c] = fminf(lbx[c], l[c]); lbx[3 + c] = fmaxf(lbx[3 + c], l[3 + c]); }
            const float cost = box_area(lbx) * (float)k + ra[k] * (float)(n - k);
            if (cost < best) { best = cost; best_axis = ax; best_k = k; }
        }
    }
    sah_axis = best_axis; sah_lb = b->leaf_aabb;
    qsort(idx, (size_t)n, sizeof(int), sah_cmp);
    free(tmp); free(ra);
    const int L = sah_build(b, idx, best_k, next_node), R = sah_build(b, idx + best_k, n - best_k, next_node);
    b->left[me] = L; b->right[me] = R;
    float *nb = b->node_aabb + 6 * me;
    const float *a = L >= 0 ? b->node_aabb + 6 * L : b->leaf_aabb + 6 * (~L), *c = R >= 0 ? b->node_aabb + 6 * R : b->leaf_aabb + 6 * (~R);
    for (int k = 0; k < 3; ++k) { nb[k] = fminf(a[k], c[k]); nb[3 + k] = fmaxf(a[3 + k], c[3 + k]); }
    return me;
}

orc_bvh *orc_bvh_create(const rt_scene_desc *sc) {
    orc_bvh *b = (orc_bvh *)calloc(1, sizeof *b);
    const int n = sc->n_objects, m = n > 0 ? n : 1;
    b->sc = sc; b->n = n;
    b->prim = (int32_t *)malloc(sizeof(int32_t) * (size_t)m);
    b->left = (int32_t *)malloc(sizeof(int32_t) * (size_t)m); b->right = (int32_t *)malloc(sizeof(int32_t) * (size_t)m);
    b->node_aabb = (float *)malloc(sizeof(float) * 6 * (size_t)m); b->leaf_aabb = (float *)malloc(sizeof(float) * 6 * (size_t)m);
    int32_t info[4] = {0, 0, 0, 0};
    orc_lbvh_build_ex(sc, NULL, b->prim, b->left, b->right, NULL, b->node_aabb, b->leaf_aabb, NULL, info);
    b->nt = info[0]; b->ng = info[1]; b->depth = info[2];
    if (getenv("ORC_BVH_SAH") && b->nt >= 2) {
        int *idx = (int *)malloc(sizeof(int) * (size_t)b->nt), next = 0;
        for (int i = 0; i < b->nt; ++i) idx[i] = i;
        sah_build(b, idx, b->nt, &next);
        free(idx);
    }
    return b;
}

void orc_bvh_destroy(orc_bvh *b) {
    if (!b) return;
    free(b->prim); free(b->left); free(b->right); free(b->node_aabb); free(b->leaf_aabb); free(b);
}

#define BOX_REL 9.5367431640625e-07f
#define BOX_ABS 4.76837158203125e-07f
#define BOX_OOD_MAX 1e18f
#define BOX_PAR 1e10f

static void count_leaf(orc_trav_counters *c, int type, int hit) {
    if (!c) return;
    switch (type) {
    case RT_OBJ_SPHERE: if (hit) c->sph_acc++; else c->sph_rej++; break;
    case RT_OBJ_XY_RECT: case RT_OBJ_XZ_RECT: case RT_OBJ_YZ_RECT: c->rect++; break;
    case RT_OBJ_CYLINDER: if (hit) c->cyl_acc++; else c->cyl_rej++; break;
    case RT_OBJ_BOX: c->boxprim++; break;
    default: c->tri++; break;
    }
}

/* one exact leaf test with the tie rule of the linear walk (rt_traverse.cuh:leaf_visit) */
static int leaf_visit(const orc_bvh *b, int slot, v3 o, v3 d, float inv_a, float tmin, float *best_t, int *best_obj, orc_hit *rec,
                      orc_trav_counters *cnt) {
    const int i = b->prim[slot];
    orc_hit tmp;
    const int h = orc_hit_object(b->sc, i, o, d, inv_a, tmin, *best_t, &tmp);
    count_leaf(cnt, b->sc->objects[i].type, h);
    if (!h) return 0;
    int take = tmp.t < *best_t || *best_obj < 0;
    if (!take) take = i > *best_obj;            /* t == best_t: the later object in the scene file wins (object.cuh:29-33) */
    if (take) { *best_t = tmp.t; *best_obj = i; *rec = tmp; }
    return take;
}

int orc_bvh_hit(const orc_bvh *b, v3 o, v3 d, float tmin, float tmax, int any_hit, orc_hit *rec, orc_trav_counters *cnt) {
    float best_t = tmax; int best_obj = -1;
    rec->hit = 0;
    if (b->n == 0) return 0;
    const float a = dot3(d, d), inv_a = 1.0f / a;
    for (int g = 0; g < b->ng; ++g)
        if (leaf_visit(b, b->nt + g, o, d, inv_a, tmin, &best_t, &best_obj, rec, cnt) && any_hit) return 1;
    if (b->nt == 0) return best_obj >= 0;
    float dd[3] = {d.x, d.y, d.z}, oo3[3] = {o.x, o.y, o.z};
    float oodn[3], oodf[3], oon[3], oof[3]; int neg[3];
    for (int k = 0; k < 3; ++k) {
        float ood = 1.0f / dd[k];
        const int par = !(fabsf(ood) <= BOX_OOD_MAX);
        if (par) ood = copysignf(BOX_OOD_MAX, dd[k]);
        const float oo = -(oo3[k] * ood);
        const float ca = fmaf(BOX_ABS, fabsf(oo), par ? BOX_PAR : 0.0f);
        const float GN = 1.0f - BOX_REL, GF = 1.0f + BOX_REL;
        oodn[k] = ood * GN; oodf[k] = ood * GF;
        oon[k] = fmaf(oo, GN, -ca); oof[k] = fmaf(oo, GF, ca);
        neg[k] = signbit(dd[k]) != 0;
    }
    struct { int node; float t; } stack[128];
    int sp = 0, cur = 0, done = 0;
    while (!done) {
        if (cur < 0) {
            if (leaf_visit(b, ~cur, o, d, inv_a, tmin, &best_t, &best_obj, rec, cnt) && any_hit) return 1;
            cur = 0; done = 1;
            while (sp > 0) { --sp; if (stack[sp].t <= best_t) { cur = stack[sp].node; done = 0; break; } }
            if (done) break;
        }
        if (cur >= 0) {
            int child[2]; float tn[2], tf[2];
            if (b->nt >= 2) { child[0] = b->left[cur]; child[1] = b->right[cur]; }
            else { child[0] = ~0; child[1] = 0x7fffffff; }      /* single leaf: the device packs (leaf 0, none) */
            for (int s = 0; s < 2; ++s) {
                if (child[s] == 0x7fffffff) { tn[s] = INFINITY; tf[s] = -INFINITY; continue; }
                const float *bx = child[s] >= 0 ? b->node_aabb + 6 * child[s] : b->leaf_aabb + 6 * (~child[s]);
                float n3[3], f3[3];
                for (int k = 0; k < 3; ++k) {
                    const float pn = neg[k] ? bx[3 + k] : bx[k], pf = neg[k] ? bx[k] : bx[3 + k];
                    n3[k] = fmaf(pn, oodn[k], oon[k]); f3[k] = fmaf(pf, oodf[k], oof[k]);
                }
                tn[s] = fmaxf(fmaxf(n3[0], n3[1]), fmaxf(n3[2], tmin));
                tf[s] = fminf(fminf(f3[0], f3[1]), fminf(f3[2], best_t));
            }
            if (cnt) { cnt->box += 2; cnt->nodes++; }
            const int hl = tn[0] <= tf[0], hr = tn[1] <= tf[1];
            if (hl && hr) {
                const int lfirst = tn[0] <= tn[1];
                if (sp < 128) { stack[sp].node = lfirst ? child[1] : child[0]; stack[sp].t = lfirst ? tn[1] : tn[0]; ++sp; }
                else if (cnt) cnt->overflow++;
                if (cnt && sp > cnt->max_stack) cnt->max_stack = sp;
                cur = lfirst ? child[0] : child[1];
            } else if (hl || hr) {
                cur = hl ? child[0] : child[1];
            } else {
                cur = 0; done = 1;
                while (sp > 0) { --sp; if (stack[sp].t <= best_t) { cur = stack[sp].node; done = 0; break; } }
            }
        }
    }
    return best_obj >= 0;
}
