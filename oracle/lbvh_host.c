/* lbvh_host.c — host twin of the GPU LBVH build (Morton codes -> radix sort -> Karras hierarchy -> refit).
 * TEST INFRASTRUCTURE ONLY (see rt_oracle.h).
 *
 * The reference has no BVH in its C++/CUDA trees (hittable_list is a linear walk, object.cuh:23-37;
 * aabb.hpp is dead code), so there is nothing to pin this against upstream: the contract is that the
 * GPU build is BIT-EXACT with this serial statement of the same algorithm (north star, kernel group 1).
 *   Morton:   63-bit, 21 bits per axis, centroid normalised to the centroid bounds (per axis; one scale for all axes in trees of more than 4096 leaves)
 *   sort:     stable LSD radix, 8 passes x 8-bit digits, on (morton, object index) pairs
 *   topology: Karras 2012, delta(i,j) = clz(m_i ^ m_j), ties broken by 64 + clz(i ^ j)
 *   boxes:    leaf = object box padded by 2^-10 * (largest object extent) + 2^-20 * (largest scene extent);
 *             internal = union of children
 *   giants:   an object whose largest box extent exceeds a quarter of the scene's (the r = 1000 ground sphere of the RTIOW
 *             scene, cmake-cpu-version/main.cpp:125-172; the 200 x 200 rects of blue2.json) would make every box on its
 *             ancestor chain cover the whole scene.  Up to ORC_MAX_GIANTS such objects stay OUT of the tree: they get the key
 *             2^63 (behind every 63-bit Morton code; the stable sort keeps them in scene-file order), occupy the last
 *             n_giants leaf slots, and the traversal tests them before it walks the tree (the "last index wins" tie rule of
 *             object.cuh:29-33 does not depend on the order of the tests).  Morton codes are normalised to the centroid
 *             bounds of the other objects.  A giant sphere's box (read only by the camera-beam tile lists) is padded by the
 *             float error bound of sphere::hit for rays that start at the camera instead of 2^-10 of its extent.
 */
#include "orc_internal.h"
#include <float.h>

#ifndef ORC_MAX_GIANTS
#define ORC_MAX_GIANTS 16
#endif
#define ORC_GIANT_KEY 0x8000000000000000ULL
#define ORC_MORTON_CUBE_MIN 4096

void orc_object_aabb(const rt_scene_desc *sc, int i, float bmin[3], float bmax[3]) {
    const rt_object_desc *ob = &sc->objects[i];
    const float *p = ob->p;
    const rt_xform_desc *xf = ob->xform >= 0 ? &sc->xforms[ob->xform] : NULL;
    float lo[3], hi[3];
    int corners = 0;
    switch (ob->type) {
    case RT_OBJ_SPHERE: {
        float r = fabsf(p[3]);
        for (int k = 0; k < 3; ++k) { bmin[k] = p[k] - r; bmax[k] = p[k] + r; }
        if (!xf) return;
        for (int k = 0; k < 3; ++k) { lo[k] = bmin[k]; hi[k] = bmax[k]; }      /* instance: world box of the object-space box's corners */
        corners = 1;
        break;
    }
    case RT_OBJ_XY_RECT: case RT_OBJ_XZ_RECT: case RT_OBJ_YZ_RECT: {
        int a = ob->type == RT_OBJ_YZ_RECT ? 1 : 0, b = ob->type == RT_OBJ_XY_RECT ? 1 : 2;
        int k = ob->type == RT_OBJ_XY_RECT ? 2 : ob->type == RT_OBJ_XZ_RECT ? 1 : 0;
        bmin[a] = fminf(p[0], p[1]); bmax[a] = fmaxf(p[0], p[1]);
        bmin[b] = fminf(p[2], p[3]); bmax[b] = fmaxf(p[2], p[3]);
        bmin[k] = p[4] - 1e-4f; bmax[k] = p[4] + 1e-4f;
        if (!xf) return;
        for (int q = 0; q < 3; ++q) { lo[q] = bmin[q]; hi[q] = bmax[q]; }
        corners = 1;
        break;
    }
    case RT_OBJ_CYLINDER: {
        float r = fabsf(p[0]);
        lo[0] = -r; lo[1] = -r; lo[2] = fminf(p[1], p[2]); hi[0] = r; hi[1] = r; hi[2] = fmaxf(p[1], p[2]);
        corners = 1;
        break;
    }
    case RT_OBJ_BOX:
        for (int k = 0; k < 3; ++k) { lo[k] = fminf(p[k], p[3 + k]); hi[k] = fmaxf(p[k], p[3 + k]); }
        corners = 1;
        break;
    case RT_OBJ_TRIANGLE:
        for (int k = 0; k < 3; ++k) {
            bmin[k] = fminf(p[k], fminf(p[3 + k], p[6 + k]));
            bmax[k] = fmaxf(p[k], fmaxf(p[3 + k], p[6 + k]));
        }
        return;
    default:
        for (int k = 0; k < 3; ++k) { bmin[k] = 0.f; bmax[k] = 0.f; }
        return;
    }
    if (corners && !xf) { for (int k = 0; k < 3; ++k) { bmin[k] = lo[k]; bmax[k] = hi[k]; } return; }
    for (int k = 0; k < 3; ++k) { bmin[k] = INFINITY; bmax[k] = -INFINITY; }
    for (int c = 0; c < 8; ++c) {
        v3 q = orc_xform_point(xf->m, V((c & 1) ? hi[0] : lo[0], (c & 2) ? hi[1] : lo[1], (c & 4) ? hi[2] : lo[2]));
        float qq[3] = {q.x, q.y, q.z};
        for (int k = 0; k < 3; ++k) { bmin[k] = fminf(bmin[k], qq[k]); bmax[k] = fmaxf(bmax[k], qq[k]); }
    }
}

static uint64_t expand21(uint32_t v) {   /* spread the low 21 bits so that two zero bits follow each bit */
    uint64_t x = v & 0x1fffffu;
    x = (x | x << 32) & 0x1f00000000ffffULL;
    x = (x | x << 16) & 0x1f0000ff0000ffULL;
    x = (x | x << 8) & 0x100f00f00f00f00fULL;
    x = (x | x << 4) & 0x10c30c30c30c30c3ULL;
    x = (x | x << 2) & 0x1249249249249249ULL;
    return x;
}

static uint32_t quant21(float c, float cmin, float ext) {
    float scale = ext > 0.0f ? 2097152.0f / ext : 0.0f;
    float q = (c - cmin) * scale;
    q = fminf(fmaxf(q, 0.0f), 2097151.0f);
    return (uint32_t)q;
}

void orc_radix_sort_pairs(uint64_t *keys, int32_t *vals, int n) {
    uint64_t *k2 = (uint64_t *)malloc(sizeof(uint64_t) * (size_t)(n > 0 ? n : 1));
    int32_t *v2 = (int32_t *)malloc(sizeof(int32_t) * (size_t)(n > 0 ? n : 1));
    for (int pass = 0; pass < 8; ++pass) {
        size_t cnt[257] = {0};
        int sh = pass * 8;
        for (int i = 0; i < n; ++i) cnt[((keys[i] >> sh) & 255) + 1]++;
        for (int d = 0; d < 256; ++d) cnt[d + 1] += cnt[d];
        for (int i = 0; i < n; ++i) { size_t d = (keys[i] >> sh) & 255; k2[cnt[d]] = keys[i]; v2[cnt[d]] = vals[i]; cnt[d]++; }
        memcpy(keys, k2, sizeof(uint64_t) * (size_t)n); memcpy(vals, v2, sizeof(int32_t) * (size_t)n);
    }
    free(k2); free(v2);
}

static inline int delta(const uint64_t *m, int n, int i, int j) {
    if (j < 0 || j >= n) return -1;
    uint64_t x = m[i] ^ m[j];
    if (x) return __builtin_clzll(x);
    return 64 + __builtin_clz((uint32_t)(i ^ j));     /* i != j here */
}

/* Giant-sphere box padding: largest distance m by which a ray starting within D of the centre can miss the sphere while the float
 * discriminant of sphere::hit (object.cuh:51-58) still comes out >= 0, and by which an accepted hit point can lie off the surface:
 * m^2 + 2 r m <= K (D^2 + r^2) with K = 2^-18 (64 ulp of the largest term, a loose bound on the ~20 roundings involved). */
static float giant_sphere_pad(float r, float D) {
    float x = 3.814697265625e-06f * (D * D + r * r);
    return x / (sqrtf(r * r + x) + r);
}

int orc_lbvh_build_ex(const rt_scene_desc *sc, uint64_t *morton_sorted, int32_t *prim_sorted, int32_t *left, int32_t *right,
                      int32_t *parent, float *node_aabb, float *leaf_aabb, float *scene_aabb, int32_t *info) {
    const int n = sc->n_objects;
    if (info) { info[0] = 0; info[1] = 0; info[2] = 0; info[3] = 0; }
    if (n <= 0) return 0;
    float *bmin = (float *)malloc(sizeof(float) * 3 * (size_t)n), *bmax = (float *)malloc(sizeof(float) * 3 * (size_t)n);
    float smin[3] = {INFINITY, INFINITY, INFINITY}, smax[3] = {-INFINITY, -INFINITY, -INFINITY};
    for (int i = 0; i < n; ++i) {
        orc_object_aabb(sc, i, bmin + 3 * i, bmax + 3 * i);
        for (int k = 0; k < 3; ++k) { smin[k] = fminf(smin[k], bmin[3 * i + k]); smax[k] = fmaxf(smax[k], bmax[3 * i + k]); }
    }
    if (scene_aabb) for (int k = 0; k < 3; ++k) { scene_aabb[k] = smin[k]; scene_aabb[3 + k] = smax[k]; }
    const float sext = fmaxf(smax[0] - smin[0], fmaxf(smax[1] - smin[1], smax[2] - smin[2]));
    /* giants: largest extent > sext / 4; none when more than ORC_MAX_GIANTS qualify (a scene made of large pieces) */
    unsigned char *giant = (unsigned char *)calloc((size_t)n, 1);
    int ng = 0;
    for (int i = 0; i < n; ++i) {
        float e = fmaxf(bmax[3 * i] - bmin[3 * i], fmaxf(bmax[3 * i + 1] - bmin[3 * i + 1], bmax[3 * i + 2] - bmin[3 * i + 2]));
        giant[i] = e > 0.25f * sext;
        ng += giant[i];
    }
    if (ng > ORC_MAX_GIANTS) { memset(giant, 0, (size_t)n); ng = 0; }
    const int nt = n - ng;                      /* leaves of the tree */
    float cmin[3] = {INFINITY, INFINITY, INFINITY}, cmax[3] = {-INFINITY, -INFINITY, -INFINITY};
    for (int i = 0; i < n; ++i) {
        if (giant[i]) continue;
        for (int k = 0; k < 3; ++k) {
            float c = 0.5f * (bmin[3 * i + k] + bmax[3 * i + k]);
            cmin[k] = fminf(cmin[k], c); cmax[k] = fmaxf(cmax[k], c);
        }
    }
    uint64_t *keys = (uint64_t *)malloc(sizeof(uint64_t) * (size_t)n);
    int32_t *vals = (int32_t *)malloc(sizeof(int32_t) * (size_t)n);
    /* trees of more than ORC_MORTON_CUBE_MIN leaves (meshes): one scale for the three axes, a cubical Morton grid (the cells of a flat
     * surface are not thin slabs); smaller scenes: every axis normalised to its own extent (rt_lbvh.cu:k_morton has the measurements) */
    const float cext = fmaxf(cmax[0] - cmin[0], fmaxf(cmax[1] - cmin[1], cmax[2] - cmin[2]));
    for (int i = 0; i < n; ++i) {
        vals[i] = i;
        if (giant[i]) { keys[i] = ORC_GIANT_KEY; continue; }
        uint32_t q[3];
        for (int k = 0; k < 3; ++k) {
            float c = 0.5f * (bmin[3 * i + k] + bmax[3 * i + k]);
            q[k] = quant21(c, cmin[k], nt > ORC_MORTON_CUBE_MIN ? cext : cmax[k] - cmin[k]);
        }
        keys[i] = expand21(q[0]) << 2 | expand21(q[1]) << 1 | expand21(q[2]);
    }
    orc_radix_sort_pairs(keys, vals, n);
    if (morton_sorted) memcpy(morton_sorted, keys, sizeof(uint64_t) * (size_t)n);
    if (prim_sorted) memcpy(prim_sorted, vals, sizeof(int32_t) * (size_t)n);

    float pad_abs = sext * 9.5367431640625e-07f;                     /* 2^-20 */
    float *lb = (float *)malloc(sizeof(float) * 6 * (size_t)n);
    for (int j = 0; j < n; ++j) {
        int i = vals[j];
        float e = fmaxf(bmax[3 * i] - bmin[3 * i], fmaxf(bmax[3 * i + 1] - bmin[3 * i + 1], bmax[3 * i + 2] - bmin[3 * i + 2]));
        float pad = fmaf(e, 9.765625e-04f, pad_abs);                  /* 2^-10 */
        if (giant[i] && sc->objects[i].type == RT_OBJ_SPHERE) {        /* camera rays only: error bound instead of 2^-10 e */
            const float *p = sc->objects[i].p, *cf = sc->camera.lookfrom;
            float dx = cf[0] - p[0], dy = cf[1] - p[1], dz = cf[2] - p[2];
            float D = sqrtf(dx * dx + dy * dy + dz * dz) + 0.5f * fabsf(sc->camera.aperture);
            pad = giant_sphere_pad(fabsf(p[3]), D) + pad_abs;
        }
        for (int k = 0; k < 3; ++k) { lb[6 * j + k] = bmin[3 * i + k] - pad; lb[6 * j + 3 + k] = bmax[3 * i + k] + pad; }
    }
    if (leaf_aabb) memcpy(leaf_aabb, lb, sizeof(float) * 6 * (size_t)n);

    int depth = nt >= 1 ? 1 : 0;
    if (nt >= 2) {
        int32_t *L = (int32_t *)malloc(sizeof(int32_t) * (size_t)(nt - 1)), *R = (int32_t *)malloc(sizeof(int32_t) * (size_t)(nt - 1));
        int32_t *P = (int32_t *)malloc(sizeof(int32_t) * (size_t)(2 * nt - 1));
        P[0] = -1;
        for (int i = 0; i < nt - 1; ++i) {     /* Karras 2012, Algorithm "construct internal nodes" */
            int d = delta(keys, nt, i, i + 1) - delta(keys, nt, i, i - 1) >= 0 ? 1 : -1;
            int dmin = delta(keys, nt, i, i - d);
            int lmax = 2;
            while (delta(keys, nt, i, i + lmax * d) > dmin) lmax *= 2;
            int l = 0;
            for (int t = lmax / 2; t >= 1; t /= 2)
                if (delta(keys, nt, i, i + (l + t) * d) > dmin) l += t;
            int j = i + l * d;
            int dnode = delta(keys, nt, i, j);
            int s = 0, t = l;
            do {
                t = (t + 1) >> 1;
                if (delta(keys, nt, i, i + (s + t) * d) > dnode) s += t;
            } while (t > 1);
            int gamma = i + s * d + (d < 0 ? -1 : 0);
            int lo = i < j ? i : j, hi = i < j ? j : i;
            L[i] = lo == gamma ? ~gamma : gamma;
            R[i] = hi == gamma + 1 ? ~(gamma + 1) : gamma + 1;
            if (L[i] >= 0) P[L[i]] = i; else P[nt - 1 + ~L[i]] = i;
            if (R[i] >= 0) P[R[i]] = i; else P[nt - 1 + ~R[i]] = i;
        }
        /* refit bottom-up: process nodes in an order where children come first (iterate until all done) */
        float *nb = (float *)malloc(sizeof(float) * 6 * (size_t)(nt - 1));
        unsigned char *done = (unsigned char *)calloc((size_t)(nt - 1), 1);
        int remaining = nt - 1;
        while (remaining > 0) {
            for (int i = nt - 2; i >= 0; --i) {
                if (done[i]) continue;
                if ((L[i] >= 0 && !done[L[i]]) || (R[i] >= 0 && !done[R[i]])) continue;
                const float *a = L[i] >= 0 ? nb + 6 * L[i] : lb + 6 * (~L[i]);
                const float *b = R[i] >= 0 ? nb + 6 * R[i] : lb + 6 * (~R[i]);
                for (int k = 0; k < 3; ++k) { nb[6 * i + k] = fminf(a[k], b[k]); nb[6 * i + 3 + k] = fmaxf(a[3 + k], b[3 + k]); }
                done[i] = 1; --remaining;
            }
        }
        /* depth = the largest number of internal nodes between the root and a leaf (the traversal stack never holds more) */
        for (int j = 0; j < nt; ++j) {
            int dpt = 0;
            for (int c = P[nt - 1 + j]; c >= 0; c = P[c]) ++dpt;
            if (dpt > depth) depth = dpt;
        }
        if (left) memcpy(left, L, sizeof(int32_t) * (size_t)(nt - 1));
        if (right) memcpy(right, R, sizeof(int32_t) * (size_t)(nt - 1));
        if (parent) memcpy(parent, P, sizeof(int32_t) * (size_t)(2 * nt - 1));
        if (node_aabb) memcpy(node_aabb, nb, sizeof(float) * 6 * (size_t)(nt - 1));
        free(L); free(R); free(P); free(nb); free(done);
    } else if (parent) parent[0] = -1;
    if (info) { info[0] = nt; info[1] = ng; info[2] = depth; }
    free(bmin); free(bmax); free(keys); free(vals); free(lb); free(giant);
    return 0;
}

int orc_lbvh_build(const rt_scene_desc *sc, uint64_t *morton_sorted, int32_t *prim_sorted, int32_t *left, int32_t *right,
                   int32_t *parent, float *node_aabb, float *leaf_aabb, float *scene_aabb) {
    return orc_lbvh_build_ex(sc, morton_sorted, prim_sorted, left, right, parent, node_aabb, leaf_aabb, scene_aabb, NULL);
}
