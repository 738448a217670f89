// rt_scene.cuh — flat device-side scene tables (SoA) and the per-primitive / per-material device functions.
//
// Replaces the reference's pointer graph of virtual objects on the device heap
// (gpu-version/object.cuh, material.cuh, texture.cuh, built by move_to_device, main.cu:374-446).
// Leaf slot order is the Morton order produced by the LBVH build; hit records carry only (t, slot)
// and everything else (p, normal, uv, material) is recomputed in the shade stage.
#pragma once
#include "rt_math.cuh"
#include "../../include/rt.h"

struct __align__(16) BvhNode {   // 64 B: both child boxes + child links; one node = four 16-byte loads
    float4 a;   // l.min.x l.min.y l.min.z l.max.x
    float4 b;   // l.max.y l.max.z r.min.x r.min.y
    float4 c;   // r.min.z r.max.x r.max.y r.max.z
    int4 d;     // left, right, unused, unused.   child >= 0: internal node; child < 0: leaf, ~child = type<<28 | slot
};
// The same node laid out for the extend stage's sign-ordered slab test.  Per box and axis the two planes are stored
// as (lo, hi, hi, lo): an 8-byte load at offset 0 yields (near, far) for a ray travelling in +axis, at offset 8 for -axis,
// so the traversal needs no per-plane min/max or select — the ray's direction signs are folded into the load address.
// 104 bytes = 13 x 8: an ODD number of 8-byte bank pairs, so the same field of nodes i and j lies in the same shared-memory
// bank pair only when i = j (mod 16).  With the 128-byte stride this struct had first, every node's field sat in the SAME bank
// pair and the 8-byte loads of a warp whose lanes visit different nodes were fully serialised: ncu (r02a) had 51-61 % of the
// shared-memory wavefronts of k_extend at bounces 2-3 as bank conflicts and the shared-memory pipe at 77-82 % of its peak.
struct __align__(8) BvhNodeS {
    float2 lx[2], ly[2], lz[2];   // left child box: per axis (lo, hi), (hi, lo)
    float2 rx[2], ry[2], rz[2];   // right child box
    int2 link;                    // left, right (leaf: ~code as BvhNode::d; inner: byte offset of the node = index * 104)
};
static_assert(sizeof(BvhNodeS) == 104, "BvhNodeS must be 13 x 8 bytes");
// Trees that do not fit into shared memory (the 1 M-triangle mesh: 64 MB of 64-byte nodes) are walked through L1/L2, where the
// walk is bound by the latency of dependent node fetches.  BvhNode4 = a binary node at EVEN depth with its four grandchildren
// (a leaf child keeps its own slot): one 128-byte line per visit, half as many dependent fetches per ray.  Child boxes are
// stored per plane (lo.x of the four children, ...), so the ray's direction signs pick whole 16-byte loads as near / far planes.
// Indexed by the binary node index (odd-depth entries stay unused: no compaction pass).
struct __align__(16) BvhNode4 {
    float4 plane[6];    // lo.x lo.y lo.z hi.x hi.y hi.z, one component per child; an empty slot holds (+inf, -inf): never hit
    int4 link;          // >= 0: BvhNode4 index, < 0: leaf, ~link = type<<28 | slot
    int4 pad;
};
#define RT_LEAF_SLOT_MASK 0x0fffffff
#define RT_LEAF_NONE (~(7 << 28))   // a leaf code that is never hit (its box is empty)

struct DMaterial { int type; int texture; float fuzz; float ior; float4 albedo; };   // 32 B; albedo.w != 0: the texture chain reads (u,v)
struct DTexture { int type; int even; int odd; int image; float4 color; };            // 32 B
struct DImage { cudaTextureObject_t tex; int width, height; };

#ifndef RT_TILE_K
#define RT_TILE_K 64        // candidates per 8x4 tile; 16 covered every RTIOW tile but left 40 % of the mesh tiles to full traversal (64: 9 %)
#endif
#define RT_MORTON_CUBE_MIN 4096 // trees with more leaves quantise their Morton codes on a cubical grid (rt_lbvh.cu:k_morton)
#define RT_MAX_GIANTS 16        // objects kept out of the tree and tested first by every ray (rt_lbvh.cu:k_giants_mark)
struct DCamera { float3 origin, llc, horizontal, vertical, u, v, w; float lens_radius; };

struct DScene {
    int n_prims, n_nodes;                  // all leaf slots; inner nodes of the tree (1 when the tree has fewer than two leaves)
    int n_tree, n_giants;                  // leaf slots [0, n_tree) form the tree, [n_tree, n_prims) hold the giants
    const int *giant_code;                 // their leaf codes (type << 28 | slot), device array (a by-value array indexed at run time
                                           // would force a local-memory copy of this whole struct in every kernel)
    const float *leaf_aabb;                // padded box per leaf slot (tile lists: boxes of the giants)
    int width, height, max_depth, sampling, bg_mode, defocus;      // sampling: the family (RT_SAMPLING_REFERENCE_GPU / _CPU)
    int direct;                            // RT_SAMPLING_DIRECT: direct draws instead of the reference's rejection loops
    float3 background;
    DCamera cam;
    const BvhNode *nodes;                  // AOV pass and build tests
    const BvhNodeS *nodes_s;               // extend stage, tree staged in shared memory
    const BvhNode4 *nodes4;                // extend stage, tree read from global memory (null when the tree fits shared memory)
    const float4 *primA, *primB, *primC;   // geometry by leaf slot
    const int *prim_obj;                   // slot -> object index in the scene file (closest-hit tie rule)
    const int *prim_mat;                   // slot -> material index
    const float4 *xforms;                  // 6 float4 per transform: rows of m (3x4) then rows of m_inv (3x4)
    const DMaterial *mats;
    // next-event estimation (RT_RENDER_NEE): sampled lights in scene-file order
    int n_lights;
    int nee_mis;                           // RT_RENDER_NEE_MIS: balance-heuristic weights on light samples and on emission found by scattered rays
    const int *light_slot;                 // light -> leaf slot
    const int *light_type;                 // light -> RT_OBJ_*
    const float *light_area;               // light -> surface area
    const int *obj_light;                  // object index -> light index or -1
    // primary-ray candidate lists (rt_kernels.cu:k_tile_lists): per 8x4 pixel tile the leaves its camera beam can touch
    int sphere_xf;                         // 1: some sphere carries an instance transform (else the sphere tests skip the primB.w flag load: a uniform branch)
    int only_type;                         // RT_OBJ_* when every primitive has that type (k_extend specialisation), else -1
    const DMaterial *slot_mat;             // material record per leaf slot (scenes of <= 65 536 leaves; else nullptr): one load instead of prim_mat -> mats
    int tile_mode;                         // 0: none, 1: every path of a pass is covered, 2: covered where tile_cnt >= 0
    const int *tile_list, *tile_cnt;       // [n_tiles * RT_TILE_K] leaf codes, [n_tiles] count or -1 (too many: fix-up pass)
    const int *tile_empty_prefix;          // [n_tiles + 1] number of tiles with count 0 before tile t (their paths skip the queues), or null
    const DTexture *texs;
    const DImage *images;
};

struct HitRec {
    float t, u, v;
    float3 p, n;
    int front;
};

// ---------------- transforms (vec3.cuh:331-370) ----------------
__device__ __forceinline__ float3 xf_point(const float4 *m, float3 p) {
    float4 r0 = m[0], r1 = m[1], r2 = m[2];
    return v3(__fmaf_rn(r0.z, p.z, __fmaf_rn(r0.y, p.y, r0.x * p.x)) + r0.w,
              __fmaf_rn(r1.z, p.z, __fmaf_rn(r1.y, p.y, r1.x * p.x)) + r1.w,
              __fmaf_rn(r2.z, p.z, __fmaf_rn(r2.y, p.y, r2.x * p.x)) + r2.w);
}
__device__ __forceinline__ float3 xf_vec(const float4 *m, float3 p) {
    float4 r0 = m[0], r1 = m[1], r2 = m[2];
    return v3(__fmaf_rn(r0.z, p.z, __fmaf_rn(r0.y, p.y, r0.x * p.x)),
              __fmaf_rn(r1.z, p.z, __fmaf_rn(r1.y, p.y, r1.x * p.x)),
              __fmaf_rn(r2.z, p.z, __fmaf_rn(r2.y, p.y, r2.x * p.x)));
}
__device__ __forceinline__ float3 xf_normal(const float4 *mi, float3 p) {   // transpose of the inverse
    float4 r0 = mi[0], r1 = mi[1], r2 = mi[2];
    return v3(__fmaf_rn(r2.x, p.z, __fmaf_rn(r1.x, p.y, r0.x * p.x)),
              __fmaf_rn(r2.y, p.z, __fmaf_rn(r1.y, p.y, r0.y * p.x)),
              __fmaf_rn(r2.z, p.z, __fmaf_rn(r1.z, p.y, r0.z * p.x)));
}

// ---------------- closest-hit distance per primitive type (extend stage: t only) ----------------
// Each returns true and the accepted root in t when the reference's hit() would (tmin <= t <= tmax semantics
// exactly as object.cuh spells them per type).

// sphere::hit, object.cuh:47-66.  inv_a = 1/dot(d,d) hoisted per ray.
__device__ __forceinline__ bool sphere_t(float4 A, float3 o, float3 d, float a, float inv_a, float tmin, float tmax, float &t) {
    float3 oc = o - f4xyz(A);
    float hb = dot3(oc, d);
    float cc = __fmaf_rn(-A.w, A.w, dot3(oc, oc));
    float disc = __fmaf_rn(hb, hb, -(a * cc));
    if (disc < 0.0f) return false;
    float sq = sqrtf(disc);
    float root = (-hb - sq) * inv_a;
    if (root < tmin || tmax < root) {
        root = (-hb + sq) * inv_a;
        if (root < tmin || tmax < root) return false;
    }
    t = root;
    return true;
}

// xy/xz/yz_rect::hit, object.cuh:105-115 (type: RT_OBJ_XY_RECT..RT_OBJ_YZ_RECT)
__device__ __forceinline__ void rect_axes(int type, float3 v, float &a, float &b, float &k) {
    if (type == RT_OBJ_XY_RECT) { a = v.x; b = v.y; k = v.z; }
    else if (type == RT_OBJ_XZ_RECT) { a = v.x; b = v.z; k = v.y; }
    else { a = v.y; b = v.z; k = v.x; }
}
__device__ __forceinline__ bool rect_t(int type, float4 A, float kk, float3 o, float3 d, float tmin, float tmax, float &t) {
    float oa, ob, ok, da, db, dk;
    rect_axes(type, o, oa, ob, ok); rect_axes(type, d, da, db, dk);
    float tt = (kk - ok) / dk;
    if (tt < tmin || tt > tmax) return false;
    float a = __fmaf_rn(tt, da, oa), b = __fmaf_rn(tt, db, ob);
    if (a < A.x || a > A.y || b < A.z || b > A.w) return false;
    t = tt;
    return true;
}

// cylinder::hit + quadratic, object.cuh:199-214,233-268.  oo/od: ray in object space.
__device__ __forceinline__ bool cylinder_t(float radius, float zmin, float zmax, float3 oo, float3 od, float tmin, float tmax, float &t) {
    float a = __fmaf_rn(od.y, od.y, od.x * od.x);
    float b = 2.0f * __fmaf_rn(od.y, oo.y, od.x * oo.x);
    float c = __fmaf_rn(-radius, radius, __fmaf_rn(oo.y, oo.y, oo.x * oo.x));
    float delta = __fmaf_rn(b, b, -(4.0f * a * c));
    if (delta < 0.0f) return false;
    float sd = sqrtf(delta);
    float t0 = -0.5f * (b - sd) / a, t1 = -0.5f * (b + sd) / a;
    if (t0 > t1) { float tmp = t0; t0 = t1; t1 = tmp; }
    if (t0 > tmax || t1 < tmin) return false;
    float tt = t0;
    if (t0 < tmin) { tt = t1; if (tt > tmax) return false; }
    float z = __fmaf_rn(tt, od.z, oo.z);
    if (z < zmin || z > zmax) {
        if (tt == t1) return false;
        tt = t1;
        if (tt > tmax || tt < tmin) return false;
        z = __fmaf_rn(tt, od.z, oo.z);
        if (z < zmin || z > zmax) return false;
    }
    t = tt;
    return true;
}

// BOX (extension): slab test in object space; returns the entry root, else the exit root; axis of the face hit.
__device__ __forceinline__ bool box_t(float3 lo, float3 hi, float3 oo, float3 od, float tmin, float tmax, float &t, int &axis) {
    float L[3] = {lo.x, lo.y, lo.z}, H[3] = {hi.x, hi.y, hi.z}, O[3] = {oo.x, oo.y, oo.z}, D[3] = {od.x, od.y, od.z};
    float tn = -INFINITY, tf = INFINITY; int an = 0, af = 0;
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        float t0 = (L[k] - O[k]) / D[k], t1 = (H[k] - O[k]) / D[k];
        if (D[k] == 0.0f) { if (O[k] < L[k] || O[k] > H[k]) return false; continue; }
        if (t0 > t1) { float tmp = t0; t0 = t1; t1 = tmp; }
        if (t0 > tn) { tn = t0; an = k; }
        if (t1 < tf) { tf = t1; af = k; }
    }
    if (tn > tf) return false;
    float tt = tn; int ax = an;
    if (tt < tmin || tmax < tt) { tt = tf; ax = af; if (tt < tmin || tmax < tt) return false; }
    t = tt; axis = ax;
    return true;
}

// TRIANGLE (extension): Moeller-Trumbore, two-sided.  A = v0, B = e1, C = e2.
__device__ __forceinline__ bool triangle_t(float3 v0, float3 e1, float3 e2, float3 o, float3 d, float tmin, float tmax,
                                           float &t, float &bu, float &bv) {
    float3 pv = v3(__fmaf_rn(d.y, e2.z, -(d.z * e2.y)), __fmaf_rn(d.z, e2.x, -(d.x * e2.z)), __fmaf_rn(d.x, e2.y, -(d.y * e2.x)));
    float det = dot3(e1, pv);
    if (det == 0.0f) return false;
    float inv = 1.0f / det;
    float3 tv = o - v0;
    float u = dot3(tv, pv) * inv;
    if (u < 0.0f || u > 1.0f) return false;
    float3 qv = v3(__fmaf_rn(tv.y, e1.z, -(tv.z * e1.y)), __fmaf_rn(tv.z, e1.x, -(tv.x * e1.z)), __fmaf_rn(tv.x, e1.y, -(tv.y * e1.x)));
    float v = dot3(d, qv) * inv;
    if (v < 0.0f || u + v > 1.0f) return false;
    float tt = dot3(e2, qv) * inv;
    if (tt < tmin || tmax < tt) return false;
    // A grazing ray (|det| tiny) carries a barycentric error of ~1e-7 |o - v0| |d x e2| / |det|: the float test above can accept a point
    // well outside the triangle — and outside its leaf box, where the tree (rightly) never looks: the 1 M-triangle terrain showed one
    // such ray in ~1e9, found by the tile candidate lists and missed by the traversal.  A hit counts only if its point lies in the
    // triangle's own box widened by 2^-11 of its largest extent, which is inside the leaf box (2^-10 of the extent + 2^-20 of the
    // scene) whatever the rounding: linear walk, tile lists and tree then agree by construction (oracle/orc_hit.c states the same).
    {
        const float lx = v0.x + fminf(0.0f, fminf(e1.x, e2.x)), hx = v0.x + fmaxf(0.0f, fmaxf(e1.x, e2.x));
        const float ly = v0.y + fminf(0.0f, fminf(e1.y, e2.y)), hy = v0.y + fmaxf(0.0f, fmaxf(e1.y, e2.y));
        const float lz = v0.z + fminf(0.0f, fminf(e1.z, e2.z)), hz = v0.z + fmaxf(0.0f, fmaxf(e1.z, e2.z));
        const float s = 4.8828125e-04f * fmaxf(hx - lx, fmaxf(hy - ly, hz - lz));
        const float3 q = vfma(tt, d, o);
        if (!(q.x >= lx - s && q.x <= hx + s && q.y >= ly - s && q.y <= hy + s && q.z >= lz - s && q.z <= hz + s)) return false;
    }
    t = tt; bu = u; bv = v;
    return true;
}

// One leaf test for the extend stage.  code = ~child = type<<28 | slot.
// ONLY >= 0: the scene holds primitives of that one type (spheres: the RTIOW scene) — no dispatch, the other tests are not compiled in.
template <int ONLY = -1>
__device__ __forceinline__ bool leaf_t(const DScene &sc, int code, float3 o, float3 d, float a, float inv_a, float tmin, float tmax, float &t) {
    const int type = (unsigned)code >> 28, slot = code & RT_LEAF_SLOT_MASK;
    const float4 A = __ldg(&sc.primA[slot]);
    if (ONLY == RT_OBJ_SPHERE) return sphere_t(A, o, d, a, inv_a, tmin, tmax, t);     // (ONLY: no sphere carries a transform)
    switch (type) {
    case RT_OBJ_SPHERE: {
        // rotate / translate instance (vec3.cuh:303-427 applied as object.cuh:225-231 does for the cylinder): the ray goes to
        // object space, t is unchanged (rigid transform, direction not normalised)
        const int xi = sc.sphere_xf ? __float_as_int(__ldg(&sc.primB[slot]).w) : -1;
        if (xi < 0) return sphere_t(A, o, d, a, inv_a, tmin, tmax, t);
        const float4 *xf = sc.xforms + 6 * xi;
        const float3 od = xf_vec(xf + 3, d);
        const float a2 = dot3(od, od);
        return sphere_t(A, xf_point(xf + 3, o), od, a2, 1.0f / a2, tmin, tmax, t);
    }
    case RT_OBJ_XY_RECT: case RT_OBJ_XZ_RECT: case RT_OBJ_YZ_RECT: {
        const float4 B = __ldg(&sc.primB[slot]);
        const int xi = __float_as_int(B.w);
        if (xi < 0) return rect_t(type, A, B.x, o, d, tmin, tmax, t);
        const float4 *xf = sc.xforms + 6 * xi;
        return rect_t(type, A, B.x, xf_point(xf + 3, o), xf_vec(xf + 3, d), tmin, tmax, t);
    }
    case RT_OBJ_CYLINDER: {
        const float4 *xf = sc.xforms + 6 * __float_as_int(A.w);
        return cylinder_t(A.x, A.y, A.z, xf_point(xf + 3, o), xf_vec(xf + 3, d), tmin, tmax, t);
    }
    case RT_OBJ_BOX: {
        const float4 B = __ldg(&sc.primB[slot]);
        const int xi = __float_as_int(A.w);
        float3 oo = o, od = d;
        if (xi >= 0) { const float4 *xf = sc.xforms + 6 * xi; oo = xf_point(xf + 3, o); od = xf_vec(xf + 3, d); }
        int axis;
        return box_t(f4xyz(A), f4xyz(B), oo, od, tmin, tmax, t, axis);
    }
    case RT_OBJ_TRIANGLE: {
        float bu, bv;
        return triangle_t(f4xyz(A), f4xyz(__ldg(&sc.primB[slot])), f4xyz(__ldg(&sc.primC[slot])), o, d, tmin, tmax, t, bu, bv);
    }
    }
    return false;
}

__device__ __forceinline__ void set_face_normal(HitRec &r, float3 d, float3 outward) {   // hittable.cuh:20-23
    r.front = dot3(d, outward) < 0.0f;
    r.n = r.front ? outward : -outward;
}

// Rebuild the full hit record of the reference (p, normal, front_face, u, v) from (t, slot) — shade stage.
// want_uv: sphere/cylinder uv cost an acos/atan2; only image textures read them.
// A, B: the slot's primA / primB records, loaded by the caller (the shade stage issues these loads before anything else it does)
// ONLY = RT_OBJ_SPHERE: every primitive of the scene is a sphere without an instance transform (DScene::only_type) — the other cases are
// not compiled into the caller, which is what lets the shade kernel of such scenes live in fewer registers.
template <int ONLY = -1>
__device__ __forceinline__ void hit_record(const DScene &sc, int type, int slot, float3 o, float3 d, float t, bool want_uv, HitRec &r,
                                           const float4 A, const float4 B) {
    r.t = t; r.u = 0.0f; r.v = 0.0f;
    if (ONLY == RT_OBJ_SPHERE) type = RT_OBJ_SPHERE;
    switch (type) {
    case RT_OBJ_SPHERE: {   // object.cuh:68-72,87-93
        const int xi = (ONLY == RT_OBJ_SPHERE || !sc.sphere_xf) ? -1 : __float_as_int(B.w);
        float3 n;
        if (xi < 0) {
            r.p = vfma(t, d, o);
            n = (1.0f / A.w) * (r.p - f4xyz(A));
            set_face_normal(r, d, n);
        } else {            // instance: hit point and normal in object space (the uv follow the rotated sphere), then to world space
            const float4 *xf = sc.xforms + 6 * xi;
            const float3 op = vfma(t, xf_vec(xf + 3, d), xf_point(xf + 3, o));
            n = (1.0f / A.w) * (op - f4xyz(A));
            r.p = xf_point(xf, op);
            set_face_normal(r, d, xf_normal(xf + 3, n));
        }
        if (want_uv) {
            float theta = rt_acosf(-n.y);
            float phi = rt_atan2f(-n.z, n.x) + RT_PI_F;
            r.u = phi * RT_INV_2PI_F; r.v = theta * RT_INV_PI_F;
        }
        break;
    }
    case RT_OBJ_XY_RECT: case RT_OBJ_XZ_RECT: case RT_OBJ_YZ_RECT: {   // object.cuh:113-121
        const int xi = __float_as_int(B.w);
        const float4 *xf = sc.xforms + 6 * (xi >= 0 ? xi : 0);
        float3 oo = o, od = d;
        if (xi >= 0) { oo = xf_point(xf + 3, o); od = xf_vec(xf + 3, d); }
        float oa, ob, ok, da, db, dk;
        rect_axes(type, oo, oa, ob, ok); rect_axes(type, od, da, db, dk);
        float a = __fmaf_rn(t, da, oa), b = __fmaf_rn(t, db, ob);
        r.u = (a - A.x) / (A.y - A.x); r.v = (b - A.z) / (A.w - A.z);
        float3 n = type == RT_OBJ_XY_RECT ? v3(0, 0, 1) : type == RT_OBJ_XZ_RECT ? v3(0, 1, 0) : v3(1, 0, 0);
        if (xi < 0) {
            r.front = dk < 0.0f;
            r.n = r.front ? n : -n;
            r.p = vfma(t, d, o);
        } else {
            r.p = xf_point(xf, vfma(t, od, oo));
            set_face_normal(r, d, xf_normal(xf + 3, n));
        }
        break;
    }
    case RT_OBJ_CYLINDER: {   // object.cuh:270-289
        const float4 *xf = sc.xforms + 6 * __float_as_int(A.w);
        float3 oo = xf_point(xf + 3, o), od = xf_vec(xf + 3, d);
        float3 op = vfma(t, od, oo);
        float len = sqrtf(__fmaf_rn(op.y, op.y, op.x * op.x));
        float3 on = v3(op.x / len, op.y / len, 0.0f / len);
        r.p = xf_point(xf, op);
        set_face_normal(r, d, xf_normal(xf + 3, on));
        if (want_uv) {
            float phi = rt_atan2f(op.y, op.x) + RT_2PI_F;
            r.u = phi * RT_INV_4PI_F; r.v = (op.z - A.y) / (A.z - A.y);
        }
        break;
    }
    case RT_OBJ_BOX: {
        const int xi = __float_as_int(A.w);
        const float4 *xf = sc.xforms + 6 * (xi >= 0 ? xi : 0);
        float3 oo = o, od = d;
        if (xi >= 0) { oo = xf_point(xf + 3, o); od = xf_vec(xf + 3, d); }
        // re-derive the face: the axis whose slab plane the point lies on (same decision as the oracle: recompute box_t)
        float tt; int ax = 0;
        box_t(f4xyz(A), f4xyz(B), oo, od, t, t, tt, ax);
        float3 op = vfma(t, od, oo);
        float P[3] = {op.x, op.y, op.z}, L[3] = {A.x, A.y, A.z}, H[3] = {B.x, B.y, B.z}, nn[3] = {0, 0, 0};
        nn[ax] = (P[ax] - L[ax]) < (H[ax] - P[ax]) ? -1.0f : 1.0f;
        int ua = (ax + 1) % 3, va = (ax + 2) % 3;
        r.u = (P[ua] - L[ua]) / (H[ua] - L[ua]); r.v = (P[va] - L[va]) / (H[va] - L[va]);
        float3 on = v3(nn[0], nn[1], nn[2]);
        if (xi >= 0) { r.p = xf_point(xf, op); on = xf_normal(xf + 3, on); } else r.p = op;
        set_face_normal(r, d, on);
        break;
    }
    default: {   // RT_OBJ_TRIANGLE
        float3 v0 = f4xyz(A), e1 = f4xyz(B), e2 = f4xyz(__ldg(&sc.primC[slot]));
        float tt, bu = 0.0f, bv = 0.0f;
        triangle_t(v0, e1, e2, o, d, t, t, tt, bu, bv);
        float3 n = v3(__fmaf_rn(e1.y, e2.z, -(e1.z * e2.y)), __fmaf_rn(e1.z, e2.x, -(e1.x * e2.z)), __fmaf_rn(e1.x, e2.y, -(e1.y * e2.x)));
        if (want_uv) {      // per-vertex texture coordinates (OBJ `vt`), 16-bit unorm pairs in the w components; without `vt` the
                            // vertices carry (0,0), (1,0), (0,1) and the sums below give (bu, bv)
            const unsigned q0 = __float_as_uint(A.w), q1 = __float_as_uint(B.w), q2 = __float_as_uint(__ldg(&sc.primC[slot]).w);
            const float w0 = (1.0f - bu) - bv;
            r.u = __fmaf_rn(bv, (float)(q2 & 0xffffu) / 65535.0f, __fmaf_rn(bu, (float)(q1 & 0xffffu) / 65535.0f, w0 * ((float)(q0 & 0xffffu) / 65535.0f)));
            r.v = __fmaf_rn(bv, (float)(q2 >> 16) / 65535.0f, __fmaf_rn(bu, (float)(q1 >> 16) / 65535.0f, w0 * ((float)(q0 >> 16) / 65535.0f)));
        }
        r.p = vfma(t, d, o);
        set_face_normal(r, d, vunit(n));
        break;
    }
    }
}
__device__ __forceinline__ void hit_record(const DScene &sc, int type, int slot, float3 o, float3 d, float t, bool want_uv, HitRec &r) {
    hit_record(sc, type, slot, o, d, t, want_uv, r, __ldg(&sc.primA[slot]), __ldg(&sc.primB[slot]));
}
