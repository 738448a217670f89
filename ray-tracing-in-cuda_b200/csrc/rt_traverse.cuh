// rt_traverse.cuh — closest-hit traversal of the 2-wide LBVH (replaces hittable_list::hit, object.cuh:23-37).
//
// Result contract = the reference's linear walk: the accepted hit is the smallest root over all objects, and on an
// exact tie in t the object with the LARGER scene-file index wins ("last index wins": the walk tests
// `t_max < root`, so an equal root replaces the earlier one).  Box tests only prune; they are conservative
// (one FFMA per slab plane, t = plane * (1/d) - o/d, evaluated with ray constants that already carry the error margin
// |t| * 2^-20 + 2^-21 * |o/d| of their axis: near planes rounded down, far planes rounded up), so no hit the linear walk
// would accept can be culled — tests compare against the brute-force oracle and expect zero mismatching pixels.
// Leaf tests are the exact reference arithmetic (rt_scene.cuh).
//
// Two node layouts, one algorithm:
//   SIGNED = true   BvhNodeS (104 B): plane pairs stored (lo, hi, hi, lo) per box and axis; the ray's direction signs pick
//                   the 8-byte half that reads (near, far), so a box costs 6 FFMA + 2 three-input min/max + 2 min/max.
//                   Used when the whole tree is staged in shared memory.
//   SIGNED = false  BvhNode (64 B): both boxes in three float4; near/far planes picked by the sign predicates.  Half the bytes per visit —
//                   used when the tree is read from global memory (L1/L2), where bytes matter more than ALU slots.
#pragma once
#include "rt_wavefront.cuh"
#include <cfloat>

struct TravCounters { unsigned box, sph_rej, sph_acc, rect, cyl_rej, cyl_acc, tri, boxprim, nodes, overflow; };

template <bool COUNT>
__device__ __forceinline__ void count_leaf(TravCounters &c, int code, bool hit) {
    if (!COUNT) return;
    switch ((unsigned)code >> 28) {
    case RT_OBJ_SPHERE: if (hit) c.sph_acc++; else c.sph_rej++; break;
    case RT_OBJ_XY_RECT: case RT_OBJ_XZ_RECT: case RT_OBJ_YZ_RECT: c.rect++; break;
    case RT_OBJ_CYLINDER: if (hit) c.cyl_acc++; else c.cyl_rej++; break;
    case RT_OBJ_BOX: c.boxprim++; break;
    default: c.tri++; break;
    }
}

#define RT_BOX_REL 9.5367431640625e-07f      // 2^-20: reciprocal + fma rounding, relative to t
#define RT_BOX_ABS 4.76837158203125e-07f     // 2^-21 * |o/d| per axis: rounding of the hoisted o/d term and of the scaled copies
#define RT_BOX_OOD_MAX 1e18f                 // |1/d| clamp: beyond it the ray counts as parallel to the slab pair
#define RT_BOX_PAR 1e10f                     // absolute margin of a parallel axis (t units; larger than any scene distance)
#define RT_TRAV_DONE ((int)0x80000000)

// One exact leaf test with the closest-hit bookkeeping of the linear walk (object.cuh:23-37): a root equal to the current
// best replaces it when the object comes later in the scene file.  Returns true when the hit was taken.
template <bool COUNT, int ONLY = -1>
__device__ __forceinline__ bool leaf_visit(const DScene &sc, int code, float3 o, float3 d, float a, float inv_a, float tmin,
                                           float &best_t, int &best_code, TravCounters &cnt) {
    float t;
    const bool h = leaf_t<ONLY>(sc, code, o, d, a, inv_a, tmin, best_t, t);
    count_leaf<COUNT>(cnt, code, h);
    if (!h) return false;
    bool take = t < best_t || best_code < 0;
    if (!take)   // t == best_t: the later object in the scene file wins (object.cuh:29-33)
        take = __ldg(&sc.prim_obj[code & RT_LEAF_SLOT_MASK]) > __ldg(&sc.prim_obj[best_code & RT_LEAF_SLOT_MASK]);
    if (take) { best_t = t; best_code = code; }
    return take;
}

__device__ __forceinline__ float2 lds_f2(unsigned addr) {
    float2 v;
    asm("ld.shared.v2.f32 {%0, %1}, [%2];" : "=f"(v.x), "=f"(v.y) : "r"(addr));
    return v;
}
__device__ __forceinline__ void sts_i2(unsigned addr, int2 v) {
    asm volatile("st.shared.v2.s32 [%0], {%1, %2};" ::"r"(addr), "r"(v.x), "r"(v.y));
}
__device__ __forceinline__ int2 lds_i2(unsigned addr) {
    int2 v;
    asm("ld.shared.v2.s32 {%0, %1}, [%2];" : "=r"(v.x), "=r"(v.y) : "r"(addr));
    return v;
}
// stack accesses: volatile, so that the compiler keeps them in program order with respect to each other
__device__ __forceinline__ int2 lds_stack(unsigned addr) {
    int2 v;
    asm volatile("ld.shared.v2.s32 {%0, %1}, [%2];" : "=r"(v.x), "=r"(v.y) : "r"(addr));
    return v;
}

// nodes: BvhNodeS array in SHARED memory (SIGNED) or BvhNode array in global memory (!SIGNED).
// Returns best_t / best_code (-1 = miss).  ANY_HIT: stop at the first accepted hit (occlusion rays).
//
// Conservative slabs without per-node work: every plane is evaluated twice-over-constant, never with a per-node margin —
//   near planes with  (ood (1-r), oo (1-r) - c|oo|)   so that  t_near' <= exact t_near,
//   far  planes with  (ood (1+r), oo (1+r) + c|oo|)   so that  t_far'  >= exact t_far      (r = 2^-20, c = 2^-21),
// per axis (an axis the ray is almost parallel to has a huge |oo|; its own margin grows, the other axes keep theirs).
// |1/d| is clamped to 1e18; a clamped axis (d = 0 or denormal: the ray is parallel to that slab pair) gets an absolute
// margin of 1e10 instead, which keeps the slab pair as an inside/outside test without NaNs — before, 0 * inf dropped the
// slab and such a ray visited every node along its projection (one ray in ~10^4 of the centre column of a camera placed at
// x = 0: 15 000 node visits, a multi-millisecond tail on the 1 M-triangle mesh).
//
// Traversal stack: the first SD levels live in SHARED memory (sstack = shared-window address of this thread's column, one
// 8-byte entry per level, levels RT_SSTACK_STRIDE bytes apart: bank = thread, so any mix of levels across the lanes of a warp
// is conflict-free), deeper levels in the per-thread local-memory array.  ncu (r01f) had the pop loop at 17 % of the warp
// samples of k_extend, waiting on local-memory loads at ~6 active lanes; mean stack depth on the reference scenes is < SD.
// SD = 0 (k_aov, k_tile_lists callers): local memory only.
template <bool SIGNED, bool COUNT, bool ANY_HIT, int SD = 0, int ONLY = -1>
__device__ __forceinline__ void traverse(const DScene &sc, const char *__restrict__ nodes, float3 o, float3 d, float tmin,
                                         float tmax, float &best_t, int &best_code, TravCounters &cnt, unsigned sstack = 0) {
    best_t = tmax; best_code = -1;
    if (sc.n_prims == 0) return;
    const float a = dot3(d, d), inv_a = 1.0f / a;
    // giants first (rt_lbvh.cu:k_giants_mark): the few objects that span the scene are not in the tree; testing them up front also
    // hands the walk a tight best_t from its first node.  The tie rule does not depend on the order of the tests.
    for (int g = 0; g < sc.n_giants; ++g)
        if (leaf_visit<COUNT, ONLY>(sc, __ldg(&sc.giant_code[g]), o, d, a, inv_a, tmin, best_t, best_code, cnt) && ANY_HIT) return;
    if (sc.n_tree == 0) return;
    float3 ood = v3(__frcp_rn(d.x), __frcp_rn(d.y), __frcp_rn(d.z));
    const bool px = !(fabsf(ood.x) <= RT_BOX_OOD_MAX), py = !(fabsf(ood.y) <= RT_BOX_OOD_MAX), pz = !(fabsf(ood.z) <= RT_BOX_OOD_MAX);
    if (px) ood.x = copysignf(RT_BOX_OOD_MAX, d.x);
    if (py) ood.y = copysignf(RT_BOX_OOD_MAX, d.y);
    if (pz) ood.z = copysignf(RT_BOX_OOD_MAX, d.z);
    const float3 oo = v3(-(o.x * ood.x), -(o.y * ood.y), -(o.z * ood.z));
    const float3 ca = v3(__fmaf_rn(RT_BOX_ABS, fabsf(oo.x), px ? RT_BOX_PAR : 0.0f), __fmaf_rn(RT_BOX_ABS, fabsf(oo.y), py ? RT_BOX_PAR : 0.0f),
                         __fmaf_rn(RT_BOX_ABS, fabsf(oo.z), pz ? RT_BOX_PAR : 0.0f));
    const float GN = 1.0f - RT_BOX_REL, GF = 1.0f + RT_BOX_REL;
    const float3 oodn = v3(ood.x * GN, ood.y * GN, ood.z * GN), oodf = v3(ood.x * GF, ood.y * GF, ood.z * GF);
    const float3 oon = v3(__fmaf_rn(oo.x, GN, -ca.x), __fmaf_rn(oo.y, GN, -ca.y), __fmaf_rn(oo.z, GN, -ca.z));
    const float3 oof = v3(__fmaf_rn(oo.x, GF, ca.x), __fmaf_rn(oo.y, GF, ca.y), __fmaf_rn(oo.z, GF, ca.z));
    const bool ngx = __float_as_int(d.x) < 0, ngy = __float_as_int(d.y) < 0, ngz = __float_as_int(d.z) < 0;
    // SIGNED: 8 where the ray travels in the negative axis direction selects the (near, far) half of a plane quadruple
    unsigned al = 96, ax = ngx ? 8 : 0, ay = ngy ? 8 : 0, az = ngz ? 8 : 0;
    if (SIGNED) {
        const unsigned sbase = (unsigned)__cvta_generic_to_shared(nodes);
        ax += sbase; ay += sbase + 16; az += sbase + 32; al += sbase;
        asm("" : "+r"(al));        // opaque: keeps the compiler from re-deriving the shared window base every iteration
    }
    int2 stack[RT_TRAV_STACK];
    int sp = 0, cur = 0;
    // next subtree that can still contain a closer (or equally close) hit; none left: done
#define RT_POP()                                                                 \
    for (;;) {                                                                   \
        if (sp == 0) { cur = RT_TRAV_DONE; break; }                              \
        --sp;                                                                    \
        int2 e_;                                                                 \
        if (SD > 0 && sp < SD) e_ = lds_stack(sstack + (unsigned)sp * RT_SSTACK_STRIDE); \
        else e_ = stack[sp - SD];                                                \
        if (__int_as_float(e_.y) <= best_t) { cur = e_.x; break; }               \
    }
    while (cur != RT_TRAV_DONE) {
        // leaf step first, inner step second: a lane that holds a leaf tests it, pops, and still takes part in this
        // iteration's inner-node step (a diverged warp executes both halves anyway)
        if (cur < 0) {
            if (leaf_visit<COUNT, ONLY>(sc, ~cur, o, d, a, inv_a, tmin, best_t, best_code, cnt) && ANY_HIT) return;
            RT_POP()
        }
        // "while-while": after the leaf step every lane descends to its NEXT leaf (or to the end of its walk) before any lane tests a
        // leaf again, so the leaf step above runs with more lanes (RT_TRAV_DONE is negative: it ends this loop too).  Against the
        // one-step "if-if" form (RT_IF_IF tuning builds): RTIOW extend 21.7 -> 20.7 ms per 128 spp, blue2 16.8 -> 16.0, sample_scene equal.
#ifdef RT_IF_IF
        if (cur >= 0) {
#else
        while (cur >= 0) {
#endif
            float tnl, tfl, tnr, tfr;
            int2 nd;
            if (SIGNED) {
                const unsigned c = (unsigned)cur;                 // byte offset of the node
                const float2 lx = lds_f2(ax + c), ly = lds_f2(ay + c), lz = lds_f2(az + c);
                const float2 rx = lds_f2(ax + c + 48), ry = lds_f2(ay + c + 48), rz = lds_f2(az + c + 48);
                nd = lds_i2(al + c);
                tnl = fmaxf(fmaxf(__fmaf_rn(lx.x, oodn.x, oon.x), __fmaf_rn(ly.x, oodn.y, oon.y)), fmaxf(__fmaf_rn(lz.x, oodn.z, oon.z), tmin));
                tfl = fminf(fminf(__fmaf_rn(lx.y, oodf.x, oof.x), __fmaf_rn(ly.y, oodf.y, oof.y)), fminf(__fmaf_rn(lz.y, oodf.z, oof.z), best_t));
                tnr = fmaxf(fmaxf(__fmaf_rn(rx.x, oodn.x, oon.x), __fmaf_rn(ry.x, oodn.y, oon.y)), fmaxf(__fmaf_rn(rz.x, oodn.z, oon.z), tmin));
                tfr = fminf(fminf(__fmaf_rn(rx.y, oodf.x, oof.x), __fmaf_rn(ry.y, oodf.y, oof.y)), fminf(__fmaf_rn(rz.y, oodf.z, oof.z), best_t));
            } else {
                const BvhNode *nb = reinterpret_cast<const BvhNode *>(nodes) + cur;
                const float4 na = nb->a, nbb = nb->b, nc = nb->c;
                const int4 nd4 = nb->d;
                nd = make_int2(nd4.x, nd4.y);
                // left box: min (na.x na.y na.z) max (na.w nbb.x nbb.y); right box: min (nbb.z nbb.w nc.x) max (nc.y nc.z nc.w);
                // the direction signs pick the (near, far) plane of every pair
                tnl = fmaxf(fmaxf(__fmaf_rn(ngx ? na.w : na.x, oodn.x, oon.x), __fmaf_rn(ngy ? nbb.x : na.y, oodn.y, oon.y)),
                            fmaxf(__fmaf_rn(ngz ? nbb.y : na.z, oodn.z, oon.z), tmin));
                tfl = fminf(fminf(__fmaf_rn(ngx ? na.x : na.w, oodf.x, oof.x), __fmaf_rn(ngy ? na.y : nbb.x, oodf.y, oof.y)),
                            fminf(__fmaf_rn(ngz ? na.z : nbb.y, oodf.z, oof.z), best_t));
                tnr = fmaxf(fmaxf(__fmaf_rn(ngx ? nc.y : nbb.z, oodn.x, oon.x), __fmaf_rn(ngy ? nc.z : nbb.w, oodn.y, oon.y)),
                            fmaxf(__fmaf_rn(ngz ? nc.w : nc.x, oodn.z, oon.z), tmin));
                tfr = fminf(fminf(__fmaf_rn(ngx ? nbb.z : nc.y, oodf.x, oof.x), __fmaf_rn(ngy ? nbb.w : nc.z, oodf.y, oof.y)),
                            fminf(__fmaf_rn(ngz ? nc.x : nc.w, oodf.z, oof.z), best_t));
            }
            if (COUNT) { cnt.box += 2; cnt.nodes++; }
            const bool hl = tnl <= tfl;
            const bool hr = tnr <= tfr;
            if (hl && hr) {
                const bool lfirst = tnl <= tnr;
                const int2 far_ = make_int2(lfirst ? nd.y : nd.x, __float_as_int(lfirst ? tnr : tnl));
                RT_CHECK(sp >= 0 && sp - SD < RT_TRAV_STACK);
                if (SD > 0 && sp < SD) { sts_i2(sstack + (unsigned)sp * RT_SSTACK_STRIDE, far_); ++sp; }
                else if (sp - SD < RT_TRAV_STACK) { stack[sp - SD] = far_; ++sp; }
                else if (COUNT) cnt.overflow++;
                cur = lfirst ? nd.x : nd.y;
            } else if (hl || hr) {
                cur = hl ? nd.x : nd.y;
            } else {
                RT_POP()
            }
        }
    }
#undef RT_POP
}

// ---- four-wide traversal for trees read from global memory (rt_scene.cuh: BvhNode4) ----
// Same result contract and the same conservative plane arithmetic as traverse(); a visit fetches one 128-byte node (the ray's
// direction signs pick the near / far plane of every pair as a whole 16-byte load), tests four boxes, orders the children it
// hits with a five-comparator network and pushes all but the nearest (farthest first).
__device__ __forceinline__ void cswap(float &ta, int &la, float &tb, int &lb) {
    const bool sw = tb < ta;
    const float t0 = sw ? tb : ta, t1 = sw ? ta : tb;
    const int l0 = sw ? lb : la, l1 = sw ? la : lb;
    ta = t0; tb = t1; la = l0; lb = l1;
}

template <bool COUNT, bool ANY_HIT, int SD = 0, int ONLY = -1>
__device__ __forceinline__ void traverse4(const DScene &sc, const BvhNode4 *__restrict__ nodes, float3 o, float3 d, float tmin, float tmax,
                                          float &best_t, int &best_code, TravCounters &cnt, unsigned sstack = 0) {
    best_t = tmax; best_code = -1;
    if (sc.n_prims == 0) return;
    const float a = dot3(d, d), inv_a = 1.0f / a;
    for (int g = 0; g < sc.n_giants; ++g)
        if (leaf_visit<COUNT, ONLY>(sc, __ldg(&sc.giant_code[g]), o, d, a, inv_a, tmin, best_t, best_code, cnt) && ANY_HIT) return;
    if (sc.n_tree == 0) return;
    float3 ood = v3(__frcp_rn(d.x), __frcp_rn(d.y), __frcp_rn(d.z));
    const bool px = !(fabsf(ood.x) <= RT_BOX_OOD_MAX), py = !(fabsf(ood.y) <= RT_BOX_OOD_MAX), pz = !(fabsf(ood.z) <= RT_BOX_OOD_MAX);
    if (px) ood.x = copysignf(RT_BOX_OOD_MAX, d.x);
    if (py) ood.y = copysignf(RT_BOX_OOD_MAX, d.y);
    if (pz) ood.z = copysignf(RT_BOX_OOD_MAX, d.z);
    const float3 oo = v3(-(o.x * ood.x), -(o.y * ood.y), -(o.z * ood.z));
    const float3 ca = v3(__fmaf_rn(RT_BOX_ABS, fabsf(oo.x), px ? RT_BOX_PAR : 0.0f), __fmaf_rn(RT_BOX_ABS, fabsf(oo.y), py ? RT_BOX_PAR : 0.0f),
                         __fmaf_rn(RT_BOX_ABS, fabsf(oo.z), pz ? RT_BOX_PAR : 0.0f));
    const float GN = 1.0f - RT_BOX_REL, GF = 1.0f + RT_BOX_REL;
    const float3 oodn = v3(ood.x * GN, ood.y * GN, ood.z * GN), oodf = v3(ood.x * GF, ood.y * GF, ood.z * GF);
    const float3 oon = v3(__fmaf_rn(oo.x, GN, -ca.x), __fmaf_rn(oo.y, GN, -ca.y), __fmaf_rn(oo.z, GN, -ca.z));
    const float3 oof = v3(__fmaf_rn(oo.x, GF, ca.x), __fmaf_rn(oo.y, GF, ca.y), __fmaf_rn(oo.z, GF, ca.z));
    // plane indices within the node: near = lo (0..2) for a ray travelling in +axis, hi (3..5) otherwise; far the other one
    const int nx = __float_as_int(d.x) < 0 ? 3 : 0, ny = __float_as_int(d.y) < 0 ? 4 : 1, nz = __float_as_int(d.z) < 0 ? 5 : 2;
    const int fx = 3 - nx, fy = 5 - ny, fz = 7 - nz;
    int2 stack[RT_TRAV_STACK];
    int sp = 0, cur = 0;
#define RT_POP4()                                                                \
    for (;;) {                                                                   \
        if (sp == 0) { cur = RT_TRAV_DONE; break; }                              \
        --sp;                                                                    \
        int2 e_;                                                                 \
        if (SD > 0 && sp < SD) e_ = lds_stack(sstack + (unsigned)sp * RT_SSTACK_STRIDE); \
        else e_ = stack[sp - SD];                                                \
        if (__int_as_float(e_.y) <= best_t) { cur = e_.x; break; }               \
    }
#define RT_PUSH4(L_, T_)                                                         \
    {                                                                            \
        const int2 e_ = make_int2(L_, __float_as_int(T_));                       \
        RT_CHECK(sp >= 0 && sp - SD < RT_TRAV_STACK);                            \
        if (SD > 0 && sp < SD) { sts_i2(sstack + (unsigned)sp * RT_SSTACK_STRIDE, e_); ++sp; } \
        else if (sp - SD < RT_TRAV_STACK) { stack[sp - SD] = e_; ++sp; }         \
        else if (COUNT) cnt.overflow++;                                          \
    }
    while (cur != RT_TRAV_DONE) {
        if (cur < 0) {
            if (leaf_visit<COUNT, ONLY>(sc, ~cur, o, d, a, inv_a, tmin, best_t, best_code, cnt) && ANY_HIT) return;
            RT_POP4()
        }
        if (cur >= 0) {     // one step per iteration: the while-while form of traverse() measured 37 % SLOWER here (latency-bound node fetches:
                            // lanes waiting at a leaf keep no loads in flight)
            const float4 *__restrict__ nd = reinterpret_cast<const float4 *>(nodes + cur);
            const float4 pnx = nd[nx], pny = nd[ny], pnz = nd[nz], pfx = nd[fx], pfy = nd[fy], pfz = nd[fz];
            const int4 lk = *reinterpret_cast<const int4 *>(nd + 6);
            float t0, t1, t2, t3;
#define RT_BOX4(C_, T_)                                                                                                             \
            {                                                                                                                       \
                const float tn_ = fmaxf(fmaxf(__fmaf_rn(pnx.C_, oodn.x, oon.x), __fmaf_rn(pny.C_, oodn.y, oon.y)),                   \
                                        fmaxf(__fmaf_rn(pnz.C_, oodn.z, oon.z), tmin));                                              \
                const float tf_ = fminf(fminf(__fmaf_rn(pfx.C_, oodf.x, oof.x), __fmaf_rn(pfy.C_, oodf.y, oof.y)),                   \
                                        fminf(__fmaf_rn(pfz.C_, oodf.z, oof.z), best_t));                                            \
                T_ = tn_ <= tf_ ? tn_ : INFINITY;                                                                                   \
            }
            RT_BOX4(x, t0) RT_BOX4(y, t1) RT_BOX4(z, t2) RT_BOX4(w, t3)
#undef RT_BOX4
            if (COUNT) { cnt.box += 4; cnt.nodes++; }
            int l0 = lk.x, l1 = lk.y, l2 = lk.z, l3 = lk.w;
            // full order of the four entry distances (singling out only the nearest and pushing the rest in slot order measured equal:
            // 65.6 vs 65.5 ms — the walk is not bound by this network)
            cswap(t0, l0, t1, l1); cswap(t2, l2, t3, l3); cswap(t0, l0, t2, l2); cswap(t1, l1, t3, l3); cswap(t1, l1, t2, l2);
            if (t0 == INFINITY) { RT_POP4() }
            else {
                if (t3 != INFINITY) RT_PUSH4(l3, t3)
                if (t2 != INFINITY) RT_PUSH4(l2, t2)
                if (t1 != INFINITY) RT_PUSH4(l1, t1)
                cur = l0;
            }
        }
    }
#undef RT_POP4
#undef RT_PUSH4
}
