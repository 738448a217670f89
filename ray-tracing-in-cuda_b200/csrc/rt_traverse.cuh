// rt_traverse.cuh — closest-hit traversal of the 2-wide LBVH (replaces hittable_list::hit, object.cuh:23-37).
//
// Result contract = the reference's linear walk: the accepted hit is the smallest root over all objects, and on an
// exact tie in t the object with the LARGER scene-file index wins ("last index wins": the walk tests
// `t_max < root`, so an equal root replaces the earlier one).  Box tests only prune; they are made
// conservative (far plane inflated by 4 ulp, stack-cull deflated) so that no hit the linear walk would
// accept can be culled — tests compare against the brute-force oracle and expect zero mismatching pixels.
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

// nodes: the node array to read (shared-memory copy of the whole tree when it fits, else global memory).
// Returns best_t / best_code (-1 = miss).  ANY_HIT: stop at the first accepted hit (shadow rays).
template <bool COUNT, bool ANY_HIT>
__device__ __forceinline__ void traverse(const DScene &sc, const BvhNode *__restrict__ nodes, float3 o, float3 d, float tmin,
                                         float tmax, float &best_t, int &best_code, TravCounters &cnt) {
    best_t = tmax; best_code = -1;
    if (sc.n_prims == 0) return;
    const float3 ood = v3(1.0f / d.x, 1.0f / d.y, 1.0f / d.z);
    const float a = dot3(d, d), inv_a = 1.0f / a;
    int stack_node[RT_TRAV_STACK];
    float stack_t[RT_TRAV_STACK];
    int sp = 0;
    int cur = 0;
    for (;;) {
        if (cur >= 0) {
            const float4 na = nodes[cur].a, nb = nodes[cur].b, nc = nodes[cur].c;
            const int4 nd = nodes[cur].d;
            if (COUNT) { cnt.box += 2; cnt.nodes++; }
            // left box: min (na.x na.y na.z) max (na.w nb.x nb.y); right box: min (nb.z nb.w nc.x) max (nc.y nc.z nc.w)
            float x0 = (na.x - o.x) * ood.x, x1 = (na.w - o.x) * ood.x;
            float y0 = (na.y - o.y) * ood.y, y1 = (nb.x - o.y) * ood.y;
            float z0 = (na.z - o.z) * ood.z, z1 = (nb.y - o.z) * ood.z;
            float tnl = fmaxf(fmaxf(fminf(x0, x1), fminf(y0, y1)), fmaxf(fminf(z0, z1), tmin));
            float tfl = fminf(fminf(fmaxf(x0, x1), fmaxf(y0, y1)), fminf(fmaxf(z0, z1), best_t));
            x0 = (nb.z - o.x) * ood.x; x1 = (nc.y - o.x) * ood.x;
            y0 = (nb.w - o.y) * ood.y; y1 = (nc.z - o.y) * ood.y;
            z0 = (nc.x - o.z) * ood.z; z1 = (nc.w - o.z) * ood.z;
            float tnr = fmaxf(fmaxf(fminf(x0, x1), fminf(y0, y1)), fmaxf(fminf(z0, z1), tmin));
            float tfr = fminf(fminf(fmaxf(x0, x1), fmaxf(y0, y1)), fminf(fmaxf(z0, z1), best_t));
            const bool hl = tnl <= tfl * 1.0000005f, hr = tnr <= tfr * 1.0000005f;
            if (hl && hr) {
                const bool lfirst = tnl <= tnr;
                const int nearc = lfirst ? nd.x : nd.y, farc = lfirst ? nd.y : nd.x;
                if (sp < RT_TRAV_STACK) { stack_node[sp] = farc; stack_t[sp] = lfirst ? tnr : tnl; ++sp; }
                else if (COUNT) cnt.overflow++;
                cur = nearc;
                continue;
            }
            if (hl) { cur = nd.x; continue; }
            if (hr) { cur = nd.y; continue; }
        } else {
            const int code = ~cur;
            float t;
            const bool h = leaf_t(sc, code, o, d, a, inv_a, tmin, best_t, t);
            count_leaf<COUNT>(cnt, code, h);
            if (h) {
                bool take = t < best_t || best_code < 0;
                if (!take) {   // t == best_t: the later object in the scene file wins (object.cuh:29-33)
                    take = __ldg(&sc.prim_obj[code & RT_LEAF_SLOT_MASK]) > __ldg(&sc.prim_obj[best_code & RT_LEAF_SLOT_MASK]);
                }
                if (take) { best_t = t; best_code = code; if (ANY_HIT) return; }
            }
        }
        // pop, skipping subtrees that start beyond the current closest hit
        for (;;) {
            if (sp == 0) return;
            --sp;
            cur = stack_node[sp];
            if (stack_t[sp] * 0.9999995f <= best_t) break;
        }
    }
}
