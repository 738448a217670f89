// rt_wavefront.cuh — queue layout shared by the wavefront kernels and the host-side pass loop.
#pragma once
#include "rt_shade.cuh"

// Checked build (tools/build_variant.py checked -DRT_CHECKED=1): device-side asserts on every queue / staging / stack index the
// kernels compute.  compute-sanitizer is closed on this GPU pool ("runs under it have left GPUs needing a reset"), so the small
// renders of tools/sanitize_run.py run against this build instead; log under profiles/.
#ifdef RT_CHECKED
#include <cassert>
#define RT_CHECK(cond) assert(cond)
#else
#define RT_CHECK(cond) ((void)0)
#endif
#define RT_NUM_BINS 5          // miss, lambertian, metal, dielectric, diffuse_light
#define RT_TRAV_STACK 64
#ifndef RT_SSTACK
#define RT_SSTACK 8            // traversal stack levels kept in shared memory by k_extend (rt_traverse.cuh); deeper ones: local memory
#endif
#ifndef RT_EXT_BLOCK
#define RT_EXT_BLOCK 512       // k_extend CTA size
#endif
#define RT_SSTACK_STRIDE (RT_EXT_BLOCK * 8)     // bytes between two levels of the shared-memory stack (one int2 per thread and level)
#define RT_SSTACK_BYTES (RT_SSTACK * RT_SSTACK_STRIDE)
#define RT_BLOCK 256

// event counters (rt_stats, SURVEY.md §8(d) flop model)
enum { CTR_BOX = 0, CTR_SPH_REJECT, CTR_SPH_ACCEPT, CTR_RECT, CTR_CYL_REJECT, CTR_CYL_ACCEPT, CTR_TRI, CTR_BOXPRIM,
       CTR_SHADE, CTR_GENERATE, CTR_NODE_VISITS, CTR_STACK_OVERFLOW, CTR_GEN_LEAF, CTR_COUNT = 16 };

// SoA ray queues: every per-ray field is one float4 array, read and written with coalesced 16-byte accesses.
struct Queues {
    float4 *ray_o[2];      // origin.xyz, path id (int bits)            — ping-pong between bounces
    float4 *ray_d[2];      // direction.xyz (unnormalised, as the reference keeps it), pixel index (Philox key word)
    float4 *thr[2];        // throughput rgb, sample index within the pass
    int2 *hits;            // (t bits, leaf code or -1) per queue entry, written by extend, read by shade
    float4 *radiance;      // per path id: the radiance this sample contributes (written exactly once; NEE: accumulated)
    int *qcount;           // [max_depth + 1] queue length per bounce
    int *workctr;          // [max_depth][2] dynamic work-fetch cursors (extend, shade)
    // shadow stage (RT_RENDER_NEE): occlusion rays written by shade, consumed by the extend kernel in any-hit mode (launch_shadow) right after it
    float4 *sh_o, *sh_d, *sh_c;   // origin.xyz + path id | direction y - x | radiance carried if unoccluded
    int *shcount;                 // [max_depth] shadow-queue length per bounce
    unsigned long long *counters;   // [CTR_COUNT]
    long long cap;
};
