// rt_kernels.cu — the wavefront path-tracing kernels for sm_100a:
//   tile_lists (once per scene) ; generate (+ bounce-0 hits) -> [ extend (LBVH traversal) -> shade (+ stream compaction)
//   (-> shadow) ] x max_depth (from bounce 8 — 2 on all-sphere scenes — one tail launch finishes the surviving paths) -> accumulate (per sample chunk)
//   -> fold chunks ; tonemap
// They replace the reference's megakernel (gpu-version/main.cu:72-105 `render` + 17-70 `ray_color`), which runs one
// thread per pixel, serial over samples, with virtual dispatch through device-heap pointers.
//
// Design for B200 (148 SMs, 228 KB smem/SM, 126 MB L2, 180 GB HBM3e):
//   * persistent grids (SM count x resident CTAs); extend warps pull batches from a device-side cursor, shade CTAs own
//     256-entry chunks — queue lengths never travel to the host, nearly-empty late bounces cost one tiny launch each;
//   * SoA float4 ray queues (coalesced 16-byte accesses / TMA bulk copies), hit records of 8 bytes, 256 M paths per pass and stream;
//   * the whole BVH is staged in shared memory in a sign-ordered 104-byte layout when tree + stack fit 112 KB (RTIOW: 484 nodes = 50 KB),
//     else four-wide 128-byte nodes are read via L1/L2; objects that span the scene ("giants") stay out of the tree and are tested first; the first traversal-stack levels live in shared memory too; primary rays skip
//     the tree through per-tile camera-beam candidate lists, and tiles whose beam touches nothing skip the queues altogether;
//   * shade: TMA bulk loads/stores through shared staging, ballot/popc compaction, one global atomic per chunk, optional
//     in-CTA sort by material class (RT_RENDER_SORT);
//   * Philox4x32-10 keyed (pixel, sample, bounce): no RNG state in memory; camera paths carry no throughput record;
//   * radiance is written once per path and summed per pixel in sample order by `accumulate` (no float atomics).
#include "rt_kernels.h"
#include "rt_traverse.cuh"
#include <cstdlib>
#include <mutex>

__device__ __forceinline__ unsigned lanemask_lt() { unsigned m; asm("mov.u32 %0, %%lanemask_lt;" : "=r"(m)); return m; }
__device__ __forceinline__ int lane_id() { return threadIdx.x & 31; }

// One warp fetches `chunk` consecutive work items from a device cursor; returns the first index to all lanes.
__device__ __forceinline__ int warp_fetch(int *cursor, int chunk) {
    int base = 0;
    if (lane_id() == 0) base = atomicAdd(cursor, chunk);
    return __shfl_sync(0xffffffffu, base, 0);
}

template <bool COUNT>
__device__ __forceinline__ void flush_counter(unsigned long long *counters, int which, unsigned v) {
    if (!COUNT) return;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if (lane_id() == 0 && v) atomicAdd(&counters[which], (unsigned long long)v);
}

// ------------------------------------------------------------------------------------------------
// pass_init: zero the per-bounce cursors/counters of a pass and publish the length of the bounce-0 queue
// ------------------------------------------------------------------------------------------------
__global__ void k_pass_init(Queues q, int max_depth, int n_queue0) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i <= max_depth) q.qcount[i] = i == 0 ? n_queue0 : 0;
    if (i < max_depth * 2) q.workctr[i] = 0;
    if (q.shcount && i < max_depth) q.shcount[i] = 0;
}

// ------------------------------------------------------------------------------------------------
// generate: one camera path per (pixel, sample) — main.cu:95-99.  Path id = s_local * n_pixels + pixel, so that
// `accumulate` reads the radiance of one sample index as one contiguous, coalesced plane.
// ------------------------------------------------------------------------------------------------
// HITS: bounce 0 needs no traversal.  All samples of the 32 pixels of a tile leave the camera inside one thin beam, and
// k_tile_lists has already found every leaf that beam can touch (2.5 on average for the RTIOW scene, never more than 8),
// so the thread tests those few leaves with the exact leaf arithmetic right here and writes the bounce-0 hit record next
// to the ray: the bounce-0 extend launch disappears.  Paths it cannot cover (margin pixels outside the 8x4 tiling, tiles
// with more than RT_TILE_K candidates) get the RT_HIT_TODO mark and are traversed by a fix-up extend launch.
#define RT_HIT_TODO ((int)0x80000001)
template <bool HITS, bool COUNT, bool DIRECT, int ONLY = -1>      // ONLY = RT_OBJ_SPHERE: all-sphere scene, candidate tests without the type dispatch
__global__ void __launch_bounds__(RT_BLOCK) k_generate(DScene sc, Queues q, unsigned seed, int spp_begin, int n_paths, int zero_radiance) {
    // one thread per (tile-ordered) pixel slot and group of samples: the pixel arithmetic is done once, the sample planes are
    // written with stride n_pixels (each plane coalesced across the warp).  The grid covers every pixel slot in x.
    const int np = sc.width * sc.height, n_s = n_paths / np;
    const float div_w = (float)(sc.width - 1), div_h = (float)(sc.height - 1);      // divisors of camera.cuh's (i + r) / (W - 1)
    const int groups = gridDim.y;
    const int n_tiled = (sc.width & ~7) * (sc.height & ~3);
    TravCounters cnt = {};
    // HITS: a tile whose beam touches no leaf box at all (tile_cnt == 0: sky, backdrop) has nothing but misses.  Its paths
    // never enter the queues: with a constant background nothing is written here (k_accumulate adds the constant, sample by
    // sample), with the sky gradient only the radiance record is written.  k_shade skips such entries of the bounce-0 queue
    // (whole chunks through sc.tile_empty_prefix), so their ray / hit records are never read either.  One warp = one tile.
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    int pixel = 0, x = 0, y = 0, n_cand = -1;
    const int *cand = nullptr;
    if (k < np) {
        pixel = plane_to_pixel(sc.width, sc.height, k);
        y = pixel / sc.width; x = pixel - y * sc.width;
        if (HITS && k < n_tiled) { n_cand = __ldg(&sc.tile_cnt[k >> 5]); cand = sc.tile_list + (size_t)(k >> 5) * RT_TILE_K; }
    }
    const bool empty_tile = HITS && n_cand == 0 && sc.tile_empty_prefix != nullptr, gradient = sc.bg_mode == RT_BG_SKY_GRADIENT;
    // nothing to do at all: pixel slots past the image, and empty tiles under a constant background when no fix-up extend runs
    const bool idle = k >= np || (empty_tile && !gradient && sc.tile_mode == 1);
    for (int s_local = blockIdx.y; s_local < n_s && !idle; s_local += groups) {
        const int i = s_local * np + k;
        if (empty_tile && sc.tile_mode == 2) q.hits[i] = make_int2(__float_as_int(FLT_MAX), -1);   // the fix-up extend scans every hit record
        if (empty_tile && !gradient) continue;
        Rng rng; rng.init(seed, (unsigned)pixel, (unsigned)(spp_begin + s_local), 0u, sc.sampling);
        const float u = ((float)x + rng.next()) / div_w;
        const float v = ((float)y + rng.next()) / div_h;
        float3 o, d;
        camera_ray<DIRECT>(sc.cam, u, v, sc.defocus != 0, rng, o, d);
        if (empty_tile) {             // main.cu:61-63 with throughput 1
            const float3 bg = background(sc, d);
            q.radiance[i] = make_float4(bg.x, bg.y, bg.z, 0.0f);
            continue;
        }
        RT_CHECK(i >= 0 && (long long)i < q.cap);
        q.ray_o[0][i] = make_float4(o.x, o.y, o.z, __int_as_float(i));
        if (zero_radiance) q.radiance[i] = make_float4(0.0f, 0.0f, 0.0f, 0.0f);      // NEE: several stages add to it
        q.ray_d[0][i] = make_float4(d.x, d.y, d.z, __int_as_float(pixel));        // the Philox key word travels with the path
        // no throughput record at bounce 0: it is (1, 1, 1) and the sample index is path id / n_pixels — k_shade knows both
        if (HITS) {
            float best_t = FLT_MAX; int best_code = RT_HIT_TODO;
            if (n_cand >= 0) {
                best_code = -1;
                const float a = dot3(d, d), inv_a = 1.0f / a;
                for (int c = 0; c < n_cand; ++c) leaf_visit<COUNT, ONLY>(sc, __ldg(&cand[c]), o, d, a, inv_a, 0.001f, best_t, best_code, cnt);
            }
            q.hits[i] = make_int2(__float_as_int(best_t), best_code);
        }
    }
    if (HITS && COUNT)       // this stage's own leaf tests (one counter: they are not extend-stage work)
        flush_counter<COUNT>(q.counters, CTR_GEN_LEAF, cnt.sph_rej + cnt.sph_acc + cnt.rect + cnt.cyl_rej + cnt.cyl_acc + cnt.tri + cnt.boxprim);
}

// ------------------------------------------------------------------------------------------------
// tile_lists: per 8x4 pixel tile, the leaves whose box the tile's camera beam can touch.  The beam = every ray the camera
// can produce for the tile's pixels (jitter included): directions affine in (u, v), so their componentwise range over the
// tile is the range over its four corners; with a thin lens the origin ranges over the box of the lens disk and the
// direction range widens by the same offsets (o and d taken as independent ranges: conservative).  Interval slab test per axis:
//   t_near(d) >= min over the range, t_far(d) <= max over the range  =>  "max of the minima <= min of the maxima" holds for
// every box a single ray of the beam hits (necessary condition; boxes are never missed, some extra ones are listed).
// One thread per tile, plain stack traversal of the 64-byte nodes; runs once per uploaded scene / size override.
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ bool beam_box(float3 ol, float3 oh, float3 dl, float3 dh, float3 lo, float3 hi) {
    float tn = 0.0f, tf = INFINITY;
    const float ovl[3] = {ol.x, ol.y, ol.z}, ovh[3] = {oh.x, oh.y, oh.z}, l[3] = {dl.x, dl.y, dl.z}, h[3] = {dh.x, dh.y, dh.z};
    const float bl[3] = {lo.x, lo.y, lo.z}, bh[3] = {hi.x, hi.y, hi.z};
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        const float A = bl[k] - ovh[k], B = bh[k] - ovl[k];      // smallest lo - o and largest hi - o over the origin range
        float n, f;
        if (l[k] > 0.0f) { n = fminf(A / l[k], A / h[k]); f = fmaxf(B / l[k], B / h[k]); }
        else if (h[k] < 0.0f) { n = fminf(B / l[k], B / h[k]); f = fmaxf(A / l[k], A / h[k]); }
        else if (A <= 0.0f && B >= 0.0f) { n = -INFINITY; f = INFINITY; }
        else if (A > 0.0f) { n = h[k] > 0.0f ? A / h[k] : INFINITY; f = INFINITY; }
        else { n = l[k] < 0.0f ? B / l[k] : INFINITY; f = INFINITY; }
        // rounding of the subtractions and divisions: widen by 2^-16 relative (and a small absolute term)
        n = n - fabsf(n) * 1.52587890625e-05f - 1e-6f; f = f + fabsf(f) * 1.52587890625e-05f + 1e-6f;
        if (n == n) tn = fmaxf(tn, n);         // NaN (0 * inf, inf - inf): no constraint from this axis
        if (f == f) tf = fminf(tf, f);
    }
    return tn <= tf;
}

__global__ void __launch_bounds__(128) k_tile_lists(DScene sc, int *__restrict__ list, int *__restrict__ tile_cnt, int *__restrict__ n_bad) {
    const int tiles_x = (sc.width & ~7) >> 3, n_tiles = tiles_x * ((sc.height & ~3) >> 2);
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n_tiles) return;
    const int ty = t / tiles_x, tx = t - ty * tiles_x;
    const DCamera &cam = sc.cam;
    // u = (x + xi) / (W - 1) with xi in [0, 1]: the tile spans x in [8 tx, 8 tx + 8]; a hair wider for the rounding of u, v
    const float u0 = (float)(tx * 8) / (float)(sc.width - 1) - 1e-6f, u1 = (float)(tx * 8 + 8) / (float)(sc.width - 1) + 1e-6f;
    const float v0 = (float)(ty * 4) / (float)(sc.height - 1) - 1e-6f, v1 = (float)(ty * 4 + 4) / (float)(sc.height - 1) + 1e-6f;
    float3 dl = v3(INFINITY, INFINITY, INFINITY), dh = v3(-INFINITY, -INFINITY, -INFINITY);
#pragma unroll
    for (int c = 0; c < 4; ++c) {
        const float u = (c & 1) ? u1 : u0, v = (c & 2) ? v1 : v0;
        const float3 d = vfma(v, cam.vertical, vfma(u, cam.horizontal, cam.llc)) - cam.origin;
        dl = v3(fminf(dl.x, d.x), fminf(dl.y, d.y), fminf(dl.z, d.z)); dh = v3(fmaxf(dh.x, d.x), fmaxf(dh.y, d.y), fmaxf(dh.z, d.z));
    }
    // rounding of camera_ray's own evaluation (three fma / subtractions on terms of this size)
    const float3 pad = v3(1e-5f * (fabsf(cam.llc.x) + fabsf(cam.horizontal.x) + fabsf(cam.vertical.x) + fabsf(cam.origin.x)) + 1e-30f,
                          1e-5f * (fabsf(cam.llc.y) + fabsf(cam.horizontal.y) + fabsf(cam.vertical.y) + fabsf(cam.origin.y)) + 1e-30f,
                          1e-5f * (fabsf(cam.llc.z) + fabsf(cam.horizontal.z) + fabsf(cam.vertical.z) + fabsf(cam.origin.z)) + 1e-30f);
    dl = dl - pad; dh = dh + pad;
    float3 ol = cam.origin, oh = cam.origin;
    if (sc.defocus) {      // camera_ray: off = u rx + v ry with (rx, ry) in the lens disk; o = origin + off, d = d0 - off
        const float r = cam.lens_radius * 1.00001f;
        const float3 lens = v3(r * (fabsf(cam.u.x) + fabsf(cam.v.x)), r * (fabsf(cam.u.y) + fabsf(cam.v.y)), r * (fabsf(cam.u.z) + fabsf(cam.v.z)));
        ol = ol - lens; oh = oh + lens; dl = dl - lens; dh = dh + lens;
    }
    int stack[64];
    int sp = 0, n = 0;
    bool bad = false;
    int *out = list + (size_t)t * RT_TILE_K;
    for (int g = 0; g < sc.n_giants; ++g) {       // the giants are not in the tree: their padded boxes sit in the last leaf slots
        const float *bx = sc.leaf_aabb + 6 * (size_t)(sc.n_tree + g);
        if (beam_box(ol, oh, dl, dh, v3(bx[0], bx[1], bx[2]), v3(bx[3], bx[4], bx[5]))) out[n++] = __ldg(&sc.giant_code[g]);
    }
    if (sc.n_tree > 0) stack[sp++] = 0;
    while (sp > 0 && !bad) {
        const BvhNode nd = sc.nodes[stack[--sp]];
#pragma unroll
        for (int side = 0; side < 2; ++side) {
            const float3 lo = side ? v3(nd.b.z, nd.b.w, nd.c.x) : v3(nd.a.x, nd.a.y, nd.a.z);
            const float3 hi = side ? v3(nd.c.y, nd.c.z, nd.c.w) : v3(nd.a.w, nd.b.x, nd.b.y);
            const int child = side ? nd.d.y : nd.d.x;
            if (!beam_box(ol, oh, dl, dh, lo, hi)) continue;
            RT_CHECK(n <= RT_TILE_K && sp <= 64);
            if (child < 0) { if (n < RT_TILE_K) out[n++] = ~child; else bad = true; }
            else if (sp < 64) stack[sp++] = child;
            else bad = true;
        }
    }
    tile_cnt[t] = bad ? -1 : n;
    if (bad) atomicAdd(n_bad, 1);
}

// ------------------------------------------------------------------------------------------------
// extend: closest hit for every ray of the bounce's queue (replaces hittable_list::hit, object.cuh:23-37).
//
// Persistent grid; a warp takes a private batch of consecutive queue indices with one global atomic (the batch size adapts
// to the queue length, so short late-bounce queues still spread over every SM; CTAs that cannot get work leave before
// staging the tree).  The whole tree is staged in shared memory in the sign-ordered layout when it fits, else the compact
// 64-byte nodes are read through L1/L2.  Traversal itself: rt_traverse.cuh.  Output: 8 B per ray (t, leaf code).
// Measured and rejected: per-lane refill of finished lanes, postponed/batched leaf tests, a 4-wide collapsed
// tree with leaf ranges, binning rays by material here through global lists.
// ------------------------------------------------------------------------------------------------
#define RT_EXT_BATCH 256

// SHADOW = true: the same kernel walks the bounce's occlusion-ray queue (RT_RENDER_NEE) with an any-hit traversal on
// t in [0.001, 0.999] of the unnormalised segment; an unoccluded sample adds its radiance to the path's slot (one entry per
// path and bounce, so the add needs no atomic).
// max_batch < 0 (bounce 0 after a generate launch that resolved most hits itself): only entries carrying RT_HIT_TODO.
// ONLY: RT_OBJ_SPHERE when every primitive of the scene is a sphere (no leaf dispatch in the loop), else -1.
#ifdef RT_EXT_MINCTAS          // tuning builds: resident CTAs per SM the register allocation aims at
#define RT_EXT_BOUNDS __launch_bounds__(RT_EXT_BLOCK, RT_EXT_MINCTAS)
#else
#define RT_EXT_BOUNDS __launch_bounds__(RT_EXT_BLOCK)
#endif
template <bool SMEM, bool COUNT, bool SHADOW, int ONLY = -1>
__global__ void RT_EXT_BOUNDS k_extend(DScene sc, Queues q, int bounce, int buf, int max_batch) {
    const bool fixup = max_batch < 0;
    if (fixup) max_batch = -max_batch;
    extern __shared__ float2 s_raw[];
    const int n = SHADOW ? q.shcount[bounce] : q.qcount[bounce];
    if (n == 0) return;
    const int total_warps = gridDim.x * (RT_EXT_BLOCK / 32);
    int batch = (n + total_warps - 1) / total_warps;
    batch = batch < 32 ? 32 : (batch > max_batch ? max_batch : ((batch + 31) & ~31));
    if ((long long)blockIdx.x * (RT_EXT_BLOCK / 32) * batch >= n) return;
    const char *nodes = reinterpret_cast<const char *>(sc.nodes);
    if (SMEM) {
        const float2 *src = reinterpret_cast<const float2 *>(sc.nodes_s);
        for (int i = threadIdx.x; i < sc.n_nodes * (int)(sizeof(BvhNodeS) / 8); i += blockDim.x) s_raw[i] = __ldg(&src[i]);
        __syncthreads();
        nodes = reinterpret_cast<const char *>(s_raw);
    }
    // this thread's column of the shared-memory traversal stack, behind the tree (dynamic smem = tree bytes + RT_SSTACK_BYTES)
    const unsigned sstack = (unsigned)__cvta_generic_to_shared(s_raw) + (SMEM ? (unsigned)sc.n_nodes * (unsigned)sizeof(BvhNodeS) : 0u) +
                            threadIdx.x * 8u;
    int *cursor = &q.workctr[bounce * 2 + (SHADOW ? 1 : 0)];
    const float4 *__restrict__ ray_o = SHADOW ? q.sh_o : q.ray_o[buf];
    const float4 *__restrict__ ray_d = SHADOW ? q.sh_d : q.ray_d[buf];
    TravCounters cnt = {};
    for (;;) {
        const int base = warp_fetch(cursor, batch);
        if (base >= n) break;
        const int end = min(base + batch, n);
#pragma unroll 1
        for (int i = base + lane_id(); i < end; i += 32) {
            if (!SHADOW && fixup && q.hits[i].y != RT_HIT_TODO) continue;
            const float4 o4 = ray_o[i], d4 = ray_d[i];
            float t; int code;
            RT_CHECK(i >= 0 && (long long)i < q.cap);
            if (SHADOW) {
                if (SMEM) traverse<true, COUNT, true, RT_SSTACK, ONLY>(sc, nodes, f4xyz(o4), f4xyz(d4), 0.001f, 0.999f, t, code, cnt, sstack);
                else if (sc.nodes4) traverse4<COUNT, true, RT_SSTACK, ONLY>(sc, sc.nodes4, f4xyz(o4), f4xyz(d4), 0.001f, 0.999f, t, code, cnt, sstack);
                else traverse<false, COUNT, true, RT_SSTACK, ONLY>(sc, nodes, f4xyz(o4), f4xyz(d4), 0.001f, 0.999f, t, code, cnt, sstack);
                if (code < 0) {
                    const int pid = __float_as_int(o4.w);
                    RT_CHECK(pid >= 0 && (long long)pid < q.cap);
                    const float4 c = q.sh_c[i], r = q.radiance[pid];
                    q.radiance[pid] = make_float4(r.x + c.x, r.y + c.y, r.z + c.z, 0.0f);
                }
            } else {
                if (SMEM) traverse<true, COUNT, false, RT_SSTACK, ONLY>(sc, nodes, f4xyz(o4), f4xyz(d4), 0.001f, FLT_MAX, t, code, cnt, sstack);
                else if (sc.nodes4) traverse4<COUNT, false, RT_SSTACK, ONLY>(sc, sc.nodes4, f4xyz(o4), f4xyz(d4), 0.001f, FLT_MAX, t, code, cnt, sstack);
                else traverse<false, COUNT, false, RT_SSTACK, ONLY>(sc, nodes, f4xyz(o4), f4xyz(d4), 0.001f, FLT_MAX, t, code, cnt, sstack);
                q.hits[i] = make_int2(__float_as_int(t), code);
            }
        }
    }
    if (COUNT) {
        flush_counter<COUNT>(q.counters, CTR_BOX, cnt.box); flush_counter<COUNT>(q.counters, CTR_SPH_REJECT, cnt.sph_rej);
        flush_counter<COUNT>(q.counters, CTR_SPH_ACCEPT, cnt.sph_acc); flush_counter<COUNT>(q.counters, CTR_RECT, cnt.rect);
        flush_counter<COUNT>(q.counters, CTR_CYL_REJECT, cnt.cyl_rej); flush_counter<COUNT>(q.counters, CTR_CYL_ACCEPT, cnt.cyl_acc);
        flush_counter<COUNT>(q.counters, CTR_TRI, cnt.tri); flush_counter<COUNT>(q.counters, CTR_BOXPRIM, cnt.boxprim);
        flush_counter<COUNT>(q.counters, CTR_NODE_VISITS, cnt.nodes); flush_counter<COUNT>(q.counters, CTR_STACK_OVERFLOW, cnt.overflow);
    }
}

// ------------------------------------------------------------------------------------------------
// shade: rebuild the hit record, evaluate emitted/scatter (main.cu:45-63), write the radiance of paths that end,
// and compact the survivors into the next bounce's queue.
//
// One CTA owns one chunk of RT_SH_CHUNK consecutive queue entries at a time (static round-robin, no work atomics):
//   1. the chunk's ray / throughput / hit records arrive in shared memory by four TMA bulk copies
//      (cp.async.bulk + mbarrier), double-buffered so the next chunk streams in while this one is shaded;
//   2. the entries are sorted by material class inside shared memory (ballot + popc multisplit), so a warp shades
//      one material class (miss / lambertian / metal / dielectric / light) except at the <= 4 class boundaries;
//   3. survivors are compacted into a shared-memory staging area (ballot + popc, one shared atomic per warp), the CTA
//      reserves its output range with ONE global atomic per chunk and three TMA bulk stores write full lines.
// ------------------------------------------------------------------------------------------------
#ifndef RT_SH_CHUNK
#define RT_SH_CHUNK 256
#endif
#ifndef RT_SH_CTAS_PER_SM
#define RT_SH_CTAS_PER_SM (1024 / RT_SH_CHUNK)
#endif
#define RT_SH_WARPS (RT_SH_CHUNK / 32)

// Camera-path entries of the bounce-0 queue (position = path id = sample * n_pixels + plane position), plane positions
// k0, k0 + 1, ... (wrapping at n_pixels), cnt of them: 1 when every one belongs to a tile whose beam touches nothing (generate
// wrote no record for them), 0 when none does, 2 when the chunk is mixed.
__device__ __forceinline__ int cam_range_state(const DScene &sc, int k0, int cnt, int n_pixels) {
    const int n_tiled = (sc.width & ~7) * (sc.height & ~3);
    int left = cnt, empties = 0, tiles = 0;
    bool margin = false;
    while (left > 0) {                        // at most two plane segments unless the image is smaller than a chunk
        const int k1 = min(k0 + left, n_pixels), kt = min(k1, n_tiled);
        if (k1 > n_tiled) margin = true;      // margin pixels outside the tiling are never "empty"
        if (kt > k0) {
            const int t0 = k0 >> 5, t1 = (kt + 31) >> 5;
            empties += __ldg(&sc.tile_empty_prefix[t1]) - __ldg(&sc.tile_empty_prefix[t0]); tiles += t1 - t0;
        }
        left -= k1 - k0; k0 = 0;
    }
    if (empties == 0) return 0;
    return (empties == tiles && !margin) ? 1 : 2;
}

struct __align__(128) ShadeSmem {
    float4 in_o[2][RT_SH_CHUNK], in_d[2][RT_SH_CHUNK], in_thr[2][RT_SH_CHUNK];
    int2 in_hit[2][RT_SH_CHUNK];
    float4 out_o[RT_SH_CHUNK], out_d[RT_SH_CHUNK], out_thr[RT_SH_CHUNK];
    unsigned long long full[2];       // mbarriers: "stage s has landed"
    int cls_count[8];
    int n_out, out_base;
    int cam_state[2], cam_k0[2];      // camera paths: 1 = the stage's chunk holds entries of empty tiles only (nothing was loaded),
                                      // 2 = some of them, 0 = none; plane position of the chunk's first entry
    unsigned short perm[RT_SH_CHUNK];
};

__device__ __forceinline__ unsigned smem_u32(const void *p) { return (unsigned)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(unsigned long long *bar, int count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(unsigned long long *bar, unsigned bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned long long *bar, unsigned parity) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_LOOP:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra WAIT_DONE;\n"
        "bra WAIT_LOOP;\n"
        "WAIT_DONE:\n"
        "}\n" ::"r"(smem_u32(bar)), "r"(parity) : "memory");
}
// TMA 1-D bulk copy global -> shared, completion counted in bytes on the mbarrier
__device__ __forceinline__ void tma_load_1d(void *dst_smem, const void *src_gmem, unsigned bytes, unsigned long long *bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst_smem)),
                 "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}
// TMA 1-D bulk copy shared -> global (bulk async-group completion)
__device__ __forceinline__ void tma_store_1d(void *dst_gmem, const void *src_smem, unsigned bytes) {
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst_gmem), "r"(smem_u32(src_smem)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void tma_store_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
__device__ __forceinline__ void tma_store_wait_read() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }
__device__ __forceinline__ void tma_store_wait_all() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

// with_thr = false at bounce 0: camera paths carry no throughput record (generate does not write one)
__device__ __forceinline__ void shade_issue_loads(ShadeSmem &sm, const DScene &sc, const Queues &q, int buf, int chunk, int n, int stage,
                                                  bool with_thr) {
    const int first = chunk * RT_SH_CHUNK;
    const int cnt = min(RT_SH_CHUNK, n - first);
    const unsigned b16 = (unsigned)cnt * 16u, b8 = ((unsigned)cnt * 8u + 15u) & ~15u;
    if (!with_thr && sc.tile_empty_prefix) {      // written before the arrive below (release), read after the wait (acquire)
        const int np = sc.width * sc.height, k0 = first % np;
        const int state = cam_range_state(sc, k0, cnt, np);
        sm.cam_state[stage] = state; sm.cam_k0[stage] = k0;
        if (state == 1) { mbar_expect_tx(&sm.full[stage], 0u); return; }
    }
    mbar_expect_tx(&sm.full[stage], (with_thr ? 3u : 2u) * b16 + b8);
    tma_load_1d(sm.in_o[stage], q.ray_o[buf] + first, b16, &sm.full[stage]);
    tma_load_1d(sm.in_d[stage], q.ray_d[buf] + first, b16, &sm.full[stage]);
    if (with_thr) tma_load_1d(sm.in_thr[stage], q.thr[buf] + first, b16, &sm.full[stage]);
    tma_load_1d(sm.in_hit[stage], q.hits + first, b8, &sm.full[stage]);
}

template <bool COUNT, bool NEE, bool CAM, bool DIRECT, int ONLY = -1>
__global__ void __launch_bounds__(RT_SH_CHUNK, RT_SH_CTAS_PER_SM) k_shade(DScene sc, Queues q, unsigned seed, int spp_begin, int bounce, int buf,
                                                         int last_bounce, int sort) {
    extern __shared__ __align__(128) unsigned char s_shade_raw[];
    ShadeSmem &sm = *reinterpret_cast<ShadeSmem *>(s_shade_raw);
    const int n = q.qcount[bounce];
    const int n_chunks = (n + RT_SH_CHUNK - 1) / RT_SH_CHUNK;
    if ((int)blockIdx.x >= n_chunks) return;
    const int tid = threadIdx.x, lane = tid & 31;
    int *next_count = &q.qcount[bounce + 1];
    if (tid == 0) {
        mbar_init(&sm.full[0], 1); mbar_init(&sm.full[1], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        sm.n_out = 0;
#pragma unroll
        for (int b = 0; b < 8; ++b) sm.cls_count[b] = 0;
    }
    __syncthreads();
    constexpr bool camera_paths = CAM;            // bounce 0: queue position = path id, throughput (1, 1, 1), no record in memory
    const int n_pixels = sc.width * sc.height;
    const float inv_np = 1.0f / (float)n_pixels;
    if (tid == 0) shade_issue_loads(sm, sc, q, buf, blockIdx.x, n, 0, !camera_paths);
    unsigned n_shaded = 0;
    int it = 0;
    for (int chunk = blockIdx.x; chunk < n_chunks; chunk += gridDim.x, ++it) {
        const int st = it & 1;
        const int cnt = min(RT_SH_CHUNK, n - chunk * RT_SH_CHUNK);
        if (tid == 0 && chunk + (int)gridDim.x < n_chunks) shade_issue_loads(sm, sc, q, buf, chunk + gridDim.x, n, st ^ 1, !camera_paths);
        mbar_wait(&sm.full[st], (unsigned)(it >> 1) & 1u);
        const int cam_state = (camera_paths && sc.tile_empty_prefix) ? sm.cam_state[st] : 0;
        if (cam_state == 1) { __syncthreads(); continue; }      // uniform: only empty-tile entries, nothing was loaded
        // (the barrier keeps thread 0 from re-using sm.cam_state[st] for the chunk after next before everyone has read it)
        // ---- 1. sort the chunk by material class (shared memory only) ----
        int j = tid;                                   // the entry this thread shades
        if (sort && cam_state == 0) {     // (a mixed camera chunk holds unwritten records)
            int cls = -1;
            if (tid < cnt) {
                const int code = sm.in_hit[st][tid].y;
                cls = code < 0 ? 0 : 1 + sc.mats[__ldg(&sc.prim_mat[code & RT_LEAF_SLOT_MASK])].type;
            }
            int pos_in_cls = 0;
#pragma unroll
            for (int b = 0; b < RT_NUM_BINS; ++b) {
                const unsigned m = __ballot_sync(0xffffffffu, cls == b);
                if (m) {
                    int base = 0;
                    const int leader = __ffs(m) - 1;
                    if (lane == leader) base = atomicAdd(&sm.cls_count[b], __popc(m));
                    base = __shfl_sync(0xffffffffu, base, leader);
                    if (cls == b) pos_in_cls = base + __popc(m & lanemask_lt());
                }
            }
            __syncthreads();
            if (cls >= 0) {
                int off = 0;
#pragma unroll
                for (int b = 0; b < RT_NUM_BINS; ++b) if (b < cls) off += sm.cls_count[b];
                sm.perm[off + pos_in_cls] = (unsigned short)tid;
            }
            __syncthreads();
            if (tid < cnt) j = sm.perm[tid];
            if (tid < 8) sm.cls_count[tid] = 0;        // next use is after the two barriers of step 3
        }
        // ---- 2. shade ----
        bool alive = false;
        float3 no = v3(0, 0, 0), nd = v3(0, 0, 0), nthr = v3(0, 0, 0);
        int path_id = 0;
        float key_pixel = 0.0f, key_sample = 0.0f;
        bool sh_on = false; int sh_flag = 0;
        float3 sh_x = v3(0, 0, 0), sh_w = v3(0, 0, 0), sh_f = v3(0, 0, 0);
        if (tid < cnt) {
            bool live = true;
            if (cam_state == 2) {     // an entry of an empty tile inside a mixed chunk: no record was written
                int kpos = sm.cam_k0[st] + j;
                while (kpos >= n_pixels) kpos -= n_pixels;
                live = kpos >= (sc.width & ~7) * (sc.height & ~3) || __ldg(&sc.tile_cnt[kpos >> 5]) != 0;
            }
            if (live) {
            const float4 o4 = sm.in_o[st][j], d4 = sm.in_d[st][j];
            float4 t4;
            if (camera_paths) {        // sample index = path id / n_pixels: float estimate (off by one at most), exact fix-up
                const int pid = __float_as_int(o4.w);
                int s_local = (int)((float)pid * inv_np), rem = pid - s_local * n_pixels;
                while (rem < 0) { rem += n_pixels; --s_local; }
                while (rem >= n_pixels) { rem -= n_pixels; ++s_local; }
                t4 = make_float4(1.0f, 1.0f, 1.0f, __int_as_float(s_local));
            } else t4 = sm.in_thr[st][j];
            const int2 h = sm.in_hit[st][j];
            const float3 o = f4xyz(o4), d = f4xyz(d4), thr = f4xyz(t4);
            path_id = __float_as_int(o4.w); key_pixel = d4.w; key_sample = t4.w;
            // NEE: bit 30 of the path id = "the previous vertex sampled the lights" (do not count a sampled light again)
            const bool prev_nee = NEE && (path_id & 0x40000000) != 0;
            if (NEE) path_id &= 0x3fffffff;
            float3 rad = v3(0, 0, 0);
            // The chunk is bound by the latency of its own dependent chain, so everything that does not need the
            // slot's records runs UNDER their loads: the three loads go out first (a miss reads slot 0: one cached line, no branch in
            // front of the loads), then the miss lanes' background, then Philox + the direct sampler of the hit lanes.
            const bool hit = h.y >= 0;
            const int type = hit ? (int)((unsigned)h.y >> 28) : 0, slot = hit ? (h.y & RT_LEAF_SLOT_MASK) : 0;
            const float4 pA = __ldg(&sc.primA[slot]), pB = __ldg(&sc.primB[slot]);
            const DMaterial m = sc.slot_mat ? sc.slot_mat[slot] : sc.mats[hit ? __ldg(&sc.prim_mat[slot]) : 0];     // (an empty scene has no slot 0 material index)
            if (!hit) {
                rad = thr * background(sc, d);
            } else {
                if (COUNT) n_shaded++;
                Rng rng; rng.init(seed, (unsigned)__float_as_int(d4.w), (unsigned)(spp_begin + __float_as_int(t4.w)), (unsigned)bounce + 1u, sc.sampling);
                BallDraw ahead;
                if (DIRECT) { rng.prime(); ahead = ball_direct_ahead(rng); }
                HitRec rec;
                hit_record<ONLY>(sc, type, slot, o, d, __int_as_float(h.x), m.albedo.w != 0.0f, rec, pA, pB);
                if (m.type == RT_MAT_DIFFUSE_LIGHT) {     // emitted(), material.cuh:175-178; scatter() == false
                    const int lk = prev_nee ? sc.obj_light[__ldg(&sc.prim_obj[slot])] : -1;
                    if (lk < 0) rad = thr * material_color(sc, m, rec.u, rec.v, rec.p);
                    else if (sc.nee_mis) rad = thr * (mis_hit_weight(sc, lk, d, __int_as_float(h.x), rec.n) * material_color(sc, m, rec.u, rec.v, rec.p));
                } else {
                    float3 att, sd;
                    const bool ok = scatter<DIRECT>(sc, m, d, rec, rng, att, sd, DIRECT ? &ahead : nullptr);
                    if (ok && !last_bounce) {
                        alive = true;
                        no = rec.p; nd = sd; nthr = thr * att;
                    }
                    if (NEE && m.type == RT_MAT_LAMBERTIAN && sc.n_lights > 0) {      // shadow stage feed (oracle/orc_nee.c)
                        Rng r2; r2.init(seed, (unsigned)__float_as_int(d4.w), (unsigned)(spp_begin + __float_as_int(t4.w)),
                                        ((unsigned)bounce + 1u) | 0x10000u, sc.sampling);
                        float3 f;
                        if (light_sample(sc, rec.p, rec.n, r2, sh_w, f)) { sh_on = true; sh_x = rec.p; sh_f = (thr * att) * f; }
                        sh_flag = 0x40000000;
                    }
                }
            }
            RT_CHECK(path_id >= 0 && (long long)path_id < q.cap);
            if (!alive) {
                if (NEE) { const float4 r0 = q.radiance[path_id]; rad = v3(r0.x + rad.x, r0.y + rad.y, r0.z + rad.z); }
                q.radiance[path_id] = make_float4(rad.x, rad.y, rad.z, 0.0f);
            }
            }
        }
        if (NEE) {      // occlusion rays of this chunk -> shadow queue (one global atomic per warp)
            const unsigned ms = __ballot_sync(0xffffffffu, sh_on);
            if (ms) {
                int pos = 0;
                const int leader = __ffs(ms) - 1;
                if (lane == leader) pos = atomicAdd(&q.shcount[bounce], __popc(ms));
                pos = __shfl_sync(0xffffffffu, pos, leader) + __popc(ms & lanemask_lt());
                RT_CHECK(pos >= 0 && (long long)pos < q.cap);
                if (sh_on) {
                    q.sh_o[pos] = make_float4(sh_x.x, sh_x.y, sh_x.z, __int_as_float(path_id));
                    q.sh_d[pos] = make_float4(sh_w.x, sh_w.y, sh_w.z, 0.0f);
                    q.sh_c[pos] = make_float4(sh_f.x, sh_f.y, sh_f.z, 0.0f);
                }
            }
        }
        // ---- 3. compact the survivors into the staging area; one global atomic + three TMA stores per chunk ----
        if (tid == 0) tma_store_wait_read();          // the previous chunk's stores have finished reading the staging area
        __syncthreads();
        const unsigned m = __ballot_sync(0xffffffffu, alive);
        if (m) {
            int pos = 0;
            const int leader = __ffs(m) - 1;
            if (lane == leader) pos = atomicAdd(&sm.n_out, __popc(m));
            pos = __shfl_sync(0xffffffffu, pos, leader) + __popc(m & lanemask_lt());
            RT_CHECK(pos >= 0 && pos + (alive ? 0 : 0) < RT_SH_CHUNK);
            if (alive) {
                sm.out_o[pos] = make_float4(no.x, no.y, no.z, __int_as_float(path_id | sh_flag));
                sm.out_d[pos] = make_float4(nd.x, nd.y, nd.z, key_pixel);
                sm.out_thr[pos] = make_float4(nthr.x, nthr.y, nthr.z, key_sample);
            }
        }
        fence_proxy_async();                          // make the generic-proxy writes visible to the TMA engine
        __syncthreads();
        if (tid == 0) {
            const int n_out = sm.n_out;
            if (n_out) {
                const int base = atomicAdd(next_count, n_out);
                RT_CHECK(n_out <= RT_SH_CHUNK && base >= 0 && (long long)base + n_out <= q.cap);
                tma_store_1d(q.ray_o[buf ^ 1] + base, sm.out_o, (unsigned)n_out * 16u);
                tma_store_1d(q.ray_d[buf ^ 1] + base, sm.out_d, (unsigned)n_out * 16u);
                tma_store_1d(q.thr[buf ^ 1] + base, sm.out_thr, (unsigned)n_out * 16u);
                tma_store_commit();
            }
            sm.n_out = 0;                             // next use is after the first barrier of the next step 3
        }
    }
    if (tid == 0) tma_store_wait_all();
    flush_counter<COUNT>(q.counters, CTR_SHADE, n_shaded);
}

// ------------------------------------------------------------------------------------------------
// tail: finishes every path that is still alive at bounce b0 in ONE launch (extend + shade in a per-lane loop, no queues).
//
// The wavefront ends in dozens of launches over queues that have shrunk to a few per cent of the pass (RTIOW: 3.5 % of the paths
// are alive at bounce 8, the queues stay non-empty to bounce ~45): each of those 2 x 40 launches is a few tens of microseconds of a
// mostly empty GPU.  From bounce b0 on (rt_ctx.cu chooses it), a lane owns one path and runs  closest hit -> emitted / scatter  until
// the path ends, then writes its radiance record; a finished lane takes the next entry of the bounce-b0 queue (warp-aggregated
// atomic on the work cursor) once at least RT_TAIL_REFILL lanes of its warp are idle, so that the loop does not run at one or two
// live lanes while 45-bounce stragglers finish.  Same Philox keys (pixel, sample, bounce + 1) and the same functions as k_extend /
// k_shade: the image is bit-identical to the queue schedule (tests: RT_RENDER_NO_TAIL).  Also the "queue-free" schedule for scenes
// of a dozen objects when b0 = 1.  Not used with NEE (the shadow stage needs its queue) or COUNT_EVENTS.
// Segments traced beyond bounce b0 are added to qcount[b0 + 1] (ray statistics; the shade stage of bounce b0 never runs).
// ------------------------------------------------------------------------------------------------
#ifndef RT_TAIL_REFILL
#define RT_TAIL_REFILL 4         // (1 / 4 / 12 / 24: tail 6.51 / 6.53 / 9.06 / 12.55 ms per 128-spp RTIOW pass)
#endif
template <bool SMEM, bool DIRECT, int ONLY = -1>
#ifndef RT_TAIL_CTAS
#define RT_TAIL_CTAS 2          // 64 registers with a few spills; 1 CTA (91 registers, 16 warps per SM): 9.06 vs 7.21 ms
#endif
__global__ void __launch_bounds__(RT_EXT_BLOCK, RT_TAIL_CTAS) k_tail(DScene sc, Queues q, unsigned seed, int spp_begin, int b0, int buf) {
    extern __shared__ float2 s_raw[];
    const int n = q.qcount[b0];
    if (n == 0) return;
    const char *nodes = reinterpret_cast<const char *>(sc.nodes);
    if (SMEM) {
        const float2 *src = reinterpret_cast<const float2 *>(sc.nodes_s);
        for (int i = threadIdx.x; i < sc.n_nodes * (int)(sizeof(BvhNodeS) / 8); i += blockDim.x) s_raw[i] = __ldg(&src[i]);
        __syncthreads();
        nodes = reinterpret_cast<const char *>(s_raw);
    }
    const unsigned sstack = (unsigned)__cvta_generic_to_shared(s_raw) + (SMEM ? (unsigned)sc.n_nodes * (unsigned)sizeof(BvhNodeS) : 0u) +
                            threadIdx.x * 8u;
    int *cursor = &q.workctr[b0 * 2];
    const float4 *__restrict__ in_o = q.ray_o[buf], *__restrict__ in_d = q.ray_d[buf], *__restrict__ in_t = q.thr[buf];
    const int lane = lane_id();
    TravCounters cnt = {};
    // per-lane path state
    int state = -1;                  // >= 0: path in flight (value = bounce); -1: idle, wants work; -2: the queue is exhausted
    float3 o = v3(0, 0, 0), d = v3(0, 0, 0), thr = v3(0, 0, 0);
    float key_pixel = 0.0f, key_sample = 0.0f;
    int path_id = 0;
    unsigned extra = 0;              // segments traced beyond bounce b0
    for (;;) {
        const unsigned idle = __ballot_sync(0xffffffffu, state == -1), live = __ballot_sync(0xffffffffu, state >= 0);
        if (idle && (live == 0u || __popc(idle) >= RT_TAIL_REFILL)) {
            int base = 0;
            const int leader = __ffs(idle) - 1;
            if (lane == leader) base = atomicAdd(cursor, __popc(idle));
            base = __shfl_sync(0xffffffffu, base, leader);
            if (state == -1) {
                const int i = base + __popc(idle & lanemask_lt());
                if (i < n) {
                    const float4 o4 = in_o[i], d4 = in_d[i], t4 = in_t[i];
                    o = f4xyz(o4); d = f4xyz(d4); thr = f4xyz(t4);
                    path_id = __float_as_int(o4.w); key_pixel = d4.w; key_sample = t4.w;
                    state = b0;
                } else state = -2;
            }
        }
        if (__ballot_sync(0xffffffffu, state >= 0) == 0u) {
            if (__ballot_sync(0xffffffffu, state == -1) == 0u) break;      // every lane has seen the end of the queue
            continue;
        }
        if (state >= 0) {
            float t; int code;
            if (SMEM) traverse<true, false, false, RT_SSTACK, ONLY>(sc, nodes, o, d, 0.001f, FLT_MAX, t, code, cnt, sstack);
            else if (sc.nodes4) traverse4<false, false, RT_SSTACK, ONLY>(sc, sc.nodes4, o, d, 0.001f, FLT_MAX, t, code, cnt, sstack);
            else traverse<false, false, false, RT_SSTACK, ONLY>(sc, nodes, o, d, 0.001f, FLT_MAX, t, code, cnt, sstack);
            if (state > b0) ++extra;
            float3 rad = v3(0, 0, 0);
            bool alive = false;
            if (code < 0) {
                rad = thr * background(sc, d);
            } else {
                const int type = (unsigned)code >> 28, slot = code & RT_LEAF_SLOT_MASK;
                const float4 pA = __ldg(&sc.primA[slot]), pB = __ldg(&sc.primB[slot]);
                const DMaterial m = sc.slot_mat ? sc.slot_mat[slot] : sc.mats[__ldg(&sc.prim_mat[slot])];
                Rng rng; rng.init(seed, (unsigned)__float_as_int(key_pixel), (unsigned)(spp_begin + __float_as_int(key_sample)), (unsigned)state + 1u, sc.sampling);
                BallDraw ahead;
                if (DIRECT) { rng.prime(); ahead = ball_direct_ahead(rng); }      // (as in k_shade: under the latency of the slot's records)
                HitRec rec;
                hit_record<ONLY>(sc, type, slot, o, d, t, m.albedo.w != 0.0f, rec, pA, pB);
                if (m.type == RT_MAT_DIFFUSE_LIGHT) {
                    rad = thr * material_color(sc, m, rec.u, rec.v, rec.p);
                } else {
                    float3 att, sd;
                    const bool ok = scatter<DIRECT>(sc, m, d, rec, rng, att, sd, DIRECT ? &ahead : nullptr);
                    if (ok && state != sc.max_depth - 1) { alive = true; o = rec.p; d = sd; thr = thr * att; }
                }
            }
            if (alive) ++state;
            else { q.radiance[path_id] = make_float4(rad.x, rad.y, rad.z, 0.0f); state = -1; }
        }
    }
#pragma unroll
    for (int k = 16; k > 0; k >>= 1) extra += __shfl_xor_sync(0xffffffffu, extra, k);
    if (lane == 0 && extra) atomicAdd(&q.qcount[b0 + 1], (int)extra);
}

// ------------------------------------------------------------------------------------------------
// accumulate: accum[pixel] += sum over this pass's samples, in ascending sample order (fixed order => the image is
// bit-reproducible; replaces the register sum of main.cu:95-104).  One thread per pixel, S coalesced plane reads.
// ------------------------------------------------------------------------------------------------
// const_tiles: tile_cnt of a pass whose generate skipped the empty tiles under a constant background (else nullptr) — their
// pixels get the constant added once per sample, in the same order as the sum over radiance records would.
// plan: the pass's sample planes fall into up to RT_ACC_CHUNKS runs, one per sample chunk of the fixed-order accumulation
// (rt.h: rt_sample_chunk); run j = planes [first[j], first[j+1]) is summed, in plane order, into chunk buffer buf[j].
__global__ void __launch_bounds__(RT_BLOCK) k_accumulate(const float4 *__restrict__ radiance, AccumPlan plan, int W, int H,
                                                         int n_samples, const int *__restrict__ qcount, int max_depth,
                                                         unsigned long long *total_rays, const int *__restrict__ shcount,
                                                         const int *__restrict__ const_tiles, float3 bg, int ext_first, int ext_end) {
    const int np = W * H;
    if (blockIdx.x == 0 && threadIdx.x == 0) {     // rays traced this pass = sum of the per-bounce queue lengths
        unsigned long long r = 0, sh = 0;
        // every camera path is one ray (the bounce-0 queue may hold fewer: paths that miss everything end in generate)
        r = (unsigned long long)np * (unsigned long long)n_samples;
        for (int b = 0; b < max_depth; ++b) { if (b > 0) r += (unsigned long long)qcount[b]; if (shcount) sh += (unsigned long long)shcount[b]; }
        total_rays[0] += r; total_rays[1] += sh;
        unsigned long long re = 0;      // rays that went through an extend launch: bounces [ext_first, ext_end) (the rest: generate's lists, the tail stage)
        for (int b = ext_first; b < ext_end && b < max_depth; ++b) re += (unsigned long long)qcount[b];
        total_rays[2] += re;
    }
    for (int k = blockIdx.x * blockDim.x + threadIdx.x; k < np; k += gridDim.x * blockDim.x) {
        const int p = plane_to_pixel(W, H, k);
        const bool constant = const_tiles && k < (W & ~7) * (H & ~3) && __ldg(&const_tiles[k >> 5]) == 0;
#pragma unroll
        for (int j = 0; j < RT_ACC_CHUNKS; ++j) {
            if (j >= plan.n) break;
            float *__restrict__ accum = plan.buf[j];
            float r = 0.0f, g = 0.0f, b = 0.0f;
            if (!plan.clear[j]) { r = accum[3 * p]; g = accum[3 * p + 1]; b = accum[3 * p + 2]; }
            const int s0 = plan.first[j], s1 = plan.first[j + 1];
            if (constant) {
                for (int s = s0; s < s1; ++s) { r += bg.x; g += bg.y; b += bg.z; }
            } else
            for (int s = s0; s < s1; ++s) {
                const float4 c = radiance[(long long)s * np + k];
                r += c.x; g += c.y; b += c.z;
            }
            accum[3 * p] = r; accum[3 * p + 1] = g; accum[3 * p + 2] = b;
        }
    }
}

// ------------------------------------------------------------------------------------------------
// primary_only: scenes whose objects are ALL emitters (the frames of gpu-version/dna.py: 60 emissive spheres + 30 emissive
// cylinders) end every path at its first vertex, and with complete tile candidate lists (tile_mode 1) the first hit needs no tree.
// Then generate -> shade<CAM> -> accumulate is one kernel with no queue at all: a thread owns a pixel, loops over the samples in
// order, and adds each sample's radiance (emitter colour | background) to the running sum of the sample chunk it belongs to.  Same
// functions, same Philox keys and the same order of float additions as the three-kernel path: bit-identical image
// (test_empty_tiles_skip_the_queues compares with RT_RENDER_NO_TILE_LISTS), 112 B of queue traffic per sample gone.
// ------------------------------------------------------------------------------------------------
template <bool DIRECT>
__global__ void __launch_bounds__(RT_BLOCK) k_primary_only(DScene sc, AccumPlan plan, unsigned seed, int spp_begin, int n_samples,
                                                          unsigned long long *total_rays) {
    const int np = sc.width * sc.height;
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k == 0) total_rays[0] += (unsigned long long)np * (unsigned long long)n_samples;       // one ray per camera path
    if (k >= np) return;
    const int pixel = plane_to_pixel(sc.width, sc.height, k);
    const int y = pixel / sc.width, x = pixel - y * sc.width;
    const float div_w = (float)(sc.width - 1), div_h = (float)(sc.height - 1);
    const int n_cand = __ldg(&sc.tile_cnt[k >> 5]);                   // tile_mode 1: every pixel belongs to a tile with a complete list
    const int *cand = sc.tile_list + (size_t)(k >> 5) * RT_TILE_K;
    const bool gradient = sc.bg_mode == RT_BG_SKY_GRADIENT;
    TravCounters cnt = {};
#pragma unroll 1
    for (int j = 0; j < plan.n; ++j) {
        float *__restrict__ accum = plan.buf[j];
        float r = 0.0f, g = 0.0f, b = 0.0f;
        if (!plan.clear[j]) { r = accum[3 * pixel]; g = accum[3 * pixel + 1]; b = accum[3 * pixel + 2]; }
#pragma unroll 1
        for (int s = plan.first[j]; s < plan.first[j + 1]; ++s) {
            float3 rad = sc.background;
            if (n_cand != 0 || gradient) {
                Rng rng; rng.init(seed, (unsigned)pixel, (unsigned)(spp_begin + s), 0u, sc.sampling);
                const float u = ((float)x + rng.next()) / div_w;
                const float v = ((float)y + rng.next()) / div_h;
                float3 o, d;
                camera_ray<DIRECT>(sc.cam, u, v, sc.defocus != 0, rng, o, d);
                float best_t = FLT_MAX; int best_code = -1;
                const float a = dot3(d, d), inv_a = 1.0f / a;
                for (int c = 0; c < n_cand; ++c) leaf_visit<false>(sc, __ldg(&cand[c]), o, d, a, inv_a, 0.001f, best_t, best_code, cnt);
                if (best_code < 0) rad = background(sc, d);
                else {
                    const int type = (unsigned)best_code >> 28, slot = best_code & RT_LEAF_SLOT_MASK;
                    const DMaterial m = sc.slot_mat ? sc.slot_mat[slot] : sc.mats[__ldg(&sc.prim_mat[slot])];
                    HitRec rec;
                    hit_record(sc, type, slot, o, d, best_t, m.albedo.w != 0.0f, rec);
                    rad = m.type == RT_MAT_DIFFUSE_LIGHT ? material_color(sc, m, rec.u, rec.v, rec.p) : v3(0, 0, 0);
                }
            }
            r += rad.x; g += rad.y; b += rad.z;
        }
        accum[3 * pixel] = r; accum[3 * pixel + 1] = g; accum[3 * pixel + 2] = b;
    }
}

// Fixed-order combination of the sample chunks of a render (rt_ctx.cu:rt_render): image = (((c0 + c1) + c2) + ...) in ascending
// chunk order, each c_k a sequential sum over the samples of chunk k.  The chunk boundaries depend on the total sample count
// only, so the image is bit-identical whether one GPU renders all chunks or eight GPUs render one each.
__global__ void __launch_bounds__(RT_BLOCK) k_fold_chunks(float *__restrict__ accum, ChunkPtrs cp, int n_chunks, int keep, long long n) {
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        float a = keep ? accum[i] : 0.0f;
#pragma unroll
        for (int k = 0; k < RT_ACC_CHUNKS; ++k)
            if (k < n_chunks) a = (k == 0 && !keep) ? cp.p[0][i] : a + cp.p[k][i];
        accum[i] = a;
    }
}

// acc = ((p0 + p1) + p2) + ... : fixed-order combination of per-GPU partial sums
__global__ void k_sum_ordered(float *__restrict__ accum, const float *const *__restrict__ partials, int n_partials, long long n) {
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        float a = partials[0][i];
        for (int k = 1; k < n_partials; ++k) a += partials[k][i];
        accum[i] = a;
    }
}

// ------------------------------------------------------------------------------------------------
// tonemap: write_image (PNG: sum/spp, clamp, x256, no gamma — color.cuh:15-35) or write_color (PPM: sum*(1/spp),
// sqrt, clamp, x256 — color.cuh:70-95); rows flipped so that row 0 of the output is the top row.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(RT_BLOCK) k_tonemap(const float *__restrict__ accum, unsigned char *__restrict__ rgb8, int W,
                                                      int H, int spp, int gamma) {
    const int n = W * H * 3;
    const float fspp = (float)spp, scale = 1.0f / fspp;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const int c = i % 3, px = i / 3, row = px / W, col = px - row * W;
        float x = accum[3 * ((H - 1 - row) * W + col) + c];
        x = gamma ? sqrtf(x * scale) : x / fspp;
        x = x < 0.0f ? 0.0f : (x > 0.999f ? 0.999f : x);
        rgb8[i] = (unsigned char)(int)(256.0f * x);
    }
}

// ------------------------------------------------------------------------------------------------
// aov: the deterministic path — pixel-centre primary ray, first hit through the same traversal and hit-record code,
// albedo / emission / background colour; 24 floats per pixel (layout: DESIGN.md §6).
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(RT_BLOCK) k_aov(DScene sc, float *__restrict__ out) {
    const int np = sc.width * sc.height;
    TravCounters cnt = {};
    for (int pixel = blockIdx.x * blockDim.x + threadIdx.x; pixel < np; pixel += gridDim.x * blockDim.x) {
        const int y = pixel / sc.width, x = pixel - y * sc.width;
        const float u = ((float)x + 0.5f) / (float)(sc.width - 1), v = ((float)y + 0.5f) / (float)(sc.height - 1);
        Rng rng; rng.init(0, 0, 0, 0, sc.sampling);
        float3 o, d;
        camera_ray(sc.cam, u, v, false, rng, o, d);
        float t; int code;
#ifdef RT_AOV_COST      // development builds (tools/build_variant.py): channel 23 = BVH nodes visited by this pixel's ray
        cnt.nodes = 0;
        traverse<false, true, false>(sc, reinterpret_cast<const char *>(sc.nodes), o, d, 0.001f, FLT_MAX, t, code, cnt);
#else
        traverse<false, false, false>(sc, reinterpret_cast<const char *>(sc.nodes), o, d, 0.001f, FLT_MAX, t, code, cnt);
#endif
        float q[24];
#pragma unroll
        for (int k = 0; k < 24; ++k) q[k] = 0.0f;
#ifdef RT_AOV_COST
        q[23] = (float)cnt.nodes;
#endif
        if (code >= 0) {
            const int type = (unsigned)code >> 28, slot = code & RT_LEAF_SLOT_MASK;
            const int mi = sc.prim_mat[slot];
            const DMaterial m = sc.mats[mi];
            HitRec rec;
            hit_record(sc, type, slot, o, d, t, true, rec);
            q[0] = 1.0f; q[1] = rec.t; q[2] = rec.p.x; q[3] = rec.p.y; q[4] = rec.p.z;
            q[5] = rec.n.x; q[6] = rec.n.y; q[7] = rec.n.z; q[8] = (float)rec.front; q[9] = rec.u; q[10] = rec.v;
            q[11] = (float)sc.prim_obj[slot]; q[12] = (float)mi;
            float3 att = v3(0, 0, 0), E = v3(0, 0, 0);
            bool ok = false;
            if (m.type == RT_MAT_DIFFUSE_LIGHT) E = material_color(sc, m, rec.u, rec.v, rec.p);
            else {
                // fixed uniforms 0.5,0.5,0.5,... as in the oracle/reference harness: only `att` and `ok` are recorded
                FixedRng fr;
                float3 sd;
                ok = sc.direct ? scatter<true>(sc, m, d, rec, fr, att, sd) : scatter<false>(sc, m, d, rec, fr, att, sd);
            }
            q[13] = att.x; q[14] = att.y; q[15] = att.z; q[16] = E.x; q[17] = E.y; q[18] = E.z; q[22] = ok ? 1.0f : 0.0f;
        } else {
            const float3 b = background(sc, d);
            q[19] = b.x; q[20] = b.y; q[21] = b.z;
        }
#pragma unroll
        for (int k = 0; k < 24; ++k) out[(long long)pixel * 24 + k] = q[k];
    }
}

// ================================================================================================
// host launchers
// ================================================================================================
static inline int persistent_grid(const LaunchCfg &c, int per_sm) { return c.num_sms * per_sm; }
static inline void bump(const LaunchCfg &c) { if (c.launches) ++*c.launches; }

// n_queue0: length of the bounce-0 queue = all camera paths of the pass (dense: position = path id)
void launch_pass_init(const LaunchCfg &c, const Queues &q, int max_depth, int n_queue0) {
    int n = max_depth * 2 + 2;
    k_pass_init<<<(n + 255) / 256, 256, 0, c.stream>>>(q, max_depth, n_queue0);
    bump(c);
}

// exclusive prefix count of empty tiles (tile_cnt == 0): prefix[t] = number of empty tiles among [0, t); one block
__global__ void __launch_bounds__(1024) k_tile_prefix(const int *__restrict__ tile_cnt, int n_tiles, int *__restrict__ prefix) {
    __shared__ int part[1024];
    const int per = (n_tiles + 1023) / 1024, t0 = threadIdx.x * per, t1 = min(t0 + per, n_tiles);
    int c = 0;
    for (int t = t0; t < t1; ++t) c += tile_cnt[t] == 0;
    part[threadIdx.x] = c;
    __syncthreads();
    for (int o = 1; o < 1024; o <<= 1) {       // inclusive Hillis-Steele scan of the per-thread counts
        const int v = threadIdx.x >= o ? part[threadIdx.x - o] : 0;
        __syncthreads();
        part[threadIdx.x] += v;
        __syncthreads();
    }
    int run = part[threadIdx.x] - c;
    for (int t = t0; t < t1; ++t) { prefix[t] = run; run += tile_cnt[t] == 0; }
    if (threadIdx.x == 1023) prefix[n_tiles] = part[1023];
}

void launch_tile_prefix(const LaunchCfg &c, const int *tile_cnt, int n_tiles, int *prefix) {
    k_tile_prefix<<<1, 1024, 0, c.stream>>>(tile_cnt, n_tiles, prefix);
    bump(c);
}

void launch_tile_lists(const LaunchCfg &c, const DScene &sc, int *list, int *tile_cnt, int *n_bad) {
    const int n_tiles = ((sc.width & ~7) >> 3) * ((sc.height & ~3) >> 2);
    if (n_tiles <= 0) return;
    k_tile_lists<<<(n_tiles + 127) / 128, 128, 0, c.stream>>>(sc, list, tile_cnt, n_bad);
    bump(c);
}

// hits: also resolve the bounce-0 hit records from the tile candidate lists (sc.tile_mode != 0)
void launch_generate(const LaunchCfg &c, const DScene &sc, const Queues &q, unsigned seed, int spp_begin, int n_paths, bool zero_radiance,
                     bool hits, bool count) {
    const int np = sc.width * sc.height, n_s = n_paths / np;
    const int bx = (np + RT_BLOCK - 1) / RT_BLOCK;
    // enough CTAs to fill the GPU: pixel blocks x sample groups
    int gy = (c.num_sms * 16 + bx - 1) / bx;
    gy = gy < 1 ? 1 : (gy > n_s ? n_s : gy);
#define RT_GEN_ARGS <<<dim3(bx, gy), RT_BLOCK, 0, c.stream>>>(sc, q, seed, spp_begin, n_paths, zero_radiance ? 1 : 0)
    if (sc.direct) {
        if (!hits) k_generate<false, false, true> RT_GEN_ARGS;
        else if (count) k_generate<true, true, true> RT_GEN_ARGS;
        else if (sc.only_type == RT_OBJ_SPHERE) k_generate<true, false, true, RT_OBJ_SPHERE> RT_GEN_ARGS;
        else k_generate<true, false, true> RT_GEN_ARGS;
    } else {
        if (!hits) k_generate<false, false, false> RT_GEN_ARGS;
        else if (count) k_generate<true, true, false> RT_GEN_ARGS;
        else if (sc.only_type == RT_OBJ_SPHERE) k_generate<true, false, false, RT_OBJ_SPHERE> RT_GEN_ARGS;
        else k_generate<true, false, false> RT_GEN_ARGS;
    }
#undef RT_GEN_ARGS
    bump(c);
}

template <bool SMEM, bool SHADOW>
static void launch_extend_t(const LaunchCfg &c, const DScene &sc, const Queues &q, int bounce, int buf, bool count, bool fixup = false) {
    const int grid = persistent_grid(c, c.extend_blocks_per_sm);
    const size_t sm = (SMEM ? c.bvh_smem_bytes : 0) + RT_SSTACK_BYTES;
    // largest run of consecutive rays a warp takes per fetch: long traversals (hundreds of nodes in the tree) balance better
    // at the tail of a launch with short runs, a dozen-object scene is bound by the fetch atomics and wants long ones
    // (RTIOW 64 vs 256: extend 18.39 vs 18.78 ms, mesh 72.8 vs 74.8, blue2 41.5 vs 40.5)
    static const int batch_env = getenv("RT_EXT_BATCH") ? atoi(getenv("RT_EXT_BATCH")) : 0;
    const int batch0 = batch_env > 0 ? batch_env : (sc.n_nodes > 64 ? 64 : RT_EXT_BATCH);
    const int batch = fixup ? -batch0 : batch0;
    const bool spheres = SMEM && sc.only_type == RT_OBJ_SPHERE;      // the specialised loop exists for the shared-memory tree
    if (count) k_extend<SMEM, true, SHADOW><<<grid, RT_EXT_BLOCK, sm, c.stream>>>(sc, q, bounce, buf, batch);
    else if (spheres) k_extend<true, false, SHADOW, RT_OBJ_SPHERE><<<grid, RT_EXT_BLOCK, sm, c.stream>>>(sc, q, bounce, buf, batch);
    else k_extend<SMEM, false, SHADOW><<<grid, RT_EXT_BLOCK, sm, c.stream>>>(sc, q, bounce, buf, batch);
}

// fixup: bounce 0 after a generate launch that wrote the hit records itself; only entries marked RT_HIT_TODO are traversed
void launch_extend(const LaunchCfg &c, const DScene &sc, const Queues &q, int bounce, int buf, bool count, bool fixup) {
    if (c.bvh_smem_bytes) launch_extend_t<true, false>(c, sc, q, bounce, buf, count, fixup);
    else launch_extend_t<false, false>(c, sc, q, bounce, buf, count, fixup);
    bump(c);
}

// shadow stage (RT_RENDER_NEE): the extend kernel in any-hit mode over the occlusion-ray queue shade has just written
void launch_shadow(const LaunchCfg &c, const DScene &sc, const Queues &q, int bounce) {
    if (c.bvh_smem_bytes) launch_extend_t<true, true>(c, sc, q, bounce, 0, false);
    else launch_extend_t<false, true>(c, sc, q, bounce, 0, false);
    bump(c);
}

// tail stage: every path alive at bounce b0 is finished by one launch (k_tail)
void launch_tail(const LaunchCfg &c, const DScene &sc, const Queues &q, unsigned seed, int spp_begin, int b0, int buf) {
    const int grid = persistent_grid(c, RT_TAIL_CTAS);
    const size_t sm = c.bvh_smem_bytes + RT_SSTACK_BYTES;
    const bool spheres = c.bvh_smem_bytes && sc.only_type == RT_OBJ_SPHERE;
#define RT_TAIL_ARGS <<<grid, RT_EXT_BLOCK, sm, c.stream>>>(sc, q, seed, spp_begin, b0, buf)
    if (sc.direct) {
        if (spheres) k_tail<true, true, RT_OBJ_SPHERE> RT_TAIL_ARGS;
        else if (c.bvh_smem_bytes) k_tail<true, true> RT_TAIL_ARGS;
        else k_tail<false, true> RT_TAIL_ARGS;
    } else {
        if (spheres) k_tail<true, false, RT_OBJ_SPHERE> RT_TAIL_ARGS;
        else if (c.bvh_smem_bytes) k_tail<true, false> RT_TAIL_ARGS;
        else k_tail<false, false> RT_TAIL_ARGS;
    }
#undef RT_TAIL_ARGS
    bump(c);
}

// The dynamic shared-memory limit is a property of the kernel on a device, shared by every context (and host thread) that uses
// the device: it is raised ONCE per device to the opt-in maximum for every k_extend / k_shade instance, never to one scene's tree
// size (a second context uploading a smaller tree must not lower the limit under the first one's launches).
static std::once_flag g_smem_once[RT_MAX_DEVICES];
static cudaError_t g_smem_err[RT_MAX_DEVICES];
cudaError_t kernels_configure_device(int dev, size_t smem_optin) {
    if (dev < 0 || dev >= RT_MAX_DEVICES) return cudaErrorInvalidDevice;
    std::call_once(g_smem_once[dev], [&] {
        cudaError_t e = cudaSuccess;
#define SET(K, B) if (e == cudaSuccess) e = cudaFuncSetAttribute(K, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(B))
        SET((k_extend<true, false, false>), smem_optin); SET((k_extend<true, true, false>), smem_optin); SET((k_extend<true, false, true>), smem_optin);
        SET((k_extend<true, true, true>), smem_optin);
        SET((k_extend<true, false, false, RT_OBJ_SPHERE>), smem_optin); SET((k_extend<true, false, true, RT_OBJ_SPHERE>), smem_optin);
        SET((k_extend<false, false, false>), smem_optin); SET((k_extend<false, true, false>), smem_optin);
        SET((k_extend<false, false, true>), smem_optin); SET((k_extend<false, true, true>), smem_optin);
        SET((k_tail<true, true, RT_OBJ_SPHERE>), smem_optin); SET((k_tail<true, true>), smem_optin); SET((k_tail<false, true>), smem_optin);
        SET((k_tail<true, false, RT_OBJ_SPHERE>), smem_optin); SET((k_tail<true, false>), smem_optin); SET((k_tail<false, false>), smem_optin);
#define SET_SH(D_) \
        SET((k_shade<false, false, false, D_>), sizeof(ShadeSmem)); SET((k_shade<true, false, false, D_>), sizeof(ShadeSmem)); \
        SET((k_shade<false, true, false, D_>), sizeof(ShadeSmem)); SET((k_shade<true, true, false, D_>), sizeof(ShadeSmem)); \
        SET((k_shade<false, false, true, D_>), sizeof(ShadeSmem)); SET((k_shade<true, false, true, D_>), sizeof(ShadeSmem)); \
        SET((k_shade<false, true, true, D_>), sizeof(ShadeSmem)); SET((k_shade<true, true, true, D_>), sizeof(ShadeSmem));
        SET_SH(false) SET_SH(true)
        SET((k_shade<false, false, false, false, RT_OBJ_SPHERE>), sizeof(ShadeSmem)); SET((k_shade<false, false, true, false, RT_OBJ_SPHERE>), sizeof(ShadeSmem));
        SET((k_shade<false, false, false, true, RT_OBJ_SPHERE>), sizeof(ShadeSmem)); SET((k_shade<false, false, true, true, RT_OBJ_SPHERE>), sizeof(ShadeSmem));
#undef SET_SH
#undef SET
        g_smem_err[dev] = e;
    });
    return g_smem_err[dev];
}

int extend_max_blocks_per_sm(size_t smem_bytes) {
    int nb = 0;
    if (smem_bytes) cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, k_extend<true, false, false>, RT_EXT_BLOCK, smem_bytes + RT_SSTACK_BYTES);
    else cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, k_extend<false, false, false>, RT_EXT_BLOCK, RT_SSTACK_BYTES);
    return nb;
}

void launch_shade(const LaunchCfg &c, const DScene &sc, const Queues &q, unsigned seed, int spp_begin, int bounce, int buf,
                  bool last_bounce, bool count, bool sort, bool nee) {
    static const int env_sh_ctas = getenv("RT_SH_CTAS") ? atoi(getenv("RT_SH_CTAS")) : 0;          // tuning: persistent CTAs per SM
    const int grid = persistent_grid(c, env_sh_ctas > 0 && env_sh_ctas < RT_SH_CTAS_PER_SM ? env_sh_ctas : RT_SH_CTAS_PER_SM);
#define RT_SHADE_ARGS <<<grid, RT_SH_CHUNK, sizeof(ShadeSmem), c.stream>>>(sc, q, seed, spp_begin, bounce, buf, last_bounce, sort)
#define RT_SHADE_CAM(C_, N_) { if (sc.direct) { if (bounce == 0) k_shade<C_, N_, true, true> RT_SHADE_ARGS; else k_shade<C_, N_, false, true> RT_SHADE_ARGS; } \
                              else { if (bounce == 0) k_shade<C_, N_, true, false> RT_SHADE_ARGS; else k_shade<C_, N_, false, false> RT_SHADE_ARGS; } }
    // all-sphere scenes (no instance transforms): the instance without the other primitives' hit-record code (fewer registers live)
    if (!count && !nee && sc.only_type == RT_OBJ_SPHERE) {
        if (sc.direct) { if (bounce == 0) k_shade<false, false, true, true, RT_OBJ_SPHERE> RT_SHADE_ARGS; else k_shade<false, false, false, true, RT_OBJ_SPHERE> RT_SHADE_ARGS; }
        else { if (bounce == 0) k_shade<false, false, true, false, RT_OBJ_SPHERE> RT_SHADE_ARGS; else k_shade<false, false, false, false, RT_OBJ_SPHERE> RT_SHADE_ARGS; }
    }
    else if (count) { if (nee) RT_SHADE_CAM(true, true) else RT_SHADE_CAM(true, false) }
    else { if (nee) RT_SHADE_CAM(false, true) else RT_SHADE_CAM(false, false) }
#undef RT_SHADE_CAM
#undef RT_SHADE_ARGS
    bump(c);
}

// light -> leaf slot (the LBVH build decides the slots on the device): one thread per slot
__global__ void k_light_slots(int n, const int *__restrict__ prim_obj, const int *__restrict__ obj_light, int *light_slot) {
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= n) return;
    const int k = obj_light[prim_obj[j]];
    if (k >= 0) light_slot[k] = j;
}

// material record per leaf slot (small scenes): shade reads it with one load instead of prim_mat[slot] -> mats[...]
__global__ void k_slot_mats(int n, const int *__restrict__ prim_mat, const DMaterial *__restrict__ mats, DMaterial *slot_mat) {
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j < n) slot_mat[j] = mats[prim_mat[j]];
}
void launch_slot_mats(const LaunchCfg &c, int n, const int *prim_mat, const DMaterial *mats, DMaterial *slot_mat) {
    if (n <= 0) return;
    k_slot_mats<<<(n + 255) / 256, 256, 0, c.stream>>>(n, prim_mat, mats, slot_mat);
    bump(c);
}

void launch_light_slots(const LaunchCfg &c, int n, const int *prim_obj, const int *obj_light, int *light_slot) {
    if (n <= 0) return;
    k_light_slots<<<(n + 255) / 256, 256, 0, c.stream>>>(n, prim_obj, obj_light, light_slot);
    bump(c);
}

void launch_accumulate(const LaunchCfg &c, const Queues &q, const AccumPlan &plan, int W, int H, int n_samples, int max_depth,
                       unsigned long long *total_rays, const int *const_tiles, float3 bg, int ext_first, int ext_end) {
    int blocks = (W * H + RT_BLOCK - 1) / RT_BLOCK;
    k_accumulate<<<blocks, RT_BLOCK, 0, c.stream>>>(q.radiance, plan, W, H, n_samples, q.qcount, max_depth, total_rays, q.shcount,
                                                    const_tiles, bg, ext_first, ext_end);
    bump(c);
}

void launch_primary_only(const LaunchCfg &c, const DScene &sc, const AccumPlan &plan, unsigned seed, int spp_begin, int n_samples,
                         unsigned long long *total_rays) {
    const int np = sc.width * sc.height;
    if (sc.direct) k_primary_only<true><<<(np + RT_BLOCK - 1) / RT_BLOCK, RT_BLOCK, 0, c.stream>>>(sc, plan, seed, spp_begin, n_samples, total_rays);
    else k_primary_only<false><<<(np + RT_BLOCK - 1) / RT_BLOCK, RT_BLOCK, 0, c.stream>>>(sc, plan, seed, spp_begin, n_samples, total_rays);
    bump(c);
}

void launch_fold_chunks(const LaunchCfg &c, float *accum, const ChunkPtrs &cp, int n_chunks, bool keep, long long n) {
    if (n_chunks <= 0) return;
    k_fold_chunks<<<persistent_grid(c, 8), RT_BLOCK, 0, c.stream>>>(accum, cp, n_chunks, keep ? 1 : 0, n);
    bump(c);
}

void launch_sum_ordered(const LaunchCfg &c, float *accum, const float *const *partials_dev, int n_partials, long long n) {
    k_sum_ordered<<<persistent_grid(c, 8), RT_BLOCK, 0, c.stream>>>(accum, partials_dev, n_partials, n);
    bump(c);
}

void launch_tonemap(const LaunchCfg &c, const float *accum, unsigned char *rgb8, int W, int H, int spp, int gamma) {
    k_tonemap<<<persistent_grid(c, 8), RT_BLOCK, 0, c.stream>>>(accum, rgb8, W, H, spp, gamma);
    bump(c);
}

void launch_aov(const LaunchCfg &c, const DScene &sc, float *out) {
    int np = sc.width * sc.height;
    k_aov<<<(np + RT_BLOCK - 1) / RT_BLOCK, RT_BLOCK, 0, c.stream>>>(sc, out);
    bump(c);
}
