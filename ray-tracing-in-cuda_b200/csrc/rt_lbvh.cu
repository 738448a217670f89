// rt_lbvh.cu — GPU LBVH build: object boxes -> 63-bit Morton codes -> radix sort -> Karras hierarchy -> refit -> pack.
//
// Replaces the reference's linear object list (gpu-version/object.cuh:11-38 hittable_list; its aabb.hpp is dead
// code).  Integer outputs (sorted keys, sorted object indices, child/parent links) are BIT-EXACT with the serial
// host statement of the same algorithm in oracle/lbvh_host.c; boxes are numerically equal (float min/max are
// order-independent).  Every step is a data-parallel kernel; nothing is built on the host.
#include "rt_kernels.h"

// ---- order-preserving float <-> uint mapping for atomicMin/atomicMax on floats ----
__device__ __forceinline__ unsigned f2ord(float f) { unsigned b = __float_as_uint(f); return b ^ ((b >> 31) ? 0xffffffffu : 0x80000000u); }
__device__ __forceinline__ float ord2f(unsigned u) { return __uint_as_float(u ^ ((u >> 31) ? 0x80000000u : 0xffffffffu)); }

__global__ void k_bounds_init(unsigned *bounds) {
    int i = threadIdx.x;
    if (i < 12) bounds[i] = (i % 6) < 3 ? f2ord(INFINITY) : f2ord(-INFINITY);
}

__device__ __forceinline__ float3 lb_xf_point(const float4 *m, float3 p) { return xf_point(m, p); }

// object box, same operations as oracle/lbvh_host.c orc_object_aabb
__device__ void object_aabb(const DObject &ob, const float4 *xforms, float bmin[3], float bmax[3]) {
    const float *p = ob.p;
    float lo[3], hi[3];
    switch (ob.type) {
    case RT_OBJ_SPHERE: {
        float r = fabsf(p[3]);
        for (int k = 0; k < 3; ++k) { bmin[k] = p[k] - r; bmax[k] = p[k] + r; }
        if (ob.xform < 0) return;
        for (int k = 0; k < 3; ++k) { lo[k] = bmin[k]; hi[k] = bmax[k]; }      // instance: world box of the object-space box's corners
        break;
    }
    case RT_OBJ_XY_RECT: case RT_OBJ_XZ_RECT: case RT_OBJ_YZ_RECT: {
        int a = ob.type == RT_OBJ_YZ_RECT ? 1 : 0, b = ob.type == RT_OBJ_XY_RECT ? 1 : 2;
        int k = ob.type == RT_OBJ_XY_RECT ? 2 : ob.type == RT_OBJ_XZ_RECT ? 1 : 0;
        bmin[a] = fminf(p[0], p[1]); bmax[a] = fmaxf(p[0], p[1]);
        bmin[b] = fminf(p[2], p[3]); bmax[b] = fmaxf(p[2], p[3]);
        bmin[k] = p[4] - 1e-4f; bmax[k] = p[4] + 1e-4f;
        if (ob.xform < 0) return;
        for (int q = 0; q < 3; ++q) { lo[q] = bmin[q]; hi[q] = bmax[q]; }
        break;
    }
    case RT_OBJ_CYLINDER: {
        float r = fabsf(p[0]);
        lo[0] = -r; lo[1] = -r; lo[2] = fminf(p[1], p[2]); hi[0] = r; hi[1] = r; hi[2] = fmaxf(p[1], p[2]);
        break;
    }
    case RT_OBJ_BOX:
        for (int k = 0; k < 3; ++k) { lo[k] = fminf(p[k], p[3 + k]); hi[k] = fmaxf(p[k], p[3 + k]); }
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
    if (ob.xform < 0) { for (int k = 0; k < 3; ++k) { bmin[k] = lo[k]; bmax[k] = hi[k]; } return; }
    for (int k = 0; k < 3; ++k) { bmin[k] = INFINITY; bmax[k] = -INFINITY; }
    const float4 *m = xforms + 6 * ob.xform;
    for (int c = 0; c < 8; ++c) {
        float3 q = xf_point(m, v3((c & 1) ? hi[0] : lo[0], (c & 2) ? hi[1] : lo[1], (c & 4) ? hi[2] : lo[2]));
        float qq[3] = {q.x, q.y, q.z};
        for (int k = 0; k < 3; ++k) { bmin[k] = fminf(bmin[k], qq[k]); bmax[k] = fmaxf(bmax[k], qq[k]); }
    }
}

// min/max of six values per thread over the warp, then one atomic per warp and word into bounds[base .. base + 5]
__device__ __forceinline__ void warp_bounds(float v[6], unsigned *bounds, int base) {
    for (int k = 0; k < 6; ++k) {
        const bool is_min = k < 3;
        float x = v[k];
        for (int o = 16; o > 0; o >>= 1) { float y = __shfl_xor_sync(0xffffffffu, x, o); x = is_min ? fminf(x, y) : fmaxf(x, y); }
        if ((threadIdx.x & 31) == 0) { if (is_min) atomicMin(&bounds[base + k], f2ord(x)); else atomicMax(&bounds[base + k], f2ord(x)); }
    }
}

// step 1: object boxes + scene bounds
__global__ void __launch_bounds__(256) k_obj_aabb(int n, const DObject *__restrict__ objects, const float4 *__restrict__ xforms,
                                                  float *__restrict__ obmin, float *__restrict__ obmax, unsigned *bounds) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    float v[6] = {INFINITY, INFINITY, INFINITY, -INFINITY, -INFINITY, -INFINITY};
    if (i < n) {
        float bmin[3], bmax[3];
        object_aabb(objects[i], xforms, bmin, bmax);
        for (int k = 0; k < 3; ++k) { obmin[3 * i + k] = bmin[k]; obmax[3 * i + k] = bmax[k]; v[k] = bmin[k]; v[3 + k] = bmax[k]; }
    }
    warp_bounds(v, bounds, 0);
}

// step 1b: giants — objects whose largest extent exceeds a quarter of the scene's (the r = 1000 ground sphere of the RTIOW scene,
// the 200 x 200 rects of blue2.json).  In the tree, every box on such an object's ancestor chain would cover the whole scene and
// prune nothing; they stay out of it (key 2^63: the last leaf slots, in scene-file order) and the traversal tests them first.
// info[3] counts the candidates; more than RT_MAX_GIANTS of them (a scene made of large pieces) and none is taken out.
__device__ __forceinline__ float scene_extent(const unsigned *bounds) {
    return fmaxf(ord2f(bounds[3]) - ord2f(bounds[0]), fmaxf(ord2f(bounds[4]) - ord2f(bounds[1]), ord2f(bounds[5]) - ord2f(bounds[2])));
}
__device__ __forceinline__ float object_extent(const float *obmin, const float *obmax, int i) {
    return fmaxf(obmax[3 * i] - obmin[3 * i], fmaxf(obmax[3 * i + 1] - obmin[3 * i + 1], obmax[3 * i + 2] - obmin[3 * i + 2]));
}
__global__ void __launch_bounds__(256) k_giants_mark(int n, const float *__restrict__ obmin, const float *__restrict__ obmax,
                                                     const unsigned *__restrict__ bounds, unsigned char *__restrict__ giant, int *info) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    bool g = false;
    if (i < n) { g = object_extent(obmin, obmax, i) > 0.25f * scene_extent(bounds); giant[i] = g ? 1 : 0; }
    const unsigned m = __ballot_sync(0xffffffffu, g);
    if (m && (threadIdx.x & 31) == 0) atomicAdd(&info[3], __popc(m));
}
// step 1c: centroid bounds of the objects that go into the tree; thread 0 publishes n_tree / n_giants
__global__ void __launch_bounds__(256) k_centroid_bounds(int n, const float *__restrict__ obmin, const float *__restrict__ obmax,
                                                         const unsigned char *__restrict__ giant, unsigned *bounds, int *info) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const int ng = info[3] > RT_MAX_GIANTS ? 0 : info[3];
    if (i == 0) { info[0] = n - ng; info[1] = ng; }
    float v[6] = {INFINITY, INFINITY, INFINITY, -INFINITY, -INFINITY, -INFINITY};
    if (i < n && !(ng && giant[i]))
        for (int k = 0; k < 3; ++k) { const float c = 0.5f * (obmin[3 * i + k] + obmax[3 * i + k]); v[k] = c; v[3 + k] = c; }
    warp_bounds(v, bounds, 6);
}

__device__ __forceinline__ unsigned long long expand21(unsigned v) {
    unsigned long long x = v & 0x1fffffu;
    x = (x | x << 32) & 0x1f00000000ffffULL;
    x = (x | x << 16) & 0x1f0000ff0000ffULL;
    x = (x | x << 8) & 0x100f00f00f00f00fULL;
    x = (x | x << 4) & 0x10c30c30c30c30c3ULL;
    x = (x | x << 2) & 0x1249249249249249ULL;
    return x;
}
__device__ __forceinline__ unsigned quant21(float c, float cmin, float ext) {
    float scale = ext > 0.0f ? 2097152.0f / ext : 0.0f;
    float q = (c - cmin) * scale;
    q = fminf(fmaxf(q, 0.0f), 2097151.0f);
    return (unsigned)q;
}

// step 2: 63-bit Morton code of the box centroid, normalised to the centroid bounds; giants: 2^63
__global__ void __launch_bounds__(256) k_morton(int n, const float *__restrict__ obmin, const float *__restrict__ obmax,
                                                const unsigned *__restrict__ bounds, const unsigned char *__restrict__ giant,
                                                const int *__restrict__ info, unsigned long long *keys, int *vals) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    vals[i] = i;
    if (info[1] && giant[i]) { keys[i] = 0x8000000000000000ull; return; }
    // Trees of more than RT_MORTON_CUBE_MIN leaves (meshes) use ONE scale for the three axes (the largest centroid extent): a
    // cubical Morton grid.  With every axis normalised to its own extent the cells of a flat surface (the terrain mesh: 20 x 1.6 x
    // 20) are thin horizontal slabs and every third split cuts the terrain into height bands that span the whole ground plane:
    // 27.0 -> 18.1 node visits per ray on a 28 800-triangle terrain (tools/visits.py).  Scenes of a few hundred separate objects keep
    // the per-axis scale, which measures better there (RTIOW 9.1 vs 10.8, dna 7.8 vs 9.1 visits per ray: height separates the
    // large spheres from the small ones); the SAH cost of the two trees does not predict this (9.17 vs 9.34 for RTIOW), so the
    // rule is stated on the scene size, not on a cost estimate.
    const bool cube = info[0] > RT_MORTON_CUBE_MIN;
    const float ext = fmaxf(ord2f(bounds[9]) - ord2f(bounds[6]), fmaxf(ord2f(bounds[10]) - ord2f(bounds[7]), ord2f(bounds[11]) - ord2f(bounds[8])));
    unsigned q[3];
    for (int k = 0; k < 3; ++k) {
        const float c = 0.5f * (obmin[3 * i + k] + obmax[3 * i + k]);
        q[k] = quant21(c, ord2f(bounds[6 + k]), cube ? ext : ord2f(bounds[9 + k]) - ord2f(bounds[6 + k]));
    }
    keys[i] = expand21(q[0]) << 2 | expand21(q[1]) << 1 | expand21(q[2]);
}

// ---- step 3: stable LSD radix sort, 8-bit digits, (key, value) pairs ----
#define RS_TILE 1024
#define RS_ROUNDS 4

__global__ void __launch_bounds__(256) k_radix_hist(int n, int n_tiles, const unsigned long long *__restrict__ keys, int shift,
                                                    unsigned *__restrict__ hist) {
    __shared__ unsigned h[256];
    h[threadIdx.x] = 0;
    __syncthreads();
    const int base = blockIdx.x * RS_TILE;
    for (int j = 0; j < RS_ROUNDS; ++j) {
        const int i = base + j * 256 + threadIdx.x;
        if (i < n) atomicAdd(&h[(unsigned)(keys[i] >> shift) & 255u], 1u);
    }
    __syncthreads();
    hist[threadIdx.x * n_tiles + blockIdx.x] = h[threadIdx.x];    // digit-major, so one scan yields global offsets
}

// exclusive scan of `len` words by a single block (len = 256 * n_tiles)
__global__ void __launch_bounds__(1024) k_radix_scan(unsigned *data, int len) {
    __shared__ unsigned warp_sum[32];
    __shared__ unsigned carry;
    if (threadIdx.x == 0) carry = 0;
    __syncthreads();
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    for (int base = 0; base < len; base += 1024) {
        const int i = base + threadIdx.x;
        const unsigned x = i < len ? data[i] : 0u;
        unsigned s = x;
        for (int o = 1; o < 32; o <<= 1) { unsigned y = __shfl_up_sync(0xffffffffu, s, o); if (lane >= o) s += y; }
        if (lane == 31) warp_sum[w] = s;
        __syncthreads();
        if (w == 0) {
            unsigned t = warp_sum[lane], u = t;
            for (int o = 1; o < 32; o <<= 1) { unsigned y = __shfl_up_sync(0xffffffffu, u, o); if (lane >= o) u += y; }
            warp_sum[lane] = u - t;      // exclusive prefix of warp totals
        }
        __syncthreads();
        const unsigned c = carry;
        if (i < len) data[i] = c + warp_sum[w] + s - x;
        __syncthreads();
        if (threadIdx.x == 1023) carry = c + warp_sum[w] + s;
        __syncthreads();
    }
}

__global__ void __launch_bounds__(256) k_radix_scatter(int n, int n_tiles, const unsigned long long *__restrict__ keys,
                                                       const int *__restrict__ vals, int shift, const unsigned *__restrict__ offs,
                                                       unsigned long long *__restrict__ out_keys, int *__restrict__ out_vals) {
    __shared__ unsigned cnt[RS_ROUNDS * 8][256];      // per (round, warp) digit counts -> start offsets
    for (int k = threadIdx.x; k < RS_ROUNDS * 8 * 256; k += 256) (&cnt[0][0])[k] = 0;
    __syncthreads();
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const int base = blockIdx.x * RS_TILE;
    unsigned long long key[RS_ROUNDS]; int val[RS_ROUNDS]; unsigned rank[RS_ROUNDS];
    for (int j = 0; j < RS_ROUNDS; ++j) {
        const int i = base + j * 256 + threadIdx.x;
        const bool valid = i < n;
        key[j] = valid ? keys[i] : 0ull; val[j] = valid ? vals[i] : 0;
        const unsigned digit = (unsigned)(key[j] >> shift) & 255u;
        const unsigned peers = __match_any_sync(0xffffffffu, valid ? digit : (0x100u | lane));
        rank[j] = __popc(peers & ((1u << lane) - 1u));
        if (valid && rank[j] == 0) cnt[j * 8 + w][digit] = __popc(peers);
    }
    __syncthreads();
    {   // one thread per digit: turn the 32 group counts into start offsets (stable: round-major, warp, lane)
        const unsigned d = threadIdx.x;
        unsigned run = offs[d * n_tiles + blockIdx.x];
        for (int g = 0; g < RS_ROUNDS * 8; ++g) { const unsigned c = cnt[g][d]; cnt[g][d] = run; run += c; }
    }
    __syncthreads();
    for (int j = 0; j < RS_ROUNDS; ++j) {
        const int i = base + j * 256 + threadIdx.x;
        if (i < n) {
            const unsigned digit = (unsigned)(key[j] >> shift) & 255u;
            const unsigned pos = cnt[j * 8 + w][digit] + rank[j];
            out_keys[pos] = key[j]; out_vals[pos] = val[j];
        }
    }
}

// step 4: per sorted leaf slot: padded box + the geometry tables the traversal reads (Morton order = memory order)
// A giant sphere's box is read by the camera-beam tile lists only (k_tile_lists): padded by the float error bound of sphere::hit
// for rays that start within D of the centre — m^2 + 2 r m <= 2^-18 (D^2 + r^2), oracle/lbvh_host.c:giant_sphere_pad — instead of
// 2^-10 of its extent (1.95 units for the RTIOW ground sphere, which kept every sky tile of that camera non-empty).
__global__ void __launch_bounds__(256) k_leaf_setup(int n, const int *__restrict__ sorted, const DObject *__restrict__ objects,
                                                    const float *__restrict__ obmin, const float *__restrict__ obmax,
                                                    const unsigned *__restrict__ bounds, const int *__restrict__ info, float4 cam,
                                                    float *__restrict__ leaf_aabb,
                                                    float4 *primA, float4 *primB, float4 *primC, int *prim_obj, int *prim_mat) {
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= n) return;
    const int i = sorted[j];
    const float pad_abs = scene_extent(bounds) * 9.5367431640625e-07f;
    const float e = object_extent(obmin, obmax, i);
    float pad = __fmaf_rn(e, 9.765625e-04f, pad_abs);
    const DObject ob = objects[i];
    if (j >= info[0] && ob.type == RT_OBJ_SPHERE) {
        const float dx = cam.x - ob.p[0], dy = cam.y - ob.p[1], dz = cam.z - ob.p[2];
        const float D = sqrtf(dx * dx + dy * dy + dz * dz) + cam.w, r = fabsf(ob.p[3]);
        const float x = 3.814697265625e-06f * (D * D + r * r);
        pad = x / (sqrtf(r * r + x) + r) + pad_abs;
    }
    for (int k = 0; k < 3; ++k) { leaf_aabb[6 * j + k] = obmin[3 * i + k] - pad; leaf_aabb[6 * j + 3 + k] = obmax[3 * i + k] + pad; }
    const float *p = ob.p;
    float4 A = make_float4(0, 0, 0, 0), B = A, C = A;
    switch (ob.type) {
    case RT_OBJ_SPHERE: A = make_float4(p[0], p[1], p[2], p[3]); B.w = __int_as_float(ob.xform); break;
    case RT_OBJ_XY_RECT: case RT_OBJ_XZ_RECT: case RT_OBJ_YZ_RECT: A = make_float4(p[0], p[1], p[2], p[3]); B.x = p[4]; B.w = __int_as_float(ob.xform); break;
    case RT_OBJ_CYLINDER: A = make_float4(p[0], p[1], p[2], __int_as_float(ob.xform)); break;
    case RT_OBJ_BOX: A = make_float4(p[0], p[1], p[2], __int_as_float(ob.xform)); B = make_float4(p[3], p[4], p[5], 0.0f); break;
    case RT_OBJ_TRIANGLE: {
        // per-vertex texture coordinates: p[9..11] hold (u, v) as 16-bit unorm pairs; all zero = none: (0,0), (1,0), (0,1), which
        // makes the interpolated uv the barycentrics
        unsigned q0 = __float_as_uint(p[9]), q1 = __float_as_uint(p[10]), q2 = __float_as_uint(p[11]);
        if ((q0 | q1 | q2) == 0u) { q1 = 0x0000ffffu; q2 = 0xffff0000u; }
        A = make_float4(p[0], p[1], p[2], __uint_as_float(q0));
        B = make_float4(p[3] - p[0], p[4] - p[1], p[5] - p[2], __uint_as_float(q1));
        C = make_float4(p[6] - p[0], p[7] - p[1], p[8] - p[2], __uint_as_float(q2));
        break;
    }
    }
    primA[j] = A; primB[j] = B; primC[j] = C;
    prim_obj[j] = i; prim_mat[j] = ob.material;
}

// step 5: Karras 2012 — one thread per internal node
__device__ __forceinline__ int lb_delta(const unsigned long long *__restrict__ m, int n, int i, int j) {
    if (j < 0 || j >= n) return -1;
    const unsigned long long x = m[i] ^ m[j];
    if (x) return __clzll((long long)x);
    return 64 + __clz(i ^ j);
}

__global__ void __launch_bounds__(256) k_karras(const int *__restrict__ info, const unsigned long long *__restrict__ keys, int *__restrict__ left,
                                                int *__restrict__ right, int *__restrict__ parent) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const int n = info[0];                      // leaves of the tree (the giants sit behind them in the sorted order)
    if (i >= n - 1) return;
    if (i == 0) parent[0] = -1;
    const int d = lb_delta(keys, n, i, i + 1) - lb_delta(keys, n, i, i - 1) >= 0 ? 1 : -1;
    const int dmin = lb_delta(keys, n, i, i - d);
    int lmax = 2;
    while (lb_delta(keys, n, i, i + lmax * d) > dmin) lmax *= 2;
    int l = 0;
    for (int t = lmax / 2; t >= 1; t /= 2)
        if (lb_delta(keys, n, i, i + (l + t) * d) > dmin) l += t;
    const int j = i + l * d;
    const int dnode = lb_delta(keys, n, i, j);
    int s = 0, t = l;
    do {
        t = (t + 1) >> 1;
        if (lb_delta(keys, n, i, i + (s + t) * d) > dnode) s += t;
    } while (t > 1);
    const int gamma = i + s * d + (d < 0 ? -1 : 0);
    const int lo = i < j ? i : j, hi = i < j ? j : i;
    const int L = lo == gamma ? ~gamma : gamma;
    const int R = hi == gamma + 1 ? ~(gamma + 1) : gamma + 1;
    left[i] = L; right[i] = R;
    if (L >= 0) parent[L] = i; else parent[n - 1 + ~L] = i;
    if (R >= 0) parent[R] = i; else parent[n - 1 + ~R] = i;
}

// step 6: bottom-up refit; the second thread to arrive at a node computes its box (min/max are order-independent)
__global__ void __launch_bounds__(256) k_refit(int *info, const float *__restrict__ leaf_aabb, const int *__restrict__ left,
                                               const int *__restrict__ right, const int *__restrict__ parent, int *flags,
                                               float *node_aabb) {
    const int leaf = blockIdx.x * blockDim.x + threadIdx.x;
    const int n = info[0];
    const bool valid = leaf < n && n >= 2;
    // depth of the tree = the largest number of internal nodes above a leaf: bounds the traversal stack (checked at upload)
    int depth = 0;
    if (valid) for (int c = parent[n - 1 + leaf]; c >= 0; c = parent[c]) ++depth;
    depth = __reduce_max_sync(0xffffffffu, depth);
    if ((threadIdx.x & 31) == 0 && depth) atomicMax(&info[2], depth);
    if (!valid) return;
    int cur = parent[n - 1 + leaf];
    while (cur >= 0) {
        __threadfence();
        if (atomicAdd(&flags[cur], 1) == 0) return;
        const int L = left[cur], R = right[cur];
        const float *a = L >= 0 ? node_aabb + 6 * L : leaf_aabb + 6 * (~L);
        const float *b = R >= 0 ? node_aabb + 6 * R : leaf_aabb + 6 * (~R);
        for (int k = 0; k < 3; ++k) {
            node_aabb[6 * cur + k] = fminf(__ldcg(a + k), __ldcg(b + k));
            node_aabb[6 * cur + 3 + k] = fmaxf(__ldcg(a + 3 + k), __ldcg(b + 3 + k));
        }
        cur = parent[cur];
    }
}

// step 7: pack into 64-byte traversal nodes (both child boxes inline)
__global__ void __launch_bounds__(256) k_pack(int n_all, int *info, const int *__restrict__ left, const int *__restrict__ right,
                                              const float *__restrict__ node_aabb, const float *__restrict__ leaf_aabb,
                                              const int *__restrict__ sorted, const DObject *__restrict__ objects, BvhNode *nodes,
                                              BvhNodeS *nodes_s) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const int n = info[0];
    // leaf codes of the giants (slots n .. n_all - 1) for the traversal prologue; depth of a one-leaf tree
    if (i < info[1]) info[4 + i] = (objects[sorted[n + i]].type << 28) | (n + i);
    if (i == 0 && n == 1) info[2] = 1;
    const int n_nodes = n >= 2 ? n - 1 : 1;
    if (i >= n_nodes) return;
    int L, R;
    if (n >= 2) { L = left[i]; R = right[i]; } else { L = n == 1 ? ~0 : RT_LEAF_NONE; R = RT_LEAF_NONE; }
    float lb[6], rb[6];
    int lc = L, rc = R;
    for (int side = 0; side < 2; ++side) {
        const int c = side ? R : L;
        float *bx = side ? rb : lb;
        if (c == RT_LEAF_NONE) { for (int k = 0; k < 3; ++k) { bx[k] = INFINITY; bx[3 + k] = -INFINITY; } continue; }
        const float *src = c >= 0 ? node_aabb + 6 * c : leaf_aabb + 6 * (~c);
        for (int k = 0; k < 6; ++k) bx[k] = src[k];
        if (c < 0) {
            const int slot = ~c;
            const int code = (objects[sorted[slot]].type << 28) | slot;
            if (side) rc = ~code; else lc = ~code;
        }
    }
    BvhNode nd;
    nd.a = make_float4(lb[0], lb[1], lb[2], lb[3]);
    nd.b = make_float4(lb[4], lb[5], rb[0], rb[1]);
    nd.c = make_float4(rb[2], rb[3], rb[4], rb[5]);
    nd.d = make_int4(lc, rc, 0, 0);
    nodes[i] = nd;
    BvhNodeS ns;
    for (int k = 0; k < 3; ++k) {
        float2 *l = k == 0 ? ns.lx : k == 1 ? ns.ly : ns.lz, *r = k == 0 ? ns.rx : k == 1 ? ns.ry : ns.rz;
        l[0] = make_float2(lb[k], lb[3 + k]); l[1] = make_float2(lb[3 + k], lb[k]);
        r[0] = make_float2(rb[k], rb[3 + k]); r[1] = make_float2(rb[3 + k], rb[k]);
    }
    // inner children as byte offsets into the node array (the traversal adds them to a shared-memory address)
    ns.link = make_int2(lc >= 0 ? lc * (int)sizeof(BvhNodeS) : lc, rc >= 0 ? rc * (int)sizeof(BvhNodeS) : rc);
    nodes_s[i] = ns;
}

// step 8 (trees walked from global memory only): four-wide nodes = every even-depth binary node with its grandchildren
__global__ void __launch_bounds__(256) k_pack4(const int *__restrict__ info, const int *__restrict__ left, const int *__restrict__ right,
                                               const int *__restrict__ parent, const float *__restrict__ node_aabb,
                                               const float *__restrict__ leaf_aabb, const int *__restrict__ sorted,
                                               const DObject *__restrict__ objects, BvhNode4 *nodes4) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const int n = info[0];
    const int n_nodes = n >= 2 ? n - 1 : 1;
    if (i >= n_nodes) return;
    int ch[4] = {RT_LEAF_NONE, RT_LEAF_NONE, RT_LEAF_NONE, RT_LEAF_NONE};      // binary links: >= 0 inner, < 0 ~slot; RT_LEAF_NONE: empty
    int m = 0;
    if (n >= 2) {
        int depth = 0;
        for (int c = parent[i]; c >= 0; c = parent[c]) ++depth;
        if (depth & 1) return;                                 // odd depth: folded into its parent's four-wide node
        const int L = left[i], R = right[i];
        if (L < 0) ch[m++] = L; else { ch[m++] = left[L]; ch[m++] = right[L]; }
        if (R < 0) ch[m++] = R; else { ch[m++] = left[R]; ch[m++] = right[R]; }
    } else if (n == 1) ch[m++] = ~0;
    float pl[6][4]; int ln[4];
    for (int k = 0; k < 4; ++k) {
        const int c = ch[k];
        if (c == RT_LEAF_NONE) { for (int a = 0; a < 3; ++a) { pl[a][k] = INFINITY; pl[3 + a][k] = -INFINITY; } ln[k] = RT_LEAF_NONE; continue; }
        const float *src = c >= 0 ? node_aabb + 6 * c : leaf_aabb + 6 * (~c);
        for (int a = 0; a < 6; ++a) pl[a][k] = src[a];
        ln[k] = c >= 0 ? c : ~((objects[sorted[~c]].type << 28) | (~c));
    }
    BvhNode4 nd;
    for (int a = 0; a < 6; ++a) nd.plane[a] = make_float4(pl[a][0], pl[a][1], pl[a][2], pl[a][3]);
    nd.link = make_int4(ln[0], ln[1], ln[2], ln[3]); nd.pad = make_int4(0, 0, 0, 0);
    nodes4[i] = nd;
}

void lbvh_pack4(const LaunchCfg &c, const LbvhBuffers &b, BvhNode4 *nodes4) {
    const int g = b.n > 0 ? (b.n + 255) / 256 : 1;
    k_pack4<<<g, 256, 0, c.stream>>>(b.info, b.left, b.right, b.parent, b.node_aabb, b.leaf_aabb, b.vals[0], b.objects, nodes4);
    if (c.launches) ++*c.launches;
}

size_t lbvh_hist_words(int n) { return (size_t)256 * (size_t)((n + RS_TILE - 1) / RS_TILE > 0 ? (n + RS_TILE - 1) / RS_TILE : 1); }

void lbvh_build(const LaunchCfg &c, LbvhBuffers &b) {
    const int n = b.n;
    cudaStream_t s = c.stream;
    auto bump = [&](int k = 1) { if (c.launches) *c.launches += k; };
    const int g = n > 0 ? (n + 255) / 256 : 1;
    k_bounds_init<<<1, 32, 0, s>>>(b.bounds); bump();
    cudaMemsetAsync(b.info, 0, sizeof(int) * RT_LBVH_INFO_WORDS, s);
    if (n > 0) {
        k_obj_aabb<<<g, 256, 0, s>>>(n, b.objects, b.xforms, b.obj_bmin, b.obj_bmax, b.bounds); bump();
        k_giants_mark<<<g, 256, 0, s>>>(n, b.obj_bmin, b.obj_bmax, b.bounds, b.giant, b.info); bump();
        k_centroid_bounds<<<g, 256, 0, s>>>(n, b.obj_bmin, b.obj_bmax, b.giant, b.bounds, b.info); bump();
        k_morton<<<g, 256, 0, s>>>(n, b.obj_bmin, b.obj_bmax, b.bounds, b.giant, b.info, b.keys[0], b.vals[0]); bump();
        const int n_tiles = (n + RS_TILE - 1) / RS_TILE;
        int cur = 0;
        for (int pass = 0; pass < 8; ++pass) {
            k_radix_hist<<<n_tiles, 256, 0, s>>>(n, n_tiles, b.keys[cur], pass * 8, b.hist);
            k_radix_scan<<<1, 1024, 0, s>>>(b.hist, 256 * n_tiles);
            k_radix_scatter<<<n_tiles, 256, 0, s>>>(n, n_tiles, b.keys[cur], b.vals[cur], pass * 8, b.hist, b.keys[cur ^ 1], b.vals[cur ^ 1]);
            bump(3);
            cur ^= 1;
        }
        // 8 passes: the sorted data is back in buffer 0
        k_leaf_setup<<<g, 256, 0, s>>>(n, b.vals[0], b.objects, b.obj_bmin, b.obj_bmax, b.bounds, b.info, b.cam, b.leaf_aabb, b.primA, b.primB,
                                       b.primC, b.prim_obj, b.prim_mat); bump();
        if (n >= 2) {           // the tree has n_tree = info[0] <= n leaves: the kernels read it on the device
            cudaMemsetAsync(b.flags, 0, sizeof(int) * (size_t)(n - 1), s);
            k_karras<<<g, 256, 0, s>>>(b.info, b.keys[0], b.left, b.right, b.parent); bump();
            k_refit<<<g, 256, 0, s>>>(b.info, b.leaf_aabb, b.left, b.right, b.parent, b.flags, b.node_aabb); bump();
        }
    }
    k_pack<<<g, 256, 0, s>>>(n, b.info, b.left, b.right, b.node_aabb, b.leaf_aabb, b.vals[0], b.objects, b.nodes, b.nodes_s); bump();
}
