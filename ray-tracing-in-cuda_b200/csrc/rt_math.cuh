// rt_math.cuh — device math for the sm_100a path tracer.
//
// Numeric contract: IEEE binary32, round-to-nearest.  The library is compiled with
// -fmad=false, so a multiply-add is fused ONLY where __fmaf_rn / fmaf is written below; sqrt and
// division are the correctly rounded ones (-prec-sqrt=true -prec-div=true).  The CPU oracle
// (oracle/orc_internal.h) states the same operation sequence with fmaf(), which is what makes the
// sample-for-sample comparison in tests/ possible.  Traversal box tests are exempt (they only prune).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#define RT_PI_F 3.14159274101257324f
#define RT_2PI_F 6.28318548202514648f
#define RT_INV_PI_F 0.318309873342514038f
#define RT_INV_2PI_F 0.159154936671257019f
#define RT_INV_4PI_F 0.0795774683356285095f
#define RT_PIO2_F 1.57079637050628662f

__device__ __forceinline__ float3 v3(float x, float y, float z) { return make_float3(x, y, z); }
__device__ __forceinline__ float3 operator+(float3 a, float3 b) { return v3(a.x + b.x, a.y + b.y, a.z + b.z); }
__device__ __forceinline__ float3 operator-(float3 a, float3 b) { return v3(a.x - b.x, a.y - b.y, a.z - b.z); }
__device__ __forceinline__ float3 operator*(float3 a, float3 b) { return v3(a.x * b.x, a.y * b.y, a.z * b.z); }
__device__ __forceinline__ float3 operator*(float t, float3 a) { return v3(t * a.x, t * a.y, t * a.z); }
__device__ __forceinline__ float3 operator-(float3 a) { return v3(-a.x, -a.y, -a.z); }
// dot(): gpu-version/vec3.cuh:110-112, fusion pinned to two fma steps
__device__ __forceinline__ float dot3(float3 a, float3 b) { return __fmaf_rn(a.z, b.z, __fmaf_rn(a.y, b.y, a.x * b.x)); }
// a + t*b (ray::at, ray.cuh:19)
__device__ __forceinline__ float3 vfma(float t, float3 b, float3 a) {
    return v3(__fmaf_rn(t, b.x, a.x), __fmaf_rn(t, b.y, a.y), __fmaf_rn(t, b.z, a.z));
}
// unit_vector = (1/len) * v (vec3.cuh:108,120)
__device__ __forceinline__ float3 vunit(float3 a) { float r = 1.0f / sqrtf(dot3(a, a)); return r * a; }
__device__ __forceinline__ float3 f4xyz(float4 a) { return v3(a.x, a.y, a.z); }

// ---- shared-algorithm transcendentals (same polynomials and fma order as oracle/orc_math.c) ----
__device__ __forceinline__ float rt_sinf(float x) {
    float k = rintf(x * 0.636619746685028076f);
    float r = __fmaf_rn(-k, 1.57079637050628662f, x);
    r = __fmaf_rn(-k, -4.37113882867379289e-08f, r);
    r = __fmaf_rn(-k, -1.77635683940025046e-15f, r);
    int q = (int)k & 3;
    float s = r * r;
    float ps = __fmaf_rn(s, 2.72425904768169858e-06f, -1.98400259250774980e-04f);
    ps = __fmaf_rn(s, ps, 8.33333190530538559e-03f);
    ps = __fmaf_rn(s, ps, -1.66666671633720398e-01f);
    float sn = __fmaf_rn(r * s, ps, r);
    float pc = __fmaf_rn(s, -2.72958857294724998e-07f, 2.48005617322633043e-05f);
    pc = __fmaf_rn(s, pc, -1.38888880610466003e-03f);
    pc = __fmaf_rn(s, pc, 4.16666679084300995e-02f);
    float cs = __fmaf_rn(s * s, pc, __fmaf_rn(-0.5f, s, 1.0f));
    float res = (q & 1) ? cs : sn;
    return (q & 2) ? -res : res;
}

__device__ __forceinline__ float rt_asin_core(float z) {
    float s = z * z;
    float p = __fmaf_rn(s, 3.38299162685871124e-02f, 1.70643702149391174e-02f);
    p = __fmaf_rn(s, p, 3.11187468469142914e-02f);
    p = __fmaf_rn(s, p, 4.45978343486785889e-02f);
    p = __fmaf_rn(s, p, 7.50009939074516296e-02f);
    p = __fmaf_rn(s, p, 1.66666656732559204e-01f);
    return __fmaf_rn(z * s, p, z);
}

__device__ __forceinline__ float rt_acosf(float x) {
    float ax = fabsf(x);
    if (ax <= 0.5f) return RT_PIO2_F - rt_asin_core(x);
    if (ax > 1.0f) ax = 1.0f;
    float z = sqrtf(__fmaf_rn(-0.5f, ax, 0.5f));
    float a = 2.0f * rt_asin_core(z);
    return x > 0.0f ? a : RT_PI_F - a;
}

__device__ __forceinline__ float rt_atan2f(float y, float x) {
    float ax = fabsf(x), ay = fabsf(y);
    float mx = ax > ay ? ax : ay, mn = ax > ay ? ay : ax;
    float a = mx > 0.0f ? mn / mx : 0.0f;
    float s = a * a;
    float p = __fmaf_rn(s, -2.44705588556826115e-03f, 1.37503892183303833e-02f);
    p = __fmaf_rn(s, p, -3.62703576683998108e-02f);
    p = __fmaf_rn(s, p, 6.28437772393226624e-02f);
    p = __fmaf_rn(s, p, -8.67317989468574524e-02f);
    p = __fmaf_rn(s, p, 1.10379971563816071e-01f);
    p = __fmaf_rn(s, p, -1.42791107296943665e-01f);
    p = __fmaf_rn(s, p, 1.99997663497924805e-01f);
    p = __fmaf_rn(s, p, -3.33333313465118408e-01f);
    float r = __fmaf_rn(a * s, p, a);
    if (ay > ax) r = RT_PIO2_F - r;
    if (x < 0.0f) r = RT_PI_F - r;
    return y < 0.0f ? -r : r;
}

// ---- Philox4x32-10, keyed (seed, pixel) with counter (sample, bounce, block, 'RT') ----
// Replaces the reference's per-pixel XORWOW state array (gpu-version/main.cu:120-125: 48 B/pixel in global
// memory, one RMW per draw) by a stateless counter-based generator: nothing to initialise, nothing to store.
#ifndef RT_PHILOX_ROUNDS
#define RT_PHILOX_ROUNDS 10      // tuning builds only (tools/build_variant.py): the oracle states 10 rounds
#endif
__device__ __forceinline__ uint4 philox4x32_10(uint4 c, uint2 k) {
#pragma unroll
    for (int i = 0; i < RT_PHILOX_ROUNDS; ++i) {
        uint32_t h0 = __umulhi(0xD2511F53u, c.x), l0 = 0xD2511F53u * c.x;
        uint32_t h1 = __umulhi(0xCD9E8D57u, c.z), l1 = 0xCD9E8D57u * c.z;
        c = make_uint4(h1 ^ c.y ^ k.x, l1, h0 ^ c.w ^ k.y, l0);
        k.x += 0x9E3779B9u; k.y += 0xBB67AE85u;
    }
    return c;
}

struct Rng {
    uint2 key;
    uint32_t sample, bounce, block;
    uint4 out;
    int idx;
    uint32_t inc;   // 1 for GPU semantics (0,1], 0 for CPU semantics [0,1)

    __device__ __forceinline__ void init(uint32_t seed, uint32_t pixel, uint32_t s, uint32_t b, int sampling) {
        key = make_uint2(seed, pixel); sample = s; bounce = b; block = 0; idx = 4; inc = sampling == 0 ? 1u : 0u;
    }
    __device__ __forceinline__ float next() {
        if (idx == 4) { out = philox4x32_10(make_uint4(sample, bounce, block, 0x5254u), key); ++block; idx = 0; }
        uint32_t x = idx == 0 ? out.x : idx == 1 ? out.y : idx == 2 ? out.z : out.w;
        ++idx;
        return (float)((x >> 8) + inc) * 5.9604644775390625e-08f;
    }
    // block-wise access for samplers that consume whole blocks (rt_shade.cuh:random_in_unit_sphere)
    __device__ __forceinline__ uint4 next_block() {
        if (idx == 0) { idx = 4; return out; }            // the block prime() drew ahead of its use (idx == 0 occurs in no other state)
        uint4 r = philox4x32_10(make_uint4(sample, bounce, block, 0x5254u), key); ++block; return r;
    }
    // Draws block 0 NOW (same values, same order of use): the shade stage calls it before the material record has arrived, so the ten
    // Philox rounds run under the latency of that load instead of after it.
    __device__ __forceinline__ void prime() { out = philox4x32_10(make_uint4(sample, bounce, 0u, 0x5254u), key); block = 1; idx = 0; }
};
