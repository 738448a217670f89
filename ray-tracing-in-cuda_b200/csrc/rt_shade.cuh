// rt_shade.cuh — textures, materials, camera rays and miss colour as device functions.
// Same operation order as oracle/orc_render.c; reference citations per function.
#pragma once
#include "rt_scene.cuh"

// texture.cuh:24-27 (solid), 44-52 (checker); image: CUDA texture unit, point sampling, unnormalised coordinates
__device__ __forceinline__ float3 texture_value(const DScene &sc, int tex, float u, float v, float3 p) {
#pragma unroll 1
    for (int guard = 0; guard < 8; ++guard) {
        const DTexture t = sc.texs[tex];
        if (t.type == RT_TEX_SOLID) return f4xyz(t.color);
        if (t.type == RT_TEX_CHECKER) {
            float sines = rt_sinf(10.0f * p.x) * rt_sinf(10.0f * p.y) * rt_sinf(10.0f * p.z);
            tex = sines < 0.0f ? t.odd : t.even;
            continue;
        }
        if (t.type == RT_TEX_IMAGE) {
            const DImage im = sc.images[t.image];
            float uu = u < 0.0f ? 0.0f : (u > 1.0f ? 1.0f : u);
            float vv = 1.0f - (v < 0.0f ? 0.0f : (v > 1.0f ? 1.0f : v));
            int i = (int)(uu * (float)im.width), j = (int)(vv * (float)im.height);
            if (i >= im.width) i = im.width - 1;
            if (j >= im.height) j = im.height - 1;
            // texel centre (i+0.5, j+0.5): the texture unit fetches texel (i, j); uchar4 -> float in [0,1] (= b/255)
            float4 c = tex2D<float4>(im.tex, (float)i + 0.5f, (float)j + 0.5f);
            return v3(c.x, c.y, c.z);
        }
        break;
    }
    return v3(0, 0, 0);
}

// Colour of a lambertian / diffuse_light material at (u, v, p).  A material whose texture is a plain solid_color carries the colour
// itself (texture < 0, resolved at upload): one dependent table load less per shaded vertex.
__device__ __forceinline__ float3 material_color(const DScene &sc, const DMaterial &m, float u, float v, float3 p) {
    return m.texture < 0 ? f4xyz(m.albedo) : texture_value(sc, m.texture, u, v, p);
}

// replayed uniforms for the deterministic AOV pass (same sequence as the oracle and the reference harness)
struct FixedRng {
    int pos = 0;
    __device__ __forceinline__ float next() {
        const int k = pos++ & 7;
        return k < 3 ? 0.5f : (k < 6 ? 0.25f : 0.1f);
    }
};

// random_in_unit_sphere: vec3.cuh:133-141 (octant, (0,1] uniforms) | cmake-cpu-version/vec3.h:121-129 (full cube)
__device__ __forceinline__ bool ball_trial(float a, float b, float c, int sampling, float3 &p) {
    p = sampling == RT_SAMPLING_REFERENCE_GPU ? v3(a, b, c)
                                              : v3(__fmaf_rn(2.0f, a, -1.0f), __fmaf_rn(2.0f, b, -1.0f), __fmaf_rn(2.0f, c, -1.0f));
    return !(dot3(p, p) >= 1.0f);
}
template <class R>
__device__ __forceinline__ float3 random_in_unit_sphere(R &rng, int sampling) {
    for (;;) {
        float a = rng.next(), b = rng.next(), c = rng.next();
        float3 p;
        if (ball_trial(a, b, c, sampling, p)) return p;
    }
}
// Philox stream: the rejection sampler consumes WHOLE blocks and cuts every 128-bit block into two trials of three 21-bit
// uniforms (oracle/orc_render.c:random_in_unit_sphere states the same packing).  Trial 1 = the low 21 bits of words x, y, z;
// trial 2 = their high 11 bits extended by 10 bits each of word w.  21-bit uniforms convert to float exactly, 2u-1 is
// exact as well, and a trial costs half a Philox block instead of three quarters: the loop (acceptance pi/6 per trial, the
// warp waits for its slowest lane) is the largest instruction block of k_shade.  Draws after the sampler start a new block.
__device__ __forceinline__ float u21(uint32_t x, uint32_t inc) { return (float)(x + inc) * 4.76837158203125e-07f; }
template <>
__device__ __forceinline__ float3 random_in_unit_sphere<Rng>(Rng &rng, int sampling) {
    float3 p;
    for (;;) {
        const uint4 B = rng.next_block();
        if (ball_trial(u21(B.x & 0x1fffffu, rng.inc), u21(B.y & 0x1fffffu, rng.inc), u21(B.z & 0x1fffffu, rng.inc), sampling, p)) break;
        if (ball_trial(u21((B.x >> 21) | ((B.w & 0x3ffu) << 11), rng.inc), u21((B.y >> 21) | ((B.w & 0xffc00u) << 1), rng.inc),
                       u21((B.z >> 21) | ((B.w >> 9) & 0x1ff800u), rng.inc), sampling, p)) break;
    }
    rng.idx = 4;
    return p;
}

// ---- direct draws (RT_SAMPLING_DIRECT): the same distributions without the rejection loops ----
// The reference's samplers are rejection loops (vec3.cuh:133-141, vec3.h:121-129,157-165): acceptance pi/6 per trial for the ball,
// so a warp waits for its unluckiest lane and the loop ran at ~10 of 32 lanes (half the instructions of k_shade, ncu r01f).
// The reference is matched statistically only (its generators are XORWOW / mt19937, ours Philox), so the same distributions
// may be drawn directly from five 24-bit integers k1..k5 of ONE Philox block:
//   direction  z = 1 - 2 k1/2^24, angle 2 pi k2/2^24 (- pi/4: a constant turn keeps a uniform angle uniform, and the octant
//              reduction below then needs no further shift), r = sqrt(1 - z^2)              -> uniform on the unit sphere
//   radius     max(k3, k4, k5) has density 3 r^2: the radial law of the uniform ball         -> unit * radius uniform in the ball
// GPU family (vec3.cuh:133-141 samples the positive octant only): componentwise |.| of the above.
// cos / sin of the angle: quarter turns from the top two bits, the polynomial pair of rt_sinf on [-pi/4, pi/4) for the rest.
__device__ __forceinline__ void rt_sincos_turn24(uint32_t k, float &c, float &s) {
    const float th = (float)((int)(k & 0x3fffffu) - 0x200000) * 3.74507398e-07f;      // (pi/2) 2^-22 : theta in [-pi/4, pi/4)
    const float s2 = th * th;
    float ps = __fmaf_rn(s2, 2.72425904768169858e-06f, -1.98400259250774980e-04f);
    ps = __fmaf_rn(s2, ps, 8.33333190530538559e-03f);
    ps = __fmaf_rn(s2, ps, -1.66666671633720398e-01f);
    const float sn = __fmaf_rn(th * s2, ps, th);
    float pc = __fmaf_rn(s2, -2.72958857294724998e-07f, 2.48005617322633043e-05f);
    pc = __fmaf_rn(s2, pc, -1.38888880610466003e-03f);
    pc = __fmaf_rn(s2, pc, 4.16666679084300995e-02f);
    const float cs = __fmaf_rn(s2 * s2, pc, __fmaf_rn(-0.5f, s2, 1.0f));
    const uint32_t q = k >> 22;
    c = (q & 1u) ? sn : cs; s = (q & 1u) ? cs : sn;
    if (q == 1u || q == 2u) c = -c;
    if (q >= 2u) s = -s;
}
// unit: uniform on the sphere; rad: radius of a uniform ball point (only metal fuzz and the GPU family's lambertian use it)
__device__ __forceinline__ void ball_direct_k(uint32_t k1, uint32_t k2, uint32_t k3, uint32_t k4, uint32_t k5, float3 &unit, float &rad) {
    const float z = __fmaf_rn(-2.0f, (float)k1 * 5.9604644775390625e-08f, 1.0f);
    const float r = sqrtf(fmaxf(__fmaf_rn(-z, z, 1.0f), 0.0f));
    float c, s;
    rt_sincos_turn24(k2, c, s);
    unit = v3(r * c, r * s, z);
    const uint32_t m = max(k3, max(k4, k5));
    rad = (float)(m + 1u) * 5.9604644775390625e-08f;
}
__device__ __forceinline__ uint32_t k24_of(float u) { const uint32_t k = (uint32_t)(u * 16777216.0f); return k > 0xffffffu ? 0xffffffu : k; }
template <class R>
__device__ __forceinline__ void ball_direct(R &rng, float3 &unit, float &rad) {      // replayed uniforms: five draws as they come
    const float a = rng.next(), b = rng.next(), c = rng.next(), d = rng.next(), e = rng.next();
    ball_direct_k(k24_of(a), k24_of(b), k24_of(c), k24_of(d), k24_of(e), unit, rad);
}
template <>
__device__ __forceinline__ void ball_direct<Rng>(Rng &rng, float3 &unit, float &rad) {
    const uint4 B = rng.next_block();
    ball_direct_k(B.x >> 8, B.y >> 8, B.z >> 8, B.w >> 8, (B.x & 0xffu) | ((B.y & 0xffu) << 8) | ((B.z & 0xffu) << 16), unit, rad);
    rng.idx = 4;
}

__device__ __forceinline__ float3 reflect3(float3 v, float3 n) {   // vec3.cuh:121-123
    float k = 2.0f * dot3(v, n);
    return v3(__fmaf_rn(-k, n.x, v.x), __fmaf_rn(-k, n.y, v.y), __fmaf_rn(-k, n.z, v.z));
}

// material::scatter for the four material classes (material.cuh:37-52, 63-70, 130-158, 169-173).
// Returns false when the path ends here.  att: attenuation; sd: scattered direction (origin is rec.p).
// The direct sampler's unit vector and radius of a primed generator (Rng::prime), computed AHEAD of the material record: scatter() takes
// them instead of drawing (the block is the one ball_direct would have consumed, so the values are the same).
struct BallDraw { float3 unit; float rad; };
__device__ __forceinline__ BallDraw ball_direct_ahead(const Rng &rng) {
    const uint4 B = rng.out;
    BallDraw b;
    ball_direct_k(B.x >> 8, B.y >> 8, B.z >> 8, B.w >> 8, (B.x & 0xffu) | ((B.y & 0xffu) << 8) | ((B.z & 0xffu) << 16), b.unit, b.rad);
    return b;
}

template <bool DIRECT, class R>
__device__ __forceinline__ bool scatter(const DScene &sc, const DMaterial &m, float3 rd, const HitRec &rec, R &rng,
                                        float3 &att, float3 &sd, const BallDraw *ahead = nullptr) {
    // lambertian and metal draw the same random_in_unit_sphere(): one shared (convergent) rejection loop for both classes,
    // then the short class-specific tails — a warp holding a mix of the two pays for the expensive part once
    float3 r = v3(0, 0, 0);
    if (m.type == RT_MAT_LAMBERTIAN || m.type == RT_MAT_METAL) {
        if (DIRECT) {
            float rad;
            if (ahead) { r = ahead->unit; rad = ahead->rad; }
            else ball_direct(rng, r, rad);
            if (sc.sampling == RT_SAMPLING_REFERENCE_GPU) r = rad * v3(fabsf(r.x), fabsf(r.y), fabsf(r.z));     // octant ball for both classes
            else if (m.type == RT_MAT_METAL) r = rad * r;                                                        // ball | unit vector
        } else r = random_in_unit_sphere(rng, sc.sampling);
    }
    switch (m.type) {
    case RT_MAT_LAMBERTIAN: {
        if (!DIRECT && sc.sampling == RT_SAMPLING_REFERENCE_CPU) r = vunit(r);     // random_unit_vector (material.h:27)
        float3 dir = rec.n + r;
        if (fabsf(dir.x) < 1e-8f && fabsf(dir.y) < 1e-8f && fabsf(dir.z) < 1e-8f) dir = rec.n;
        sd = dir;
        att = material_color(sc, m, rec.u, rec.v, rec.p);
        return true;
    }
    case RT_MAT_METAL: {
        float3 refl = reflect3(vunit(rd), rec.n);
        sd = vfma(m.fuzz, r, refl);
        att = f4xyz(m.albedo);
        return dot3(sd, rec.n) > 0.0f;
    }
    case RT_MAT_DIELECTRIC: {
        att = v3(1.0f, 1.0f, 1.0f);
        float ratio = rec.front ? (1.0f / m.ior) : m.ior;
        float3 ud = vunit(rd);
        float cos_theta = fminf(-dot3(ud, rec.n), 1.0f);
        float sin_theta = sqrtf(__fmaf_rn(-cos_theta, cos_theta, 1.0f));
        bool do_reflect = ratio * sin_theta > 1.0f;
        if (!do_reflect) {
            float r0 = (1.0f - ratio) / (1.0f + ratio);
            r0 = r0 * r0;
            float x = 1.0f - cos_theta, x2 = x * x;
            float refl = __fmaf_rn(1.0f - r0, x2 * x2 * x, r0);
            do_reflect = refl > rng.next();
        }
        if (do_reflect) sd = reflect3(ud, rec.n);
        else {
            float3 perp = ratio * vfma(cos_theta, rec.n, ud);
            float k = -sqrtf(fabsf(1.0f - dot3(perp, perp)));
            sd = vfma(k, rec.n, perp);
        }
        return true;
    }
    default:
        att = v3(0, 0, 0); sd = v3(0, 0, 0);
        return false;
    }
}

// Position k of a sample plane -> pixel index.  Camera paths are laid out in 8x4 pixel tiles (one warp = one tile), so the
// 32 primary rays of a warp form a compact bundle instead of a 32x1 strip; pixels outside the tiled W8 x H4 region
// (right and top margins when the image size is not a multiple of the tile) follow in row order.  Pure bookkeeping:
// the Philox key is the true pixel index, so images do not depend on this layout.
__device__ __forceinline__ int plane_to_pixel(int W, int H, int k) {
    const int W8 = W & ~7, H4 = H & ~3, nt = W8 * H4;
    int x, y;
    if (k < nt) {
        const int tile = k >> 5, j = k & 31, tiles_x = W8 >> 3;
        const int ty = tile / tiles_x, tx = tile - ty * tiles_x;
        x = tx * 8 + (j & 7); y = ty * 4 + (j >> 3);
    } else {
        int r = k - nt;
        const int wr = W - W8;
        if (r < wr * H4) { y = r / wr; x = W8 + (r - y * wr); }
        else { r -= wr * H4; y = H4 + r / W; x = r - (r / W) * W; }
    }
    return y * W + x;
}

// ---- next-event estimation: one point on one light (same operation order as oracle/orc_nee.c) ----
__device__ __forceinline__ float rt_cosf(float x) { return rt_sinf(x + RT_PIO2_F); }

// Samples light k's surface with (u1, u2).  Returns false when x cannot receive light from the sampled point; else
// w = y - x (the occlusion ray, to be tested on t in [0.001, 0.999]) and f = Le(y) * cos_x * cos_y / |w|^2 * area * n_lights / pi.
template <class R>
__device__ __forceinline__ bool light_sample(const DScene &sc, float3 x, float3 n, R &rng, float3 &w, float3 &f) {
    const float u0 = rng.next(), u1 = rng.next(), u2 = rng.next();
    int k = (int)(u0 * (float)sc.n_lights);
    if (k > sc.n_lights - 1) k = sc.n_lights - 1;
    const int slot = sc.light_slot[k];
    const int type = sc.light_type[k];
    const float4 A = sc.primA[slot];
    float3 y, ny; float tu = 0.0f, tv = 0.0f;
    switch (type) {
    case RT_OBJ_SPHERE: {
        const float z = __fmaf_rn(-2.0f, u1, 1.0f);
        const float r = sqrtf(fmaxf(0.0f, __fmaf_rn(-z, z, 1.0f)));
        const float phi = RT_2PI_F * u2;
        const float3 dir = v3(r * rt_cosf(phi), r * rt_sinf(phi), z);
        ny = dir;
        y = vfma(fabsf(A.w), dir, f4xyz(A));
        const float theta = rt_acosf(-dir.y), ph = rt_atan2f(-dir.z, dir.x) + RT_PI_F;
        tu = ph * RT_INV_2PI_F; tv = theta * RT_INV_PI_F;
        const int xi = __float_as_int(sc.primB[slot].w);      // rotate / translate instance: the point and the normal go to world space
        if (xi >= 0) { const float4 *xf = sc.xforms + 6 * xi; y = xf_point(xf, y); ny = vunit(xf_normal(xf + 3, ny)); }
        break;
    }
    case RT_OBJ_XY_RECT: case RT_OBJ_XZ_RECT: case RT_OBJ_YZ_RECT: {
        const float4 B = sc.primB[slot];
        const float a = __fmaf_rn(u1, A.y - A.x, A.x), b = __fmaf_rn(u2, A.w - A.z, A.z), kk = B.x;
        if (type == RT_OBJ_XY_RECT) { y = v3(a, b, kk); ny = v3(0, 0, 1); }
        else if (type == RT_OBJ_XZ_RECT) { y = v3(a, kk, b); ny = v3(0, 1, 0); }
        else { y = v3(kk, a, b); ny = v3(1, 0, 0); }
        tu = (a - A.x) / (A.y - A.x); tv = (b - A.z) / (A.w - A.z);
        const int xi = __float_as_int(B.w);
        if (xi >= 0) { const float4 *xf = sc.xforms + 6 * xi; y = xf_point(xf, y); ny = vunit(xf_normal(xf + 3, ny)); }
        break;
    }
    case RT_OBJ_CYLINDER: {
        const float4 *xf = sc.xforms + 6 * __float_as_int(A.w);
        const float phi = RT_2PI_F * u1, z = __fmaf_rn(u2, A.z - A.y, A.y);
        const float c = rt_cosf(phi), s = rt_sinf(phi);
        const float3 op = v3(A.x * c, A.x * s, z);
        y = xf_point(xf, op);
        ny = vunit(xf_normal(xf + 3, v3(c, s, 0.0f)));
        tu = (rt_atan2f(op.y, op.x) + RT_2PI_F) * RT_INV_4PI_F; tv = (z - A.y) / (A.z - A.y);
        break;
    }
    case RT_OBJ_BOX: {      // A = lo (w: instance transform), B = hi; u1 picks the face in proportion to its area and is re-used inside it
        const float4 B = sc.primB[slot];
        const float L[3] = {A.x, A.y, A.z}, H[3] = {B.x, B.y, B.z};
        const float e[3] = {fabsf(B.x - A.x), fabsf(B.y - A.y), fabsf(B.z - A.z)};
        const float wf[3] = {e[1] * e[2], e[2] * e[0], e[0] * e[1]};
        const float v = u1 * (2.0f * ((wf[0] + wf[1]) + wf[2]));
        int fc = 0; float c = 0.0f;
        while (fc < 5 && !(v < c + wf[fc >> 1])) { c += wf[fc >> 1]; ++fc; }
        const float s1 = fminf(fmaxf((v - c) / wf[fc >> 1], 0.0f), 0.99999994f);
        const int ax = fc >> 1, ua = (ax + 1) % 3, va = (ax + 2) % 3;
        float P[3], nn[3] = {0.0f, 0.0f, 0.0f};
        P[ax] = (fc & 1) ? H[ax] : L[ax]; nn[ax] = (fc & 1) ? 1.0f : -1.0f;
        P[ua] = __fmaf_rn(s1, H[ua] - L[ua], L[ua]); P[va] = __fmaf_rn(u2, H[va] - L[va], L[va]);
        tu = (P[ua] - L[ua]) / (H[ua] - L[ua]); tv = (P[va] - L[va]) / (H[va] - L[va]);      // hit_record's box uv
        y = v3(P[0], P[1], P[2]); ny = v3(nn[0], nn[1], nn[2]);
        const int xi = __float_as_int(A.w);
        if (xi >= 0) { const float4 *xf = sc.xforms + 6 * xi; y = xf_point(xf, y); ny = vunit(xf_normal(xf + 3, ny)); }
        break;
    }
    default: {   // RT_OBJ_TRIANGLE: A = v0, B = e1, C = e2
        const float3 e1 = f4xyz(sc.primB[slot]), e2 = f4xyz(sc.primC[slot]);
        const float su = sqrtf(u1);
        const float b1 = u2 * su, b2 = 1.0f - su;
        y = vfma(b2, e2, vfma(b1, e1, f4xyz(A)));
        const float3 c = v3(__fmaf_rn(e1.y, e2.z, -(e1.z * e2.y)), __fmaf_rn(e1.z, e2.x, -(e1.x * e2.z)), __fmaf_rn(e1.x, e2.y, -(e1.y * e2.x)));
        ny = vunit(c);
        tu = b1; tv = b2;
        break;
    }
    }
    w = y - x;
    const float d2 = dot3(w, w);
    if (!(d2 > 0.0f)) return false;
    const float inv_d = 1.0f / sqrtf(d2);
    const float cs = dot3(n, w) * inv_d, cl = fabsf(dot3(ny, w)) * inv_d;
    if (!(cs > 0.0f) || !(cl > 0.0f)) return false;
    float g = ((cs * cl) / d2) * (sc.light_area[k] * ((float)sc.n_lights * RT_INV_PI_F));
    if (sc.nee_mis) {       // balance heuristic against the cosine-weighted scatter (oracle/orc_nee.c)
        const float pl = d2 / (cl * (sc.light_area[k] * (float)sc.n_lights)), pb = cs * RT_INV_PI_F;
        g = g * (pl / (pl + pb));
    }
    const DMaterial m = sc.mats[sc.prim_mat[slot]];
    f = g * material_color(sc, m, tu, tv, y);
    return true;
}

// RT_RENDER_NEE_MIS: weight of the emission a ray scattered at a lambertian vertex (d = n + unit vector, not normalised: cos_x = |d| / 2)
// finds on sampled light k at parameter t, unit normal ny (oracle/orc_nee.c:orc_mis_hit_weight)
__device__ __forceinline__ float mis_hit_weight(const DScene &sc, int k, float3 d, float t, float3 ny) {
    const float dd = dot3(d, d), len = sqrtf(dd);
    const float dist2 = (t * t) * dd;
    const float cy = fabsf(dot3(ny, d)) / len;
    const float pl = dist2 / (cy * (sc.light_area[k] * (float)sc.n_lights)), pb = (len * 0.5f) * RT_INV_PI_F;
    return pb / (pb + pl);
}

// miss colour: constant background (main.cu:61-63) or the RTIOW sky gradient (cmake-cpu-version/main.cpp:36-38)
__device__ __forceinline__ float3 background(const DScene &sc, float3 d) {
    if (sc.bg_mode == RT_BG_SKY_GRADIENT) {
        float3 ud = vunit(d);
        float t = 0.5f * (ud.y + 1.0f), w = 1.0f - t;
        return v3(__fmaf_rn(t, 0.5f, w), __fmaf_rn(t, 0.7f, w), __fmaf_rn(t, 1.0f, w));
    }
    return sc.background;
}

// camera::get_ray, camera.cuh:31-39 (no lens) / cmake-cpu-version/camera.h:32-39 (thin lens)
template <bool DIRECT = false>
__device__ __forceinline__ void camera_ray(const DCamera &cam, float s, float t, bool defocus, Rng &rng, float3 &o, float3 &d) {
    float rx = 0.0f, ry = 0.0f;
    if (defocus && DIRECT) {      // uniform disk without the loop: radius sqrt(u1), angle from u2
        const float u1 = rng.next(), u2 = rng.next();
        float c, sn;
        rt_sincos_turn24(k24_of(u2), c, sn);
        const float rr = sqrtf(u1);
        rx = cam.lens_radius * (rr * c); ry = cam.lens_radius * (rr * sn);
    } else if (defocus) {
        for (;;) {
            float px = __fmaf_rn(2.0f, rng.next(), -1.0f), py = __fmaf_rn(2.0f, rng.next(), -1.0f);
            if (__fmaf_rn(py, py, px * px) >= 1.0f) continue;
            rx = cam.lens_radius * px; ry = cam.lens_radius * py;
            break;
        }
    }
    float3 off = v3(__fmaf_rn(cam.v.x, ry, cam.u.x * rx), __fmaf_rn(cam.v.y, ry, cam.u.y * rx), __fmaf_rn(cam.v.z, ry, cam.u.z * rx));
    o = cam.origin + off;
    float3 a = vfma(t, cam.vertical, vfma(s, cam.horizontal, cam.llc));
    d = (a - cam.origin) - off;
}
