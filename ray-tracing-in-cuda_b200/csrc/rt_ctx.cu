// rt_ctx.cu — C ABI implementation (include/rt.h): context, scene upload (flatten + GPU LBVH build), the wavefront
// pass loop, download/tonemap.  Replaces jsonmain() of gpu-version/main.cu:448-522 minus file I/O.
#include "rt_ctx.h"
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <fstream>
#include <sstream>
#include <nvtx3/nvToolsExt.h>      // header-only NVTX v3: a no-op unless a profiler injects itself (ranges per stage)

static thread_local std::string g_create_err;

int rt_fail(rt_ctx *ctx, int code, const std::string &msg) {
    if (ctx) ctx->err = msg; else g_create_err = msg;
    return code;
}

#define CU(call)                                                                                            \
    do {                                                                                                    \
        cudaError_t e__ = (call);                                                                           \
        if (e__ != cudaSuccess)                                                                             \
            return rt_fail(ctx, RT_ERR_CUDA, std::string(#call) + ": " + cudaGetErrorString(e__));          \
    } while (0)

void HostScene::fix_pointers() {
    d.n_objects = (int)objects.size(); d.n_xforms = (int)xforms.size(); d.n_materials = (int)materials.size();
    d.n_textures = (int)textures.size(); d.n_images = (int)images.size();
    for (size_t i = 0; i < images.size(); ++i) images[i].rgb8 = image_data[i].data();
    d.objects = objects.data(); d.xforms = xforms.data(); d.materials = materials.data(); d.textures = textures.data();
    d.images = images.data();
}

extern "C" int rt_abi_version(void) { return RT_ABI_VERSION; }

extern "C" const char *rt_last_error(const rt_ctx *ctx) { return ctx ? ctx->err.c_str() : g_create_err.c_str(); }

extern "C" int rt_create(rt_ctx **out, int device, void *stream) {
    rt_ctx *ctx = nullptr;   // errors before the ctx exists go to the thread-local slot
    if (!out) return rt_fail(nullptr, RT_ERR_INVALID, "rt_create: out is NULL");
    *out = nullptr;
    int count = 0;
    cudaError_t e = cudaGetDeviceCount(&count);
    if (e != cudaSuccess || count == 0)
        return rt_fail(nullptr, RT_ERR_NO_DEVICE, std::string("no CUDA device: ") + cudaGetErrorString(e) +
                                                      " (this library has no CPU fallback)");
    if (device < 0 || device >= count) return rt_fail(nullptr, RT_ERR_INVALID, "rt_create: bad device ordinal");
    CU(cudaSetDevice(device));
    rt_ctx *c = new rt_ctx();
    c->device = device;
    cudaDeviceProp prop;
    e = cudaGetDeviceProperties(&prop, device);
    if (e != cudaSuccess) { delete c; return rt_fail(nullptr, RT_ERR_CUDA, cudaGetErrorString(e)); }
    c->num_sms = prop.multiProcessorCount;
    c->smem_optin = prop.sharedMemPerBlockOptin;
    if (stream) { c->stream = (cudaStream_t)stream; c->own_stream = false; }
    else {
        e = cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking);
        if (e != cudaSuccess) { delete c; return rt_fail(nullptr, RT_ERR_CUDA, cudaGetErrorString(e)); }
        c->own_stream = true;
    }
    cudaEventCreate(&c->ev0); cudaEventCreate(&c->ev1); cudaEventCreate(&c->ev_count[0]); cudaEventCreate(&c->ev_count[1]);
    for (int k = 1; k < RT_MAX_SETS; ++k) cudaStreamCreateWithFlags(&c->extra_stream[k], cudaStreamNonBlocking);
    cudaEventCreateWithFlags(&c->ev_fork, cudaEventDisableTiming);
    for (int k = 0; k < RT_MAX_SETS; ++k) cudaEventCreateWithFlags(&c->qs[k].ev_acc, cudaEventDisableTiming);
    c->cfg.stream = c->stream; c->cfg.num_sms = c->num_sms; c->cfg.launches = &c->launches;
    e = kernels_configure_device(device, c->smem_optin);
    if (e != cudaSuccess) { std::string m = std::string("kernel configuration: ") + cudaGetErrorString(e); rt_destroy(c); return rt_fail(nullptr, RT_ERR_CUDA, m); }
    *out = c;
    return RT_OK;
}

static void free_images(rt_ctx *c) {
    for (auto t : c->img_tex) cudaDestroyTextureObject(t);
    for (auto a : c->img_arrays) cudaFreeArray(a);
    c->img_tex.clear(); c->img_arrays.clear();
}

extern "C" void rt_destroy(rt_ctx *c) {
    if (!c) return;
    rt_comm_destroy(c);
    cudaSetDevice(c->device);
    cudaStreamSynchronize(c->stream);
    free_images(c);
    c->d_objects.release(); c->d_xforms.release(); c->d_mats.release(); c->d_texs.release(); c->d_images.release();
    c->d_obmin.release(); c->d_obmax.release(); c->d_node_aabb.release(); c->d_leaf_aabb.release(); c->d_bounds.release();
    c->d_hist.release(); c->d_keys[0].release(); c->d_keys[1].release(); c->d_vals[0].release(); c->d_vals[1].release();
    c->d_left.release(); c->d_right.release(); c->d_parent.release(); c->d_flags.release(); c->d_prim_obj.release();
    c->d_prim_mat.release(); c->d_slot_mat.release(); c->d_giant.release(); c->d_lbvh_info.release(); c->d_tile_prefix.release(); c->d_nodes.release(); c->d_nodes_s.release(); c->d_nodes4.release(); c->d_light_slot.release(); c->d_light_type.release(); c->d_obj_light.release(); c->d_light_area.release();
    c->d_primA.release(); c->d_primB.release(); c->d_primC.release();
    for (int k = 0; k < RT_MAX_SETS; ++k) { c->qs[k].release(); if (c->qs[k].ev_acc) cudaEventDestroy(c->qs[k].ev_acc); }
    if (c->ev_fork) cudaEventDestroy(c->ev_fork);
    for (int k = 1; k < RT_MAX_SETS; ++k) if (c->extra_stream[k]) cudaStreamDestroy(c->extra_stream[k]);
    c->d_counters.release(); c->d_tile_list.release(); c->d_tile_cnt.release(); c->d_tile_bad.release(); c->d_accum.release(); c->d_rgb8.release(); c->d_aov.release();
    c->d_partials.release(); c->d_rays.release();
    for (int k = 0; k < RT_ACC_CHUNKS; ++k) c->d_chunk[k].release();
    for (auto e : c->ev_pool) cudaEventDestroy(e);
    for (int k = 0; k < 2; ++k) if (c->ev_count[k]) cudaEventDestroy(c->ev_count[k]);
    if (c->ev0) cudaEventDestroy(c->ev0);
    if (c->ev1) cudaEventDestroy(c->ev1);
    if (c->own_stream) cudaStreamDestroy(c->stream);
    delete c;
}

// ---- host-side set-up that the reference also does on the host ----
struct H3 { float x, y, z; };
static inline H3 hsub(H3 a, H3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
static inline H3 hscale(float t, H3 a) { return {t * a.x, t * a.y, t * a.z}; }
static inline float hdot(H3 a, H3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
static inline H3 hunit(H3 a) { float r = 1.0f / sqrtf(hdot(a, a)); return hscale(r, a); }
static inline H3 hcross(H3 u, H3 v) { return {u.y * v.z - u.z * v.y, u.z * v.x - u.x * v.z, u.x * v.y - u.y * v.x}; }
static inline float3 f3(H3 a) { return make_float3(a.x, a.y, a.z); }

// camera::camera, gpu-version/camera.cuh:9-29 (float arithmetic, unfused: this file is built with -ffp-contract=off)
static void camera_setup(const rt_camera_desc &c, DCamera &cam) {
    float theta = (float)((double)c.vfov_deg * M_PI / (double)180.0f);       // rtweekend.cuh:17-19
    float h = tanf(theta / 2.0f);
    float viewport_height = 2.0f * h;
    float viewport_width = c.aspect * viewport_height;
    H3 from = {c.lookfrom[0], c.lookfrom[1], c.lookfrom[2]}, at = {c.lookat[0], c.lookat[1], c.lookat[2]};
    H3 vup = {c.vup[0], c.vup[1], c.vup[2]};
    H3 w = hunit(hsub(from, at)), u = hunit(hcross(vup, w)), v = hcross(w, u);
    H3 horizontal = hscale(c.focus_dist * viewport_width, u), vertical = hscale(c.focus_dist * viewport_height, v);
    H3 hh = hscale(1.0f / 2.0f, horizontal), vh = hscale(1.0f / 2.0f, vertical);
    H3 llc = hsub(hsub(hsub(from, hh), vh), hscale(c.focus_dist, w));
    cam.origin = f3(from); cam.llc = f3(llc); cam.horizontal = f3(horizontal); cam.vertical = f3(vertical);
    cam.u = f3(u); cam.v = f3(v); cam.w = f3(w);
    cam.lens_radius = c.aperture / 2.0f;
}

static bool texture_reads_uv(const HostScene &hs, int tex, int depth = 0) {
    if (tex < 0 || tex >= (int)hs.textures.size() || depth > 8) return false;
    const rt_texture_desc &t = hs.textures[tex];
    if (t.type == RT_TEX_IMAGE) return true;
    if (t.type == RT_TEX_CHECKER) return texture_reads_uv(hs, t.even, depth + 1) || texture_reads_uv(hs, t.odd, depth + 1);
    return false;
}

static int validate_scene(rt_ctx *ctx, const HostScene &hs) {
    const rt_scene_desc &d = hs.d;
    if (d.width < 2 || d.height < 2) return rt_fail(ctx, RT_ERR_INVALID, "scene: width and height must be >= 2");
    if (d.max_depth < 1 || d.samples_per_pixel < 1) return rt_fail(ctx, RT_ERR_INVALID, "scene: max_depth and samples_per_pixel must be >= 1");
    if ((long long)d.n_objects >= (1ll << 28)) return rt_fail(ctx, RT_ERR_INVALID, "scene: too many objects");
    for (int i = 0; i < d.n_objects; ++i) {
        const rt_object_desc &o = hs.objects[i];
        if (o.type < 0 || o.type > RT_OBJ_TRIANGLE) return rt_fail(ctx, RT_ERR_INVALID, "scene: unknown object type at index " + std::to_string(i));
        if (o.material < 0 || o.material >= d.n_materials) return rt_fail(ctx, RT_ERR_INVALID, "scene: object " + std::to_string(i) + " references a missing material");
        if (o.xform >= d.n_xforms) return rt_fail(ctx, RT_ERR_INVALID, "scene: object " + std::to_string(i) + " references a missing transform");
        if (o.type == RT_OBJ_CYLINDER && o.xform < 0) return rt_fail(ctx, RT_ERR_INVALID, "scene: cylinder without a transform");
    }
    for (int i = 0; i < d.n_materials; ++i) {
        const rt_material_desc &m = hs.materials[i];
        if (m.type < 0 || m.type > RT_MAT_DIFFUSE_LIGHT) return rt_fail(ctx, RT_ERR_INVALID, "scene: unknown material type");
        if ((m.type == RT_MAT_LAMBERTIAN || m.type == RT_MAT_DIFFUSE_LIGHT) && (m.texture < 0 || m.texture >= d.n_textures))
            return rt_fail(ctx, RT_ERR_INVALID, "scene: material " + std::to_string(i) + " references a missing texture");
    }
    for (int i = 0; i < d.n_textures; ++i) {
        const rt_texture_desc &t = hs.textures[i];
        if (t.type < 0 || t.type > RT_TEX_IMAGE) return rt_fail(ctx, RT_ERR_INVALID, "scene: unknown texture type");
        if (t.type == RT_TEX_CHECKER && (t.even < 0 || t.even >= d.n_textures || t.odd < 0 || t.odd >= d.n_textures))
            return rt_fail(ctx, RT_ERR_INVALID, "scene: checker texture references a missing texture");
        if (t.type == RT_TEX_IMAGE && (t.image < 0 || t.image >= d.n_images)) return rt_fail(ctx, RT_ERR_INVALID, "scene: image texture references a missing image");
    }
    return RT_OK;
}

// Copies the tables to the device and builds the LBVH there (replaces parse_scene's per-node cudaMalloc/cudaMemcpy,
// parser.hpp:130-561, and the single-thread move_to_device kernel, main.cu:374-446).
static int upload_to_device(rt_ctx *ctx) {
    HostScene &hs = ctx->hs;
    const rt_scene_desc &d = hs.d;
    CU(cudaSetDevice(ctx->device));
    const int n = d.n_objects;
    cudaStream_t s = ctx->stream;

    // objects, transforms, materials, textures
    static_assert(sizeof(DObject) == sizeof(rt_object_desc), "DObject must mirror rt_object_desc");
    CU(ctx->d_objects.alloc(n));
    if (n) CU(cudaMemcpyAsync(ctx->d_objects.p, hs.objects.data(), sizeof(DObject) * n, cudaMemcpyHostToDevice, s));
    std::vector<float4> xf(6 * (size_t)std::max(1, d.n_xforms));
    for (int i = 0; i < d.n_xforms; ++i)
        for (int r = 0; r < 3; ++r) {
            const float *m = hs.xforms[i].m + 4 * r, *mi = hs.xforms[i].m_inv + 4 * r;
            xf[6 * i + r] = make_float4(m[0], m[1], m[2], m[3]);
            xf[6 * i + 3 + r] = make_float4(mi[0], mi[1], mi[2], mi[3]);
        }
    CU(ctx->d_xforms.alloc(xf.size()));
    CU(cudaMemcpyAsync(ctx->d_xforms.p, xf.data(), sizeof(float4) * xf.size(), cudaMemcpyHostToDevice, s));
    std::vector<DMaterial> mats(std::max(1, d.n_materials));
    for (int i = 0; i < d.n_materials; ++i) {
        const rt_material_desc &m = hs.materials[i];
        DMaterial dm{};
        dm.type = m.type; dm.texture = m.texture;
        dm.fuzz = m.fuzz < 1.0f ? m.fuzz : 1.0f;                         // metal ctor, material.cuh:61
        dm.ior = m.ior;
        const bool uv = (m.type == RT_MAT_LAMBERTIAN || m.type == RT_MAT_DIFFUSE_LIGHT) && texture_reads_uv(hs, m.texture);
        dm.albedo = make_float4(m.albedo[0], m.albedo[1], m.albedo[2], uv ? 1.0f : 0.0f);
        // a plain solid_color texture is resolved here: the material carries the colour, the kernels skip the texture table
        if ((m.type == RT_MAT_LAMBERTIAN || m.type == RT_MAT_DIFFUSE_LIGHT) && hs.textures[m.texture].type == RT_TEX_SOLID) {
            const rt_texture_desc &t = hs.textures[m.texture];
            dm.albedo = make_float4(t.color[0], t.color[1], t.color[2], 0.0f);
            dm.texture = -1;
        }
        mats[i] = dm;
    }
    CU(ctx->d_mats.alloc(mats.size()));
    CU(cudaMemcpyAsync(ctx->d_mats.p, mats.data(), sizeof(DMaterial) * mats.size(), cudaMemcpyHostToDevice, s));
    std::vector<DTexture> texs(std::max(1, d.n_textures));
    for (int i = 0; i < d.n_textures; ++i) {
        const rt_texture_desc &t = hs.textures[i];
        DTexture dt{};
        dt.type = t.type; dt.even = t.even; dt.odd = t.odd; dt.image = t.image;
        dt.color = make_float4(t.color[0], t.color[1], t.color[2], 0.0f);
        texs[i] = dt;
    }
    CU(ctx->d_texs.alloc(texs.size()));
    CU(cudaMemcpyAsync(ctx->d_texs.p, texs.data(), sizeof(DTexture) * texs.size(), cudaMemcpyHostToDevice, s));

    // image textures -> CUDA arrays sampled by the texture unit (point filter, unnormalised coordinates, uchar4 -> [0,1])
    free_images(ctx);
    std::vector<DImage> imgs(std::max(1, d.n_images));
    for (int i = 0; i < d.n_images; ++i) {
        const rt_image_desc &im = hs.images[i];
        if (im.width < 1 || im.height < 1 || !im.rgb8) return rt_fail(ctx, RT_ERR_INVALID, "scene: empty image");
        std::vector<uchar4> px((size_t)im.width * im.height);
        for (size_t k = 0; k < px.size(); ++k) px[k] = make_uchar4(im.rgb8[3 * k], im.rgb8[3 * k + 1], im.rgb8[3 * k + 2], 255);
        cudaChannelFormatDesc fmt = cudaCreateChannelDesc<uchar4>();
        cudaArray_t arr;
        CU(cudaMallocArray(&arr, &fmt, im.width, im.height));
        ctx->img_arrays.push_back(arr);
        CU(cudaMemcpy2DToArray(arr, 0, 0, px.data(), sizeof(uchar4) * im.width, sizeof(uchar4) * im.width, im.height, cudaMemcpyHostToDevice));
        cudaResourceDesc rd{}; rd.resType = cudaResourceTypeArray; rd.res.array.array = arr;
        cudaTextureDesc td{};
        td.addressMode[0] = td.addressMode[1] = cudaAddressModeClamp;
        td.filterMode = cudaFilterModePoint; td.readMode = cudaReadModeNormalizedFloat; td.normalizedCoords = 0;
        cudaTextureObject_t tex;
        CU(cudaCreateTextureObject(&tex, &rd, &td, nullptr));
        ctx->img_tex.push_back(tex);
        imgs[i].tex = tex; imgs[i].width = im.width; imgs[i].height = im.height;
    }
    CU(ctx->d_images.alloc(imgs.size()));
    CU(cudaMemcpyAsync(ctx->d_images.p, imgs.data(), sizeof(DImage) * imgs.size(), cudaMemcpyHostToDevice, s));

    // LBVH buffers
    const size_t nn = (size_t)std::max(1, n);
    CU(ctx->d_obmin.alloc(3 * nn)); CU(ctx->d_obmax.alloc(3 * nn)); CU(ctx->d_bounds.alloc(12));
    CU(ctx->d_keys[0].alloc(nn)); CU(ctx->d_keys[1].alloc(nn)); CU(ctx->d_vals[0].alloc(nn)); CU(ctx->d_vals[1].alloc(nn));
    CU(ctx->d_hist.alloc(lbvh_hist_words(n)));
    CU(ctx->d_left.alloc(nn)); CU(ctx->d_right.alloc(nn)); CU(ctx->d_parent.alloc(2 * nn)); CU(ctx->d_flags.alloc(nn));
    CU(ctx->d_node_aabb.alloc(6 * nn)); CU(ctx->d_leaf_aabb.alloc(6 * nn));
    CU(ctx->d_nodes.alloc(nn)); CU(ctx->d_nodes_s.alloc(nn)); CU(ctx->d_primA.alloc(nn)); CU(ctx->d_primB.alloc(nn)); CU(ctx->d_primC.alloc(nn));
    CU(ctx->d_prim_obj.alloc(nn)); CU(ctx->d_prim_mat.alloc(nn)); CU(ctx->d_giant.alloc(nn)); CU(ctx->d_lbvh_info.alloc(RT_LBVH_INFO_WORDS));
    LbvhBuffers &lb = ctx->lb;
    lb.giant = ctx->d_giant.p; lb.info = ctx->d_lbvh_info.p;
    lb.cam = make_float4(d.camera.lookfrom[0], d.camera.lookfrom[1], d.camera.lookfrom[2], 0.5f * fabsf(d.camera.aperture));
    lb.n = n; lb.objects = ctx->d_objects.p; lb.xforms = ctx->d_xforms.p; lb.obj_bmin = ctx->d_obmin.p; lb.obj_bmax = ctx->d_obmax.p;
    lb.bounds = ctx->d_bounds.p; lb.keys[0] = ctx->d_keys[0].p; lb.keys[1] = ctx->d_keys[1].p; lb.vals[0] = ctx->d_vals[0].p;
    lb.vals[1] = ctx->d_vals[1].p; lb.hist = ctx->d_hist.p; lb.left = ctx->d_left.p; lb.right = ctx->d_right.p;
    lb.parent = ctx->d_parent.p; lb.flags = ctx->d_flags.p; lb.node_aabb = ctx->d_node_aabb.p; lb.leaf_aabb = ctx->d_leaf_aabb.p;
    lb.nodes = ctx->d_nodes.p; lb.nodes_s = ctx->d_nodes_s.p; lb.primA = ctx->d_primA.p; lb.primB = ctx->d_primB.p; lb.primC = ctx->d_primC.p;
    lb.prim_obj = ctx->d_prim_obj.p; lb.prim_mat = ctx->d_prim_mat.p;
    CU(cudaEventRecord(ctx->ev0, s));
    lbvh_build(ctx->cfg, lb);
    static const bool no_slot_mat = getenv("RT_NO_SLOT_MAT") != nullptr;
    if (n > 0 && n <= (1 << 16) && !no_slot_mat) {       // per-slot material records (rt_scene.cuh: DScene::slot_mat)
        CU(ctx->d_slot_mat.alloc(nn));
        launch_slot_mats(ctx->cfg, n, ctx->d_prim_mat.p, ctx->d_mats.p, ctx->d_slot_mat.p);
        ctx->ds.slot_mat = ctx->d_slot_mat.p;
    } else ctx->ds.slot_mat = nullptr;
    CU(cudaEventRecord(ctx->ev1, s));
    CU(cudaGetLastError());
    int info[RT_LBVH_INFO_WORDS] = {};
    CU(cudaMemcpyAsync(info, ctx->d_lbvh_info.p, sizeof info, cudaMemcpyDeviceToHost, s));
    CU(cudaStreamSynchronize(s));       // the host vectors above go out of scope
    float ms = 0; cudaEventElapsedTime(&ms, ctx->ev0, ctx->ev1); ctx->stats.ms_build = ms;
    ctx->bvh_depth = info[2];
    // the traversal stacks hold one entry per level at most (rt_traverse.cuh, k_tile_lists): a deeper tree (tens of thousands of
    // objects sharing one Morton cell) would drop subtrees silently, so the upload fails instead
    if (info[2] > RT_TRAV_STACK) {
        ctx->have_scene = false;
        return rt_fail(ctx, RT_ERR_INVALID, "scene: LBVH depth " + std::to_string(info[2]) + " exceeds the traversal stack (" + std::to_string(RT_TRAV_STACK) +
                                            " levels): too many objects share one Morton cell");
    }

    // sampled lights for RT_RENDER_NEE: emissive spheres / rects / cylinders / boxes / triangles in scene-file order (oracle/orc_nee.c)
    {
        std::vector<int> obj_light((size_t)std::max(1, n), -1), ltype; std::vector<float> larea;
        for (int i = 0; i < n; ++i) {
            const rt_object_desc &o = hs.objects[i];
            if (hs.materials[o.material].type != RT_MAT_DIFFUSE_LIGHT) continue;
            const float *p = o.p; float area;
            if (o.type == RT_OBJ_SPHERE) area = 4.0f * 3.14159274101257324f * (p[3] * p[3]);
            else if (o.type == RT_OBJ_XY_RECT || o.type == RT_OBJ_XZ_RECT || o.type == RT_OBJ_YZ_RECT) area = fabsf((p[1] - p[0]) * (p[3] - p[2]));
            else if (o.type == RT_OBJ_CYLINDER) area = 6.28318548202514648f * fabsf(p[0]) * fabsf(p[2] - p[1]);
            else if (o.type == RT_OBJ_TRIANGLE) {
                const float e1[3] = {p[3] - p[0], p[4] - p[1], p[5] - p[2]}, e2[3] = {p[6] - p[0], p[7] - p[1], p[8] - p[2]};
                const float c[3] = {fmaf(e1[1], e2[2], -(e1[2] * e2[1])), fmaf(e1[2], e2[0], -(e1[0] * e2[2])), fmaf(e1[0], e2[1], -(e1[1] * e2[0]))};
                area = 0.5f * sqrtf(fmaf(c[2], c[2], fmaf(c[1], c[1], c[0] * c[0])));
            } else if (o.type == RT_OBJ_BOX) {
                const float ex = fabsf(p[3] - p[0]), ey = fabsf(p[4] - p[1]), ez = fabsf(p[5] - p[2]);
                area = 2.0f * ((ey * ez + ez * ex) + ex * ey);
            } else continue;
            if (!(area > 0.0f)) continue;                      // a degenerate emitter has no surface to sample (and 1/area in the MIS weight)
            obj_light[i] = (int)ltype.size(); ltype.push_back(o.type); larea.push_back(area);
        }
        const int nl = (int)ltype.size();
        CU(ctx->d_obj_light.alloc(obj_light.size())); CU(ctx->d_light_type.alloc(std::max(1, nl)));
        CU(ctx->d_light_area.alloc(std::max(1, nl))); CU(ctx->d_light_slot.alloc(std::max(1, nl)));
        CU(cudaMemcpyAsync(ctx->d_obj_light.p, obj_light.data(), sizeof(int) * obj_light.size(), cudaMemcpyHostToDevice, s));
        if (nl) {
            CU(cudaMemcpyAsync(ctx->d_light_type.p, ltype.data(), sizeof(int) * nl, cudaMemcpyHostToDevice, s));
            CU(cudaMemcpyAsync(ctx->d_light_area.p, larea.data(), sizeof(float) * nl, cudaMemcpyHostToDevice, s));
            launch_light_slots(ctx->cfg, n, ctx->d_prim_obj.p, ctx->d_obj_light.p, ctx->d_light_slot.p);
        }
        CU(cudaStreamSynchronize(s));
        ctx->ds.n_lights = nl; ctx->ds.light_slot = ctx->d_light_slot.p; ctx->ds.light_type = ctx->d_light_type.p;
        ctx->ds.light_area = ctx->d_light_area.p; ctx->ds.obj_light = ctx->d_obj_light.p;
    }

    // device scene header
    DScene &ds = ctx->ds;
    ds.n_prims = n; ds.n_tree = info[0]; ds.n_giants = info[1]; ds.n_nodes = ds.n_tree >= 2 ? ds.n_tree - 1 : 1;
    ds.giant_code = ctx->d_lbvh_info.p + 4;
    ds.leaf_aabb = ctx->d_leaf_aabb.p;
    ds.width = d.width; ds.height = d.height; ds.max_depth = d.max_depth; ds.sampling = d.sampling & 1; ds.direct = (d.sampling & RT_SAMPLING_DIRECT) ? 1 : 0;
    ds.bg_mode = d.background_mode;
    ds.defocus = (ds.sampling == RT_SAMPLING_REFERENCE_CPU && d.camera.aperture > 0.0f) ? 1 : 0;
    ds.background = make_float3(d.background[0], d.background[1], d.background[2]);
    camera_setup(d.camera, ds.cam);
    ds.nodes = ctx->d_nodes.p; ds.nodes_s = ctx->d_nodes_s.p; ds.primA = ctx->d_primA.p; ds.primB = ctx->d_primB.p; ds.primC = ctx->d_primC.p;
    ds.prim_obj = ctx->d_prim_obj.p; ds.prim_mat = ctx->d_prim_mat.p; ds.xforms = ctx->d_xforms.p; ds.mats = ctx->d_mats.p;
    ds.texs = ctx->d_texs.p; ds.images = ctx->d_images.p;

    // extend-kernel configuration: stage the whole BVH in shared memory when it leaves room for >= 2 CTAs per SM
    const size_t bvh_bytes = sizeof(BvhNodeS) * (size_t)ds.n_nodes;
    ctx->cfg.bvh_smem_bytes = 0;
    if (bvh_bytes + RT_SSTACK_BYTES <= ctx->smem_optin && bvh_bytes + RT_SSTACK_BYTES <= 112 * 1024)      // two CTAs per SM: tree + stack each
        ctx->cfg.bvh_smem_bytes = bvh_bytes;
    int per_sm = extend_max_blocks_per_sm(ctx->cfg.bvh_smem_bytes);
    if (per_sm < 1) { ctx->cfg.bvh_smem_bytes = 0; per_sm = extend_max_blocks_per_sm(0); }
    ctx->cfg.extend_blocks_per_sm = per_sm < 1 ? 1 : per_sm;
    static const int env_ext_ctas = getenv("RT_EXT_CTAS") ? atoi(getenv("RT_EXT_CTAS")) : 0;       // tuning: persistent CTAs per SM
    if (env_ext_ctas > 0) ctx->cfg.extend_blocks_per_sm = std::min(env_ext_ctas, ctx->cfg.extend_blocks_per_sm);
    ds.nodes4 = nullptr;
    static const bool no_bvh4 = getenv("RT_NO_BVH4") != nullptr;      // A/B: two-wide 64-byte nodes
    if (!ctx->cfg.bvh_smem_bytes && ds.n_tree > 0 && !no_bvh4) {          // tree walked from global memory: four-wide nodes (rt_scene.cuh: BvhNode4)
        CU(ctx->d_nodes4.alloc((size_t)ds.n_nodes));
        lbvh_pack4(ctx->cfg, ctx->lb, ctx->d_nodes4.p);
        CU(cudaGetLastError());
        CU(cudaStreamSynchronize(s));
        ds.nodes4 = ctx->d_nodes4.p;
    }
    ds.sphere_xf = 0;
    for (int i = 0; i < d.n_objects; ++i) if (hs.objects[i].type == RT_OBJ_SPHERE && hs.objects[i].xform >= 0) { ds.sphere_xf = 1; break; }
    ds.only_type = d.n_objects > 0 ? hs.objects[0].type : -1;
    for (int i = 0; i < d.n_objects; ++i)       // the specialised sphere loop carries no instance transforms
        if (hs.objects[i].type != ds.only_type || (ds.only_type == RT_OBJ_SPHERE && hs.objects[i].xform >= 0)) { ds.only_type = -1; break; }
    // a scene whose objects are all emitters (dna.py frames) ends every path at its first vertex: one bounce is the whole job
    ctx->scene_scatters = false;
    for (int i = 0; i < d.n_objects; ++i)
        if (hs.materials[hs.objects[i].material].type != RT_MAT_DIFFUSE_LIGHT) { ctx->scene_scatters = true; break; }
    ctx->have_scene = true;
    ctx->tiles_dirty = true;
    ctx->accum_valid = false;
    return RT_OK;
}

static int rt_scene_upload_impl(rt_ctx *ctx, const rt_scene_desc *sc) {
    if (!ctx) return RT_ERR_INVALID;
    if (!sc) return rt_fail(ctx, RT_ERR_INVALID, "rt_scene_upload: scene is NULL");
    HostScene hs;
    hs.d = *sc;
    if (sc->n_objects < 0 || sc->n_xforms < 0 || sc->n_materials < 0 || sc->n_textures < 0 || sc->n_images < 0)
        return rt_fail(ctx, RT_ERR_INVALID, "rt_scene_upload: negative table size");
    if ((sc->n_objects > 0 && !sc->objects) || (sc->n_xforms > 0 && !sc->xforms) || (sc->n_materials > 0 && !sc->materials) ||
        (sc->n_textures > 0 && !sc->textures) || (sc->n_images > 0 && !sc->images))
        return rt_fail(ctx, RT_ERR_INVALID, "rt_scene_upload: a table pointer is NULL although its count is > 0");
    hs.objects.assign(sc->objects, sc->objects + sc->n_objects);
    hs.xforms.assign(sc->xforms, sc->xforms + sc->n_xforms);
    hs.materials.assign(sc->materials, sc->materials + sc->n_materials);
    hs.textures.assign(sc->textures, sc->textures + sc->n_textures);
    hs.images.assign(sc->images, sc->images + sc->n_images);
    for (int i = 0; i < sc->n_images; ++i) {
        const rt_image_desc &im = sc->images[i];
        if (!im.rgb8 || im.width < 1 || im.height < 1) return rt_fail(ctx, RT_ERR_INVALID, "rt_scene_upload: empty image");
        hs.image_data.emplace_back(im.rgb8, im.rgb8 + (size_t)im.width * im.height * 3);
    }
    hs.fix_pointers();
    int rc = validate_scene(ctx, hs);
    if (rc) return rc;
    ctx->hs = std::move(hs);
    ctx->hs.fix_pointers();
    return upload_to_device(ctx);
}
extern "C" int rt_scene_upload(rt_ctx *ctx, const rt_scene_desc *sc) {
    return rt_guard(ctx, "rt_scene_upload", [&] { return rt_scene_upload_impl(ctx, sc); });
}

static int rt_scene_upload_json_text_impl(rt_ctx *ctx, const char *text, const char *base_dir) {
    if (!ctx) return RT_ERR_INVALID;
    if (!text) return rt_fail(ctx, RT_ERR_INVALID, "rt_scene_upload_json_text: text is NULL");
    HostScene hs; std::string err;
    int rc = rt_json_to_scene(text, base_dir, hs, err);
    if (rc) return rt_fail(ctx, rc, err);
    hs.fix_pointers();
    rc = validate_scene(ctx, hs);
    if (rc) return rc;
    ctx->hs = std::move(hs);
    ctx->hs.fix_pointers();
    return upload_to_device(ctx);
}
extern "C" int rt_scene_upload_json_text(rt_ctx *ctx, const char *text, const char *base_dir) {
    return rt_guard(ctx, "rt_scene_upload_json_text", [&] { return rt_scene_upload_json_text_impl(ctx, text, base_dir); });
}

static int rt_scene_upload_json_impl(rt_ctx *ctx, const char *path) {
    if (!ctx) return RT_ERR_INVALID;
    if (!path) return rt_fail(ctx, RT_ERR_INVALID, "rt_scene_upload_json: path is NULL");
    std::ifstream f(path, std::ios::binary);
    if (!f) return rt_fail(ctx, RT_ERR_IO, std::string("cannot open scene file ") + path);
    std::stringstream ss; ss << f.rdbuf();
    std::string p(path), dir = ".";
    size_t k = p.find_last_of('/');
    if (k != std::string::npos) dir = p.substr(0, k);
    return rt_scene_upload_json_text(ctx, ss.str().c_str(), dir.c_str());
}
extern "C" int rt_scene_upload_json(rt_ctx *ctx, const char *path) {
    return rt_guard(ctx, "rt_scene_upload_json", [&] { return rt_scene_upload_json_impl(ctx, path); });
}

extern "C" void rt_make_xform(const float axis[3], double angle_deg, int has_rotate, const float translate[3], int has_translate, rt_xform_desc *out) {
    static const float z3[3] = {0, 0, 1}, o3[3] = {0, 0, 0};
    if (out) rt_build_xform(axis ? axis : z3, angle_deg, has_rotate != 0 && axis, translate ? translate : o3, has_translate != 0 && translate, out);
}

// Replaces a run of objects and / or transforms of the uploaded scene and rebuilds the device tables and the LBVH.  "Refit by
// rebuild": the whole device build is 31 launches (0.12 ms for 486 objects, ~4 ms for a million triangles), so a separate refit pass
// for rigid motion would save less than the JSON text it already avoids (animation drivers: blue.py / blue2.py move four cylinders).
static int rt_scene_update_impl(rt_ctx *ctx, const rt_object_desc *objects, int first_object, int n_objects, const rt_xform_desc *xforms,
                                int first_xform, int n_xforms) {
    if (!ctx) return RT_ERR_INVALID;
    if (!ctx->have_scene) return rt_fail(ctx, RT_ERR_NO_SCENE, "rt_scene_update: no scene uploaded");
    HostScene &hs = ctx->hs;
    if (n_objects < 0 || n_xforms < 0 || first_object < 0 || first_xform < 0 || (n_objects > 0 && !objects) || (n_xforms > 0 && !xforms) ||
        (long long)first_object + n_objects > (long long)hs.objects.size() || (long long)first_xform + n_xforms > (long long)hs.xforms.size())
        return rt_fail(ctx, RT_ERR_INVALID, "rt_scene_update: range outside the uploaded tables");
    std::vector<rt_object_desc> old_o(hs.objects.begin() + first_object, hs.objects.begin() + first_object + n_objects);
    std::vector<rt_xform_desc> old_x(hs.xforms.begin() + first_xform, hs.xforms.begin() + first_xform + n_xforms);
    std::copy(objects, objects + n_objects, hs.objects.begin() + first_object);
    std::copy(xforms, xforms + n_xforms, hs.xforms.begin() + first_xform);
    hs.fix_pointers();
    int rc = validate_scene(ctx, hs);
    if (rc) {       // leave the uploaded scene as it was
        std::copy(old_o.begin(), old_o.end(), hs.objects.begin() + first_object);
        std::copy(old_x.begin(), old_x.end(), hs.xforms.begin() + first_xform);
        return rc;
    }
    return upload_to_device(ctx);
}
extern "C" int rt_scene_update(rt_ctx *ctx, const rt_object_desc *objects, int first_object, int n_objects, const rt_xform_desc *xforms,
                               int first_xform, int n_xforms) {
    return rt_guard(ctx, "rt_scene_update", [&] { return rt_scene_update_impl(ctx, objects, first_object, n_objects, xforms, first_xform, n_xforms); });
}

extern "C" int rt_scene_info_get(rt_ctx *ctx, rt_scene_info *out) {
    if (!ctx) return RT_ERR_INVALID;
    if (!out) return rt_fail(ctx, RT_ERR_INVALID, "rt_scene_info_get: out is NULL");
    if (!ctx->have_scene) return rt_fail(ctx, RT_ERR_NO_SCENE, "no scene uploaded");
    const rt_scene_desc &d = ctx->hs.d;
    memset(out, 0, sizeof *out);
    out->width = d.width; out->height = d.height; out->max_depth = d.max_depth; out->samples_per_pixel = d.samples_per_pixel;
    out->n_objects = d.n_objects; out->n_materials = d.n_materials; out->n_textures = d.n_textures; out->n_bvh_nodes = ctx->ds.n_nodes;
    snprintf(out->output_file, sizeof out->output_file, "%s", ctx->hs.output_file.c_str());
    return RT_OK;
}

extern "C" int rt_scene_override(rt_ctx *ctx, int width, int height, int spp, int max_depth) {
    if (!ctx) return RT_ERR_INVALID;
    if (!ctx->have_scene) return rt_fail(ctx, RT_ERR_NO_SCENE, "no scene uploaded");
    rt_scene_desc &d = ctx->hs.d;
    if (width > 0) d.width = width;
    if (height > 0) d.height = height;
    if (spp > 0) d.samples_per_pixel = spp;
    if (max_depth > 0) d.max_depth = max_depth;
    if (d.width < 2 || d.height < 2) return rt_fail(ctx, RT_ERR_INVALID, "override: width and height must be >= 2");
    // the camera keeps the aspect the scene was loaded with (the reference ties it to the JSON's width/height, parser.hpp:122)
    ctx->ds.width = d.width; ctx->ds.height = d.height; ctx->ds.max_depth = d.max_depth;
    ctx->accum_valid = false;
    ctx->tiles_dirty = true;
    return RT_OK;
}

// ---- primary-ray candidate lists (k_tile_lists): once per scene / image size ----
static int ensure_tile_lists(rt_ctx *ctx) {
    DScene &ds = ctx->ds;
    if (!ctx->tiles_dirty) return RT_OK;
    ds.tile_mode = 0; ds.tile_list = nullptr; ds.tile_cnt = nullptr; ds.tile_empty_prefix = nullptr;
    ctx->tiles_dirty = false; ctx->tiles_all_empty = false;
    static const bool off = getenv("RT_NO_TILE_LISTS") != nullptr;
    const int n_tiles = ((ds.width & ~7) >> 3) * ((ds.height & ~3) >> 2);
    if (off || n_tiles == 0) return RT_OK;
    CU(ctx->d_tile_list.alloc((size_t)n_tiles * RT_TILE_K)); CU(ctx->d_tile_cnt.alloc(n_tiles)); CU(ctx->d_tile_bad.alloc(1));
    CU(ctx->d_tile_prefix.alloc((size_t)n_tiles + 1));
    CU(cudaMemsetAsync(ctx->d_tile_bad.p, 0, sizeof(int), ctx->stream));
    launch_tile_lists(ctx->cfg, ds, ctx->d_tile_list.p, ctx->d_tile_cnt.p, ctx->d_tile_bad.p);
    launch_tile_prefix(ctx->cfg, ctx->d_tile_cnt.p, n_tiles, ctx->d_tile_prefix.p);
    int bad = 0, n_empty = 0;
    CU(cudaMemcpyAsync(&bad, ctx->d_tile_bad.p, sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
    CU(cudaMemcpyAsync(&n_empty, ctx->d_tile_prefix.p + n_tiles, sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
    CU(cudaStreamSynchronize(ctx->stream));
    if (getenv("RT_TRACE")) {
        std::vector<int> h((size_t)n_tiles);
        CU(cudaMemcpy(h.data(), ctx->d_tile_cnt.p, (size_t)n_tiles * sizeof(int), cudaMemcpyDeviceToHost));
        int hist[RT_TILE_K + 2] = {};
        for (int c : h) hist[c < 0 ? RT_TILE_K + 1 : c]++;
        fprintf(stderr, "[tile lists] %d tiles, %d empty, %d over RT_TILE_K, defocus %d; candidates per tile:", n_tiles, n_empty, bad, ds.defocus);
        for (int c = 0; c <= RT_TILE_K + 1; ++c) if (hist[c]) fprintf(stderr, " %d:%d", c, hist[c]);
        fprintf(stderr, "\n");
        if (getenv("RT_TRACE_TILE")) {          // candidates of one tile (development aid)
            const int t = atoi(getenv("RT_TRACE_TILE"));
            if (t >= 0 && t < n_tiles) {
                int cand[RT_TILE_K];
                CU(cudaMemcpy(cand, ctx->d_tile_list.p + (size_t)t * RT_TILE_K, sizeof cand, cudaMemcpyDeviceToHost));
                fprintf(stderr, "[tile %d] %d candidates:", t, h[t]);
                for (int c = 0; c < h[t] && c < RT_TILE_K; ++c) {
                    int obj = -1;
                    CU(cudaMemcpy(&obj, ctx->d_prim_obj.p + (cand[c] & RT_LEAF_SLOT_MASK), sizeof(int), cudaMemcpyDeviceToHost));
                    fprintf(stderr, " type %d object %d", (unsigned)cand[c] >> 28, obj);
                }
                fprintf(stderr, "\n");
            }
        }
    }
    if (bad * 4 > n_tiles) return RT_OK;                      // most beams touch more than RT_TILE_K leaves (dense meshes): not worth it
    ds.tile_list = ctx->d_tile_list.p; ds.tile_cnt = ctx->d_tile_cnt.p;
    ds.tile_mode = (bad == 0 && n_tiles * 32 == ds.width * ds.height) ? 1 : 2;
    // tiles whose beam touches nothing: their paths skip the queues (k_generate / k_shade / k_accumulate)
    static const bool keep_empty = getenv("RT_NO_EMPTY_TILES") != nullptr;
    if (n_empty > 0 && !keep_empty) ds.tile_empty_prefix = ctx->d_tile_prefix.p;
    ctx->tiles_all_empty = ds.tile_empty_prefix && ds.tile_mode == 1 && n_empty == n_tiles;
    ctx->tiles_n_empty = ds.tile_empty_prefix ? n_empty : 0;
    return RT_OK;
}

// ---- wavefront queues ----
static int ensure_queues(rt_ctx *ctx, QueueSet &S, long long cap, int max_depth, bool nee) {
    for (int k = 0; k < 2; ++k) { CU(S.ray_o[k].alloc(cap)); CU(S.ray_d[k].alloc(cap)); CU(S.thr[k].alloc(cap)); }
    CU(S.radiance.alloc(cap)); CU(S.hits.alloc(cap + 2));   // +2: the TMA copy of an odd-sized tail rounds up to 16 B
    CU(S.qcount.alloc(max_depth + 2)); CU(S.workctr.alloc((size_t)max_depth * 2));
    CU(ctx->d_counters.alloc(CTR_COUNT));
    if (!ctx->d_rays.p) { CU(ctx->d_rays.alloc(3)); CU(cudaMemsetAsync(ctx->d_rays.p, 0, 3 * sizeof(unsigned long long), ctx->stream)); }
    S.cap = std::max(S.cap, cap);
    Queues &q = S.q;
    for (int k = 0; k < 2; ++k) { q.ray_o[k] = S.ray_o[k].p; q.ray_d[k] = S.ray_d[k].p; q.thr[k] = S.thr[k].p; }
    q.hits = S.hits.p; q.radiance = S.radiance.p; q.counters = ctx->d_counters.p; q.cap = cap;
    q.qcount = S.qcount.p; q.workctr = S.workctr.p;
    q.sh_o = nullptr; q.sh_d = nullptr; q.sh_c = nullptr; q.shcount = nullptr;
    if (nee) {
        CU(S.sh_o.alloc(cap)); CU(S.sh_d.alloc(cap)); CU(S.sh_c.alloc(cap)); CU(S.shcount.alloc(max_depth + 1));
        q.sh_o = S.sh_o.p; q.sh_d = S.sh_d.p; q.sh_c = S.sh_c.p; q.shcount = S.shcount.p;
    }
    return RT_OK;
}

static int ensure_accum(rt_ctx *ctx) {
    const size_t n = (size_t)ctx->ds.width * ctx->ds.height * 3;
    if (ctx->d_accum.n < n) { CU(ctx->d_accum.alloc(n)); ctx->accum_valid = false; }
    return RT_OK;
}

static_assert(RT_ACC_CHUNKS == RT_SAMPLE_CHUNKS, "chunk count of the kernels and of rt.h");
extern "C" void rt_sample_chunk(int total_spp, int k, int *begin, int *end) {
    if (total_spp < 0) total_spp = 0;
    if (k < 0) k = 0;
    if (k > RT_SAMPLE_CHUNKS) k = RT_SAMPLE_CHUNKS;
    const long long b = (long long)k * total_spp / RT_SAMPLE_CHUNKS, e = k < RT_SAMPLE_CHUNKS ? (long long)(k + 1) * total_spp / RT_SAMPLE_CHUNKS : total_spp;
    if (begin) *begin = (int)b;
    if (end) *end = (int)e;
}

static int rt_render_impl(rt_ctx *ctx, const rt_render_params *prm) {
    if (!ctx) return RT_ERR_INVALID;
    if (!prm) return rt_fail(ctx, RT_ERR_INVALID, "rt_render: params is NULL");
    if (!ctx->have_scene) return rt_fail(ctx, RT_ERR_NO_SCENE, "rt_render: no scene uploaded");
    CU(cudaSetDevice(ctx->device));
    const DScene &ds = ctx->ds;
    const int np = ds.width * ds.height;
    const int spp = prm->spp_count > 0 ? prm->spp_count : ctx->hs.d.samples_per_pixel;
    if (prm->spp_begin < 0) return rt_fail(ctx, RT_ERR_INVALID, "rt_render: spp_begin < 0");
    const bool count = (prm->flags & RT_RENDER_COUNT_EVENTS) != 0;
    const bool sort = (prm->flags & RT_RENDER_SORT) != 0;
    const bool nee = (prm->flags & (RT_RENDER_NEE | RT_RENDER_NEE_MIS)) != 0;
    if (nee && ds.sampling != RT_SAMPLING_REFERENCE_CPU)
        return rt_fail(ctx, RT_ERR_INVALID, "rt_render: RT_RENDER_NEE needs the reference_cpu sampling convention (cosine-weighted lambertian)");
    static const bool trace = getenv("RT_TRACE") != nullptr;      // host-side timing of this call on stderr
    const auto t_enter = std::chrono::steady_clock::now();
    const bool time_stages = !count && (prm->flags & RT_RENDER_TIME_STAGES) != 0;
    // Two passes in flight on two streams (unless a diagnostic mode wants serial stage times): the bounce loop ends in
    // dozens of short launches that cannot fill the GPU; the other pass's long early-bounce launches fill it meanwhile.
    // (measured: RTIOW -4.5 %, blue2 -3 % with two streams; a tree read from global memory — the 1 M-triangle mesh — gets 8 %
    // SLOWER, two passes thrash each other's L1/L2 working set, so it keeps one stream unless RT_STREAMS says otherwise)
    static const int env_sets = getenv("RT_STREAMS") ? std::max(1, std::min(RT_MAX_SETS, atoi(getenv("RT_STREAMS")))) : 0;
    const int want_sets = env_sets ? env_sets : (ctx->cfg.bvh_smem_bytes ? 2 : 1);
    const int n_sets = (count || time_stages || (prm->flags & RT_RENDER_SERIAL)) ? 1 : std::min(want_sets, spp);
    // wavefront size: whole sample planes, at most max_paths_per_pass paths and at most 2^30 queue entries per set
    long long max_paths = prm->max_paths_per_pass;
    if (max_paths <= 0) {
        size_t free_b = 0, total_b = 0;
        cudaMemGetInfo(&free_b, &total_b);
        const long long per_path = 2 * 48 + 16 + 8 + (nee ? 48 : 0);            // bytes per queue entry
        long long held = 0;
        for (int k = 0; k < RT_MAX_SETS; ++k) held += ctx->qs[k].cap * per_path;
        long long by_mem = (long long)((double)(free_b + held) * 0.5 / per_path) / n_sets;
        // big wavefronts amortise the ~100 short late-bounce launches of a pass: 2 x 256 M paths = 61 GB of the 180 GB
        // (RTIOW 1080p 1024 spp: 508.9 / 498.3 / 493.3 ms per step at 64 / 128 / 256 M paths per stream)
        // (576 M: one 128-sample chunk of a 1080p image — 265 M paths — stays one pass per stream)
        max_paths = std::min<long long>(by_mem, (576ll << 20) / n_sets);
    }
    max_paths = std::min<long long>(max_paths, 1ll << 30);
    // Fixed-order accumulation (rt.h: rt_sample_chunk): the samples of the render fall into RT_ACC_CHUNKS chunks whose boundaries
    // depend on the TOTAL sample count only; a chunk is summed sample by sample into its own buffer, the buffers are combined in
    // chunk order.  A pass may span several chunks: k_accumulate sums each chunk's planes into that chunk's buffer (AccumPlan).
    const int total_spp = prm->total_spp > 0 ? prm->total_spp : std::max(ctx->hs.d.samples_per_pixel, prm->spp_begin + spp);
    struct Piece { int chunk, begin, end; };
    std::vector<Piece> pieces;
    for (int k = 0; k < RT_ACC_CHUNKS; ++k) {
        int cb, ce;
        rt_sample_chunk(total_spp, k, &cb, &ce);
        if (k == RT_ACC_CHUNKS - 1) ce = std::max(ce, prm->spp_begin + spp);      // samples past the total count the last chunk
        const int pb = std::max(cb, prm->spp_begin), pe = std::min(ce, prm->spp_begin + spp);
        if (pb < pe) pieces.push_back({k, pb, pe});
    }
    // equal passes (1024 samples at a 139-sample cap: 8 x 128, not 7 x 139 + 51), at least one per stream
    const int s_cap = (int)std::max<long long>(1, std::min<long long>(spp, max_paths / np));
    const int n_pass_total = std::max((spp + s_cap - 1) / s_cap, std::min(n_sets, spp));
    const int s_pass = (spp + n_pass_total - 1) / n_pass_total;
    const long long cap = (long long)s_pass * np;
    if (cap > (1ll << 30) - 64) return rt_fail(ctx, RT_ERR_INVALID, "rt_render: image too large for one wavefront pass");
    int rc = ensure_accum(ctx); if (rc) return rc;
    rc = ensure_tile_lists(ctx); if (rc) return rc;
    // All-emitter scenes with complete tile candidate lists need no queues at all: one kernel per render (k_primary_only)
    static const bool no_primary_only = getenv("RT_NO_PRIMARY_ONLY") != nullptr;
    const bool primary_only = !ctx->scene_scatters && ds.tile_mode == 1 && !(prm->flags & RT_RENDER_NO_TILE_LISTS) && !nee && !count &&
                              !no_primary_only && ds.n_prims > 0;
    for (int k = 0; k < n_sets && !primary_only; ++k) {
        rc = ensure_queues(ctx, ctx->qs[k], cap, ds.max_depth, nee); if (rc) return rc;
        ctx->qs[k].cfg = ctx->cfg; ctx->qs[k].cfg.stream = k == 0 ? ctx->stream : ctx->extra_stream[k];
    }
    cudaStream_t s = ctx->stream;
    bool clear = !((prm->flags & RT_RENDER_KEEP_ACCUM) && ctx->accum_valid);
    if (count) CU(cudaMemsetAsync(ctx->d_counters.p, 0, sizeof(unsigned long long) * CTR_COUNT, s));
    long long passes = 0;
    float ms_stage[6] = {0, 0, 0, 0, 0, 0};
    // Tail stage (k_tail): the bounce from which the surviving paths are finished by one launch.  Trees staged in shared memory: bounce
    // 8 (RTIOW: 3.5 % of the paths are still alive there); RT_TAIL_BOUNCE overrides, 0 or RT_RENDER_NO_TAIL switches it off.  Not with
    // NEE (the shadow stage needs the queues) or COUNT_EVENTS (the counters belong to the extend stage).
    static const int env_tail = getenv("RT_TAIL_BOUNCE") ? atoi(getenv("RT_TAIL_BOUNCE")) : -1;
    // all-sphere scenes run the lean instances of k_tail (rt_kernels.cu), for which an early hand-over pays: RTIOW 128-spp pass 43.28 /
    // 43.42 / 43.62 / 43.99 ms from bounce 2 / 3 / 4 / 8; with the generic instances 8 is best (blue2 47.5 / 44.3 / 42.1 / 40.6)
    int tail_b0 = env_tail >= 0 ? env_tail : ((ds.only_type == RT_OBJ_SPHERE && ctx->cfg.bvh_smem_bytes) ? 2 : 8);
    if (nee || count || (prm->flags & RT_RENDER_NO_TAIL) || tail_b0 <= 0 || tail_b0 >= ds.max_depth) tail_b0 = -1;
    cudaEvent_t *evs = ctx->ev_count;       // COUNT_EVENTS: one synchronising event pair per stage launch (owned by the context)
    if (time_stages) {          // event pairs are recorded without synchronising and resolved in rt_download
        const size_t n_passes = (size_t)n_pass_total;
        const size_t need = ctx->pending.size() + n_passes * (2 + 3 * (size_t)ds.max_depth);
        while (ctx->ev_pool.size() < 2 * need) { cudaEvent_t e; if (cudaEventCreate(&e) != cudaSuccess) return rt_fail(ctx, RT_ERR_CUDA, "cudaEventCreate"); ctx->ev_pool.push_back(e); }
    }
    static const char *stage_name[6] = {"rt:generate", "rt:extend", "rt:shade", "rt:accumulate", "rt:shadow", "rt:tail"};
    static const bool nvtx = getenv("RT_NVTX") != nullptr;       // one NVTX range per stage launch (host-side enqueue) + one per rt_render
    if (nvtx) nvtxRangePushA("rt_render");
    auto timed = [&](int stage, auto &&fn) {
        if (nvtx) nvtxRangePushA(stage_name[stage]);
        if (count) cudaEventRecord(evs[0], s);
        size_t slot = 0;
        if (time_stages) { slot = ctx->pending.size(); ctx->pending.push_back(stage); cudaEventRecord(ctx->ev_pool[2 * slot], s); }
        fn();
        if (time_stages) cudaEventRecord(ctx->ev_pool[2 * slot + 1], s);
        if (count) { cudaEventRecord(evs[1], s); cudaEventSynchronize(evs[1]); float ms = 0; cudaEventElapsedTime(&ms, evs[0], evs[1]); ms_stage[stage] += ms; }
        if (nvtx) nvtxRangePop();
    };
    const auto t_prep = std::chrono::steady_clock::now();
    CU(cudaEventRecord(ctx->ev0, s));
    if (n_sets > 1) {           // the extra streams start after everything already queued on the caller's stream
        CU(cudaEventRecord(ctx->ev_fork, s));
        for (int k = 1; k < n_sets; ++k) CU(cudaStreamWaitEvent(ctx->extra_stream[k], ctx->ev_fork, 0));
    }
    int last_set = 0;
    const size_t n_img = (size_t)np * 3;
    ChunkPtrs cp{};
    int n_cp = 0;
    bool chunk_clear[RT_ACC_CHUNKS];
    for (const Piece &pc : pieces) {
        CU(ctx->d_chunk[pc.chunk].alloc(n_img));
        cp.p[n_cp++] = ctx->d_chunk[pc.chunk].p;
        chunk_clear[pc.chunk] = true;
    }
    if (primary_only) {
        AccumPlan plan{};
        for (const Piece &pc : pieces) {
            plan.first[plan.n] = pc.begin - prm->spp_begin; plan.first[plan.n + 1] = pc.end - prm->spp_begin;
            plan.buf[plan.n] = ctx->d_chunk[pc.chunk].p; plan.clear[plan.n] = 1;
            ++plan.n;
        }
        if (!ctx->d_rays.p) { CU(ctx->d_rays.alloc(3)); CU(cudaMemsetAsync(ctx->d_rays.p, 0, 3 * sizeof(unsigned long long), s)); }
        timed(0, [&] { launch_primary_only(ctx->cfg, ds, plan, prm->seed, prm->spp_begin, spp, ctx->d_rays.p); });
        if (ds.tile_empty_prefix) ctx->stats.paths_skipped += (long long)ctx->tiles_n_empty * 32 * spp;
        ++passes;
    } else {
        for (int done = 0; done < spp; done += s_pass) {
            const int sn = std::min(s_pass, spp - done);
            const int n_paths = sn * np;
            const int k = (int)(passes % n_sets);
            QueueSet &S = ctx->qs[k];
            const LaunchCfg &cfg = S.cfg;
            const int first_sample = prm->spp_begin + done;
            // this pass's sample planes by chunk
            AccumPlan plan{};
            for (const Piece &pc : pieces) {
                const int lo = std::max(pc.begin, first_sample), hi = std::min(pc.end, first_sample + sn);
                if (lo >= hi) continue;
                plan.first[plan.n] = lo - first_sample; plan.first[plan.n + 1] = hi - first_sample;
                plan.buf[plan.n] = ctx->d_chunk[pc.chunk].p; plan.clear[plan.n] = chunk_clear[pc.chunk] ? 1 : 0;
                chunk_clear[pc.chunk] = false;
                ++plan.n;
            }
            launch_pass_init(cfg, S.q, ds.max_depth, n_paths);
            // bounce 0: generate resolves the hits of every path whose tile has a candidate list (ds.tile_mode; k_generate)
            const int tile_mode = (prm->flags & RT_RENDER_NO_TILE_LISTS) ? 0 : ds.tile_mode;
            DScene dsp = ds;                 // this pass's view of the scene: no tile data when the candidate lists are switched off
            dsp.nee_mis = (prm->flags & RT_RENDER_NEE_MIS) ? 1 : 0;
            if (tile_mode == 0) { dsp.tile_mode = 0; dsp.tile_list = nullptr; dsp.tile_cnt = nullptr; dsp.tile_empty_prefix = nullptr; }
            timed(0, [&] { launch_generate(cfg, dsp, S.q, prm->seed, first_sample, n_paths, nee, tile_mode != 0, count); });
            int buf = 0;
            // every tile's beam misses everything (empty scene, camera looking away): no path enters the queues
            const bool no_paths = tile_mode == 1 && ctx->tiles_all_empty;
            if (dsp.tile_empty_prefix) ctx->stats.paths_skipped += (long long)ctx->tiles_n_empty * 32 * sn;
            const int depth_needed = ctx->scene_scatters ? ds.max_depth : 1;      // nothing survives the first vertex of an all-emitter scene
            for (int b = 0; b < depth_needed && !no_paths; ++b) {
                // from bounce tail_b0 on, one launch finishes every surviving path (k_tail) instead of 2 x (depth - b) queue launches
                if (b == tail_b0 && b > 0) { timed(5, [&] { launch_tail(cfg, dsp, S.q, prm->seed, first_sample, b, buf); }); break; }
                if (b > 0 || tile_mode != 1) { ++ctx->stats.launches_extend; timed(1, [&] { launch_extend(cfg, dsp, S.q, b, buf, count, b == 0 && tile_mode == 2); }); }
                timed(2, [&] { launch_shade(cfg, dsp, S.q, prm->seed, first_sample, b, buf, b == ds.max_depth - 1, count, sort, nee); });
                if (nee) timed(4, [&] { launch_shadow(cfg, dsp, S.q, b); });
                buf ^= 1;
            }
            // samples are added to a chunk buffer in ascending order whatever the stream: pass p's accumulate follows pass p-1's
            if (n_sets > 1 && passes > 0) CU(cudaStreamWaitEvent(cfg.stream, ctx->qs[last_set].ev_acc, 0));
            const int *const_tiles = (dsp.tile_empty_prefix && ds.bg_mode != RT_BG_SKY_GRADIENT) ? dsp.tile_cnt : nullptr;
            timed(3, [&] { launch_accumulate(cfg, S.q, plan, ds.width, ds.height, sn, ds.max_depth, ctx->d_rays.p, const_tiles, ds.background,
                                             (tile_mode == 0 ? 0 : 1), (tail_b0 > 0 && !no_paths) ? tail_b0 : ds.max_depth); });
            if (n_sets > 1) CU(cudaEventRecord(S.ev_acc, cfg.stream));
            last_set = k;
            ++passes;
        }
    }
    if (n_sets > 1) {           // join: the caller's stream continues after both pass streams
        // (a last pass on the caller's stream has already waited for the other stream's final accumulate)
        if (last_set != 0) CU(cudaStreamWaitEvent(s, ctx->qs[last_set].ev_acc, 0));
    }
    for (int k = 0; k < RT_ACC_CHUNKS; ++k) ctx->chunk_live[k] = 0;
    if (prm->flags & RT_RENDER_KEEP_CHUNKS) {      // multi-GPU: the chunk buffers travel to the root (rt_gather_image), which folds them
        for (const Piece &pc : pieces) ctx->chunk_live[pc.chunk] = 1;
    } else {
        launch_fold_chunks(ctx->cfg, ctx->d_accum.p, cp, n_cp, !clear, (long long)n_img);
    }
    CU(cudaEventRecord(ctx->ev1, s));
    if (nvtx) nvtxRangePop();
    if (trace) {
        const auto t_done = std::chrono::steady_clock::now();
        {       // queue lengths of the last pass of stream 0 (synchronises: a development aid for reading ncu launch lists)
            std::vector<int> qc((size_t)ds.max_depth + 1);
            cudaDeviceSynchronize();
            if (ctx->qs[0].qcount.p && cudaMemcpy(qc.data(), ctx->qs[0].qcount.p, qc.size() * sizeof(int), cudaMemcpyDeviceToHost) == cudaSuccess) {
                fprintf(stderr, "[rt_render] queue lengths by bounce:");
                for (size_t b = 0; b < qc.size() && b < 14; ++b) fprintf(stderr, " %d", qc[b]);
                fprintf(stderr, "\n");
            }
        }
        fprintf(stderr, "[rt_render] prep %.2f ms, enqueue %.2f ms, %lld passes of %d spp on %d stream(s)\n",
                std::chrono::duration<double, std::milli>(t_prep - t_enter).count(), std::chrono::duration<double, std::milli>(t_done - t_prep).count(),
                passes, s_pass, n_sets);
    }
    CU(cudaGetLastError());
    if (!(prm->flags & RT_RENDER_KEEP_CHUNKS)) ctx->accum_valid = true;
    rt_stats &st = ctx->stats;
    st.samples += (long long)spp * np; st.passes += passes;
    if (count) {
        unsigned long long c[CTR_COUNT];
        CU(cudaMemcpyAsync(c, ctx->d_counters.p, sizeof c, cudaMemcpyDeviceToHost, s));
        CU(cudaStreamSynchronize(s));
        st.n_box_tests += c[CTR_BOX]; st.n_sphere_reject += c[CTR_SPH_REJECT]; st.n_sphere_accept += c[CTR_SPH_ACCEPT];
        st.n_rect_tests += c[CTR_RECT]; st.n_cyl_reject += c[CTR_CYL_REJECT]; st.n_cyl_accept += c[CTR_CYL_ACCEPT];
        st.n_tri_tests += c[CTR_TRI]; st.n_boxprim_tests += c[CTR_BOXPRIM]; st.n_shade += c[CTR_SHADE];
        st.n_generate += (long long)spp * np; st.n_node_visits += c[CTR_NODE_VISITS];
        st.n_generate_leaf_tests += c[CTR_GEN_LEAF];
        st.ms_generate += ms_stage[0]; st.ms_extend += ms_stage[1]; st.ms_shade += ms_stage[2]; st.ms_accumulate += ms_stage[3]; st.ms_shadow += ms_stage[4]; st.ms_tail += ms_stage[5];
        if (c[CTR_STACK_OVERFLOW]) return rt_fail(ctx, RT_ERR_INVALID, "traversal stack overflow (BVH deeper than RT_TRAV_STACK)");
    }
    return RT_OK;
}
extern "C" int rt_render(rt_ctx *ctx, const rt_render_params *prm) {
    return rt_guard(ctx, "rt_render", [&] { return rt_render_impl(ctx, prm); });
}

static int rt_render_aov_impl(rt_ctx *ctx, float *out, size_t n_floats) {
    if (!ctx) return RT_ERR_INVALID;
    if (!ctx->have_scene) return rt_fail(ctx, RT_ERR_NO_SCENE, "rt_render_aov: no scene uploaded");
    CU(cudaSetDevice(ctx->device));
    const size_t need = (size_t)ctx->ds.width * ctx->ds.height * 24;
    if (!out || n_floats < need) return rt_fail(ctx, RT_ERR_INVALID, "rt_render_aov: output buffer too small");
    CU(ctx->d_aov.alloc(need));
    launch_aov(ctx->cfg, ctx->ds, ctx->d_aov.p);
    CU(cudaGetLastError());
    CU(cudaMemcpyAsync(out, ctx->d_aov.p, sizeof(float) * need, cudaMemcpyDeviceToHost, ctx->stream));
    CU(cudaStreamSynchronize(ctx->stream));
    return RT_OK;
}
extern "C" int rt_render_aov(rt_ctx *ctx, float *out, size_t n_floats) {
    return rt_guard(ctx, "rt_render_aov", [&] { return rt_render_aov_impl(ctx, out, n_floats); });
}

extern "C" int rt_accum_device_ptr(rt_ctx *ctx, void **dev_ptr, size_t *n_floats) {
    if (!ctx) return RT_ERR_INVALID;
    if (!ctx->have_scene) return rt_fail(ctx, RT_ERR_NO_SCENE, "no scene uploaded");
    CU(cudaSetDevice(ctx->device));
    int rc = ensure_accum(ctx); if (rc) return rc;
    if (dev_ptr) *dev_ptr = ctx->d_accum.p;
    if (n_floats) *n_floats = (size_t)ctx->ds.width * ctx->ds.height * 3;
    return RT_OK;
}

extern "C" int rt_accum_sum_ordered(rt_ctx *ctx, const void *const *partials, int n_partials) {
    if (!ctx) return RT_ERR_INVALID;
    if (!ctx->have_scene) return rt_fail(ctx, RT_ERR_NO_SCENE, "no scene uploaded");
    if (!partials || n_partials < 1) return rt_fail(ctx, RT_ERR_INVALID, "rt_accum_sum_ordered: no partial buffers");
    CU(cudaSetDevice(ctx->device));
    int rc = ensure_accum(ctx); if (rc) return rc;
    CU(ctx->d_partials.alloc(n_partials));
    CU(cudaMemcpyAsync(ctx->d_partials.p, partials, sizeof(void *) * n_partials, cudaMemcpyHostToDevice, ctx->stream));
    launch_sum_ordered(ctx->cfg, ctx->d_accum.p, ctx->d_partials.p, n_partials, (long long)ctx->ds.width * ctx->ds.height * 3);
    CU(cudaGetLastError());
    CU(cudaStreamSynchronize(ctx->stream));
    ctx->accum_valid = true;
    return RT_OK;
}

extern "C" int rt_synchronize(rt_ctx *ctx) {
    if (!ctx) return RT_ERR_INVALID;
    CU(cudaSetDevice(ctx->device));
    CU(cudaStreamSynchronize(ctx->stream));
    return RT_OK;
}

extern "C" int rt_stats_reset(rt_ctx *ctx) {
    if (!ctx) return RT_ERR_INVALID;
    float ms_build = ctx->stats.ms_build;
    memset(&ctx->stats, 0, sizeof ctx->stats);
    ctx->stats.ms_build = ms_build;
    ctx->launches = 0;
    ctx->pending.clear();
    if (ctx->d_rays.p) { CU(cudaSetDevice(ctx->device)); CU(cudaMemsetAsync(ctx->d_rays.p, 0, 3 * sizeof(unsigned long long), ctx->stream)); }
    return RT_OK;
}

static int rt_download_impl(rt_ctx *ctx, float *rgb_sum, size_t n_floats, unsigned char *rgb8, int total_spp, int gamma, rt_stats *stats) {
    if (!ctx) return RT_ERR_INVALID;
    if (!ctx->have_scene) return rt_fail(ctx, RT_ERR_NO_SCENE, "rt_download: no scene uploaded");
    CU(cudaSetDevice(ctx->device));
    const size_t need = (size_t)ctx->ds.width * ctx->ds.height * 3;
    if ((rgb_sum || rgb8) && !ctx->accum_valid) return rt_fail(ctx, RT_ERR_INVALID, "rt_download: nothing rendered yet");
    if (rgb_sum && n_floats < need) return rt_fail(ctx, RT_ERR_INVALID, "rt_download: rgb_sum buffer too small");
    cudaStream_t s = ctx->stream;
    if (rgb8) {
        if (total_spp < 1) total_spp = ctx->hs.d.samples_per_pixel;
        CU(ctx->d_rgb8.alloc(need));
        launch_tonemap(ctx->cfg, ctx->d_accum.p, ctx->d_rgb8.p, ctx->ds.width, ctx->ds.height, total_spp, gamma);
        CU(cudaGetLastError());
        CU(cudaMemcpyAsync(rgb8, ctx->d_rgb8.p, need, cudaMemcpyDeviceToHost, s));
    }
    if (rgb_sum) CU(cudaMemcpyAsync(rgb_sum, ctx->d_accum.p, sizeof(float) * need, cudaMemcpyDeviceToHost, s));
    if (stats && ctx->d_rays.p) { unsigned long long r[3] = {0, 0, 0}; CU(cudaMemcpyAsync(r, ctx->d_rays.p, sizeof r, cudaMemcpyDeviceToHost, s)); CU(cudaStreamSynchronize(s)); ctx->stats.rays = (long long)r[0]; ctx->stats.shadow_rays = (long long)r[1]; ctx->stats.rays_extend = (long long)r[2]; }
    CU(cudaStreamSynchronize(s));
    if (!ctx->pending.empty()) {
        float *acc[6] = {&ctx->stats.ms_generate, &ctx->stats.ms_extend, &ctx->stats.ms_shade, &ctx->stats.ms_accumulate, &ctx->stats.ms_shadow,
                         &ctx->stats.ms_tail};
        for (size_t k = 0; k < ctx->pending.size(); ++k) {
            float ms = 0;
            if (cudaEventElapsedTime(&ms, ctx->ev_pool[2 * k], ctx->ev_pool[2 * k + 1]) == cudaSuccess) *acc[ctx->pending[k]] += ms; else cudaGetLastError();
        }
        ctx->pending.clear();
    }
    if (stats) {
        float ms = 0;
        if (cudaEventElapsedTime(&ms, ctx->ev0, ctx->ev1) == cudaSuccess) ctx->stats.ms_total = ms; else cudaGetLastError();
        ctx->stats.launches = ctx->launches;
        *stats = ctx->stats;
    }
    return RT_OK;
}
extern "C" int rt_download(rt_ctx *ctx, float *rgb_sum, size_t n_floats, unsigned char *rgb8, int total_spp, int gamma, rt_stats *stats) {
    return rt_guard(ctx, "rt_download", [&] { return rt_download_impl(ctx, rgb_sum, n_floats, rgb8, total_spp, gamma, stats); });
}

static int rt_bvh_download_impl(rt_ctx *ctx, int32_t *n_prims, uint64_t *morton_sorted, int32_t *prim_sorted, int32_t *left,
                               int32_t *right, int32_t *parent, float *node_aabb, float *leaf_aabb, float *scene_aabb) {
    if (!ctx) return RT_ERR_INVALID;
    if (!ctx->have_scene) return rt_fail(ctx, RT_ERR_NO_SCENE, "rt_bvh_download: no scene uploaded");
    CU(cudaSetDevice(ctx->device));
    CU(cudaStreamSynchronize(ctx->stream));
    const int n = ctx->lb.n;
    if (n_prims) *n_prims = n;
    if (n <= 0) return RT_OK;
    if (morton_sorted) CU(cudaMemcpy(morton_sorted, ctx->d_keys[0].p, sizeof(uint64_t) * n, cudaMemcpyDeviceToHost));
    if (prim_sorted) CU(cudaMemcpy(prim_sorted, ctx->d_vals[0].p, sizeof(int) * n, cudaMemcpyDeviceToHost));
    if (leaf_aabb) CU(cudaMemcpy(leaf_aabb, ctx->d_leaf_aabb.p, sizeof(float) * 6 * n, cudaMemcpyDeviceToHost));
    if (scene_aabb) {
        unsigned b[12];
        CU(cudaMemcpy(b, ctx->d_bounds.p, sizeof b, cudaMemcpyDeviceToHost));
        for (int k = 0; k < 6; ++k) { unsigned u = b[k]; u ^= (u >> 31) ? 0x80000000u : 0xffffffffu; memcpy(&scene_aabb[k], &u, 4); }
    }
    const int nt = ctx->ds.n_tree;      // leaves of the tree; the giants occupy the sorted slots behind them
    if (nt >= 2) {
        if (left) CU(cudaMemcpy(left, ctx->d_left.p, sizeof(int) * (nt - 1), cudaMemcpyDeviceToHost));
        if (right) CU(cudaMemcpy(right, ctx->d_right.p, sizeof(int) * (nt - 1), cudaMemcpyDeviceToHost));
        if (parent) CU(cudaMemcpy(parent, ctx->d_parent.p, sizeof(int) * (2 * nt - 1), cudaMemcpyDeviceToHost));
        if (node_aabb) CU(cudaMemcpy(node_aabb, ctx->d_node_aabb.p, sizeof(float) * 6 * (nt - 1), cudaMemcpyDeviceToHost));
    } else if (parent) parent[0] = -1;
    return RT_OK;
}
extern "C" int rt_bvh_download(rt_ctx *ctx, int32_t *n_prims, uint64_t *morton_sorted, int32_t *prim_sorted, int32_t *left,
                               int32_t *right, int32_t *parent, float *node_aabb, float *leaf_aabb, float *scene_aabb) {
    return rt_guard(ctx, "rt_bvh_download", [&] { return rt_bvh_download_impl(ctx, n_prims, morton_sorted, prim_sorted, left, right, parent, node_aabb, leaf_aabb, scene_aabb); });
}

extern "C" int rt_bvh_info(rt_ctx *ctx, int32_t *n_tree, int32_t *n_giants, int32_t *depth) {
    if (!ctx) return RT_ERR_INVALID;
    if (!ctx->have_scene) return rt_fail(ctx, RT_ERR_NO_SCENE, "rt_bvh_info: no scene uploaded");
    if (n_tree) *n_tree = ctx->ds.n_tree;
    if (n_giants) *n_giants = ctx->ds.n_giants;
    if (depth) *depth = ctx->bvh_depth;
    return RT_OK;
}

extern "C" void rt_free(void *p) { free(p); }

extern "C" int rt_host_alloc(rt_ctx *ctx, void **ptr, size_t bytes) {
    if (!ctx || !ptr) return RT_ERR_INVALID;
    CU(cudaSetDevice(ctx->device));
    *ptr = nullptr;
    CU(cudaHostAlloc(ptr, bytes ? bytes : 1, cudaHostAllocDefault));
    return RT_OK;
}

extern "C" int rt_host_free(rt_ctx *ctx, void *ptr) {
    if (!ctx) return RT_ERR_INVALID;
    if (ptr) CU(cudaFreeHost(ptr));
    return RT_OK;
}
