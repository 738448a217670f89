// rt_ctx.h — the host-side context behind the opaque rt_ctx of include/rt.h.
#pragma once
#include <new>
#include <stdexcept>
#include <string>
#include <vector>
#include "rt_kernels.h"

struct HostScene {                 // owning copy of an rt_scene_desc (what rt_scene_upload received / the JSON reader produced)
    rt_scene_desc d{};
    std::vector<rt_object_desc> objects;
    std::vector<rt_xform_desc> xforms;
    std::vector<rt_material_desc> materials;
    std::vector<rt_texture_desc> textures;
    std::vector<rt_image_desc> images;
    std::vector<std::vector<uint8_t>> image_data;
    std::string output_file = "main.png";
    void fix_pointers();
};

template <class T> struct DevBuf {
    T *p = nullptr; size_t n = 0;
    cudaError_t alloc(size_t count) {
        if (count <= n && p) return cudaSuccess;
        release();
        cudaError_t e = cudaMalloc((void **)&p, sizeof(T) * (count ? count : 1));
        if (e == cudaSuccess) n = count ? count : 1; else p = nullptr;
        return e;
    }
    void release() { if (p) cudaFree(p); p = nullptr; n = 0; }
};

#define RT_MAX_SETS 4
struct QueueSet {
    Queues q{};
    DevBuf<float4> ray_o[2], ray_d[2], thr[2], radiance, sh_o, sh_d, sh_c; DevBuf<int2> hits; DevBuf<int> qcount, workctr, shcount;
    long long cap = 0;
    LaunchCfg cfg{};                    // the context's launch configuration on this set's stream
    cudaEvent_t ev_acc = nullptr;       // recorded after this set's accumulate: orders the accumulation across streams
    void release() {
        for (int k = 0; k < 2; ++k) { ray_o[k].release(); ray_d[k].release(); thr[k].release(); }
        radiance.release(); sh_o.release(); sh_d.release(); sh_c.release(); hits.release(); qcount.release(); workctr.release(); shcount.release();
        cap = 0;
    }
};

struct rt_ctx {
    int device = 0;
    cudaStream_t stream = nullptr;
    bool own_stream = false;
    int num_sms = 0;
    size_t smem_optin = 0;
    std::string err;
    long long launches = 0;

    bool have_scene = false;
    bool scene_scatters = true;      // false: every material referenced by an object is a diffuse_light (paths end at their first vertex)
    HostScene hs;
    DScene ds{};
    LaunchCfg cfg{};

    // scene tables
    DevBuf<DObject> d_objects; DevBuf<float4> d_xforms; DevBuf<DMaterial> d_mats; DevBuf<DTexture> d_texs; DevBuf<DImage> d_images;
    std::vector<cudaArray_t> img_arrays; std::vector<cudaTextureObject_t> img_tex;
    // LBVH
    LbvhBuffers lb{};
    DevBuf<float> d_obmin, d_obmax, d_node_aabb, d_leaf_aabb; DevBuf<unsigned> d_bounds, d_hist;
    DevBuf<unsigned long long> d_keys[2]; DevBuf<int> d_vals[2], d_left, d_right, d_parent, d_flags, d_prim_obj, d_prim_mat, d_lbvh_info; DevBuf<DMaterial> d_slot_mat;
    DevBuf<unsigned char> d_giant;
    int bvh_depth = 0;
    DevBuf<BvhNode> d_nodes; DevBuf<BvhNodeS> d_nodes_s; DevBuf<BvhNode4> d_nodes4; DevBuf<float4> d_primA, d_primB, d_primC;
    // sampled lights (next-event estimation)
    DevBuf<int> d_light_slot, d_light_type, d_obj_light; DevBuf<float> d_light_area;
    // wavefront: two independent queue sets, one per stream, so that the short late-bounce launches of one pass overlap
    // the long early-bounce launches of the next (rt_render)
    QueueSet qs[RT_MAX_SETS];
    cudaStream_t extra_stream[RT_MAX_SETS] = {};   // pass streams 1.. (ours; pass stream 0 is `stream`)
    cudaEvent_t ev_fork = nullptr;
    DevBuf<unsigned long long> d_counters, d_rays;
    // primary-ray candidate lists per 8x4 tile (rebuilt after an upload or a size override)
    DevBuf<int> d_tile_list, d_tile_cnt, d_tile_bad, d_tile_prefix; bool tiles_dirty = true, tiles_all_empty = false; int tiles_n_empty = 0;
    DevBuf<float> d_chunk[RT_ACC_CHUNKS];   // per-sample-chunk sums of the current render (folded into d_accum in chunk order)
    int chunk_live[RT_ACC_CHUNKS] = {};    // chunks rendered by the last rt_render with RT_RENDER_KEEP_CHUNKS (multi-GPU gather)
    DevBuf<float> d_accum; DevBuf<unsigned char> d_rgb8; DevBuf<float> d_aov; DevBuf<const float *> d_partials;
    // stats
    rt_stats stats{};
    cudaEvent_t ev0 = nullptr, ev1 = nullptr, ev_count[2] = {nullptr, nullptr};
    std::vector<cudaEvent_t> ev_pool;   // RT_RENDER_TIME_STAGES: (start, stop) pairs, resolved in rt_download
    std::vector<int> pending;           // stage id per recorded pair
    bool accum_valid = false;
    // multi-GPU (rt_multi.cu): NCCL communicator of this context's rank, bound at run time
    void *comm = nullptr; int comm_ranks = 1, comm_rank = 0; bool comm_owned = false;
    int shard_total = 0;               // total sample count of the last rt_render_shard
};

int rt_fail(rt_ctx *ctx, int code, const std::string &msg);
// No C++ exception may cross the extern "C" boundary (through ctypes that would be std::terminate in the host process): entry points
// that allocate host memory run their body through this guard, which maps an exception to an RT_ERR_* code + message.
template <class F> int rt_guard(rt_ctx *ctx, const char *what, F &&f) {
    try { return f(); }
    catch (const std::bad_alloc &) { return rt_fail(ctx, RT_ERR_INVALID, std::string(what) + ": out of host memory (size fields too large?)"); }
    catch (const std::exception &e) { return rt_fail(ctx, RT_ERR_INVALID, std::string(what) + ": " + e.what()); }
    catch (...) { return rt_fail(ctx, RT_ERR_INVALID, std::string(what) + ": unknown C++ exception"); }
}
void rt_build_xform(const float axis[3], double angle_deg, bool has_rotate, const float t3[3], bool has_translate, rt_xform_desc *out);
int rt_json_to_scene(const char *text, const char *base_dir, HostScene &out, std::string &err);
int rt_png_decode_file(const char *path, std::vector<uint8_t> &rgb, int &w, int &h, std::string &err);
