// rt_multi.cu — multi-GPU behind the C ABI (include/rt.h: rt_comm_*, rt_render_shard, rt_gather_image, rt_group_*).
//
// Replaces the reference's process-per-GPU scheme (gpu-version/blue.py:23-32: one `parallel_compute` per device, frames dealt by
// hand) for still images: the samples of ONE image are dealt to the GPUs in whole sample chunks (rt_sample_chunk), every GPU
// renders its chunks into per-chunk float buffers, and the root gathers the chunk buffers over NVLink (NCCL send / receive, one
// group call on the render streams) and combines them in chunk order (k_fold_chunks).  No float reduction happens inside NCCL, so
// the image is bit-identical for 1, 2, 4 or 8 GPUs: the same eight chunk sums are added in the same order wherever they were
// computed.  Volume: 8 x W x H x 12 bytes at most (1080p: 199 MB, < 0.5 ms at NVLink 5 rates) after >= 50 ms of rendering.
//
// NCCL is bound at run time (dlopen of libnccl.so.2: inside a torch process that is torch's own copy, already loaded), so the
// library has no link-time dependency on it and single-GPU users never touch it.
#include "rt_ctx.h"
#include <dlfcn.h>
#include <mutex>
#include <thread>

namespace {
typedef struct ncclComm *ncclComm_t;
struct ncclUniqueId { char internal[128]; };
enum { ncclSuccess = 0, ncclFloat32 = 7 };

struct NcclApi {
    void *h = nullptr;
    int (*GetUniqueId)(ncclUniqueId *) = nullptr;
    int (*CommInitRank)(ncclComm_t *, int, ncclUniqueId, int) = nullptr;
    int (*CommInitAll)(ncclComm_t *, int, const int *) = nullptr;
    int (*CommDestroy)(ncclComm_t) = nullptr;
    int (*Send)(const void *, size_t, int, int, ncclComm_t, cudaStream_t) = nullptr;
    int (*Recv)(void *, size_t, int, int, ncclComm_t, cudaStream_t) = nullptr;
    int (*GroupStart)() = nullptr;
    int (*GroupEnd)() = nullptr;
    const char *(*GetErrorString)(int) = nullptr;
    std::string err;
};

NcclApi &nccl() {
    static NcclApi api;
    static std::once_flag once;
    std::call_once(once, [] {
        const char *env = getenv("RT_NCCL_LIB");
        const char *names[] = {env, "libnccl.so.2", "libnccl.so"};
        for (const char *n : names) {
            if (!n) continue;
            api.h = dlopen(n, RTLD_NOW | RTLD_GLOBAL);
            if (api.h) break;
        }
        if (!api.h) {
            const char *why = dlerror();             // (the call clears the message: read it once)
            api.err = std::string("cannot load libnccl.so.2 (set RT_NCCL_LIB): ") + (why ? why : "");
            return;
        }
#define SYM(field, name) *(void **)(&api.field) = dlsym(api.h, name); if (!api.field) api.err = std::string("libnccl: missing symbol ") + name
        SYM(GetUniqueId, "ncclGetUniqueId"); SYM(CommInitRank, "ncclCommInitRank"); SYM(CommInitAll, "ncclCommInitAll");
        SYM(CommDestroy, "ncclCommDestroy"); SYM(Send, "ncclSend"); SYM(Recv, "ncclRecv"); SYM(GroupStart, "ncclGroupStart");
        SYM(GroupEnd, "ncclGroupEnd"); SYM(GetErrorString, "ncclGetErrorString");
#undef SYM
    });
    return api;
}

int nccl_fail(rt_ctx *ctx, const char *what, int rc) {
    NcclApi &a = nccl();
    return rt_fail(ctx, RT_ERR_CUDA, std::string(what) + ": " + (a.GetErrorString ? a.GetErrorString(rc) : "NCCL error"));
}
#define NC(ctx, call) do { int rc__ = (call); if (rc__ != ncclSuccess) return nccl_fail(ctx, #call, rc__); } while (0)
#define CUM(ctx, call) do { cudaError_t e__ = (call); if (e__ != cudaSuccess) return rt_fail(ctx, RT_ERR_CUDA, std::string(#call) + ": " + cudaGetErrorString(e__)); } while (0)
}   // namespace

// chunks [k0, k1) of rank `rank` out of n_ranks: contiguous, as even as 8 chunks allow; ranks past the eighth get none
extern "C" void rt_rank_chunks(int n_ranks, int rank, int *k0, int *k1) {
    if (n_ranks < 1) n_ranks = 1;
    if (rank < 0) rank = 0;
    if (rank >= n_ranks) rank = n_ranks - 1;
    const int a = (int)((long long)rank * RT_SAMPLE_CHUNKS / n_ranks), b = (int)((long long)(rank + 1) * RT_SAMPLE_CHUNKS / n_ranks);
    if (k0) *k0 = a;
    if (k1) *k1 = b;
}

extern "C" int rt_comm_unique_id(void *id, size_t bytes) {
    if (!id || bytes < sizeof(ncclUniqueId)) return rt_fail(nullptr, RT_ERR_INVALID, "rt_comm_unique_id: need a 128-byte buffer");
    NcclApi &a = nccl();
    if (!a.err.empty()) return rt_fail(nullptr, RT_ERR_IO, a.err);
    ncclUniqueId u;
    NC(nullptr, a.GetUniqueId(&u));
    memcpy(id, &u, sizeof u);
    return RT_OK;
}

extern "C" int rt_comm_init(rt_ctx *ctx, const void *id, int n_ranks, int rank) {
    if (!ctx) return RT_ERR_INVALID;
    if (!id || n_ranks < 1 || rank < 0 || rank >= n_ranks) return rt_fail(ctx, RT_ERR_INVALID, "rt_comm_init: bad arguments");
    NcclApi &a = nccl();
    if (!a.err.empty()) return rt_fail(ctx, RT_ERR_IO, a.err);
    if (ctx->comm) return rt_fail(ctx, RT_ERR_INVALID, "rt_comm_init: the context already has a communicator");
    CUM(ctx, cudaSetDevice(ctx->device));
    ncclUniqueId u;
    memcpy(&u, id, sizeof u);
    ncclComm_t c = nullptr;
    NC(ctx, a.CommInitRank(&c, n_ranks, u, rank));
    ctx->comm = c; ctx->comm_ranks = n_ranks; ctx->comm_rank = rank; ctx->comm_owned = true;
    return RT_OK;
}

extern "C" int rt_comm_destroy(rt_ctx *ctx) {
    if (!ctx) return RT_ERR_INVALID;
    if (ctx->comm && ctx->comm_owned) { cudaSetDevice(ctx->device); cudaStreamSynchronize(ctx->stream); nccl().CommDestroy((ncclComm_t)ctx->comm); }
    ctx->comm = nullptr; ctx->comm_ranks = 1; ctx->comm_rank = 0; ctx->comm_owned = false;
    return RT_OK;
}

// This rank's whole chunks of the image: samples [chunk k0 begin, chunk k1-1 end) of prm->spp_count (or the scene's) samples.
extern "C" int rt_render_shard(rt_ctx *ctx, const rt_render_params *prm) {
    if (!ctx) return RT_ERR_INVALID;
    if (!prm) return rt_fail(ctx, RT_ERR_INVALID, "rt_render_shard: params is NULL");
    if (!ctx->have_scene) return rt_fail(ctx, RT_ERR_NO_SCENE, "rt_render_shard: no scene uploaded");
    if (prm->spp_begin != 0) return rt_fail(ctx, RT_ERR_INVALID, "rt_render_shard: spp_begin must be 0 (the shard is derived from the rank)");
    const int total = prm->spp_count > 0 ? prm->spp_count : ctx->hs.d.samples_per_pixel;
    int k0, k1, b, e, dummy;
    rt_rank_chunks(ctx->comm_ranks, ctx->comm_rank, &k0, &k1);
    rt_sample_chunk(total, k0, &b, &dummy);
    rt_sample_chunk(total, k1, &e, &dummy);          // k1 == 8 yields (total, total)
    ctx->shard_total = total;
    for (int k = 0; k < RT_ACC_CHUNKS; ++k) ctx->chunk_live[k] = 0;
    if (e <= b) return RT_OK;                        // more ranks than samples: nothing to do here
    rt_render_params p = *prm;
    p.spp_begin = b; p.spp_count = e - b; p.total_spp = total; p.flags |= RT_RENDER_KEEP_CHUNKS;
    p.flags &= ~RT_RENDER_KEEP_ACCUM;
    return rt_render(ctx, &p);
}

// Enqueues this rank's part of the gather on its stream (inside an NCCL group opened by the caller) and returns the chunks that
// end up on the root.
// prepare = true: allocate the root's receive buffers (before the NCCL group is opened), nothing else.
static int gather_enqueue(rt_ctx *ctx, int root, int total, bool *root_chunks, bool prepare) {
    NcclApi &a = nccl();
    const size_t n_img = (size_t)ctx->ds.width * ctx->ds.height * 3;
    CUM(ctx, cudaSetDevice(ctx->device));
    for (int r = 0; r < ctx->comm_ranks; ++r) {
        int k0, k1;
        rt_rank_chunks(ctx->comm_ranks, r, &k0, &k1);
        for (int k = k0; k < k1; ++k) {
            int b, e;
            rt_sample_chunk(total, k, &b, &e);
            if (e <= b) continue;                    // empty chunk (fewer samples than chunks)
            if (root_chunks) root_chunks[k] = true;
            if (r == root) continue;                 // the root's own chunks are already in place
            if (ctx->comm_rank == root) {
                if (prepare) CUM(ctx, ctx->d_chunk[k].alloc(n_img));
                else NC(ctx, a.Recv(ctx->d_chunk[k].p, n_img, ncclFloat32, r, (ncclComm_t)ctx->comm, ctx->stream));
            } else if (ctx->comm_rank == r && !prepare) {
                if (!ctx->chunk_live[k]) return rt_fail(ctx, RT_ERR_INVALID, "rt_gather_image: chunk not rendered (call rt_render_shard first)");
                NC(ctx, a.Send(ctx->d_chunk[k].p, n_img, ncclFloat32, root, (ncclComm_t)ctx->comm, ctx->stream));
            }
        }
    }
    return RT_OK;
}

static int fold_on_root(rt_ctx *ctx, const bool *root_chunks) {
    const size_t n_img = (size_t)ctx->ds.width * ctx->ds.height * 3;
    CUM(ctx, cudaSetDevice(ctx->device));
    if (ctx->d_accum.n < n_img) CUM(ctx, ctx->d_accum.alloc(n_img));
    ChunkPtrs cp{};
    int n = 0;
    for (int k = 0; k < RT_ACC_CHUNKS; ++k) if (root_chunks[k]) cp.p[n++] = ctx->d_chunk[k].p;
    launch_fold_chunks(ctx->cfg, ctx->d_accum.p, cp, n, false, (long long)n_img);
    CUM(ctx, cudaGetLastError());
    ctx->accum_valid = true;
    return RT_OK;
}

extern "C" int rt_gather_image(rt_ctx *ctx, int root) {
    if (!ctx) return RT_ERR_INVALID;
    if (!ctx->have_scene) return rt_fail(ctx, RT_ERR_NO_SCENE, "rt_gather_image: no scene uploaded");
    if (root < 0 || root >= ctx->comm_ranks) return rt_fail(ctx, RT_ERR_INVALID, "rt_gather_image: bad root");
    if (ctx->shard_total <= 0) return rt_fail(ctx, RT_ERR_INVALID, "rt_gather_image: nothing rendered with rt_render_shard");
    bool root_chunks[RT_ACC_CHUNKS] = {};
    if (ctx->comm_ranks > 1) {
        if (!ctx->comm) return rt_fail(ctx, RT_ERR_INVALID, "rt_gather_image: no communicator (rt_comm_init)");
        NcclApi &a = nccl();
        int rc = gather_enqueue(ctx, root, ctx->shard_total, nullptr, true);
        if (rc) return rc;
        NC(ctx, a.GroupStart());
        rc = gather_enqueue(ctx, root, ctx->shard_total, root_chunks, false);
        NC(ctx, a.GroupEnd());
        if (rc) return rc;
    } else {
        for (int k = 0; k < RT_ACC_CHUNKS; ++k) root_chunks[k] = ctx->chunk_live[k] != 0;
    }
    if (ctx->comm_rank == root) return fold_on_root(ctx, root_chunks);
    return RT_OK;
}

// ---------------- one process, several devices ----------------
struct rt_group {
    std::vector<rt_ctx *> ctx;
    std::string err;
};

static int group_fail(rt_group *g, int code, const std::string &m) { if (g) g->err = m; return code; }

extern "C" const char *rt_group_last_error(const rt_group *g) { return g ? g->err.c_str() : "rt_group is NULL"; }

extern "C" void rt_group_destroy(rt_group *g) {
    if (!g) return;
    NcclApi &a = nccl();
    for (rt_ctx *c : g->ctx) {
        if (c && c->comm && a.CommDestroy) { cudaSetDevice(c->device); cudaStreamSynchronize(c->stream); a.CommDestroy((ncclComm_t)c->comm); c->comm = nullptr; }
        rt_destroy(c);
    }
    delete g;
}

static int rt_group_create_impl(rt_group **out, const int *devices, int n_devices);
extern "C" int rt_group_create(rt_group **out, const int *devices, int n_devices) {
    if (!out) return RT_ERR_INVALID;
    *out = nullptr;
    return rt_guard(nullptr, "rt_group_create", [&] { return rt_group_create_impl(out, devices, n_devices); });
}
static int rt_group_create_impl(rt_group **out, const int *devices, int n_devices) {
    if (!devices || n_devices < 1 || n_devices > 64) return rt_fail(nullptr, RT_ERR_INVALID, "rt_group_create: bad device list");
    rt_group *g = new rt_group();
    for (int i = 0; i < n_devices; ++i) {
        rt_ctx *c = nullptr;
        const int rc = rt_create(&c, devices[i], nullptr);
        if (rc) { const std::string m = rt_last_error(nullptr); rt_group_destroy(g); return rt_fail(nullptr, rc, m); }
        c->comm_ranks = n_devices; c->comm_rank = i;
        g->ctx.push_back(c);
    }
    if (n_devices > 1) {
        NcclApi &a = nccl();
        if (!a.err.empty()) { const std::string m = a.err; rt_group_destroy(g); return rt_fail(nullptr, RT_ERR_IO, m); }
        std::vector<ncclComm_t> comms((size_t)n_devices);
        const int rc = a.CommInitAll(comms.data(), n_devices, devices);
        if (rc != ncclSuccess) { const std::string m = std::string("ncclCommInitAll: ") + a.GetErrorString(rc); rt_group_destroy(g); return rt_fail(nullptr, RT_ERR_CUDA, m); }
        for (int i = 0; i < n_devices; ++i) { g->ctx[i]->comm = comms[i]; g->ctx[i]->comm_owned = false; }
    }
    *out = g;
    return RT_OK;
}

extern "C" int rt_group_size(const rt_group *g) { return g ? (int)g->ctx.size() : 0; }
extern "C" rt_ctx *rt_group_ctx(rt_group *g, int i) { return (g && i >= 0 && i < (int)g->ctx.size()) ? g->ctx[i] : nullptr; }

// the same call on every context, one host thread per device (uploads synchronise; eight of them in a row would serialise)
template <class F>
static int group_each(rt_group *g, const char *what, F &&fn) {
    std::vector<int> rc(g->ctx.size(), 0);
    std::vector<std::thread> th;
    bool spawned = true;
    try { for (size_t i = 1; i < g->ctx.size(); ++i) th.emplace_back([&, i] { rc[i] = fn(g->ctx[i]); }); }
    catch (...) { spawned = false; }              // (std::system_error: no more threads) the ones that started still get joined
    if (spawned) rc[0] = fn(g->ctx[0]);
    for (auto &t : th) t.join();
    if (!spawned) return group_fail(g, RT_ERR_INVALID, std::string(what) + ": cannot start a host thread per device");
    for (size_t i = 0; i < rc.size(); ++i)
        if (rc[i]) return group_fail(g, rc[i], std::string(what) + " on device " + std::to_string(g->ctx[i]->device) + ": " + rt_last_error(g->ctx[i]));
    return RT_OK;
}

extern "C" int rt_group_scene_upload(rt_group *g, const rt_scene_desc *scene) {
    if (!g) return RT_ERR_INVALID;
    return group_each(g, "rt_scene_upload", [&](rt_ctx *c) { return rt_scene_upload(c, scene); });
}
extern "C" int rt_group_scene_upload_json(rt_group *g, const char *path) {
    if (!g) return RT_ERR_INVALID;
    return group_each(g, "rt_scene_upload_json", [&](rt_ctx *c) { return rt_scene_upload_json(c, path); });
}
extern "C" int rt_group_scene_upload_json_text(rt_group *g, const char *text, const char *base_dir) {
    if (!g) return RT_ERR_INVALID;
    return group_each(g, "rt_scene_upload_json_text", [&](rt_ctx *c) { return rt_scene_upload_json_text(c, text, base_dir); });
}
extern "C" int rt_group_scene_override(rt_group *g, int width, int height, int spp, int max_depth) {
    if (!g) return RT_ERR_INVALID;
    for (rt_ctx *c : g->ctx) { const int rc = rt_scene_override(c, width, height, spp, max_depth); if (rc) return group_fail(g, rc, rt_last_error(c)); }
    return RT_OK;
}

// Every device renders its chunks (asynchronous enqueue, one device after the other from this thread), then ONE NCCL group moves the
// chunk buffers to device 0, which folds them in chunk order.  Returns without synchronising; rt_group_download synchronises.
extern "C" int rt_group_render(rt_group *g, const rt_render_params *prm) {
    if (!g) return RT_ERR_INVALID;
    if (!prm) return group_fail(g, RT_ERR_INVALID, "rt_group_render: params is NULL");
    for (rt_ctx *c : g->ctx) { const int rc = rt_render_shard(c, prm); if (rc) return group_fail(g, rc, rt_last_error(c)); }
    bool root_chunks[RT_ACC_CHUNKS] = {};
    rt_ctx *root = g->ctx[0];
    if (g->ctx.size() > 1) {
        NcclApi &a = nccl();
        int rc = gather_enqueue(root, 0, root->shard_total, nullptr, true);
        if (rc) return group_fail(g, rc, rt_last_error(root));
        if (a.GroupStart() != ncclSuccess) return group_fail(g, RT_ERR_CUDA, "ncclGroupStart");
        for (rt_ctx *c : g->ctx) { rc = gather_enqueue(c, 0, c->shard_total, c == root ? root_chunks : nullptr, false); if (rc) { g->err = rt_last_error(c); break; } }
        const int rg = a.GroupEnd();
        if (rc) return rc;
        if (rg != ncclSuccess) return group_fail(g, RT_ERR_CUDA, std::string("ncclGroupEnd: ") + a.GetErrorString(rg));
    } else {
        for (int k = 0; k < RT_ACC_CHUNKS; ++k) root_chunks[k] = root->chunk_live[k] != 0;
    }
    const int rc = fold_on_root(root, root_chunks);
    if (rc) return group_fail(g, rc, rt_last_error(root));
    return RT_OK;
}

// Image from device 0 (the root); stats->rays / samples / launches are summed over the devices, ms_total is the largest.
extern "C" int rt_group_download(rt_group *g, float *rgb_sum, size_t n_floats, unsigned char *rgb8, int total_spp, int gamma, rt_stats *stats) {
    if (!g) return RT_ERR_INVALID;
    rt_stats total{};
    for (size_t i = 1; i < g->ctx.size(); ++i) {
        rt_stats st{};
        const int rc = rt_download(g->ctx[i], nullptr, 0, nullptr, total_spp, gamma, &st);
        if (rc) return group_fail(g, rc, rt_last_error(g->ctx[i]));
        total.samples += st.samples; total.rays += st.rays; total.shadow_rays += st.shadow_rays; total.launches += st.launches; total.passes += st.passes;
        total.ms_total = std::max(total.ms_total, st.ms_total);
    }
    rt_stats st0{};
    const int rc = rt_download(g->ctx[0], rgb_sum, n_floats, rgb8, total_spp, gamma, &st0);
    if (rc) return group_fail(g, rc, rt_last_error(g->ctx[0]));
    if (stats) {
        *stats = st0;
        stats->samples += total.samples; stats->rays += total.rays; stats->shadow_rays += total.shadow_rays; stats->launches += total.launches;
        stats->passes += total.passes; stats->ms_total = std::max(st0.ms_total, total.ms_total);
    }
    return RT_OK;
}
