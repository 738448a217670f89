// rt_kernels.h — host-callable launchers of the device kernels (rt_kernels.cu, rt_lbvh.cu).
#pragma once
#include "rt_wavefront.cuh"

struct LaunchCfg {
    cudaStream_t stream;
    int num_sms;
    int extend_blocks_per_sm;     // resident CTAs per SM for the extend kernel (depends on the smem BVH copy)
    size_t bvh_smem_bytes;        // 0: traverse from global memory
    long long *launches;          // counter of kernels launched (rt_stats.launches)
};

void launch_pass_init(const LaunchCfg &c, const Queues &q, int max_depth, int n_queue0);
void launch_tile_lists(const LaunchCfg &c, const DScene &sc, int *list, int *tile_cnt, int *n_bad);
void launch_tile_prefix(const LaunchCfg &c, const int *tile_cnt, int n_tiles, int *prefix);
void launch_generate(const LaunchCfg &c, const DScene &sc, const Queues &q, unsigned seed, int spp_begin, int n_paths, bool zero_radiance,
                     bool hits, bool count);
void launch_extend(const LaunchCfg &c, const DScene &sc, const Queues &q, int bounce, int buf, bool count, bool fixup);
void launch_shade(const LaunchCfg &c, const DScene &sc, const Queues &q, unsigned seed, int spp_begin, int bounce, int buf,
                  bool last_bounce, bool count, bool sort, bool nee);
void launch_tail(const LaunchCfg &c, const DScene &sc, const Queues &q, unsigned seed, int spp_begin, int b0, int buf);
void launch_light_slots(const LaunchCfg &c, int n, const int *prim_obj, const int *obj_light, int *light_slot);
void launch_slot_mats(const LaunchCfg &c, int n, const int *prim_mat, const DMaterial *mats, DMaterial *slot_mat);
void launch_shadow(const LaunchCfg &c, const DScene &sc, const Queues &q, int bounce);
#define RT_ACC_CHUNKS 8            // sample chunks of the fixed-order accumulation (rt.h: rt_sample_chunk)
struct ChunkPtrs { const float *p[RT_ACC_CHUNKS]; };
// one pass's sample planes by chunk: run j = planes [first[j], first[j + 1]) goes to buf[j] (cleared first when clear[j])
struct AccumPlan { int n; int first[RT_ACC_CHUNKS + 1]; int clear[RT_ACC_CHUNKS]; float *buf[RT_ACC_CHUNKS]; };
void launch_accumulate(const LaunchCfg &c, const Queues &q, const AccumPlan &plan, int W, int H, int n_samples, int max_depth,
                       unsigned long long *total_rays, const int *const_tiles, float3 bg, int ext_first, int ext_end);
void launch_primary_only(const LaunchCfg &c, const DScene &sc, const AccumPlan &plan, unsigned seed, int spp_begin, int n_samples,
                         unsigned long long *total_rays);
void launch_fold_chunks(const LaunchCfg &c, float *accum, const ChunkPtrs &cp, int n_chunks, bool keep, long long n);
void launch_sum_ordered(const LaunchCfg &c, float *accum, const float *const *partials_dev, int n_partials, long long n);
void launch_tonemap(const LaunchCfg &c, const float *accum, unsigned char *rgb8, int W, int H, int spp, int gamma);
void launch_aov(const LaunchCfg &c, const DScene &sc, float *out);
int extend_max_blocks_per_sm(size_t smem_bytes);
#define RT_MAX_DEVICES 64
cudaError_t kernels_configure_device(int dev, size_t smem_optin);   // once per device: dynamic shared-memory limits of every kernel instance

// ---- LBVH build (rt_lbvh.cu) ----
struct DObject { int type, material, xform, pad; float p[12]; };   // = rt_object_desc

#define RT_LBVH_INFO_WORDS (4 + RT_MAX_GIANTS)
struct LbvhBuffers {                // all device pointers, sized for n objects
    int n;
    const DObject *objects;
    const float4 *xforms;           // 6 float4 per transform
    float *obj_bmin, *obj_bmax;     // 3 floats per object (unpadded boxes)
    unsigned *bounds;               // 12 ordered-uint words: scene min/max, centroid min/max
    unsigned long long *keys[2];    // Morton keys, ping-pong
    int *vals[2];                   // object indices, ping-pong
    unsigned *hist;                 // radix histograms: 256 * n_tiles
    int *left, *right, *parent;     // Karras topology (n-1, n-1, 2n-1)
    int *flags;                     // refit arrival counters (n-1)
    float *node_aabb, *leaf_aabb;   // 6 floats per internal node / per sorted leaf
    unsigned char *giant;           // per object: kept out of the tree (k_giants_mark)
    int *info;                      // RT_LBVH_INFO_WORDS ints: [0] leaves of the tree, [1] giants, [2] tree depth, [3] giant candidates,
                                    // [4 ..] leaf codes of the giants
    float4 cam;                     // camera origin + lens radius (padding of a giant sphere's box, read by the tile lists only)
    // outputs consumed by traversal
    BvhNode *nodes;
    BvhNodeS *nodes_s;
    float4 *primA, *primB, *primC;
    int *prim_obj, *prim_mat;
};
void lbvh_build(const LaunchCfg &c, LbvhBuffers &b);
void lbvh_pack4(const LaunchCfg &c, const LbvhBuffers &b, BvhNode4 *nodes4);      // after lbvh_build: four-wide nodes for global-memory trees
size_t lbvh_hist_words(int n);
