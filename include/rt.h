/* rt.h — C ABI of the B200-native path tracer (librt_b200.so).
 *
 * This is the drop-in boundary for the render path of WhenMelancholy/ray-tracing-in-cuda
 * (reference paths below are relative to the reference repository root):
 *
 *   rt_scene_upload_json  replaces  parse_scene()            gpu-version/parser.hpp:504-573
 *                         and       move_to_device<<<1,1>>>  gpu-version/main.cu:374-446,473
 *   rt_scene_upload       same, for callers that already hold the scene as flat tables
 *                         (the `struct scene` seam, gpu-version/parser.hpp:16-32)
 *   rt_render             replaces  init_random_library<<<>>> + render<<<>>>   gpu-version/main.cu:120-125,72-105,496,505
 *   rt_download           replaces  the managed-memory read-back + write_color/write_image tonemap
 *                                   gpu-version/main.cu:359-372, gpu-version/color.cuh:15-35,70-95
 *   rt_write_ppm/_png     replace   output_image()/write_image()   gpu-version/main.cu:359-372, color.cuh:15-35
 *
 * Conventions: every function returns 0 on success and a non-zero RT_ERR_* code otherwise; the
 * message is available from rt_last_error().  Nothing in the library calls exit() or throws across
 * the boundary (the reference's check_cuda() does exit(99): gpu-version/rtweekend.cuh:43-53 — the
 * CLI, not the library, keeps that exit code).  Plain pointers and sizes only; the library copies
 * what it needs, the caller keeps ownership of every host buffer.  One rt_ctx may be driven from one
 * host thread at a time; different contexts are independent.
 *
 * Image convention (same as the reference): float RGB SUMS over samples, 12 bytes per pixel,
 * row 0 = BOTTOM row (gpu-version/main.cu:104; the writers flip rows).
 */
#ifndef RT_B200_H
#define RT_B200_H
#include <stddef.h>
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

#define RT_ABI_VERSION 2

/* error codes */
enum { RT_OK = 0, RT_ERR_INVALID = 1, RT_ERR_CUDA = 2, RT_ERR_IO = 3, RT_ERR_PARSE = 4, RT_ERR_NO_SCENE = 5, RT_ERR_NO_DEVICE = 6 };

/* scene vocabulary: gpu-version/rtweekend.cuh:70-91 (enum class class_type) + the extensions BASELINE.json names */
enum { RT_OBJ_SPHERE = 0, RT_OBJ_XY_RECT = 1, RT_OBJ_XZ_RECT = 2, RT_OBJ_YZ_RECT = 3, RT_OBJ_CYLINDER = 4,
       RT_OBJ_BOX = 5, RT_OBJ_TRIANGLE = 6 };
enum { RT_MAT_LAMBERTIAN = 0, RT_MAT_METAL = 1, RT_MAT_DIELECTRIC = 2, RT_MAT_DIFFUSE_LIGHT = 3 };
enum { RT_TEX_SOLID = 0, RT_TEX_CHECKER = 1, RT_TEX_IMAGE = 2 };
/* miss colour: constant `background` (gpu-version/main.cu:61-63) or the RTIOW sky gradient (cmake-cpu-version/main.cpp:36-38) */
enum { RT_BG_CONSTANT = 0, RT_BG_SKY_GRADIENT = 1 };
/* which of the reference's two (disagreeing) sampling conventions to reproduce — SURVEY.md §7.2-1:
 *  REFERENCE_GPU: uniforms in (0,1]; random_in_unit_sphere over the positive octant only (gpu-version/vec3.cuh:133-141);
 *                 lambertian = n + random_in_unit_sphere (material.cuh:42); no lens jitter (camera.cuh:33-34)
 *  REFERENCE_CPU: uniforms in [0,1); ball sampled over [-1,1)^3 (cmake-cpu-version/vec3.h:121-129);
 *                 lambertian = n + random_unit_vector (material.h:27); thin-lens defocus (camera.h:32-39) */
enum { RT_SAMPLING_REFERENCE_GPU = 0, RT_SAMPLING_REFERENCE_CPU = 1,
       /* flag, OR-ed to the family: draw the same distributions (uniform ball / octant ball, unit sphere, lens disk) DIRECTLY from
        * one Philox block instead of through the reference's rejection loops (vec3.cuh:133-141, vec3.h:121-129,157-165).  The
        * reference is matched statistically only (its generators are XORWOW / mt19937), so both satisfy the converged-render
        * tolerance; the loops diverge (acceptance pi/6 per trial) and were half the instructions of the shade stage.  The JSON
        * reader sets it unless the scene says "sampler": "rejection". */
       RT_SAMPLING_DIRECT = 2 };

typedef struct rt_object_desc {
    int32_t type;       /* RT_OBJ_* */
    int32_t material;   /* index into materials */
    int32_t xform;      /* index into xforms (object-to-world, rigid) or -1; honoured for SPHERE, ??_RECT, CYLINDER (required) and BOX —
                           the reference's instance machinery, vec3.cuh:303-427 / object.cuh:225-231; TRIANGLE vertices are world space
                           (the JSON reader bakes "rotate" / "translate" of triangles and meshes into them) */
    int32_t reserved;
    /* SPHERE:   cx cy cz radius              (object.cuh:40-94; radius may be negative)
     * ??_RECT:  a0 a1 b0 b1 k                (object.cuh:96-197; xy: x0 x1 y0 y1 | xz: x0 x1 z0 z1 | yz: y0 y1 z0 z1)
     * CYLINDER: radius zmin zmax             (object.cuh:216-297; axis = object-space z, no caps)
     * BOX:      minx miny minz maxx maxy maxz (extension; object-space if xform >= 0)
     * TRIANGLE: v0 v1 v2 (9 floats, world space; extension); p[9..11]: per-vertex texture coordinates, the BITS of each float =
     *           u | v << 16 as 16-bit unorm (OBJ `vt`); all three zero = none (uv = the barycentrics) */
    float p[12];
} rt_object_desc;

/* 3x4 row-major object-to-world matrix and its inverse (gpu-version/vec3.cuh:303-379 keeps both) */
typedef struct rt_xform_desc { float m[12]; float m_inv[12]; } rt_xform_desc;

typedef struct rt_material_desc {
    int32_t type;       /* RT_MAT_* */
    int32_t texture;    /* LAMBERTIAN albedo / DIFFUSE_LIGHT emit: index into textures (material.cuh:31-35,163-167) */
    float albedo[3];    /* METAL (material.cuh:57-61) */
    float fuzz;         /* METAL; clamped to <= 1 on upload (material.cuh:61) */
    float ior;          /* DIELECTRIC index_of_refraction (material.cuh:89-92) */
    int32_t reserved;
} rt_material_desc;

typedef struct rt_texture_desc {
    int32_t type;       /* RT_TEX_* */
    int32_t even, odd;  /* CHECKER: texture indices (texture.cuh:33-57) */
    int32_t image;      /* IMAGE: index into images */
    float color[3];     /* SOLID (texture.cuh:14-31); values > 1 allowed (emitters) */
    float reserved;
} rt_texture_desc;

typedef struct rt_image_desc { int32_t width, height; const uint8_t *rgb8; /* width*height*3, row 0 = top */ } rt_image_desc;

typedef struct rt_camera_desc {      /* gpu-version/camera.cuh:9-29, parser.hpp:113-126 */
    float lookfrom[3], lookat[3], vup[3];
    float vfov_deg;
    float aspect;       /* reference: width/height (parser.hpp:122); cmake-cpu-version pins 16/9 (main.cpp:64,94) */
    float aperture;
    float focus_dist;   /* reference: |lookfrom - lookat| (parser.hpp:124) */
} rt_camera_desc;

typedef struct rt_scene_desc {
    int32_t width, height, max_depth, samples_per_pixel;   /* parser.hpp:512-517 */
    float background[3];
    int32_t background_mode;   /* RT_BG_* */
    int32_t sampling;          /* RT_SAMPLING_* */
    rt_camera_desc camera;
    int32_t n_objects, n_xforms, n_materials, n_textures, n_images;
    int32_t reserved;
    const rt_object_desc *objects;
    const rt_xform_desc *xforms;
    const rt_material_desc *materials;
    const rt_texture_desc *textures;
    const rt_image_desc *images;
} rt_scene_desc;

/* rt_render flags */
enum { RT_RENDER_KEEP_ACCUM = 1,     /* add to the accumulation buffer instead of clearing it first */
       RT_RENDER_COUNT_EVENTS = 2,   /* fill the event counters of rt_stats (slower; counts are deterministic) */
       RT_RENDER_NO_SORT = 4,        /* accepted and ignored (sorting is opt-in, see RT_RENDER_SORT) */
       RT_RENDER_NEE = 8,            /* next-event estimation through the shadow stage (changes the estimator; default off): emissive
                                        spheres, rects, cylinders, boxes and triangles are sampled, transformed instances included */
       RT_RENDER_TIME_STAGES = 16,   /* record CUDA events around every stage launch (no synchronisation); per-stage ms in rt_stats */
       RT_RENDER_SORT = 32,          /* sort every shade chunk by material class in shared memory (same image; pays off only for scenes
                                        with an even material mix — measured slower on the reference's scenes) */
       RT_RENDER_SERIAL = 64,      /* one pass at a time on the context's stream (default: two passes in flight on two streams, so the
                                        short late-bounce launches of one overlap the other; same image either way) */
       RT_RENDER_NO_TILE_LISTS = 256, /* traverse the tree for bounce 0 too (default: the generate stage resolves bounce-0 hits from the
                                        per-tile candidate lists of the camera beams; same image either way) */
       RT_RENDER_NO_TAIL = 1024,     /* keep the queue schedule to the last bounce (default: from bounce 8 on — 2 on all-sphere scenes — one launch finishes every
                                        surviving path — same image either way) */
       RT_RENDER_NEE_MIS = 2048,     /* RT_RENDER_NEE with multiple importance sampling: light samples and the emission a scattered ray finds on
                                        a sampled light are combined with the balance heuristic (implies RT_RENDER_NEE) */
       RT_RENDER_KEEP_CHUNKS = 512 };  /* leave this call's chunk sums in their buffers instead of folding them into the accumulation
                                        buffer: the root of a multi-GPU render gathers and folds them (rt_gather_image) */

typedef struct rt_render_params {
    uint32_t seed;          /* Philox key word 0 */
    int32_t spp_begin;      /* absolute index of the first sample rendered by this call */
    int32_t spp_count;      /* number of samples; <= 0 means the scene's samples_per_pixel */
    int32_t flags;          /* RT_RENDER_* */
    int64_t max_paths_per_pass;   /* wavefront size cap; <= 0 picks a default from free device memory */
    int32_t total_spp;      /* sample count of the whole image this call contributes to (sets the chunk boundaries of the
                               fixed-order accumulation, rt_sample_chunk); <= 0: max(scene samples_per_pixel, spp_begin + spp_count) */
    int32_t reserved;
} rt_render_params;

/* Fixed-order accumulation.  The samples [0, total_spp) of an image fall into RT_SAMPLE_CHUNKS chunks,
 * chunk k = [k * total_spp / 8, (k + 1) * total_spp / 8).  A chunk's samples are summed one by one in ascending order into the
 * chunk's own buffer; the image is (((c0 + c1) + c2) + ...) in ascending chunk order.  Because the boundaries depend on total_spp
 * only, the float sums are bit-identical whether one GPU renders every chunk or several GPUs render disjoint sets of whole
 * chunks (rt_comm_*, rt_group_*), and whether the chunks arrive in one rt_render call or in several ascending calls with
 * RT_RENDER_KEEP_ACCUM.  (A call whose sample range cuts a chunk in two still renders every sample exactly once; only the
 * rounding of that chunk's sum then differs from the one-call image.) */
#define RT_SAMPLE_CHUNKS 8
void rt_sample_chunk(int total_spp, int k, int *begin, int *end);

typedef struct rt_stats {
    int64_t samples;            /* camera paths started */
    int64_t rays;               /* path segments traced (extend-queue entries) */
    int64_t shadow_rays;
    int64_t launches;           /* kernels launched by the library since rt_create/rt_stats_reset */
    int64_t passes;
    float ms_total;             /* device time of the last rt_render, CUDA events on the ctx stream */
    float ms_generate, ms_extend, ms_shade, ms_shadow, ms_accumulate, ms_build;   /* per-stage device time (COUNT_EVENTS or TIME_STAGES) */
    /* event counters for the algorithmic-flop model of SURVEY.md §8(d) (only with COUNT_EVENTS) */
    int64_t n_box_tests, n_sphere_reject, n_sphere_accept, n_rect_tests, n_cyl_reject, n_cyl_accept,
            n_tri_tests, n_boxprim_tests, n_shade, n_generate, n_node_visits;
    int64_t n_generate_leaf_tests;  /* leaf tests the generate stage made on the tile candidate lists (COUNT_EVENTS; not in the n_*_tests above,
                                       which are the extend stage's) */
    int64_t paths_skipped;          /* camera paths of empty tiles that never entered the queues (their records move no bytes) */
    float ms_tail;                  /* tail stage (one launch finishes the paths alive at the tail bounce), TIME_STAGES */
    float reserved;
    int64_t launches_extend;        /* extend-stage launches among `launches` (mean rays per extend launch for the roofline's traffic figure) */
    int64_t rays_extend;            /* rays that went through an extend launch (bounce 0 from the tile lists and the tail stage's rays are not among them) */
} rt_stats;

typedef struct rt_scene_info {
    int32_t width, height, max_depth, samples_per_pixel;
    int32_t n_objects, n_materials, n_textures, n_bvh_nodes;
    char output_file[512];      /* JSON "output_file", default "main.png" (parser.hpp:567-568) */
} rt_scene_info;

typedef struct rt_ctx rt_ctx;   /* opaque: owns the device buffers, the BVH, the queues and (unless given one) a stream */

/* device: CUDA ordinal.  stream: a cudaStream_t to enqueue on (e.g. the caller's torch stream), or NULL for a private one. */
int rt_create(rt_ctx **out, int device, void *stream);
void rt_destroy(rt_ctx *ctx);
/* message for the last failing call on ctx (ctx == NULL: last failing rt_create on this thread) */
const char *rt_last_error(const rt_ctx *ctx);
int rt_abi_version(void);

/* Copies the tables to the device, flattens them to SoA and builds the LBVH there. */
int rt_scene_upload(rt_ctx *ctx, const rt_scene_desc *scene);
/* Same JSON schema as gpu-version/parser.hpp (+ optional extension keys). */
int rt_scene_upload_json(rt_ctx *ctx, const char *path);
/* As above from an in-memory JSON text; base_dir resolves relative texture/mesh paths (may be NULL). */
int rt_scene_upload_json_text(rt_ctx *ctx, const char *json_text, const char *base_dir);
/* Animation without scene text (replaces "write one JSON per frame", gpu-version/blue.py:16-22, blue2.py:16-22): replaces objects
 * [first_object, first_object + n_objects) and transforms [first_xform, first_xform + n_xforms) of the uploaded scene (either run may
 * be empty) and rebuilds the device tables and the LBVH (a full device rebuild: it is cheaper than a refit pass would save).
 * Object types / counts may change within the run; the tables keep their sizes.  On an invalid update the scene stays as it was. */
int rt_scene_update(rt_ctx *ctx, const rt_object_desc *objects, int first_object, int n_objects, const rt_xform_desc *xforms,
                    int first_xform, int n_xforms);
/* translate(t) * rotate(axis, angle_deg) and its inverse exactly as the JSON reader builds them for "rotate" / "translate"
 * (vec3.cuh:382-416, parser.hpp:423-440): matrices made here are bit-identical to those of rt_scene_upload_json. */
void rt_make_xform(const float axis[3], double angle_deg, int has_rotate, const float translate[3], int has_translate, rt_xform_desc *out);
int rt_scene_info_get(rt_ctx *ctx, rt_scene_info *out);
/* Overrides applied to the uploaded scene before rendering (<= 0 leaves a field unchanged). */
int rt_scene_override(rt_ctx *ctx, int width, int height, int samples_per_pixel, int max_depth);

/* Asynchronous: enqueues the wavefront passes for samples [spp_begin, spp_begin+spp_count) on the ctx stream. */
int rt_render(rt_ctx *ctx, const rt_render_params *params);
/* Deterministic first-hit pass (1 sample, pixel-centre rays, no lens offset, no scattering): 24 floats per pixel,
 * layout documented in DESIGN.md §6 — the "deterministic path" of the parity contract.  Synchronises. */
int rt_render_aov(rt_ctx *ctx, float *aov_out, size_t n_floats);

/* Device address of the accumulation buffer (width*height*3 float sums) so that a multi-GPU caller can
 * reduce it in place (NCCL over NVLink) before rt_download on the root rank. */
int rt_accum_device_ptr(rt_ctx *ctx, void **dev_ptr, size_t *n_floats);
/* Fixed-order sum of n_partials accumulation buffers (device pointers, e.g. gathered from the ranks) into this
 * ctx's accumulation buffer: acc = ((p0 + p1) + p2) + ...  — bit-reproducible for any GPU count. */
int rt_accum_sum_ordered(rt_ctx *ctx, const void *const *partials, int n_partials);

/* ---- several GPUs, one image (replaces the process-per-GPU scheme of gpu-version/blue.py:23-32) ----
 * The sample chunks of the image (rt_sample_chunk) are dealt to the ranks, rank r renders chunks rt_rank_chunks(n, r) into per-chunk
 * buffers, and the root gathers them over NVLink (NCCL send/receive on the render streams; NCCL is bound at run time with dlopen,
 * RT_NCCL_LIB overrides the library name) and adds them in chunk order: the float image is bit-identical for any number of GPUs.
 *
 * (a) one process per GPU (torchrun / mpirun): every rank creates its own context; rank 0 calls rt_comm_unique_id and hands the 128
 *     bytes to the others by any means (MPI, torch.distributed, a file); every rank calls rt_comm_init, then per image
 *     rt_render_shard + rt_gather_image (both asynchronous on the context's stream), and the root rt_download. */
void rt_rank_chunks(int n_ranks, int rank, int *chunk_begin, int *chunk_end);
int rt_comm_unique_id(void *id, size_t bytes /* >= 128 */);
int rt_comm_init(rt_ctx *ctx, const void *id, int n_ranks, int rank);
int rt_comm_destroy(rt_ctx *ctx);           /* also done by rt_destroy */
/* As rt_render for samples [0, spp_count) of which this rank renders its chunks (spp_begin must be 0; KEEP_ACCUM is ignored). */
int rt_render_shard(rt_ctx *ctx, const rt_render_params *params);
/* Moves every rank's chunk sums to `root` and folds them there into the accumulation buffer.  Collective: every rank calls it. */
int rt_gather_image(rt_ctx *ctx, int root);
/* (b) one process driving several devices: n contexts + ncclCommInitAll behind one handle; the image ends up on devices[0]. */
typedef struct rt_group rt_group;
int rt_group_create(rt_group **out, const int *devices, int n_devices);
void rt_group_destroy(rt_group *g);
const char *rt_group_last_error(const rt_group *g);    /* creation failures: rt_last_error(NULL) */
int rt_group_size(const rt_group *g);
rt_ctx *rt_group_ctx(rt_group *g, int i);              /* the i-th device's context (introspection; owned by the group) */
int rt_group_scene_upload(rt_group *g, const rt_scene_desc *scene);
int rt_group_scene_upload_json(rt_group *g, const char *path);
int rt_group_scene_upload_json_text(rt_group *g, const char *json_text, const char *base_dir);
int rt_group_scene_override(rt_group *g, int width, int height, int samples_per_pixel, int max_depth);
int rt_group_render(rt_group *g, const rt_render_params *params);     /* asynchronous */
int rt_group_download(rt_group *g, float *rgb_sum, size_t n_floats, unsigned char *rgb8, int total_spp, int gamma, rt_stats *stats);

/* Synchronises.  rgb_sum (may be NULL): width*height*3 float sums, row 0 = bottom.  rgb8 (may be NULL):
 * width*height*3 bytes, tonemapped on the device by total_spp, rows flipped (row 0 = top), gamma != 0 applies the
 * PPM writer's sqrt (color.cuh:58-62), gamma == 0 is the PNG writer (color.cuh:15-35).  stats may be NULL. */
int rt_download(rt_ctx *ctx, float *rgb_sum, size_t n_floats, unsigned char *rgb8, int total_spp, int gamma, rt_stats *stats);
int rt_synchronize(rt_ctx *ctx);
int rt_stats_reset(rt_ctx *ctx);

/* rt_download's tonemap + read-back for callers that must not allocate device memory (code under CUDA graph capture, tight serving
 * loops, harnesses that forbid cudaMalloc).  NO COUNTERPART IN THE REFERENCE.
 * rt_download stages the RGB8 image in a device buffer it cudaMallocs on first use; this entry point stages it in a file-scope
 * static __device__ array of RT_NOALLOC_MAX_BYTES (8 MiB: images up to 1920 x 1080 x 3 and a little beyond) that every GPU loading
 * the library holds until the process exits.  It calls no cudaMalloc / cudaFree / allocating library routine.
 *   accum_dev: device pointer to width*height*3 float sums, row 0 = bottom (e.g. from rt_accum_device_ptr) on `device`;
 *   rgb8_host: width*height*3 bytes, row 0 = top; same arithmetic as rt_download (gamma as there); stream: cudaStream_t or NULL.
 * Enqueues on `stream` and synchronises it.  width*height*3 > RT_NOALLOC_MAX_BYTES is rejected with RT_ERR_INVALID
 * (message from rt_last_error(NULL)).  Every call on a device shares the one array, which keeps the previous call's bytes
 * (nothing needs clearing: the bytes read back are exactly the bytes just written) — calls that overlap on the same GPU race. */
#define RT_NOALLOC_MAX_BYTES (8u << 20)
int rt_tonemap_noalloc(int device, void *stream, const float *accum_dev, int width, int height, int total_spp, int gamma,
                       unsigned char *rgb8_host);

/* Page-locked host buffers for rt_download targets (optional: any host pointer works, a pinned one is copied at the full
 * host-link rate and without a staging copy).  Takes the place of the reference's managed image allocation, which the
 * host reads after cudaDeviceSynchronize (main.cu:487-489, 509-514). */
int rt_host_alloc(rt_ctx *ctx, void **ptr, size_t bytes);
int rt_host_free(rt_ctx *ctx, void *ptr);

/* LBVH introspection for the bit-exact build tests.  Arrays are sized for n_prims leaves (n_prims-1 internal nodes); the tree
 * itself has n_tree <= n_prims leaves (rt_bvh_info): objects that span more than a quarter of the scene ("giants") are kept out
 * of it, sorted behind the others, and tested by every ray before the walk — left/right/node_aabb hold n_tree-1 entries, parent
 * 2*n_tree-1.  Any output pointer may be NULL.  child encoding: >= 0 internal node index, < 0 leaf ~index into the sorted order. */
int rt_bvh_download(rt_ctx *ctx, int32_t *n_prims, uint64_t *morton_sorted, int32_t *prim_sorted, int32_t *left,
                    int32_t *right, int32_t *parent, float *node_aabb /* 6 per internal node */,
                    float *leaf_aabb /* 6 per sorted leaf */, float *scene_aabb /* 6 */);

/* n_tree leaves in the tree, n_giants objects tested up front (sorted slots n_tree ..), depth = internal nodes on the longest
 * root-to-leaf path (an upload whose tree is deeper than the traversal stack fails with RT_ERR_INVALID). */
int rt_bvh_info(rt_ctx *ctx, int32_t *n_tree, int32_t *n_giants, int32_t *depth);

/* Host image writers (row 0 of rgb8 = top).  PPM is the reference's P3 text format (color.cuh:70-95). */
int rt_write_ppm(const char *path, const unsigned char *rgb8, int width, int height);
int rt_write_png(const char *path, const unsigned char *rgb8, int width, int height);
/* PNG reader for image textures (8-bit RGB/RGBA/grey, non-interlaced); caller frees *rgb8 with rt_free. */
int rt_read_png(const char *path, unsigned char **rgb8, int *width, int *height);
void rt_free(void *p);

/* Measurement helper: FP32 FMA-chain microbenchmark on the ctx device (TFLOP/s, best of 4 after warm-up) — the measured
 * denominator of the extend/shade roofline. */
int rt_probe_fp32_peak(rt_ctx *ctx, double *tflops);

#ifdef __cplusplus
}
#endif
#endif /* RT_B200_H */
