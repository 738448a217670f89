"""ctypes mirror of include/rt.h (the C ABI of librt_b200.so).

Structure layouts follow include/rt.h field for field; the reference seam each one replaces is cited
there (gpu-version/parser.hpp:16-32 `struct scene`, main.cu:72-74 `render`).  This module only declares
types and loads the shared library; it contains no rendering logic and no CPU fallback: if the CUDA
library is missing, `load_library()` raises.
"""
import ctypes as C
import os

RT_OBJ_SPHERE, RT_OBJ_XY_RECT, RT_OBJ_XZ_RECT, RT_OBJ_YZ_RECT, RT_OBJ_CYLINDER, RT_OBJ_BOX, RT_OBJ_TRIANGLE = range(7)
RT_MAT_LAMBERTIAN, RT_MAT_METAL, RT_MAT_DIELECTRIC, RT_MAT_DIFFUSE_LIGHT = range(4)
RT_TEX_SOLID, RT_TEX_CHECKER, RT_TEX_IMAGE = range(3)
RT_BG_CONSTANT, RT_BG_SKY_GRADIENT = range(2)
RT_OK, RT_ERR_INVALID, RT_ERR_CUDA, RT_ERR_IO, RT_ERR_PARSE, RT_ERR_NO_SCENE, RT_ERR_NO_DEVICE = range(7)
RT_SAMPLING_REFERENCE_GPU, RT_SAMPLING_REFERENCE_CPU = range(2)
RT_SAMPLING_DIRECT = 2       # flag: direct draws instead of the reference's rejection loops
RT_RENDER_KEEP_ACCUM, RT_RENDER_COUNT_EVENTS, RT_RENDER_NO_SORT, RT_RENDER_NEE, RT_RENDER_TIME_STAGES, RT_RENDER_SORT = 1, 2, 4, 8, 16, 32
RT_RENDER_SERIAL = 64
RT_RENDER_NO_TILE_LISTS = 256
RT_RENDER_KEEP_CHUNKS = 512
RT_RENDER_NO_TAIL = 1024
RT_RENDER_NEE_MIS = 2048
RT_SAMPLE_CHUNKS = 8
RT_NOALLOC_MAX_BYTES = 8 << 20      # static scratch of rt_tonemap_noalloc


class ObjectDesc(C.Structure):
    _fields_ = [("type", C.c_int32), ("material", C.c_int32), ("xform", C.c_int32), ("reserved", C.c_int32),
                ("p", C.c_float * 12)]


class XformDesc(C.Structure):
    _fields_ = [("m", C.c_float * 12), ("m_inv", C.c_float * 12)]


class MaterialDesc(C.Structure):
    _fields_ = [("type", C.c_int32), ("texture", C.c_int32), ("albedo", C.c_float * 3), ("fuzz", C.c_float),
                ("ior", C.c_float), ("reserved", C.c_int32)]


class TextureDesc(C.Structure):
    _fields_ = [("type", C.c_int32), ("even", C.c_int32), ("odd", C.c_int32), ("image", C.c_int32),
                ("color", C.c_float * 3), ("reserved", C.c_float)]


class ImageDesc(C.Structure):
    _fields_ = [("width", C.c_int32), ("height", C.c_int32), ("rgb8", C.POINTER(C.c_uint8))]


class CameraDesc(C.Structure):
    _fields_ = [("lookfrom", C.c_float * 3), ("lookat", C.c_float * 3), ("vup", C.c_float * 3),
                ("vfov_deg", C.c_float), ("aspect", C.c_float), ("aperture", C.c_float), ("focus_dist", C.c_float)]


class SceneDesc(C.Structure):
    _fields_ = [("width", C.c_int32), ("height", C.c_int32), ("max_depth", C.c_int32), ("samples_per_pixel", C.c_int32),
                ("background", C.c_float * 3), ("background_mode", C.c_int32), ("sampling", C.c_int32),
                ("camera", CameraDesc),
                ("n_objects", C.c_int32), ("n_xforms", C.c_int32), ("n_materials", C.c_int32),
                ("n_textures", C.c_int32), ("n_images", C.c_int32), ("reserved", C.c_int32),
                ("objects", C.POINTER(ObjectDesc)), ("xforms", C.POINTER(XformDesc)),
                ("materials", C.POINTER(MaterialDesc)), ("textures", C.POINTER(TextureDesc)),
                ("images", C.POINTER(ImageDesc))]


class RenderParams(C.Structure):
    _fields_ = [("seed", C.c_uint32), ("spp_begin", C.c_int32), ("spp_count", C.c_int32), ("flags", C.c_int32),
                ("max_paths_per_pass", C.c_int64), ("total_spp", C.c_int32), ("reserved", C.c_int32)]


class Stats(C.Structure):
    _fields_ = [("samples", C.c_int64), ("rays", C.c_int64), ("shadow_rays", C.c_int64), ("launches", C.c_int64),
                ("passes", C.c_int64), ("ms_total", C.c_float), ("ms_generate", C.c_float), ("ms_extend", C.c_float),
                ("ms_shade", C.c_float), ("ms_shadow", C.c_float), ("ms_accumulate", C.c_float), ("ms_build", C.c_float),
                ("n_box_tests", C.c_int64), ("n_sphere_reject", C.c_int64), ("n_sphere_accept", C.c_int64),
                ("n_rect_tests", C.c_int64), ("n_cyl_reject", C.c_int64), ("n_cyl_accept", C.c_int64),
                ("n_tri_tests", C.c_int64), ("n_boxprim_tests", C.c_int64), ("n_shade", C.c_int64),
                ("n_generate", C.c_int64), ("n_node_visits", C.c_int64),
                ("n_generate_leaf_tests", C.c_int64), ("paths_skipped", C.c_int64), ("ms_tail", C.c_float), ("reserved", C.c_float),
                ("launches_extend", C.c_int64), ("rays_extend", C.c_int64)]

    def as_dict(self):
        return {k: getattr(self, k) for k, _ in self._fields_}


class SceneInfo(C.Structure):
    _fields_ = [("width", C.c_int32), ("height", C.c_int32), ("max_depth", C.c_int32), ("samples_per_pixel", C.c_int32),
                ("n_objects", C.c_int32), ("n_materials", C.c_int32), ("n_textures", C.c_int32),
                ("n_bvh_nodes", C.c_int32), ("output_file", C.c_char * 512)]


# every symbol include/rt.h declares (tests check that the library exports all of them)
RT_SYMBOLS = ["rt_create", "rt_destroy", "rt_last_error", "rt_abi_version", "rt_scene_upload", "rt_scene_upload_json",
              "rt_scene_upload_json_text", "rt_scene_info_get", "rt_scene_override", "rt_render", "rt_render_aov",
              "rt_accum_device_ptr", "rt_accum_sum_ordered", "rt_download", "rt_synchronize", "rt_stats_reset",
              "rt_bvh_download", "rt_bvh_info", "rt_sample_chunk", "rt_tonemap_noalloc", "rt_scene_update", "rt_make_xform", "rt_rank_chunks", "rt_comm_unique_id", "rt_comm_init", "rt_comm_destroy",
              "rt_render_shard", "rt_gather_image", "rt_group_create", "rt_group_destroy", "rt_group_last_error", "rt_group_size", "rt_group_ctx",
              "rt_group_scene_upload", "rt_group_scene_upload_json", "rt_group_scene_upload_json_text", "rt_group_scene_override",
              "rt_group_render", "rt_group_download", "rt_write_ppm", "rt_write_png", "rt_read_png", "rt_free", "rt_probe_fp32_peak", "rt_host_alloc",
              "rt_host_free"]

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("RT_B200_LIB") or os.path.join(_HERE, "librt_b200.so")   # env override: tuning builds only
_lib = None


def load_library(path=None):
    """Load librt_b200.so (built in-tree by __graft_entry__.build()).  Raises if it is missing: there is no fallback."""
    global _lib
    if _lib is not None and path is None:
        return _lib
    p = path or LIB_PATH
    if not os.path.exists(p):
        raise RuntimeError(f"{p} not found: build it with `python __graft_entry__.py build` "
                           "(the render path has no CPU fallback)")
    lib = C.CDLL(p)
    P = C.POINTER
    vp = C.c_void_p
    lib.rt_create.argtypes = [P(vp), C.c_int, vp]
    lib.rt_destroy.argtypes = [vp]; lib.rt_destroy.restype = None
    lib.rt_last_error.argtypes = [vp]; lib.rt_last_error.restype = C.c_char_p
    lib.rt_abi_version.argtypes = []
    lib.rt_scene_upload.argtypes = [vp, P(SceneDesc)]
    lib.rt_scene_upload_json.argtypes = [vp, C.c_char_p]
    lib.rt_scene_upload_json_text.argtypes = [vp, C.c_char_p, C.c_char_p]
    lib.rt_scene_info_get.argtypes = [vp, P(SceneInfo)]
    lib.rt_scene_override.argtypes = [vp, C.c_int, C.c_int, C.c_int, C.c_int]
    lib.rt_render.argtypes = [vp, P(RenderParams)]
    lib.rt_rank_chunks.argtypes = [C.c_int, C.c_int, P(C.c_int), P(C.c_int)]; lib.rt_rank_chunks.restype = None
    lib.rt_comm_unique_id.argtypes = [vp, C.c_size_t]
    lib.rt_comm_init.argtypes = [vp, vp, C.c_int, C.c_int]
    lib.rt_comm_destroy.argtypes = [vp]
    lib.rt_render_shard.argtypes = [vp, P(RenderParams)]
    lib.rt_gather_image.argtypes = [vp, C.c_int]
    lib.rt_group_create.argtypes = [P(vp), P(C.c_int), C.c_int]
    lib.rt_group_destroy.argtypes = [vp]; lib.rt_group_destroy.restype = None
    lib.rt_group_last_error.argtypes = [vp]; lib.rt_group_last_error.restype = C.c_char_p
    lib.rt_group_size.argtypes = [vp]
    lib.rt_group_ctx.argtypes = [vp, C.c_int]; lib.rt_group_ctx.restype = vp
    lib.rt_group_scene_upload.argtypes = [vp, P(SceneDesc)]
    lib.rt_group_scene_upload_json.argtypes = [vp, C.c_char_p]
    lib.rt_group_scene_upload_json_text.argtypes = [vp, C.c_char_p, C.c_char_p]
    lib.rt_group_scene_override.argtypes = [vp, C.c_int, C.c_int, C.c_int, C.c_int]
    lib.rt_group_render.argtypes = [vp, P(RenderParams)]
    lib.rt_group_download.argtypes = [vp, P(C.c_float), C.c_size_t, P(C.c_ubyte), C.c_int, C.c_int, P(Stats)]
    lib.rt_scene_update.argtypes = [vp, P(ObjectDesc), C.c_int, C.c_int, P(XformDesc), C.c_int, C.c_int]
    lib.rt_make_xform.argtypes = [P(C.c_float), C.c_double, C.c_int, P(C.c_float), C.c_int, P(XformDesc)]; lib.rt_make_xform.restype = None
    lib.rt_tonemap_noalloc.argtypes = [C.c_int, vp, vp, C.c_int, C.c_int, C.c_int, C.c_int, P(C.c_ubyte)]
    lib.rt_sample_chunk.argtypes = [C.c_int, C.c_int, P(C.c_int), P(C.c_int)]; lib.rt_sample_chunk.restype = None
    lib.rt_render_aov.argtypes = [vp, P(C.c_float), C.c_size_t]
    lib.rt_accum_device_ptr.argtypes = [vp, P(vp), P(C.c_size_t)]
    lib.rt_accum_sum_ordered.argtypes = [vp, P(vp), C.c_int]
    lib.rt_download.argtypes = [vp, P(C.c_float), C.c_size_t, P(C.c_ubyte), C.c_int, C.c_int, P(Stats)]
    lib.rt_synchronize.argtypes = [vp]
    lib.rt_stats_reset.argtypes = [vp]
    lib.rt_host_alloc.argtypes = [vp, P(C.c_void_p), C.c_size_t]
    lib.rt_host_free.argtypes = [vp, C.c_void_p]
    lib.rt_bvh_download.argtypes = [vp, P(C.c_int32), P(C.c_uint64), P(C.c_int32), P(C.c_int32), P(C.c_int32),
                                    P(C.c_int32), P(C.c_float), P(C.c_float), P(C.c_float)]
    lib.rt_bvh_info.argtypes = [vp, P(C.c_int32), P(C.c_int32), P(C.c_int32)]
    lib.rt_write_ppm.argtypes = [C.c_char_p, P(C.c_ubyte), C.c_int, C.c_int]
    lib.rt_write_png.argtypes = [C.c_char_p, P(C.c_ubyte), C.c_int, C.c_int]
    lib.rt_read_png.argtypes = [C.c_char_p, P(P(C.c_ubyte)), P(C.c_int), P(C.c_int)]
    lib.rt_free.argtypes = [vp]; lib.rt_free.restype = None
    lib.rt_probe_fp32_peak.argtypes = [vp, P(C.c_double)]
    for name in RT_SYMBOLS:
        if getattr(lib, name).restype is C.c_int:
            pass
    if path is None:
        _lib = lib
    return lib
