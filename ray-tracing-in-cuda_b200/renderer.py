"""Host-side mirror of the reference's render-path interface, on top of the C ABI (include/rt.h).

Reference seams (gpu-version/): `parse_scene(filename)` parser.hpp:504, the `render<<<>>>` launch
main.cu:505, `output_image` / `write_image` main.cu:359-372 + color.cuh:15-35, and the whole of
`jsonmain` main.cu:448-522 (= `render_scene_file`).  All pixels come from librt_b200.so; there is no
CPU fallback here — a missing library or device raises RtError.
"""
import ctypes as C
import os

import numpy as np

from . import capi


class RtError(RuntimeError):
    pass


class Renderer:
    """One rt_ctx = one GPU.  `stream` may be a raw cudaStream_t (e.g. torch.cuda.current_stream().cuda_stream)."""

    def __init__(self, device=0, stream=None, lib_path=None):
        self.lib = capi.load_library(lib_path)
        self.ctx = C.c_void_p()
        rc = self.lib.rt_create(C.byref(self.ctx), int(device), C.c_void_p(stream) if stream else None)
        if rc != 0:
            msg = self.lib.rt_last_error(None)
            raise RtError(f"rt_create failed ({rc}): {msg.decode() if msg else ''}")
        self.device = device

    def _check(self, rc, what):
        if rc != 0:
            msg = self.lib.rt_last_error(self.ctx)
            raise RtError(f"{what} failed ({rc}): {msg.decode() if msg else ''}")

    def close(self):
        if getattr(self, "ctx", None):
            for ptr in getattr(self, "_pinned", []):
                self.lib.rt_host_free(self.ctx, ptr)
            self._pinned = []
            self.lib.rt_destroy(self.ctx)
            self.ctx = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- scene in (parse_scene + move_to_device) ----
    def upload_json(self, path):
        self._check(self.lib.rt_scene_upload_json(self.ctx, os.fsencode(path)), "rt_scene_upload_json")
        return self.info()

    def upload_json_text(self, text, base_dir=None):
        self._check(self.lib.rt_scene_upload_json_text(self.ctx, text.encode(), os.fsencode(base_dir) if base_dir else None),
                    "rt_scene_upload_json_text")
        return self.info()

    def upload(self, scene_desc):
        """scene_desc: capi.SceneDesc (or an object with a `.desc` attribute holding one)."""
        d = getattr(scene_desc, "desc", scene_desc)
        self._check(self.lib.rt_scene_upload(self.ctx, C.byref(d)), "rt_scene_upload")
        return self.info()

    def scene_update(self, objects=None, first_object=0, xforms=None, first_xform=0):
        """rt_scene_update: replace a run of objects (capi.ObjectDesc) and / or transforms (capi.XformDesc), rebuild on the device."""
        objects, xforms = list(objects or []), list(xforms or [])
        OA = (capi.ObjectDesc * max(1, len(objects)))(*objects)
        XA = (capi.XformDesc * max(1, len(xforms)))(*xforms)
        self._check(self.lib.rt_scene_update(self.ctx, OA, int(first_object), len(objects), XA, int(first_xform), len(xforms)), "rt_scene_update")

    def make_xform(self, rotate=None, translate=None):
        """translate * rotate exactly as the JSON reader builds it (rt_make_xform)."""
        x = capi.XformDesc()
        ax = (C.c_float * 3)(*(rotate["axis"] if rotate else (0, 0, 1)))
        tr = (C.c_float * 3)(*(translate if translate is not None else (0, 0, 0)))
        self.lib.rt_make_xform(ax, float(rotate["angle"]) if rotate else 0.0, int(rotate is not None), tr, int(translate is not None), C.byref(x))
        return x

    def override(self, width=0, height=0, spp=0, max_depth=0):
        self._check(self.lib.rt_scene_override(self.ctx, width or 0, height or 0, spp or 0, max_depth or 0), "rt_scene_override")
        return self.info()

    def info(self):
        i = capi.SceneInfo()
        self._check(self.lib.rt_scene_info_get(self.ctx, C.byref(i)), "rt_scene_info_get")
        return i

    # ---- render ----
    def render(self, seed=0x2023, spp_begin=0, spp_count=0, flags=0, max_paths_per_pass=0, total_spp=0):
        p = capi.RenderParams(seed, spp_begin, spp_count, flags, max_paths_per_pass, total_spp, 0)
        self._check(self.lib.rt_render(self.ctx, C.byref(p)), "rt_render")

    # ---- several GPUs, one image: one context per rank (rt_comm_*, rt_render_shard, rt_gather_image) ----
    @staticmethod
    def comm_unique_id(lib_path=None):
        """128 bytes from rank 0, to be handed to every rank (torch.distributed / MPI / a file) for comm_init."""
        lib = capi.load_library(lib_path)
        buf = C.create_string_buffer(128)
        rc = lib.rt_comm_unique_id(buf, 128)
        if rc != 0:
            msg = lib.rt_last_error(None)
            raise RtError(f"rt_comm_unique_id failed ({rc}): {msg.decode() if msg else ''}")
        return buf.raw

    def comm_init(self, unique_id, n_ranks, rank):
        self._check(self.lib.rt_comm_init(self.ctx, C.c_char_p(bytes(unique_id)), int(n_ranks), int(rank)), "rt_comm_init")

    def render_shard(self, seed=0x2023, spp_count=0, flags=0, max_paths_per_pass=0):
        """This rank's whole sample chunks of an image of spp_count (default: the scene's) samples; asynchronous."""
        p = capi.RenderParams(seed, 0, spp_count, flags, max_paths_per_pass, 0, 0)
        self._check(self.lib.rt_render_shard(self.ctx, C.byref(p)), "rt_render_shard")

    def gather_image(self, root=0):
        """Collective: every rank's chunk sums travel to `root`, which adds them in chunk order (bit-identical for any rank count)."""
        self._check(self.lib.rt_gather_image(self.ctx, int(root)), "rt_gather_image")

    def render_aov(self):
        i = self.info()
        out = np.zeros((i.height, i.width, 24), np.float32)
        self._check(self.lib.rt_render_aov(self.ctx, out.ctypes.data_as(C.POINTER(C.c_float)), out.size), "rt_render_aov")
        return out

    def synchronize(self):
        self._check(self.lib.rt_synchronize(self.ctx), "rt_synchronize")

    def accum_device_ptr(self):
        p = C.c_void_p(); n = C.c_size_t()
        self._check(self.lib.rt_accum_device_ptr(self.ctx, C.byref(p), C.byref(n)), "rt_accum_device_ptr")
        return p.value, n.value

    def accum_sum_ordered(self, device_ptrs):
        arr = (C.c_void_p * len(device_ptrs))(*device_ptrs)
        self._check(self.lib.rt_accum_sum_ordered(self.ctx, arr, len(device_ptrs)), "rt_accum_sum_ordered")

    # ---- image out ----
    def pinned_empty(self, shape, dtype):
        """numpy array over page-locked host memory (rt_host_alloc); freed when the renderer is closed."""
        n = int(np.prod(shape)) * np.dtype(dtype).itemsize
        p = C.c_void_p()
        self._check(self.lib.rt_host_alloc(self.ctx, C.byref(p), n), "rt_host_alloc")
        if not hasattr(self, "_pinned"):
            self._pinned = []
        self._pinned.append(p)
        return np.ctypeslib.as_array((C.c_ubyte * max(n, 1)).from_address(p.value))[:n].view(dtype).reshape(shape)

    def download(self, total_spp=0, gamma=0, want_sum=True, want_rgb8=True, out_sum=None, out_rgb8=None):
        """Returns (rgb_sum[H,W,3] float32 with row 0 = bottom, rgb8[H,W,3] uint8 with row 0 = top, stats dict).
        out_sum / out_rgb8: caller-owned target arrays of those shapes (e.g. from pinned_empty), reused across calls."""
        i = self.info()
        if want_sum:
            rgb_sum = out_sum if out_sum is not None else np.empty((i.height, i.width, 3), np.float32)
            assert rgb_sum.shape == (i.height, i.width, 3) and rgb_sum.dtype == np.float32 and rgb_sum.flags.c_contiguous
        else:
            rgb_sum = None
        if want_rgb8:
            rgb8 = out_rgb8 if out_rgb8 is not None else np.empty((i.height, i.width, 3), np.uint8)
            assert rgb8.shape == (i.height, i.width, 3) and rgb8.dtype == np.uint8 and rgb8.flags.c_contiguous
        else:
            rgb8 = None
        st = capi.Stats()
        self._check(self.lib.rt_download(
            self.ctx, rgb_sum.ctypes.data_as(C.POINTER(C.c_float)) if want_sum else None, rgb_sum.size if want_sum else 0,
            rgb8.ctypes.data_as(C.POINTER(C.c_ubyte)) if want_rgb8 else None, int(total_spp or i.samples_per_pixel),
            int(gamma), C.byref(st)), "rt_download")
        return rgb_sum, rgb8, st.as_dict()

    def probe_fp32_peak(self):
        v = C.c_double()
        self._check(self.lib.rt_probe_fp32_peak(self.ctx, C.byref(v)), "rt_probe_fp32_peak")
        return v.value

    def stats_reset(self):
        self._check(self.lib.rt_stats_reset(self.ctx), "rt_stats_reset")

    def bvh(self):
        n = C.c_int32()
        self._check(self.lib.rt_bvh_download(self.ctx, C.byref(n), None, None, None, None, None, None, None, None), "rt_bvh_download")
        n = n.value
        m = max(n - 1, 1)
        r = dict(morton=np.zeros(max(n, 1), np.uint64), prim=np.zeros(max(n, 1), np.int32), left=np.zeros(m, np.int32),
                 right=np.zeros(m, np.int32), parent=np.zeros(max(2 * n - 1, 1), np.int32),
                 node_aabb=np.zeros((m, 6), np.float32), leaf_aabb=np.zeros((max(n, 1), 6), np.float32),
                 scene_aabb=np.zeros(6, np.float32))
        P = C.POINTER
        nn = C.c_int32()
        self._check(self.lib.rt_bvh_download(
            self.ctx, C.byref(nn), r["morton"].ctypes.data_as(P(C.c_uint64)), r["prim"].ctypes.data_as(P(C.c_int32)),
            r["left"].ctypes.data_as(P(C.c_int32)), r["right"].ctypes.data_as(P(C.c_int32)),
            r["parent"].ctypes.data_as(P(C.c_int32)), r["node_aabb"].ctypes.data_as(P(C.c_float)),
            r["leaf_aabb"].ctypes.data_as(P(C.c_float)), r["scene_aabb"].ctypes.data_as(P(C.c_float))), "rt_bvh_download")
        nt, ng, depth = C.c_int32(), C.c_int32(), C.c_int32()
        self._check(self.lib.rt_bvh_info(self.ctx, C.byref(nt), C.byref(ng), C.byref(depth)), "rt_bvh_info")
        # the tree holds n_tree leaves (n_tree - 1 internal nodes); the giants occupy the last n_giants sorted slots
        m = max(nt.value - 1, 0)
        r.update(n=n, n_tree=nt.value, n_giants=ng.value, depth=depth.value, morton=r["morton"][:n], prim=r["prim"][:n],
                 leaf_aabb=r["leaf_aabb"][:n], left=r["left"][:m], right=r["right"][:m], node_aabb=r["node_aabb"][:m],
                 parent=r["parent"][:max(2 * nt.value - 1, 0)])
        return r


class Group:
    """One process driving several GPUs (rt_group_*): scene replicated, the sample chunks of an image dealt to the devices,
    chunk sums gathered on devices[0] over NVLink and added in chunk order."""

    def __init__(self, devices, lib_path=None):
        self.lib = capi.load_library(lib_path)
        self.g = C.c_void_p()
        arr = (C.c_int * len(devices))(*devices)
        rc = self.lib.rt_group_create(C.byref(self.g), arr, len(devices))
        if rc != 0:
            msg = self.lib.rt_last_error(None)
            raise RtError(f"rt_group_create failed ({rc}): {msg.decode() if msg else ''}")
        self.devices = list(devices)

    def _check(self, rc, what):
        if rc != 0:
            msg = self.lib.rt_group_last_error(self.g)
            raise RtError(f"{what} failed ({rc}): {msg.decode() if msg else ''}")

    def close(self):
        if getattr(self, "g", None):
            self.lib.rt_group_destroy(self.g)
            self.g = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def upload_json(self, path):
        self._check(self.lib.rt_group_scene_upload_json(self.g, os.fsencode(path)), "rt_group_scene_upload_json")

    def upload_json_text(self, text, base_dir=None):
        self._check(self.lib.rt_group_scene_upload_json_text(self.g, text.encode(), os.fsencode(base_dir) if base_dir else None),
                    "rt_group_scene_upload_json_text")

    def upload(self, scene_desc):
        d = getattr(scene_desc, "desc", scene_desc)
        self._check(self.lib.rt_group_scene_upload(self.g, C.byref(d)), "rt_group_scene_upload")

    def scene_update(self, objects=None, first_object=0, xforms=None, first_xform=0):
        """rt_scene_update: replace a run of objects (capi.ObjectDesc) and / or transforms (capi.XformDesc), rebuild on the device."""
        objects, xforms = list(objects or []), list(xforms or [])
        OA = (capi.ObjectDesc * max(1, len(objects)))(*objects)
        XA = (capi.XformDesc * max(1, len(xforms)))(*xforms)
        self._check(self.lib.rt_scene_update(self.ctx, OA, int(first_object), len(objects), XA, int(first_xform), len(xforms)), "rt_scene_update")

    def make_xform(self, rotate=None, translate=None):
        """translate * rotate exactly as the JSON reader builds it (rt_make_xform)."""
        x = capi.XformDesc()
        ax = (C.c_float * 3)(*(rotate["axis"] if rotate else (0, 0, 1)))
        tr = (C.c_float * 3)(*(translate if translate is not None else (0, 0, 0)))
        self.lib.rt_make_xform(ax, float(rotate["angle"]) if rotate else 0.0, int(rotate is not None), tr, int(translate is not None), C.byref(x))
        return x

    def override(self, width=0, height=0, spp=0, max_depth=0):
        self._check(self.lib.rt_group_scene_override(self.g, width or 0, height or 0, spp or 0, max_depth or 0), "rt_group_scene_override")

    def info(self):
        i = capi.SceneInfo()
        if self.lib.rt_scene_info_get(self.lib.rt_group_ctx(self.g, 0), C.byref(i)) != 0:
            raise RtError("rt_scene_info_get failed")
        return i

    def render(self, seed=0x2023, spp_count=0, flags=0, max_paths_per_pass=0):
        p = capi.RenderParams(seed, 0, spp_count, flags, max_paths_per_pass, 0, 0)
        self._check(self.lib.rt_group_render(self.g, C.byref(p)), "rt_group_render")

    def download(self, total_spp=0, gamma=0, want_sum=True, want_rgb8=True):
        i = self.info()
        rgb_sum = np.zeros((i.height, i.width, 3), np.float32) if want_sum else None
        rgb8 = np.zeros((i.height, i.width, 3), np.uint8) if want_rgb8 else None
        st = capi.Stats()
        self._check(self.lib.rt_group_download(self.g, rgb_sum.ctypes.data_as(C.POINTER(C.c_float)) if want_sum else None,
                                               rgb_sum.size if want_sum else 0,
                                               rgb8.ctypes.data_as(C.POINTER(C.c_ubyte)) if want_rgb8 else None, total_spp, gamma, C.byref(st)),
                    "rt_group_download")
        return rgb_sum, rgb8, st.as_dict()


def write_ppm(path, rgb8):
    lib = capi.load_library()
    a = np.ascontiguousarray(rgb8, np.uint8)
    if lib.rt_write_ppm(os.fsencode(path), a.ctypes.data_as(C.POINTER(C.c_ubyte)), a.shape[1], a.shape[0]) != 0:
        raise RtError(f"cannot write {path}")


def write_png(path, rgb8):
    lib = capi.load_library()
    a = np.ascontiguousarray(rgb8, np.uint8)
    if lib.rt_write_png(os.fsencode(path), a.ctypes.data_as(C.POINTER(C.c_ubyte)), a.shape[1], a.shape[0]) != 0:
        raise RtError(f"cannot write {path}")


def read_png(path):
    lib = capi.load_library()
    p = C.POINTER(C.c_ubyte)(); w = C.c_int(); h = C.c_int()
    if lib.rt_read_png(os.fsencode(path), C.byref(p), C.byref(w), C.byref(h)) != 0:
        raise RtError(f"cannot read {path}")
    a = np.ctypeslib.as_array(p, shape=(h.value, w.value, 3)).copy()
    lib.rt_free(p)
    return a


def parse_scene(filename, device=0, stream=None):
    """parse_scene(filename) of gpu-version/parser.hpp:504 — returns a Renderer holding the uploaded scene."""
    r = Renderer(device, stream)
    r.upload_json(filename)
    return r


def render_scene_file(filename, device=0, seed=0x2023, write_outputs=True, out_dir=None):
    """jsonmain() of gpu-version/main.cu:448-522: parse, render, write main.ppm (gamma) + output_file PNG (no gamma)."""
    r = parse_scene(filename, device)
    info = r.info()
    r.render(seed=seed)
    rgb_sum, png8, stats = r.download(gamma=0)
    if write_outputs:
        _, ppm8, _ = r.download(gamma=1, want_sum=False)
        base = out_dir or os.getcwd()
        write_ppm(os.path.join(base, "main.ppm"), ppm8)
        out = info.output_file.decode() or "main.png"
        write_png(out if os.path.isabs(out) else os.path.join(base, out), png8)
    r.close()
    return rgb_sum, png8, stats
