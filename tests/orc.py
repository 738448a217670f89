"""Test-side binding of the CPU oracle (oracle/liboracle.so) and an independent JSON -> flat-table loader.

TEST INFRASTRUCTURE.  The product has its own C++ JSON reader (rt_scene_upload_json); this Python
loader follows gpu-version/parser.hpp:504-573 independently so that tests can (a) feed the oracle and
(b) cross-check the C++ reader by uploading the same tables through rt_scene_upload.
"""
import ctypes as C
import json
import math
import os
import subprocess

import numpy as np

from rt_b200 import capi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")
GOLDEN = os.path.join(ROOT, "tests", "golden")
_lib = None


def lib():
    global _lib
    if _lib is None:
        path = os.path.join(ORACLE_DIR, "liboracle.so")
        if not os.path.exists(path):
            subprocess.check_call(["make", "-C", ORACLE_DIR, "liboracle.so"])
        L = C.CDLL(path)
        P = C.POINTER
        fp = P(C.c_float)
        L.orc_render.argtypes = [P(capi.SceneDesc), C.c_uint32, C.c_int, C.c_int, fp, C.c_int, P(C.c_int64)]
        L.orc_render_ex.argtypes = [P(capi.SceneDesc), C.c_uint32, C.c_int, C.c_int, fp, C.c_int, C.c_int, P(C.c_int64), P(C.c_int64)]
        L.orc_aov.argtypes = [P(capi.SceneDesc), fp]
        L.orc_tonemap.argtypes = [fp, C.c_int, C.c_int, C.c_int, C.c_int, P(C.c_ubyte)]
        L.orc_tonemap.restype = None
        L.orc_make_xform.argtypes = [fp, C.c_double, C.c_int, fp, C.c_int, P(capi.XformDesc)]
        L.orc_make_xform.restype = None
        L.orc_philox4x32_10.argtypes = [P(C.c_uint32), P(C.c_uint32), P(C.c_uint32)]
        L.orc_philox4x32_10.restype = None
        L.orc_uniforms.argtypes = [C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32, C.c_int, C.c_int, fp]
        L.orc_uniforms.restype = None
        L.orc_ball_samples.argtypes = [C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32, C.c_int, C.c_int, fp]
        L.orc_ball_samples.restype = None
        for k in ("sphere", "rect", "cyl", "camera", "xform"):
            f = getattr(L, "orc_kat_" + k); f.argtypes = [fp, fp, C.c_int]; f.restype = None
        L.orc_kat_scatter.argtypes = [fp, fp, C.c_int, C.c_int]; L.orc_kat_scatter.restype = None
        for k in ("orc_sinf", "orc_acosf"):
            f = getattr(L, k); f.argtypes = [C.c_float]; f.restype = C.c_float
        L.orc_atan2f.argtypes = [C.c_float, C.c_float]; L.orc_atan2f.restype = C.c_float
        L.orc_lbvh_build.argtypes = [P(capi.SceneDesc), P(C.c_uint64), P(C.c_int32), P(C.c_int32), P(C.c_int32),
                                     P(C.c_int32), fp, fp, fp]
        L.orc_lbvh_build_ex.argtypes = [P(capi.SceneDesc), P(C.c_uint64), P(C.c_int32), P(C.c_int32), P(C.c_int32),
                                        P(C.c_int32), fp, fp, fp, P(C.c_int32)]
        L.orc_render_bvh_strided.argtypes = [P(capi.SceneDesc), C.c_uint32, fp, C.c_int, C.c_int, P(C.c_int64)]
        L.orc_render_pixels.argtypes = [P(capi.SceneDesc), C.c_uint32, fp, C.c_int, P(C.c_int32), C.c_int, C.c_int]
        L.orc_render_range.argtypes = [P(capi.SceneDesc), C.c_uint32, C.c_int, C.c_int, C.c_int, fp, C.c_int, P(C.c_int64)]
        L.orc_render_bvh.argtypes = [P(capi.SceneDesc), C.c_uint32, C.c_int, C.c_int, fp, C.c_int, C.c_int, P(C.c_int64), P(C.c_int64)]
        L.orc_radix_sort_pairs.argtypes = [P(C.c_uint64), P(C.c_int32), C.c_int]
        L.orc_radix_sort_pairs.restype = None
        _lib = L
    return _lib


def fptr(a):
    return a.ctypes.data_as(C.POINTER(C.c_float))


class Scene:
    """Owns the ctypes arrays behind a capi.SceneDesc."""

    def __init__(self):
        self.desc = capi.SceneDesc()
        self.keep = []
        self.output_file = "main.png"

    @property
    def ptr(self):
        return C.byref(self.desc)


_OBJ = {"sphere": capi.RT_OBJ_SPHERE, "xy_rect": capi.RT_OBJ_XY_RECT, "xz_rect": capi.RT_OBJ_XZ_RECT,
        "yz_rect": capi.RT_OBJ_YZ_RECT, "cylinder": capi.RT_OBJ_CYLINDER, "box": capi.RT_OBJ_BOX,
        "triangle": capi.RT_OBJ_TRIANGLE}
_MAT = {"lambertian": capi.RT_MAT_LAMBERTIAN, "metal": capi.RT_MAT_METAL, "dielectric": capi.RT_MAT_DIELECTRIC,
        "diffuse_light": capi.RT_MAT_DIFFUSE_LIGHT}


def read_image_rgb8(path):
    from PIL import Image
    im = np.asarray(Image.open(path).convert("RGB"), dtype=np.uint8)
    return np.ascontiguousarray(im)


def make_xform(o):
    """translate(t) * rotate(axis, angle) of an object's "rotate" / "translate" keys, through the oracle's orc_make_xform."""
    xf = capi.XformDesc()
    ax = np.array(o.get("rotate", {}).get("axis", [0, 0, 1]), np.float32)
    tr = np.array(o.get("translate", [0, 0, 0]), np.float32)
    lib().orc_make_xform(fptr(ax), float(o.get("rotate", {}).get("angle", 0.0)), int("rotate" in o), fptr(tr), int("translate" in o), C.byref(xf))
    return xf


def bake(m, p):
    """m * p for a 3x4 row-major matrix in float32, left to right (csrc/rt_json.cpp: Loader::bake)."""
    f = np.float32
    x, y, z = f(p[0]), f(p[1]), f(p[2])
    return [float(f(f(f(f(m[4 * r]) * x) + f(f(m[4 * r + 1]) * y)) + f(f(m[4 * r + 2]) * z)) + f(m[4 * r + 3])) for r in range(3)]


def pack_uv(u, v):
    """(u, v) clamped to [0, 1] as two 16-bit unorm values in one 32-bit word (rt.h: TRIANGLE p[9..11])."""
    f = np.float32
    u = min(max(f(u), f(0)), f(1)); v = min(max(f(v), f(0)), f(1))
    return int(f(f(u * f(65535.0)) + f(0.5))) | (int(f(f(v * f(65535.0)) + f(0.5))) << 16)


def expand_meshes(objs, base_dir):
    """`mesh` objects -> triangles, in place in the object order (same rules as csrc/rt_json.cpp load_obj: v' = v*scale + t
    in float32 — with "rotate": v' = M (v*scale), M = translate * rotate —, 1-based / negative indices, v/vt/vn tokens, `vt` ->
    per-vertex uv, fan triangulation)."""
    out = []
    for o in objs:
        if o["type"] != "mesh":
            out.append(o)
            continue
        sc_, tr = np.float32(o.get("scale", 1.0)), np.array(o.get("translate", [0, 0, 0]), np.float32)
        rot = make_xform(o).m if "rotate" in o else None
        path = o["file"] if os.path.isabs(o["file"]) else os.path.join(base_dir or ".", o["file"])
        v, vt = [], []
        for line in open(path):
            t = line.split()
            if not t:
                continue
            if t[0] == "v":
                if rot is not None:
                    v.append(bake(rot, [np.float32(np.float32(float(t[1 + k])) * sc_) for k in range(3)]))
                else:
                    v.append([float(np.float32(np.float32(float(t[1 + k])) * sc_) + tr[k]) for k in range(3)])
            elif t[0] == "vt":
                vt.append((float(t[1]), float(t[2]) if len(t) > 2 else 0.0))
            elif t[0] == "f":
                toks = [tok.split("/") for tok in t[1:]]
                idx = [int(tk[0]) for tk in toks]
                idx = [len(v) + i if i < 0 else i - 1 for i in idx]
                tidx = [(int(tk[1]) if len(tk) > 1 and tk[1] else None) for tk in toks]
                tidx = [None if j is None else (len(vt) + j if j < 0 else j - 1) for j in tidx]
                for k in range(1, len(idx) - 1):
                    tri = {"type": "triangle", "v0": v[idx[0]], "v1": v[idx[k]], "v2": v[idx[k + 1]], "material": o.get("material", 0)}
                    tt = [tidx[0], tidx[k], tidx[k + 1]]
                    if all(j is not None for j in tt):
                        tri["uv"] = [list(vt[j]) for j in tt]
                    out.append(tri)
    return out


def load_scene(src, width=None, height=None, spp=None, max_depth=None, base_dir=None):
    """src: path to a JSON scene or an already parsed dict (schema of gpu-version/parser.hpp + extensions)."""
    if isinstance(src, (str, os.PathLike)):
        base_dir = base_dir or os.path.dirname(os.path.abspath(src))
        with open(src) as f:
            d = json.load(f)
    else:
        d = src
    L = lib()
    sc = Scene()
    D = sc.desc
    D.width = int(width or d["width"]); D.height = int(height or d["height"])
    D.max_depth = int(max_depth or d["max_depth"]); D.samples_per_pixel = int(spp or d["samples_per_pixel"])
    for k in range(3):
        D.background[k] = d["background"][k]
    D.background_mode = capi.RT_BG_SKY_GRADIENT if d.get("background_mode") == "sky_gradient" else capi.RT_BG_CONSTANT
    D.sampling = capi.RT_SAMPLING_REFERENCE_CPU if d.get("sampling") == "reference_cpu" else capi.RT_SAMPLING_REFERENCE_GPU
    if d.get("sampler") != "rejection":
        D.sampling |= capi.RT_SAMPLING_DIRECT
    cam = d["camera"]
    for k in range(3):
        D.camera.lookfrom[k] = cam["lookfrom"][k]; D.camera.lookat[k] = cam["lookat"][k]; D.camera.vup[k] = cam["vup"][k]
    D.camera.vfov_deg = cam["vfov"]
    # parser.hpp:122 — aspect from the JSON's own width/height (not from overrides), as float division
    D.camera.aspect = cam.get("aspect_ratio", float(np.float32(d["width"]) / np.float32(d["height"])))
    D.camera.aperture = cam["aperture"]
    lf = np.array(cam["lookfrom"], np.float32) - np.array(cam["lookat"], np.float32)
    fd = np.sqrt(np.float32(lf[0] * lf[0] + lf[1] * lf[1]) + lf[2] * lf[2], dtype=np.float32)   # parser.hpp:124
    D.camera.focus_dist = cam.get("focus_dist", float(fd))
    sc.output_file = d.get("output_file", "main.png")

    objs = expand_meshes(d["object"]["data"], base_dir); mats = d["material"]["data"]; texs = d["texture"]["data"]
    xforms = []
    OA = (capi.ObjectDesc * max(1, len(objs)))()
    for i, o in enumerate(objs):
        q = OA[i]
        q.type = _OBJ[o["type"]]; q.material = int(o.get("material", 0)); q.xform = -1
        t = o["type"]
        if t == "sphere":
            vals = list(o["center"]) + [o["radius"]]
        elif t == "xy_rect":
            vals = [o["x0"], o["x1"], o["y0"], o["y1"], o["k"]]
        elif t == "xz_rect":
            vals = [o["x0"], o["x1"], o["z0"], o["z1"], o["k"]]
        elif t == "yz_rect":
            vals = [o["y0"], o["y1"], o["z0"], o["z1"], o["k"]]
        elif t == "cylinder":
            vals = [o["radius"], o["zmin"], o["zmax"]]
        elif t == "box":
            vals = list(o["min"]) + list(o["max"])
        elif t == "triangle":
            vs = [list(o["v0"]), list(o["v1"]), list(o["v2"])]
            if "rotate" in o or "translate" in o:          # baked into the vertices (csrc/rt_json.cpp)
                m = make_xform(o).m
                vs = [bake(m, [np.float32(c) for c in vv]) for vv in vs]
            vals = vs[0] + vs[1] + vs[2]
        for k, v in enumerate(vals):
            q.p[k] = v
        if t == "triangle" and "uv" in o:
            bits = C.cast(q.p, C.POINTER(C.c_uint32))        # raw bits: a float round trip could alter NaN patterns
            for c in range(3):
                bits[9 + c] = pack_uv(*o["uv"][c])
        if t != "triangle" and (t == "cylinder" or "rotate" in o or "translate" in o):
            q.xform = len(xforms); xforms.append(make_xform(o))
    XA = (capi.XformDesc * max(1, len(xforms)))(*xforms)
    MA = (capi.MaterialDesc * max(1, len(mats)))()
    for i, m in enumerate(mats):
        q = MA[i]
        q.type = _MAT[m["type"]]
        q.texture = int(m.get("texture", 0)) if m["type"] in ("lambertian", "diffuse_light") else 0
        if m["type"] == "metal":
            for k in range(3):
                q.albedo[k] = m["albedo"][k]
            q.fuzz = m["fuzz"]
        if m["type"] == "dielectric":
            q.ior = m["index_of_refraction"]
    images = []
    TA = (capi.TextureDesc * max(1, len(texs)))()
    for i, t in enumerate(texs):
        q = TA[i]
        if t["type"] == "solid_color":
            q.type = capi.RT_TEX_SOLID
            for k in range(3):
                q.color[k] = t["color"][k]
        elif t["type"] == "checker":
            q.type = capi.RT_TEX_CHECKER; q.even = int(t["even"]); q.odd = int(t["odd"])
        elif t["type"] == "image":
            q.type = capi.RT_TEX_IMAGE; q.image = len(images)
            path = t["file"] if os.path.isabs(t["file"]) else os.path.join(base_dir or ".", t["file"])
            images.append(read_image_rgb8(path))
    IA = (capi.ImageDesc * max(1, len(images)))()
    for i, im in enumerate(images):
        IA[i].height, IA[i].width = im.shape[0], im.shape[1]
        IA[i].rgb8 = im.ctypes.data_as(C.POINTER(C.c_uint8))
    D.n_objects, D.n_xforms, D.n_materials, D.n_textures, D.n_images = len(objs), len(xforms), len(mats), len(texs), len(images)
    D.objects = C.cast(OA, C.POINTER(capi.ObjectDesc)); D.xforms = C.cast(XA, C.POINTER(capi.XformDesc))
    D.materials = C.cast(MA, C.POINTER(capi.MaterialDesc)); D.textures = C.cast(TA, C.POINTER(capi.TextureDesc))
    D.images = C.cast(IA, C.POINTER(capi.ImageDesc))
    sc.keep = [OA, XA, MA, TA, IA, images]
    return sc


def render(sc, seed=0x2023, spp_begin=0, spp_count=0, threads=None):
    W, H = sc.desc.width, sc.desc.height
    out = np.zeros((H, W, 3), np.float32)
    nr = C.c_int64(0)
    rc = lib().orc_render(sc.ptr, seed, spp_begin, spp_count, fptr(out), threads or os.cpu_count() or 1, C.byref(nr))
    assert rc == 0
    return out, nr.value


def render_bvh_strided(sc, seed, stride, threads=None):
    """Host-twin render of the pixels whose index is a multiple of `stride` (the other pixels stay 0); returns (image, rays)."""
    out = np.zeros((sc.desc.height, sc.desc.width, 3), np.float32)
    nr = C.c_int64(0)
    assert lib().orc_render_bvh_strided(sc.ptr, seed, fptr(out), threads or os.cpu_count() or 1, stride, C.byref(nr)) == 0
    return out, nr.value


def render_pixels(sc, seed, pixels, use_bvh=False, threads=None):
    """Oracle render of the listed pixel indices only (brute-force walk unless use_bvh); returns the (H, W, 3) image, zero elsewhere."""
    out = np.zeros((sc.desc.height, sc.desc.width, 3), np.float32)
    px = np.ascontiguousarray(sorted(int(p) for p in pixels), np.int32)
    assert lib().orc_render_pixels(sc.ptr, seed, fptr(out), threads or os.cpu_count() or 1, px.ctypes.data_as(C.POINTER(C.c_int32)), len(px), int(use_bvh)) == 0
    return out


def render_range(sc, seed, spp_begin, spp_count, total_spp, threads=None):
    """One rank's share [spp_begin, spp_begin + spp_count) of an image of total_spp samples."""
    out = np.zeros((sc.desc.height, sc.desc.width, 3), np.float32)
    nr = C.c_int64(0)
    rc = lib().orc_render_range(sc.ptr, seed, spp_begin, spp_count, total_spp, fptr(out), threads or os.cpu_count() or 1, C.byref(nr))
    assert rc == 0
    return out, nr.value


def render_ex(sc, seed=0x2023, spp_begin=0, spp_count=0, threads=None, flags=0):
    """orc_render with rt_render flags (RT_RENDER_NEE); returns (image, rays, shadow_rays)."""
    W, H = sc.desc.width, sc.desc.height
    out = np.zeros((H, W, 3), np.float32)
    nr, ns = C.c_int64(0), C.c_int64(0)
    rc = lib().orc_render_ex(sc.ptr, seed, spp_begin, spp_count, fptr(out), threads or os.cpu_count() or 1, flags, C.byref(nr), C.byref(ns))
    assert rc == 0, rc
    return out, nr.value, ns.value


COUNTER_NAMES = ("n_box_tests", "n_node_visits", "n_sphere_reject", "n_sphere_accept", "n_rect_tests", "n_cyl_reject", "n_cyl_accept",
                 "n_tri_tests", "n_boxprim_tests", "max_stack")


def render_bvh(sc, seed=0x2023, spp_begin=0, spp_count=0, threads=None, flags=0):
    """orc_render through the host twin of the device traversal (oracle/orc_trav.c); returns (image, rays, counters dict)."""
    W, H = sc.desc.width, sc.desc.height
    out = np.zeros((H, W, 3), np.float32)
    nr = C.c_int64(0)
    cnt = (C.c_int64 * 10)()
    rc = lib().orc_render_bvh(sc.ptr, seed, spp_begin, spp_count, fptr(out), threads or os.cpu_count() or 1, flags, C.byref(nr), cnt)
    assert rc == 0, rc
    return out, nr.value, dict(zip(COUNTER_NAMES, [int(x) for x in cnt]))


def aov(sc):
    out = np.zeros((sc.desc.height, sc.desc.width, 24), np.float32)
    assert lib().orc_aov(sc.ptr, fptr(out)) == 0
    return out


def tonemap(rgb_sum, spp, gamma):
    H, W, _ = rgb_sum.shape
    out = np.zeros((H, W, 3), np.uint8)
    lib().orc_tonemap(fptr(np.ascontiguousarray(rgb_sum, np.float32)), W, H, spp, gamma,
                      out.ctypes.data_as(C.POINTER(C.c_ubyte)))
    return out


def lbvh(sc):
    n = sc.desc.n_objects
    m = max(n - 1, 1)
    r = dict(morton=np.zeros(max(n, 1), np.uint64), prim=np.zeros(max(n, 1), np.int32), left=np.zeros(m, np.int32),
             right=np.zeros(m, np.int32), parent=np.zeros(max(2 * n - 1, 1), np.int32), node_aabb=np.zeros((m, 6), np.float32),
             leaf_aabb=np.zeros((max(n, 1), 6), np.float32), scene_aabb=np.zeros(6, np.float32))
    P = C.POINTER
    info = np.zeros(4, np.int32)
    lib().orc_lbvh_build_ex(sc.ptr, r["morton"].ctypes.data_as(P(C.c_uint64)), r["prim"].ctypes.data_as(P(C.c_int32)),
                            r["left"].ctypes.data_as(P(C.c_int32)), r["right"].ctypes.data_as(P(C.c_int32)),
                            r["parent"].ctypes.data_as(P(C.c_int32)), fptr(r["node_aabb"]), fptr(r["leaf_aabb"]),
                            fptr(r["scene_aabb"]), info.ctypes.data_as(P(C.c_int32)))
    return trim_bvh(r, n, int(info[0]), int(info[1]), int(info[2]))


def trim_bvh(r, n, n_tree, n_giants, depth):
    """Keeps the valid part of the LBVH arrays: n_tree leaves form the tree (n_tree - 1 internal nodes), the giants sit in the
    last n_giants leaf slots; parent = n_tree - 1 internal entries then n_tree leaf entries."""
    m = max(n_tree - 1, 0)
    r = dict(r)
    r.update(n=n, n_tree=n_tree, n_giants=n_giants, depth=depth, morton=r["morton"][:n], prim=r["prim"][:n], leaf_aabb=r["leaf_aabb"][:n],
             left=r["left"][:m], right=r["right"][:m], node_aabb=r["node_aabb"][:m], parent=r["parent"][:max(2 * n_tree - 1, 0)])
    return r


def kat(kind, inp, sampling=capi.RT_SAMPLING_REFERENCE_GPU):
    widths = {"sphere": (12, 11), "rect": (14, 11), "cyl": (18, 11), "scatter": (32, 11), "camera": (14, 6), "xform": (10, 12)}
    wi, wo = widths[kind]
    inp = np.ascontiguousarray(inp, np.float32).reshape(-1, wi)
    out = np.zeros((inp.shape[0], wo), np.float32)
    f = getattr(lib(), "orc_kat_" + kind)
    if kind == "scatter":
        f(fptr(inp), fptr(out), inp.shape[0], sampling)
    else:
        f(fptr(inp), fptr(out), inp.shape[0])
    return out
