"""Procedural scenes handed to the library as flat tables (rt_scene_upload), for configurations the reference names
but ships no file for: BASELINE.json config 4, "loaded triangle-mesh model (~1M triangles) under GPU LBVH".

The reference's only mesh path is taichi-version/main.py:23-54 (an OBJ that is missing from the tree plus 21x21 point
grids, ~800 triangles per frame — a displaced grid); this module generates the same kind of surface at any size from a
seed, as numpy arrays laid out exactly like rt_object_desc, so a 1M-triangle scene is built in well under a second
and no file has to travel.  JSON scenes can also pull an OBJ through the `mesh` object type (csrc/rt_json.cpp).
"""
import ctypes as C

import numpy as np

from . import capi

OBJECT_DTYPE = np.dtype([("type", "<i4"), ("material", "<i4"), ("xform", "<i4"), ("reserved", "<i4"), ("p", "<f4", (12,))])
assert OBJECT_DTYPE.itemsize == C.sizeof(capi.ObjectDesc)


def displaced_grid(nx, nz, seed=2023, extent=10.0, amplitude=0.6):
    """(nx+1) x (nz+1) vertices over [-extent, extent]^2, heights = three fixed waves + seeded noise; 2*nx*nz triangles.
    Returns float32 arrays v0, v1, v2 of shape (2*nx*nz, 3)."""
    rng = np.random.default_rng(seed)
    x = np.linspace(-extent, extent, nx + 1, dtype=np.float32)
    z = np.linspace(-extent, extent, nz + 1, dtype=np.float32)
    X, Z = np.meshgrid(x, z, indexing="ij")
    Y = amplitude * (np.sin(0.9 * X) * np.cos(0.7 * Z) + 0.5 * np.sin(2.3 * X + 1.0) * np.sin(1.9 * Z)
                     + 0.25 * np.cos(4.1 * X - 2.2 * Z))
    Y = (Y + 0.05 * amplitude * rng.standard_normal(Y.shape)).astype(np.float32)
    P = np.stack([X, Y, Z], axis=-1).astype(np.float32)
    a, b, c, d = P[:-1, :-1], P[1:, :-1], P[1:, 1:], P[:-1, 1:]
    v0 = np.concatenate([a.reshape(-1, 3), a.reshape(-1, 3)])
    v1 = np.concatenate([c.reshape(-1, 3), d.reshape(-1, 3)])
    v2 = np.concatenate([b.reshape(-1, 3), c.reshape(-1, 3)])
    return v0, v1, v2


class TableScene:
    """Owns the numpy / ctypes arrays behind a capi.SceneDesc (`.desc`), ready for Renderer.upload()."""

    def __init__(self, objects, materials, textures, camera, width, height, spp, max_depth=50, background=(0.5, 0.7, 1.0),
                 sky_gradient=True, sampling=capi.RT_SAMPLING_REFERENCE_GPU | capi.RT_SAMPLING_DIRECT):
        self.objects = np.ascontiguousarray(objects, dtype=OBJECT_DTYPE)
        D = self.desc = capi.SceneDesc()
        D.width, D.height, D.max_depth, D.samples_per_pixel = width, height, max_depth, spp
        for k in range(3):
            D.background[k] = background[k]
            D.camera.lookfrom[k], D.camera.lookat[k], D.camera.vup[k] = camera["lookfrom"][k], camera["lookat"][k], camera.get("vup", (0, 1, 0))[k]
        D.background_mode = capi.RT_BG_SKY_GRADIENT if sky_gradient else capi.RT_BG_CONSTANT
        D.sampling = sampling
        D.camera.vfov_deg = camera["vfov"]
        D.camera.aspect = camera.get("aspect_ratio", width / height)
        D.camera.aperture = camera.get("aperture", 0.0)
        lf = np.asarray(camera["lookfrom"], np.float32) - np.asarray(camera["lookat"], np.float32)
        D.camera.focus_dist = camera.get("focus_dist", float(np.sqrt((lf * lf).sum(dtype=np.float32))))
        self.MA = (capi.MaterialDesc * max(1, len(materials)))()
        for i, m in enumerate(materials):
            q = self.MA[i]
            q.type = m["type"]; q.texture = m.get("texture", 0); q.fuzz = m.get("fuzz", 0.0); q.ior = m.get("ior", 1.5)
            for k in range(3):
                q.albedo[k] = m.get("albedo", (0, 0, 0))[k]
        self.TA = (capi.TextureDesc * max(1, len(textures)))()
        for i, t in enumerate(textures):
            q = self.TA[i]
            q.type = t.get("type", capi.RT_TEX_SOLID); q.even = t.get("even", 0); q.odd = t.get("odd", 0)
            for k in range(3):
                q.color[k] = t.get("color", (0, 0, 0))[k]
        self.XA = (capi.XformDesc * 1)()
        self.IA = (capi.ImageDesc * 1)()
        D.n_objects, D.n_xforms, D.n_materials, D.n_textures, D.n_images = len(self.objects), 0, len(materials), len(textures), 0
        D.objects = self.objects.ctypes.data_as(C.POINTER(capi.ObjectDesc))
        D.xforms = C.cast(self.XA, C.POINTER(capi.XformDesc)); D.materials = C.cast(self.MA, C.POINTER(capi.MaterialDesc))
        D.textures = C.cast(self.TA, C.POINTER(capi.TextureDesc)); D.images = C.cast(self.IA, C.POINTER(capi.ImageDesc))

    @property
    def ptr(self):
        return C.byref(self.desc)


def mesh_scene(nx=708, nz=708, width=3840, height=2160, spp=1024, seed=2023, max_depth=50):
    """Config 4: a displaced-grid terrain of 2*nx*nz triangles (708 x 708 -> 1,002,528) with a checker-free lambertian
    surface, three large spheres (metal, glass, lambertian) resting on it, one emissive rectangle overhead, sky gradient."""
    v0, v1, v2 = displaced_grid(nx, nz, seed)
    nt = len(v0)
    extra = 4
    obj = np.zeros(nt + extra, OBJECT_DTYPE)
    obj["xform"] = -1
    obj["type"][:nt] = capi.RT_OBJ_TRIANGLE
    obj["material"][:nt] = 0
    obj["p"][:nt, 0:3], obj["p"][:nt, 3:6], obj["p"][:nt, 6:9] = v0, v1, v2
    spheres = [((-3.0, 1.6, 0.0, 1.2), 1), ((0.0, 1.7, 1.5, 1.2), 2), ((3.2, 1.6, -0.5, 1.2), 3)]
    for k, (s, m) in enumerate(spheres):
        obj["type"][nt + k] = capi.RT_OBJ_SPHERE; obj["material"][nt + k] = m; obj["p"][nt + k, 0:4] = s
    obj["type"][nt + 3] = capi.RT_OBJ_XZ_RECT; obj["material"][nt + 3] = 4
    obj["p"][nt + 3, 0:5] = (-4.0, 4.0, -4.0, 4.0, 7.0)
    textures = [{"color": (0.55, 0.5, 0.35)}, {"color": (0.4, 0.2, 0.1)}, {"color": (6.0, 6.0, 5.5)}]
    materials = [{"type": capi.RT_MAT_LAMBERTIAN, "texture": 0}, {"type": capi.RT_MAT_METAL, "albedo": (0.7, 0.6, 0.5), "fuzz": 0.05},
                 {"type": capi.RT_MAT_DIELECTRIC, "ior": 1.5}, {"type": capi.RT_MAT_LAMBERTIAN, "texture": 1},
                 {"type": capi.RT_MAT_DIFFUSE_LIGHT, "texture": 2}]
    camera = {"lookfrom": (0.0, 6.0, 14.0), "lookat": (0.0, 0.5, 0.0), "vfov": 40.0, "aspect_ratio": width / height}
    return TableScene(obj, materials, textures, camera, width, height, spp, max_depth)
