#!/usr/bin/env python3
"""Writes sample_scene_ext.json + its two assets (ext_texture.png, ext_cube.obj): the scene BASELINE.json's config 3
describes ("AABB boxes, rotate_y instances, image textures and multiple diffuse_light emitters") but the reference's
sample_scene.json does not contain.  Uses only this repository's schema extensions; the base
layout (camera, ground sphere, three unit spheres) follows gpu-version/sample_scene.json.  Deterministic: no RNG."""
import json
import os

import numpy as np
from PIL import Image

HERE = os.path.dirname(os.path.abspath(__file__))

# 64 x 48 brick-like pattern (RGB8)
y, x = np.mgrid[0:48, 0:64]
row = y // 8
mortar = ((y % 8) == 0) | (((x + 16 * (row % 2)) % 32) == 0)
img = np.zeros((48, 64, 3), np.uint8)
img[..., 0] = np.where(mortar, 200, 150 + (x * 3 + y * 5) % 60)
img[..., 1] = np.where(mortar, 196, 60 + (x * 7 + y) % 40)
img[..., 2] = np.where(mortar, 190, 40 + (x + y * 3) % 30)
Image.fromarray(img).save(os.path.join(HERE, "ext_texture.png"))

with open(os.path.join(HERE, "ext_cube.obj"), "w") as f:          # unit cube, quads (fan-triangulated by the loaders) + one n-gon
    f.write("# unit cube centred at the origin\n")
    for v in [(-.5, -.5, -.5), (.5, -.5, -.5), (.5, .5, -.5), (-.5, .5, -.5), (-.5, -.5, .5), (.5, -.5, .5), (.5, .5, .5), (-.5, .5, .5)]:
        f.write("v %g %g %g\n" % v)
    f.write("vt 0 0\nvn 0 0 1\n")
    for face in ["1 2 3 4", "5/1/1 8/1/1 7/1/1 6/1/1", "1 5 6 2", "2 6 7 3", "3 7 8 4", "-4 -8 -1 -5"]:
        f.write("f " + face + "\n")

with open(os.path.join(HERE, "ext_pyramid.obj"), "w") as f:       # square pyramid with texture coordinates (v/vt faces, one quad)
    f.write("# pyramid: unit base in the plane y = 0, apex at y = 1\n")
    for v in [(-.5, 0, -.5), (.5, 0, -.5), (.5, 0, .5), (-.5, 0, .5), (0, 1, 0)]:
        f.write("v %g %g %g\n" % v)
    for vt in [(0, 0), (1, 0), (1, 1), (0, 1), (0.5, 0.5), (0.25, 0.75)]:
        f.write("vt %g %g\n" % vt)
    for face in ["1/1 2/2 5/5", "2/2 3/3 5/5", "3/3 4/4 5/5", "4/4 1/1 5/6", "4/4 3/3 2/2 1/1"]:
        f.write("f " + face + "\n")

scene = {
    "background": [0.02, 0.03, 0.05], "max_depth": 50, "samples_per_pixel": 4096, "width": 1600, "height": 900,
    "output_file": "./build/output/sample_scene_ext.png",
    "camera": {"lookfrom": [0, 1.6, 5.5], "lookat": [0, 0.6, -1], "vup": [0, 1, 0], "vfov": 40, "aperture": 0.0},
    "texture": {"data": [
        {"type": "solid_color", "color": [0.1, 0.2, 0.5]},          # 0
        {"type": "solid_color", "color": [0.9, 0.9, 0.9]},          # 1
        {"type": "checker", "even": 0, "odd": 1},                   # 2 ground
        {"type": "image", "file": "ext_texture.png"},               # 3
        {"type": "solid_color", "color": [6, 5.5, 5]},              # 4 warm emitter
        {"type": "solid_color", "color": [2, 3, 8]},                # 5 blue emitter
        {"type": "solid_color", "color": [7, 2, 2]},                # 6 red emitter
        {"type": "solid_color", "color": [0.7, 0.7, 0.3]},          # 7
    ]},
    "material": {"data": [
        {"type": "lambertian", "texture": 2},                        # 0 checker ground
        {"type": "lambertian", "texture": 3},                        # 1 image-textured
        {"type": "metal", "albedo": [0.8, 0.8, 0.85], "fuzz": 0.1},  # 2
        {"type": "dielectric", "index_of_refraction": 1.5},          # 3
        {"type": "diffuse_light", "texture": 4},                     # 4
        {"type": "diffuse_light", "texture": 5},                     # 5
        {"type": "diffuse_light", "texture": 6},                     # 6
        {"type": "lambertian", "texture": 7},                        # 7
    ]},
    "object": {"data": [
        {"type": "sphere", "center": [0, -100.5, -1], "radius": 100, "material": 0},
        {"type": "box", "min": [-0.6, 0, -0.6], "max": [0.6, 1.8, 0.6], "material": 1,
         "rotate": {"axis": [0, 1, 0], "angle": 18}, "translate": [-1.6, -0.5, -1.8]},                 # Cornell-style tall box
        {"type": "box", "min": [-0.5, 0, -0.5], "max": [0.5, 0.9, 0.5], "material": 2,
         "rotate": {"axis": [0, 1, 0], "angle": -15}, "translate": [1.5, -0.5, -0.6]},                 # short box
        {"type": "sphere", "center": [0, 0, 0], "radius": 0.6, "material": 1,
         "rotate": {"axis": [0.3, 1, 0.2], "angle": 40}, "translate": [0, 0.1, -1]},                      # image texture through the uv of a ROTATED sphere
        {"type": "sphere", "center": [0.2, 0.0, 0.9], "radius": 0.5, "material": 3},
        {"type": "sphere", "center": [0.2, 0.0, 0.9], "radius": -0.42, "material": 3},                     # hollow glass
        {"type": "xz_rect", "x0": -1.5, "x1": 1.5, "z0": -2.5, "z1": 0.5, "k": 3.2, "material": 4},       # emitter 1
        {"type": "sphere", "center": [-2.6, 0.2, 0.6], "radius": 0.3, "material": 5},                      # emitter 2
        {"type": "cylinder", "radius": 0.08, "zmin": -0.9, "zmax": 0.9, "material": 6,
         "rotate": {"axis": [0, 1, 0], "angle": 70}, "translate": [2.4, 1.3, -1.4]},                      # emitter 3
        {"type": "mesh", "file": "ext_cube.obj", "material": 7, "scale": 0.55, "translate": [-0.5, -0.2, 0.9]},
        {"type": "triangle", "v0": [-3.5, -0.5, -3.5], "v1": [3.5, -0.5, -3.5], "v2": [0, 3.0, -3.8], "material": 2},
        {"type": "yz_rect", "y0": -0.5, "y1": 2.5, "z0": -3.5, "z1": 1.0, "k": -3.4, "material": 7},
        # rotate / translate on the other object types: an image-textured rect turned about y, a `vt` mesh (scaled, rotated,
        # translated), a rotated triangle with explicit uv
        {"type": "xy_rect", "x0": -0.5, "x1": 0.5, "y0": -0.4, "y1": 0.4, "k": 0, "material": 1,
         "rotate": {"axis": [0, 1, 0], "angle": -35}, "translate": [2.3, 0.3, 0.9]},
        {"type": "mesh", "file": "ext_pyramid.obj", "material": 1, "scale": 0.7,
         "rotate": {"axis": [0, 1, 0], "angle": 25}, "translate": [-2.2, -0.5, 1.6]},
        {"type": "triangle", "v0": [-0.4, 0, 0], "v1": [0.4, 0, 0], "v2": [0, 0.7, 0], "material": 1, "uv": [[0, 0], [1, 0], [0.5, 1]],
         "rotate": {"axis": [1, 0, 0], "angle": -20}, "translate": [1.0, 0.9, -2.2]},
    ]},
}
json.dump(scene, open(os.path.join(HERE, "sample_scene_ext.json"), "w"), indent=1)
print("wrote sample_scene_ext.json, ext_texture.png, ext_cube.obj, ext_pyramid.obj")
