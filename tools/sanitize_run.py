#!/usr/bin/env python3
"""Small renders that instantiate every k_shade / k_extend / k_generate template (CAM, NEE, SORT, SHADOW, fix-up, COUNT, both samplers,
shared- and global-memory trees) for compute-sanitizer: python tools/sanitize_run.py"""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as ge
rt = ge.load_package()
from rt_b200 import capi, scenes
G = os.path.join(ROOT, "tests", "golden")
r = rt.Renderer(0)
def go(name, W, H, spp, flags=0, sampler=None, sampling=None, **kw):
    d = json.load(open(os.path.join(G, name + ".json")))
    if sampler: d["sampler"] = sampler
    if sampling: d["sampling"] = sampling
    r.upload_json_text(json.dumps(d), G); r.override(width=W, height=H, spp=spp)
    r.render(seed=3, flags=flags, **kw); r.download()
    print(name, W, H, spp, flags, sampler, "ok")
go("rtiow_486", 96, 54, 3)                                        # sphere-only extend, tile_mode 2 at this size (fix-up launch), direct sampler
go("rtiow_486", 256, 144, 2, sampler="rejection")                 # tile_mode 1, rejection loops
go("blue2", 64, 36, 3, capi.RT_RENDER_SORT)                       # mixed primitives, in-CTA sort
go("blue2", 67, 35, 3, capi.RT_RENDER_COUNT_EVENTS | capi.RT_RENDER_NO_TILE_LISTS)   # odd size, COUNT instances
go("sample_scene_ext", 64, 36, 3, capi.RT_RENDER_NEE, sampling="reference_cpu")     # NEE shade + shadow (any-hit) instances, image texture
go("sample_scene_ext", 64, 36, 2, capi.RT_RENDER_NEE | capi.RT_RENDER_COUNT_EVENTS, sampling="reference_cpu")
go("dna_000", 64, 36, 2, max_paths_per_pass=64 * 36)              # all-emitter scene, several passes on two streams
go("basic_scene", 64, 36, 2)
ts = scenes.mesh_scene(40, 40, width=64, height=36, spp=2, seed=5, max_depth=6)      # global-memory tree needs > ~1000 nodes
r.upload(ts); r.render(seed=1); r.download(); r.render(seed=1, flags=capi.RT_RENDER_COUNT_EVENTS); r.download()
print("mesh ok")
r.render_aov()
r.close()
