#!/usr/bin/env python3
"""GPU self-consistency at production size: bounce 0 from the tile candidate lists vs from the tree (RT_RENDER_NO_TILE_LISTS), many samples.
Both must give the same image and ray count; a difference means one of the two is not conservative (this check found the grazing-triangle
acceptance bug of round 2: 3 pixels in 2.1 G paths on the 1 M-triangle mesh).  Usage: python tools/tile_vs_traversal.py [spp]"""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import __graft_entry__ as ge, bench
rt = ge.load_package()
from rt_b200 import scenes, capi
spp = int(sys.argv[1]) if len(sys.argv) > 1 else 256
r = rt.Renderer(0)
for wl in ("mesh1m_4k_1024spp", "rtiow_1080p_1024spp", "blue2_1080p_256spp", "sample_scene_ext_1600x900_512spp", "dna000_1080p_1024spp"):
    scene, W, H, _ = bench.WORKLOADS[wl]
    bench.WORKLOADS[wl] = (scene, W, H, spp)
    upload, _ = bench.make_upload(rt, wl)
    upload(r)
    r.stats_reset(); r.render(seed=0x2023); a, _, sa = r.download()
    r.stats_reset(); r.render(seed=0x2023, flags=capi.RT_RENDER_NO_TILE_LISTS); b, _, sb = r.download()
    diff = (a != b).any(-1)
    print(wl, f"{W}x{H}x{spp}: differing pixels", int(diff.sum()), "rays", sa["rays"], sb["rays"], flush=True)
r.close()
