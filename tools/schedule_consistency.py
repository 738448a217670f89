#!/usr/bin/env python3
"""GPU self-consistency at production size: the image must not depend on the schedule.  Per workload the default render is compared,
bit for bit and ray for ray, with  RT_RENDER_NO_TAIL (every bounce through the queues), RT_RENDER_SERIAL (one pass stream),
RT_RENDER_NO_TILE_LISTS (bounce 0 through the tree), a small pass size (many passes across sample chunks) and rejection-free
schedule combinations of those.  Usage: python tools/schedule_consistency.py [spp]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import __graft_entry__ as ge, bench
rt = ge.load_package()
from rt_b200 import capi
spp = int(sys.argv[1]) if len(sys.argv) > 1 else 96
r = rt.Renderer(0)
bad = 0
for wl in ("rtiow_1080p_1024spp", "blue2_1080p_256spp", "sample_scene_ext_1600x900_512spp", "dna000_1080p_1024spp", "mesh1m_4k_1024spp"):
    scene, W, H, _ = bench.WORKLOADS[wl]
    bench.WORKLOADS[wl] = (scene, W, H, spp)
    upload, _ = bench.make_upload(rt, wl)
    upload(r)
    r.stats_reset(); r.render(seed=0x2023); ref, _, s0 = r.download()
    variants = {"no_tail": dict(flags=capi.RT_RENDER_NO_TAIL), "serial": dict(flags=capi.RT_RENDER_SERIAL),
                "no_tile_lists": dict(flags=capi.RT_RENDER_NO_TILE_LISTS), "small_passes": dict(max_paths_per_pass=W * H * 7),
                "all": dict(flags=capi.RT_RENDER_NO_TAIL | capi.RT_RENDER_SERIAL | capi.RT_RENDER_NO_TILE_LISTS, max_paths_per_pass=W * H * 5)}
    for name, kw in variants.items():
        r.stats_reset(); r.render(seed=0x2023, **kw); img, _, s1 = r.download()
        d = int((img != ref).any(-1).sum())
        ok = d == 0 and s1["rays"] == s0["rays"]
        bad += not ok
        print(wl, f"{W}x{H}x{spp}", name, "differing pixels", d, "rays", s1["rays"], "==" if s1["rays"] == s0["rays"] else "!=", s0["rays"], "passes", s1["passes"], flush=True)
r.close()
print("RESULT", "ok" if bad == 0 else f"{bad} variant(s) differ")
sys.exit(1 if bad else 0)
