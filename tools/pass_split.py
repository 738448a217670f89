#!/usr/bin/env python3
"""How should ONE rank's share of an 8-GPU render (128 of 1024 spp at 1080p) be cut into passes?  ms per variant (GPU box)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as ge, bench
rt = ge.load_package()
from rt_b200 import capi
r = rt.Renderer(0)
upload, _ = bench.make_upload(rt, "rtiow_1080p_1024spp")
upload(r)
np_ = 1920 * 1080
for spp_pass in (128, 64, 43, 32, 22, 16):
    ms = []
    for k in range(5):
        r.render(seed=0x2023, spp_begin=0, spp_count=128, total_spp=1024, max_paths_per_pass=spp_pass * np_, flags=capi.RT_RENDER_KEEP_CHUNKS)
        _, _, st = r.download(want_sum=False, want_rgb8=False)
        ms.append(st["ms_total"])
    print("128 spp in passes of <=", spp_pass, "spp:", round(min(ms[1:]), 2), "ms,", st["passes"] // 1, "passes total counted", flush=True)
r.close()
