#!/usr/bin/env python3
"""Stage time as a function of max_depth (shows where the bounce loop spends its time). Usage: depth_sweep.py [workload] [spp]"""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as ge
import bench
rt = ge.load_package()
from rt_b200 import capi
workload = sys.argv[1] if len(sys.argv) > 1 else "rtiow_1080p_1024spp"
scene, W, H, spp = bench.WORKLOADS[workload]
spp = int(sys.argv[2]) if len(sys.argv) > 2 else 16
r = rt.Renderer(0)
bench.WORKLOADS[workload] = (scene, W, H, spp)
upload, _ = bench.make_upload(rt, workload)
upload(r)
prev = None
for depth in (1, 2, 3, 4, 6, 8, 12, 16, 24, 50):
    r.override(max_depth=depth)
    for _ in range(2):
        r.render(seed=0x2023)
    r.synchronize(); r.stats_reset()
    reps = 3
    for _ in range(reps):
        r.render(seed=0x2023, flags=capi.RT_RENDER_TIME_STAGES)
    _, _, st = r.download(want_sum=False, want_rgb8=False)
    row = {"depth": depth, "rays": st["rays"] // reps, "ms_extend": round(st["ms_extend"] / reps, 3), "ms_shade": round(st["ms_shade"] / reps, 3),
           "ms_total": round(st["ms_total"], 3)}
    if prev:
        dr = row["rays"] - prev["rays"]
        row["d_rays_M"] = round(dr / 1e6, 2)
        row["ext_ns_per_kray"] = round((row["ms_extend"] - prev["ms_extend"]) * 1e6 / max(dr, 1) * 1e3, 1)
        row["shade_ns_per_kray"] = round((row["ms_shade"] - prev["ms_shade"]) * 1e6 / max(dr, 1) * 1e3, 1)
    print(json.dumps(row)); prev = row
r.close()
