#!/usr/bin/env python3
"""Per-stage device time and traversal counters of one workload (development aid; not the headline bench).
Usage: python tools/stage_bench.py [workload] [spp] [reps] [flags]"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as ge  # noqa: E402
import bench  # noqa: E402

rt = ge.load_package()
from rt_b200 import capi  # noqa: E402

workload = sys.argv[1] if len(sys.argv) > 1 else "rtiow_1080p_1024spp"
scene, W, H, spp = bench.WORKLOADS[workload]
spp = int(sys.argv[2]) if len(sys.argv) > 2 else 64
reps = int(sys.argv[3]) if len(sys.argv) > 3 else 3
extra = int(sys.argv[4]) if len(sys.argv) > 4 else 0
r = rt.Renderer(0)
bench.WORKLOADS[workload] = (scene, W, H, spp)
upload, _ = bench.make_upload(rt, workload)
upload(r)
r.render(seed=0x2023, flags=capi.RT_RENDER_COUNT_EVENTS | extra, spp_count=max(1, spp // 8))
_, _, c = r.download(want_sum=False, want_rgb8=False)
rays = c["rays"]
out = {"workload": workload, "spp": spp, "rays_per_sample": rays / c["samples"], "node_visits_per_ray": c["n_node_visits"] / rays,
       "sphere_tests_per_ray": (c["n_sphere_reject"] + c["n_sphere_accept"]) / rays,
       "other_prim_tests_per_ray": (c["n_rect_tests"] + c["n_cyl_reject"] + c["n_cyl_accept"] + c["n_tri_tests"] + c["n_boxprim_tests"]) / rays}
for _ in range(2):
    r.render(seed=0x2023, flags=extra)
r.synchronize()
r.stats_reset()
for _ in range(reps):
    r.render(seed=0x2023, flags=capi.RT_RENDER_TIME_STAGES | extra)
_, _, st = r.download(want_sum=False, want_rgb8=False)
tot = st["ms_generate"] + st["ms_extend"] + st["ms_shade"] + st["ms_accumulate"] + st["ms_tail"]
out.update({"ms_generate": st["ms_generate"] / reps, "ms_extend": st["ms_extend"] / reps, "ms_shade": st["ms_shade"] / reps,
            "ms_accumulate": st["ms_accumulate"] / reps, "ms_tail": st["ms_tail"] / reps, "ms_sum": tot / reps, "launches": st["launches"] / reps,
            "Msamples_per_s": W * H * spp / (tot / reps * 1e-3) / 1e6, "Grays_per_s_extend": st["rays"] / reps / (max(st["ms_extend"], 1e-9) / reps * 1e-3) / 1e9,
            "Grays_per_s_shade": st["rays"] / reps / (max(st["ms_shade"], 1e-9) / reps * 1e-3) / 1e9})
# whole render, one launch of rt_render per repetition: two passes in flight (default) vs one at a time
for name, fl in (("ms_render_overlapped", 0), ("ms_render_serial", getattr(capi, "RT_RENDER_SERIAL", 0))):
    ms = []
    for _ in range(reps + 1):
        r.render(seed=0x2023, flags=fl | extra)
        _, _, st = r.download(want_sum=False, want_rgb8=False)
        ms.append(st["ms_total"])
    out[name] = min(ms[1:])
print(json.dumps(out))
r.close()
