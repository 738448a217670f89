#!/usr/bin/env python3
"""Small driver for ncu captures: uploads a workload scene and renders it N times through the C ABI (no timing, no
oracle).  Usage: python tools/profile_run.py [workload] [renders] [spp]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as ge  # noqa: E402
import bench  # noqa: E402

rt = ge.load_package()
workload = sys.argv[1] if len(sys.argv) > 1 else "rtiow_c1_1200x800_10spp"
renders = int(sys.argv[2]) if len(sys.argv) > 2 else 2
scene, W, H, spp = bench.WORKLOADS[workload]
if len(sys.argv) > 3:
    spp = int(sys.argv[3])
r = rt.Renderer(0)
bench.WORKLOADS[workload] = (scene, W, H, spp)
upload, _ = bench.make_upload(rt, workload)
upload(r)
for _ in range(renders):
    r.render(seed=0x2023)
_, _, st = r.download(want_sum=False, want_rgb8=False)
print({k: st[k] for k in ("samples", "rays", "launches", "passes", "ms_total")})
r.close()
