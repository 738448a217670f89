#!/usr/bin/env python3
"""Wall-clock breakdown of one end-to-end step (upload + BVH build, render, download). Usage: e2e_breakdown.py [workload]"""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as ge
import bench
rt = ge.load_package()
workload = sys.argv[1] if len(sys.argv) > 1 else "rtiow_1080p_1024spp"
upload, _ = bench.make_upload(rt, workload)
r = rt.Renderer(0)
for it in range(3):
    t0 = time.perf_counter(); upload(r); r.synchronize()
    t1 = time.perf_counter(); r.render(seed=0x2023); t1b = time.perf_counter(); r.synchronize()
    t2 = time.perf_counter(); out = r.download(gamma=0)
    t3 = time.perf_counter()
    print({"upload_ms": round((t1 - t0) * 1e3, 2), "render_enqueue_ms": round((t1b - t1) * 1e3, 2), "render_ms": round((t2 - t1) * 1e3, 2),
           "download_ms": round((t3 - t2) * 1e3, 2), "device_ms": round(out[2]["ms_total"], 2)})
r.close()
