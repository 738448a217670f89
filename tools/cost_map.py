#!/usr/bin/env python3
"""Traversal cost per primary ray (BVH nodes visited), by image row. Needs a library built with -DRT_AOV_COST=1:
  python tools/build_variant.py cost -DRT_AOV_COST=1;  RT_B200_LIB=build_variants/cost.so python tools/cost_map.py [workload]"""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as ge
import bench
rt = ge.load_package()
workload = sys.argv[1] if len(sys.argv) > 1 else "mesh1m_1080p_64spp"
upload, _ = bench.make_upload(rt, workload)
r = rt.Renderer(0)
upload(r)
a = r.render_aov()
c = a[..., 23]
print("nodes visited per primary ray: mean %.1f  median %.0f  p99 %.0f  p99.9 %.0f  max %.0f" % (c.mean(), np.median(c), np.percentile(c, 99), np.percentile(c, 99.9), c.max()))
H = c.shape[0]
for y0 in range(0, H, H // 18):
    rows = c[y0:y0 + H // 18]
    print("rows %4d-%4d  mean %8.1f  max %8.0f  hit %.2f" % (y0, y0 + rows.shape[0] - 1, rows.mean(), rows.max(), a[y0:y0 + H // 18, :, 0].mean()))
# cost of the 32-ray groups (8x4 tiles): the warp runs as long as its slowest ray
W = c.shape[1]
t = c[:H // 4 * 4, :W // 8 * 8].reshape(H // 4, 4, W // 8, 8).max(axis=(1, 3))
print("per 8x4 tile: mean of max %.1f  max %.0f  -> lane efficiency bound %.2f" % (t.mean(), t.max(), c.mean() / t.mean()))
r.close()
