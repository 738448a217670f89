#!/usr/bin/env python3
"""Plain estimator vs next-event estimation on the extended sample scene (CPU sampling convention): time per render and
noise (RMSE between two seeds) at equal spp. Usage: nee_bench.py [spp] [width] [height]"""
import json, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as ge
rt = ge.load_package()
from rt_b200 import capi
spp = int(sys.argv[1]) if len(sys.argv) > 1 else 64
W = int(sys.argv[2]) if len(sys.argv) > 2 else 1600
H = int(sys.argv[3]) if len(sys.argv) > 3 else 900
G = os.path.join(ROOT, "tests", "golden")
d = json.load(open(os.path.join(G, "sample_scene_ext.json")))
d["sampling"] = "reference_cpu"; d["background"] = [0.02, 0.02, 0.03]
r = rt.Renderer(0)
r.upload_json_text(json.dumps(d), base_dir=G)
r.override(width=W, height=H, spp=spp)
out = {}
for name, fl in (("plain", 0), ("nee", capi.RT_RENDER_NEE)):
    imgs = []
    for seed in (1, 2, 1, 2):
        r.stats_reset(); r.render(seed=seed, flags=fl); img, _, st = r.download(); imgs.append(img / spp)
    out[name] = {"ms": round(st["ms_total"], 2), "rays": st["rays"], "shadow_rays": st["shadow_rays"],
                 "rmse_between_seeds": float(np.sqrt(np.mean((imgs[2] - imgs[3]) ** 2))), "mean": float(imgs[2].mean())}
for name, fl in (("plain", 0), ("nee", capi.RT_RENDER_NEE)):
    r.stats_reset(); r.render(seed=1, flags=fl | capi.RT_RENDER_TIME_STAGES); _, _, st = r.download(want_sum=False, want_rgb8=False)
    out[name]["stages_ms"] = {k: round(st["ms_" + k], 2) for k in ("generate", "extend", "shade", "shadow", "accumulate")}
print(json.dumps(out))
r.close()
