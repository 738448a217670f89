#!/usr/bin/env python3
"""CPU: node visits per ray of the host twin of the traversal (oracle/orc_trav.c) on the shipped scenes (tree-quality experiments)."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import __graft_entry__ as ge
ge.load_package()
import numpy as np, orc
from rt_b200 import scenes
cases = [("rtiow_486", 192, 108, 4), ("blue2", 96, 54, 4), ("sample_scene", 96, 54, 4), ("dna_000", 96, 54, 4), ("sample_scene_ext", 96, 54, 4)]
for name, W, H, spp in cases:
    sc = orc.load_scene(f"{orc.GOLDEN}/{name}.json", width=W, height=H, spp=spp)
    b = orc.lbvh(sc)
    img, n, c = orc.render_bvh(sc, seed=1)
    print(name, "n_tree", b["n_tree"], "giants", b["n_giants"], "depth", b["depth"], "nodes/ray %.2f" % (c["n_node_visits"] / n),
          "leaf tests/ray %.2f" % (sum(c[k] for k in c if k.startswith("n_") and k not in ("n_box_tests", "n_node_visits")) / n), "maxstack", c["max_stack"])
ts = scenes.mesh_scene(120, 120, width=96, height=54, spp=2, seed=5, max_depth=8)
b = orc.lbvh(ts); img, n, c = orc.render_bvh(ts, seed=1)
print("mesh120", "n_tree", b["n_tree"], "giants", b["n_giants"], "depth", b["depth"], "nodes/ray %.2f" % (c["n_node_visits"] / n), "tri/ray %.2f" % (c["n_tri_tests"] / n))
