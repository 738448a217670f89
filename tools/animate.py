#!/usr/bin/env python3
"""Animation driver: the reference's dna.py / blue.py / blue2.py in one process per GPU, frames dealt round-robin.
  python tools/animate.py blue2 [--frames N] [--spp S] [--width W --height H] [--out DIR] [--gpus G]
With --gpus G > 1 start it under torchrun (one rank per GPU); there is no collective — ranks only share the frame rule."""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as ge  # noqa: E402

rt = ge.load_package()
from rt_b200 import animation, multi_gpu  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("name", choices=sorted(animation.FRAME_RULES))
ap.add_argument("--frames", type=int, default=0)
ap.add_argument("--spp", type=int, default=0)
ap.add_argument("--width", type=int, default=0)
ap.add_argument("--height", type=int, default=0)
ap.add_argument("--out", default="")
ap.add_argument("--scene-dir", default=os.path.join(ROOT, "tests", "golden"))
args = ap.parse_args()
rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
base_file, fn, n_ref = animation.FRAME_RULES[args.name]
base = json.load(open(os.path.join(args.scene_dir, base_file)))
frames = multi_gpu.frames_for_rank(args.frames or n_ref, rank, world)
if args.out:
    os.makedirs(args.out, exist_ok=True)
r = rt.Renderer(local)
res = animation.render_animation(r, base, fn, frames, args.out or None, args.width, args.height, args.spp, base_dir=args.scene_dir)
res.update(rank=rank, world=world, samples_per_s=res["samples"] / max(res["seconds"], 1e-9))
print(json.dumps(res))
r.close()
