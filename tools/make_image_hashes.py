#!/usr/bin/env python3
"""Collects the image_sha256 of every bench line under gpurun_out/ (workloads.jsonl, bench*.json) into tests/golden/image_hashes.json —
the committed per-workload hashes bench.py checks its end-to-end image against (the render is bit-reproducible on any GPU count).
Refuses to overwrite a hash with a different one unless --force (a changed image must be a decision, not an accident)."""
import glob, json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
path = os.path.join(ROOT, "tests", "golden", "image_hashes.json")
known = json.load(open(path)) if os.path.exists(path) else {}
force = "--force" in sys.argv
for f in sorted(glob.glob(os.path.join(ROOT, "gpurun_out", "*.json*"))):
    for line in open(f):
        try:
            d = json.loads(line)
        except Exception:
            continue
        if not isinstance(d, dict) or "image_sha256" not in d:
            continue
        wl, h = d["config"]["workload"], d["image_sha256"]
        if wl in known and known[wl] != h and not force:
            print(f"{wl}: {f} has {h[:16]}, committed {known[wl][:16]} — kept the committed one (use --force)")
            continue
        known[wl] = h
json.dump(dict(sorted(known.items())), open(path, "w"), indent=1)
print(json.dumps(known, indent=1))
