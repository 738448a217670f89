#!/usr/bin/env python3
"""bench.py at several wavefront sizes (paths per pass and stream). Usage: sweep_paths.py [workload] sizes..."""
import json, os, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
wl = sys.argv[1]
for mp in sys.argv[2:]:
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--workload", wl, "--steps", "2", "--warmup", "2", "--no-cpu-baseline",
                          "--max-paths", mp], capture_output=True, text=True)
    try:
        d = json.loads(out.stdout.strip().splitlines()[-1])
        print(mp, "resident ms", round(d["ms_per_step"], 1), "e2e ms", round(d["e2e"]["ms_per_step"], 1), "launches", d["gpu_launches"])
    except Exception as e:
        print(mp, "failed", e, out.stderr[-300:])
