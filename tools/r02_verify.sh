#!/bin/bash
# What the driver runs at round end, in one visit: GPU tests, smoke, default bench, reference arm.
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q 2>&1 | tail -2
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
python bench.py > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err; echo "bench rc=$?"
python - <<'PY'
import json
d = json.loads(open("gpurun_out/bench_default.json").read().strip().splitlines()[-1])
print("ms/step", d["ms_per_step"], "e2e", d["e2e"]["ms_per_step"], "roofline", d["roofline"]["frac"], d["roofline_hbm"]["frac"], "image", d["image_check"],
      "launches", d["gpu_launches"], "cpu", d["cpu_baseline"]["value"], "clocks", d["clocks"]["sm_mhz"], d["clocks"]["reasons"])
PY
python bench.py --impl reference --steps 2 --warmup 1 2>/dev/null | tail -c 700
