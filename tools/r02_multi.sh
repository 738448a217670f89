#!/bin/bash
# multi-GPU visit: group test, CLI --gpus, bench at N ranks (N = number of GPUs of the box)
N=$(python -c "import torch; print(torch.cuda.device_count())")
mkdir -p gpurun_out
python -m pytest tests -m gpu -q -k "group or shards" 2>&1 | tail -5
(cd gpurun_out && mkdir -p build/output && ../ray-tracing-in-cuda_b200/parallel_compute -f ../tests/golden/blue2.json --gpus $N --spp 64 --width 640 --height 360; ../ray-tracing-in-cuda_b200/parallel_compute -f ../tests/golden/blue2.json --spp 64 --width 640 --height 360 --no-ppm; rm -rf build main.ppm)
python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench_1gpu.json 2> gpurun_out/bench_1gpu.err; tail -c 600 gpurun_out/bench_1gpu.json; tail -3 gpurun_out/bench_1gpu.err
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus $N --steps 3 --warmup 3 > gpurun_out/bench_${N}gpu.json 2> gpurun_out/bench_${N}gpu.err; tail -c 600 gpurun_out/bench_${N}gpu.json; tail -5 gpurun_out/bench_${N}gpu.err
python - <<'PY'
import json,glob
for f in sorted(glob.glob("gpurun_out/bench_*gpu.json")):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1]); print(f, d["n_gpus"], round(d["ms_per_step"],2), round(d["e2e"]["ms_per_step"],2), d["image_sha256"][:16], d["image_check"])
    except Exception as e: print(f, "ERR", e)
PY
