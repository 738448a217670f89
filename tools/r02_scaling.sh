#!/bin/bash
# Per-scene scaling lines (BASELINE.json configs + the headline) at N GPUs of this box: tools/r02_scaling.sh N  ->  gpurun_out/scaling_N.jsonl
# N = 1 runs bench.py directly, N > 1 through torchrun (one rank per GPU; the library's NCCL gather, bit-identical image hash).
N=${1:-1}
mkdir -p gpurun_out; out=gpurun_out/scaling_$N.jsonl; [ -n "$2" ] && out=gpurun_out/scaling_subset_$N.jsonl; : > $out
port=29600
run() {  # workload steps warmup
  port=$((port + 1))
  if [ "$N" = "1" ]; then timeout 600 python bench.py --workload $1 --steps $2 --warmup $3 --no-cpu-baseline >> $out 2>> gpurun_out/scaling_$N.err
  else timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $port bench.py --gpus $N --workload $1 --steps $2 --warmup $3 --no-cpu-baseline >> $out 2>> gpurun_out/scaling_$N.err; fi
}
if [ -n "$2" ]; then shift; for w in "$@"; do run $w 3 3; done      # a subset: tools/r02_scaling.sh N workload ...
else
run rtiow_1080p_1024spp 3 3
run rtiow_c1_1200x800_10spp 5 3
run dna000_1080p_1024spp 3 3
run sample_scene_1600x900_4096spp 2 3
run sample_scene_ext_1600x900_4096spp 2 3
run mesh1m_4k_1024spp 1 1
run blue2_anim_240f_1080p_256spp 1 1
fi
python - <<PY
import json
for l in open("$out"):
    try: d = json.loads(l)
    except Exception: continue
    print(d["n_gpus"], d["config"]["workload"], "ms", round(d["ms_per_step"], 2), "e2e", round(d["e2e"].get("ms_per_step", 0), 2), "Gs/s", round(d["value"] / 1e9, 3),
          "Mrays/s", round(d.get("mrays_per_s", 0)), d.get("image_check"), (d.get("clocks") or {}).get("reasons"))
PY
grep -v "OMP_NUM_THREADS\|^\*\*\*\|^$" gpurun_out/scaling_$N.err | tail -3
