#!/bin/bash
# One bench line per named workload (BASELINE.json configs) -> gpurun_out/workloads.jsonl
mkdir -p gpurun_out; : > gpurun_out/workloads.jsonl
for wl in rtiow_c1_1200x800_10spp basic_scene_1080p_1024spp dna000_1080p_1024spp blue2_1080p_256spp sample_scene_1600x900_512spp sample_scene_ext_1600x900_512spp; do
  python bench.py --workload $wl --steps 3 --warmup 3 --no-cpu-baseline >> gpurun_out/workloads.jsonl 2>> gpurun_out/workloads.err
done
python bench.py --workload mesh1m_4k_1024spp --steps 1 --warmup 1 --no-cpu-baseline >> gpurun_out/workloads.jsonl 2>> gpurun_out/workloads.err
python bench.py --workload blue2_anim_16f_1080p_64spp --steps 1 --warmup 1 --no-cpu-baseline >> gpurun_out/workloads.jsonl 2>> gpurun_out/workloads.err
python - <<'PY'
import json
for l in open('gpurun_out/workloads.jsonl'):
    try: d = json.loads(l)
    except Exception: continue
    print(d['config']['workload'], 'ms', round(d['ms_per_step'], 2), 'Gsamples/s', round(d['value'] / 1e9, 3), 'Mrays/s', round(d.get('mrays_per_s', 0)), 'e2e ms', round(d['e2e'].get('ms_per_step', 0), 2))
PY
tail -3 gpurun_out/workloads.err
