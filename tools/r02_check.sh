#!/bin/bash
# Round-2 GPU visit: parity tests, then per-stage times of the main workloads.
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -15 gpurun_out/pytest_gpu.log
for wl in rtiow_1080p_1024spp blue2_1080p_256spp sample_scene_1600x900_512spp dna000_1080p_1024spp; do
  python tools/stage_bench.py $wl 64 3 2>&1 | tail -1 | tee -a gpurun_out/stage.jsonl
done
python tools/stage_bench.py mesh1m_1080p_64spp 16 2 2>&1 | tail -1 | tee -a gpurun_out/stage.jsonl
