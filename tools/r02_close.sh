#!/bin/bash
# Closing numbers of round 2 with the final build: default bench line, every named workload, launch list of the headline scene.
mkdir -p gpurun_out
python bench.py > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err; echo "bench rc=$?"; python -c "
import json; d=json.loads(open('gpurun_out/bench_default.json').read().strip().splitlines()[-1]); print(d['ms_per_step'], d['e2e'], d['roofline']['frac'], d['roofline_hbm']['frac'], d['cpu_baseline'], d['image_check'], d['gpu_launches'])"
bash tools/all_workloads.sh
python tools/profile_run.py rtiow_1080p_1024spp 1 32 > gpurun_out/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv python tools/profile_run.py rtiow_1080p_1024spp 1 32 > gpurun_out/ncu1.log 2>&1
python tools/ncu_summary.py launches gpurun_out/launches.csv 2>/dev/null | tail -25
