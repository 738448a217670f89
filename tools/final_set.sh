#!/bin/bash
# The round's closing measurement set on one GPU box: parity tests (PSNR lines kept), smoke, headline bench + reference arm,
# every named workload, the same-box comparison with the reference's GPU program, ncu launch list + full capture.
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q -s -k converged > gpurun_out/psnr.log 2>&1; grep "psnr=" gpurun_out/psnr.log
timeout 1400 bash tools/gpu_check.sh > gpurun_out/gpu_check.log 2>&1; tail -3 gpurun_out/pytest_gpu.log; tail -1 gpurun_out/smoke.log
python -c "import json; d=json.load(open('gpurun_out/bench.json')); print('bench', d['ms_per_step'], d['e2e']['ms_per_step'], d['roofline']['frac'], d['roofline_hbm']['frac'], d['clocks'])"
timeout 900 bash tools/all_workloads.sh 2>&1 | tail -10
timeout 300 python tools/ref_gpu_compare.py > gpurun_out/ref_gpu_compare.log 2>&1
python -c "
import json
for c in json.load(open('gpurun_out/ref_gpu_compare.json'))['cases']:
    print(c['scene'], 'reference GPU Msamples/s', round((c.get('reference_samples_per_s') or 0)/1e6,1), 'ours', round(c['ours_samples_per_s']/1e6,1), 'x', round(c.get('speedup_render',0),1))
"
