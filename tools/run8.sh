#!/bin/bash
# 8-GPU bench lines: headline, config 3 at 4096 spp (both scenes), the 1 M-triangle mesh, the 240-frame animation
mkdir -p gpurun_out
run() { timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port $1 bench.py --gpus 8 --steps $3 --warmup $4 --workload $2 --no-cpu-baseline > gpurun_out/$5 2>> gpurun_out/run8.err; tail -1 gpurun_out/$5 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['config']['workload'], 'ms', round(d['ms_per_step'],2), 'e2e', round(d['e2e'].get('ms_per_step',0),2), 'Gs/s', round(d['value']/1e9,2), 'Mrays/s', round(d.get('mrays_per_s',0)))"; }
run 29541 rtiow_1080p_1024spp 3 3 bench_8gpu.json
run 29542 sample_scene_1600x900_4096spp 3 3 c3_sample_scene_8gpu.json
run 29543 sample_scene_ext_1600x900_4096spp 3 3 c3_sample_scene_ext_8gpu.json
run 29544 mesh1m_4k_1024spp 1 1 mesh4k_8gpu.json
run 29545 blue2_anim_240f_1080p_256spp 1 1 anim240_8gpu.json
tail -3 gpurun_out/run8.err
