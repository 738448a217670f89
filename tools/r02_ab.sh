#!/bin/bash
# A/B of environment switches on one box: tools/r02_ab.sh "ENV1=.. ENV2=.." "ENVB=.." ...   (each argument = one variant's env)
mkdir -p gpurun_out
for v in "$@"; do
  for wl in rtiow_1080p_1024spp blue2_1080p_256spp sample_scene_1600x900_512spp; do
    echo "== [$v] $wl" | tee -a gpurun_out/ab.log
    env $v python tools/stage_bench.py $wl 64 3 2>&1 | tail -1 | python -c "import json,sys; d=json.loads(sys.stdin.read()); print({k:round(d[k],2) for k in ('ms_generate','ms_extend','ms_shade','ms_tail','launches','ms_sum','ms_render_overlapped','ms_render_serial')})" | tee -a gpurun_out/ab.log
  done
done
