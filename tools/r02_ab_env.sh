#!/bin/bash
# A/B of one environment switch: tools/r02_ab_env.sh VAR [workloads...]   (stage times per 128 spp, switch unset vs set to 1)
VAR=$1; shift
for wl in ${@:-rtiow_1080p_1024spp blue2_1080p_256spp sample_scene_1600x900_512spp}; do
  for on in "" 1; do
    echo "$wl $VAR=${on:-unset} $(env ${on:+$VAR=1} timeout 120 python tools/stage_bench.py $wl 128 3 | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print({k:round(d[k],2) for k in ('ms_generate','ms_extend','ms_shade','ms_tail','ms_sum','ms_render_overlapped')})")"
  done
done
