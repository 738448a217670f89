#!/bin/bash
# A/B of library variants: stage times of the given workloads
for wl in ${WLS:-rtiow_1080p_1024spp blue2_1080p_256spp sample_scene_1600x900_512spp}; do
for v in "" $VARIANTS; do
  echo "$wl lib=${v:-default} $(RT_B200_LIB=$v python tools/stage_bench.py $wl ${SPP:-128} 3 | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print({k:round(d[k],2) for k in ('ms_generate','ms_extend','ms_shade','ms_tail','ms_sum','ms_render_overlapped')})")"
done; done
