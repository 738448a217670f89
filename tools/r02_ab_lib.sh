#!/bin/bash
# A/B of the working-tree library against build_variants/<name>.so: tools/r02_ab_lib.sh name [workloads...]  (stage times per 128 spp)
V=build_variants/$1.so; shift
for wl in ${@:-rtiow_1080p_1024spp blue2_1080p_256spp sample_scene_1600x900_512spp}; do
  for lib in "" $V; do
    echo "$wl lib=${lib:-working tree} $(RT_B200_LIB=$lib timeout 120 python tools/stage_bench.py $wl 128 3 | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print({k:round(d[k],2) for k in ('ms_generate','ms_extend','ms_shade','ms_tail','ms_sum','ms_render_overlapped')})")"
  done
done
