#!/bin/bash
# A/B of library variants: stage times of the given workloads
for wl in ${WLS:-mesh1m_1080p_64spp}; do
for v in "" $VARIANTS; do
  echo "$wl lib=${v:-default} $(RT_B200_LIB=$v timeout 300 python tools/stage_bench.py $wl ${SPP:-16} 3 | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print({k:round(d[k],2) for k in ('ms_generate','ms_extend','ms_shade','ms_tail','ms_sum','ms_render_overlapped')})")"
done; done
