#!/bin/bash
# A/B of library variants on the mesh: stage times per 16 spp at 1080p
timeout 600 python -m pytest tests -m gpu -x -q -k "mesh or triangle or lbvh or counters" 2>&1 | tail -2
for v in "" $VARIANTS; do
  echo "lib=${v:-default} $(RT_B200_LIB=$v python tools/stage_bench.py mesh1m_1080p_64spp 16 3 | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print({k:round(d[k],2) for k in ('ms_generate','ms_extend','ms_shade','ms_tail','ms_sum','ms_render_overlapped')})")"
done
