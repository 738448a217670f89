#!/bin/bash
# A/B of an environment variable on the same box: tools/ab_env.sh VAR workload spp v1 v2 ...
var=$1; wl=$2; spp=$3; shift 3
for v in "$@"; do
  echo "$var=$v"
  env $var=$v python tools/stage_bench.py $wl $spp 2 | python -c "import json,sys; d=json.loads(sys.stdin.read()); print({k:round(d[k],2) for k in ('ms_generate','ms_extend','ms_shade','ms_sum','ms_render_overlapped','ms_render_serial')})"
done
