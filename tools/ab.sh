#!/bin/bash
# A/B of render flags on the same box: tools/ab.sh workload spp flagsA flagsB ...
wl=$1; spp=$2; shift 2
for f in "$@"; do
  echo "flags=$f"; python tools/stage_bench.py $wl $spp 3 $f | python -c "import json,sys; d=json.loads(sys.stdin.read()); print({k:round(d[k],2) for k in ('ms_generate','ms_extend','ms_shade','ms_sum','ms_render_overlapped','ms_render_serial')})"
done
