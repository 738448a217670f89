#!/bin/bash
# A/B of library builds on the same box: tools/ab_lib.sh workload spp variantA variantB ...   ("default" = the in-tree library)
wl=$1; spp=$2; shift 2
for v in "$@"; do
  echo "lib=$v"
  if [ "$v" = default ]; then unset RT_B200_LIB; else export RT_B200_LIB=build_variants/$v.so; fi
  python tools/stage_bench.py $wl $spp 3 | python -c "import json,sys; d=json.loads(sys.stdin.read()); print({k:round(d[k],2) for k in ('ms_generate','ms_extend','ms_shade','ms_sum','ms_render_overlapped','ms_render_serial')})"
done
