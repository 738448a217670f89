#!/bin/bash
# ncu: launch list of one RTIOW render + one full capture of the first extend/shade launches (bounces 0-3) and of the tail launch
mkdir -p gpurun_out
WL=${1:-rtiow_1080p_1024spp}; SPP=${2:-32}
RT_TRACE=1 python tools/profile_run.py $WL 1 $SPP > gpurun_out/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv python tools/profile_run.py $WL 1 $SPP > gpurun_out/ncu1.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:'k_extend|k_shade|k_tail' -c 7 -o gpurun_out/prof_extend_shade -f python tools/profile_run.py $WL 1 $SPP > gpurun_out/ncu2.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:'k_tail|k_generate' -c 2 -o gpurun_out/prof_tail -f python tools/profile_run.py $WL 1 $SPP > gpurun_out/ncu3.log 2>&1
cat gpurun_out/plain.log | tail -4
