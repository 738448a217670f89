#!/bin/bash
# one full ncu capture of the launches matching a kernel regex: tools/r02_ncu_one.sh <regex> <count> <name> [workload] [spp]
mkdir -p gpurun_out
ncu --set full --clock-control none --import-source on -k regex:"$1" -c ${2:-3} -o gpurun_out/prof_$3 -f python tools/profile_run.py ${4:-rtiow_1080p_1024spp} 1 ${5:-32} > gpurun_out/ncu_$3.log 2>&1
tail -1 gpurun_out/ncu_$3.log
