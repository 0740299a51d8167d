#!/bin/bash
# quick ncu capture of one enumerate + one records launch on the perf.py workload.  usage: scripts/prof2.sh <tag> [cfg]
TAG=${1:-x}; CFG=${2:-cfg3}
timeout 600 ncu --set full --clock-control none --import-source on -k regex:'enumerate_kernel|records_kernel' -s 2 -c 2 -o gpurun_out/prof_$TAG -f python scripts/perf.py $CFG > gpurun_out/ncu_$TAG.log 2>&1
tail -2 gpurun_out/ncu_$TAG.log
