#!/bin/bash
# ncu evidence of one tree: full capture of the enumerate and store kernels on the perf.py workload (cfg3 shape, 10 Mbp
# genome so that the replays stay cheap), then the launch list of the bench command.  usage: scripts/prof_round.sh <tag>
TAG=${1:-x}
python scripts/perf.py cfg3 > gpurun_out/perf_$TAG.log 2>&1 || exit 1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:'enumerate_kernel|records_kernel' -s 2 -c 3 -o gpurun_out/prof_$TAG -f python scripts/perf.py cfg3 > gpurun_out/ncu_$TAG.log 2>&1
tail -2 gpurun_out/ncu_$TAG.log
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --skip-e2e --no-parity --genome 10000000"
$CMD > gpurun_out/prof_plain_$TAG.json 2> gpurun_out/prof_plain_$TAG.err || exit 1
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_$TAG.csv $CMD > gpurun_out/ncu_launch_$TAG.log 2>&1
tail -1 gpurun_out/ncu_launch_$TAG.log | cut -c1-300
