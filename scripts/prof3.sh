#!/bin/bash
# DRAM traffic of the store kernels at the bench's full table size (few metrics: one replay pass).  usage: scripts/prof3.sh <tag>
TAG=${1:-x}
timeout 500 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_sector_hit_rate.pct,smsp__inst_executed.sum,sm__warps_active.avg.pct_of_peak_sustained_active \
  --clock-control none -k regex:'records_kernel' -s 6 -c 4 --csv --log-file gpurun_out/full_table_$TAG.csv \
  python bench.py --steps 2 --warmup 3 --no-cpu-baseline --skip-e2e > gpurun_out/ncu_full_table_$TAG.log 2>&1
tail -3 gpurun_out/ncu_full_table_$TAG.log; tail -30 gpurun_out/full_table_$TAG.csv | cut -c1-250
