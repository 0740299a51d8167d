#!/bin/bash
# store kernels at the bench's FULL table size: duration, DRAM bytes, L2 atomic / reduction sectors, LSU atomic requests
# (a handful of metrics = few replay passes; the 20 GB table is saved / restored per pass).  usage: scripts/prof_store.sh <tag>
TAG=${1:-x}
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --skip-e2e --no-parity"
$CMD > gpurun_out/store_plain_$TAG.json 2> gpurun_out/store_plain_$TAG.err || exit 1
M=gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_sectors.sum,lts__t_sectors_op_atom.sum,lts__t_sectors_op_red.sum,lts__t_sector_hit_rate.pct,l1tex__t_set_accesses_pipe_lsu_mem_global_op_atom.sum,l1tex__t_set_accesses_pipe_lsu_mem_global_op_red.sum,l1tex__t_requests_pipe_lsu_mem_global_op_ld.sum,smsp__inst_executed.sum,smsp__issue_active.avg.pct_of_peak_sustained_active,sm__warps_active.avg.pct_of_peak_sustained_active
timeout 900 ncu --metrics $M --clock-control none -k regex:records_kernel -s 5 -c 4 --csv --log-file gpurun_out/store_full_$TAG.csv $CMD > gpurun_out/store_ncu_$TAG.log 2>&1
tail -2 gpurun_out/store_ncu_$TAG.log
