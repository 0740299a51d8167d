#!/bin/bash
# ncu evidence of the final tree at the bench's own size (cfg 3, 100 Mbp): the launch list of the bench command and one
# full capture of the dominant kernel (enumerate_kernel; its DRAM traffic does not depend on the table).
mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --skip-e2e --no-parity"
$CMD > gpurun_out/prof_final_plain.json 2> gpurun_out/prof_final_plain.err || exit 1
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches_final.csv $CMD > gpurun_out/ncu_launch_final.log 2>&1
tail -1 gpurun_out/ncu_launch_final.log | cut -c1-200
timeout 900 ncu --set full --clock-control none --import-source on -k regex:enumerate_kernel -s 6 -c 1 -o gpurun_out/prof_final_enum -f $CMD > gpurun_out/ncu_final_enum.log 2>&1
tail -2 gpurun_out/ncu_final_enum.log | cut -c1-200
ncu -i gpurun_out/prof_final_enum.ncu-rep --page raw --csv > gpurun_out/prof_final_enum_raw.csv 2>/dev/null
ls -la gpurun_out/prof_final_enum.ncu-rep
