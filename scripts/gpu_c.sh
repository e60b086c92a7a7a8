#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 300 python bench.py --workload c3 --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/plain_c3.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum,smsp__inst_executed.sum,smsp__issue_active.avg.pct_of_peak_sustained_active,sm__warps_active.avg.pct_of_peak_sustained_active,smsp__thread_inst_executed_per_inst_executed.ratio --clock-control none -s 2660 -c 8 --csv --log-file gpurun_out/launches_c3_staged.csv \
   python bench.py --workload c3 --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_c3.log 2>&1
tail -c 300 gpurun_out/plain_c3.log
