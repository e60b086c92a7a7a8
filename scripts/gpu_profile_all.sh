#!/bin/bash
# Round-end evidence: launch list of the bench command + one full capture per staged kernel (steady state).
set -u
mkdir -p gpurun_out
CMD="python bench.py --workload c3 --steps 2 --warmup 3 --no-cpu-baseline"
timeout 300 $CMD > gpurun_out/plain_c3.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -s 2400 -c 400 --csv --log-file gpurun_out/launches_c3_full.csv $CMD > gpurun_out/ncu_launches.log 2>&1
for K in stage_sat stage_solve stage_prepare stage_manifold; do
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:$K -s 665 -c 1 -f -o gpurun_out/prof_$K $CMD > gpurun_out/ncu_full_$K.log 2>&1
done
CMD2="python bench.py --workload c2 --steps 3 --warmup 3 --no-cpu-baseline --worlds 262144"
timeout 300 $CMD2 > gpurun_out/plain_c2.log 2>&1 && \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:step_single -s 4 -c 1 -f -o gpurun_out/prof_step_single $CMD2 > gpurun_out/ncu_full_c2.log 2>&1
ls -la gpurun_out/*.ncu-rep
