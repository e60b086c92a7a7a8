#!/bin/bash
set -u
mkdir -p gpurun_out
K=${1:-stage_sat}
timeout 300 python bench.py --workload c3 --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/plain_c3.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:$K -s 665 -c 1 -f -o gpurun_out/prof_$K \
   python bench.py --workload c3 --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_full_$K.log 2>&1
ls -la gpurun_out/prof_$K.ncu-rep
