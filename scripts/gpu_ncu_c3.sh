#!/bin/bash
# ncu --set full of the multi-body step kernel on a reduced C3 batch (ncu replays each launch ~40x)
set -u
mkdir -p gpurun_out
W=${1:-20000}
timeout 300 python bench.py --workload c3 --steps 3 --warmup 3 --no-cpu-baseline --worlds $W > gpurun_out/plain_c3s.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:step_world -s 4 -c 1 -f -o gpurun_out/prof_c3 \
   python bench.py --workload c3 --steps 3 --warmup 3 --no-cpu-baseline --worlds $W > gpurun_out/ncu_full_c3.log 2>&1
tail -2 gpurun_out/plain_c3s.log | cut -c1-200
ls -la gpurun_out/prof_c3.ncu-rep
