#!/bin/bash
# full captures of the C4 stage kernels (single chain of launches, steady state)
set -u
mkdir -p gpurun_out
CMD="python bench.py --workload c4 --steps 2 --warmup 3 --no-cpu-baseline"
for K in ${1:-stage_prepare stage_sat}; do
  XPBD_SUB_BATCHES=1 timeout 900 ncu --set full --clock-control none --import-source on -k regex:$K -s 465 -c 1 -f -o gpurun_out/prof_c4_$K $CMD > gpurun_out/ncu_full_c4_$K.log 2>&1
done
ls -la gpurun_out/prof_c4_*.ncu-rep
