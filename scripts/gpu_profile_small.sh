#!/bin/bash
# One full ncu capture per small staged kernel (steady state, C3).  KERNELS overrides the list.
set -u
mkdir -p gpurun_out
CMD="python bench.py --workload c3 --steps 2 --warmup 3 --no-cpu-baseline"
timeout 300 $CMD > gpurun_out/plain_c3.log 2>&1 || { tail -5 gpurun_out/plain_c3.log; exit 1; }
for K in ${KERNELS:-stage_solve stage_prepare stage_manifold}; do
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:$K -s 665 -c 1 -f -o gpurun_out/prof_$K $CMD > gpurun_out/ncu_full_$K.log 2>&1
done
ls -la gpurun_out/*.ncu-rep
