#!/bin/bash
# Evidence run: launch list of the bench command + one full ncu capture per kernel in steady state (graphs off so that the
# chains are plain launches ncu can address with -k / -s; C3 at frame 31 of the bench scene, C4 at frame 30 = the peak of the
# landing pile).  usage: bash scripts/gpu_profile_all.sh [tag] [c2]   (default tag r02; c2: also the single-body kernel)
set -u
TAG=${1:-r02}
mkdir -p gpurun_out
export XPBD_GRAPH=0
CMD="python bench.py --workload c3 --steps 2 --warmup 3 --no-cpu-baseline --no-extras"
timeout 300 $CMD > gpurun_out/plain_c3.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -s 9920 -c 640 --csv --log-file gpurun_out/${TAG}_launches_c3_bench.csv $CMD > gpurun_out/ncu_launches.log 2>&1
for K in stage_sat stage_solve stage_prepare stage_manifold; do
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:$K -s 2500 -c 1 -f -o gpurun_out/${TAG}_c3_$K $CMD > gpurun_out/ncu_full_$K.log 2>&1; echo "c3 $K ncu exit $?"
done
CMD4="python bench.py --workload c4 --steps 12 --warmup 3 --no-cpu-baseline --no-extras"
timeout 300 $CMD4 > gpurun_out/plain_c4.log 2>&1
for K in stage_sat stage_prepare stage_solve stage_manifold; do
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:$K -s 1810 -c 1 -f -o gpurun_out/${TAG}_c4_$K $CMD4 > gpurun_out/ncu_full_c4_$K.log 2>&1; echo "c4 $K ncu exit $?"
done
if [ "${2:-}" = "c2" ]; then
CMD2="python bench.py --workload c2 --steps 3 --warmup 3 --no-cpu-baseline --no-extras --worlds 262144"
timeout 300 $CMD2 > gpurun_out/plain_c2.log 2>&1 && \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:step_single -s 4 -c 1 -f -o gpurun_out/${TAG}_c2_step_single $CMD2 > gpurun_out/ncu_full_c2.log 2>&1; echo "c2 ncu exit $?"
fi
ls -la gpurun_out/*.ncu-rep
