#!/bin/bash
# One full ncu capture of the C3 SAT kernel in steady state (graphs off so that -k / -s address plain launches).
set -u
mkdir -p gpurun_out
NAME=${1:-r02_stage_sat_thread}
CMD="python bench.py --workload c3 --steps 2 --warmup 3 --no-cpu-baseline --no-extras"
XPBD_GRAPH=0 timeout 900 ncu --set full --clock-control none --import-source on -k regex:stage_sat -s ${2:-2800} -c 1 -f -o gpurun_out/prof_$NAME $CMD > gpurun_out/ncu_full_sat.log 2>&1; echo "ncu exit $?"
( timeout 300 python -m pytest tests/test_abi_harness.py -m gpu -x -q 2>&1 | tail -3 )
