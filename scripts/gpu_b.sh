#!/bin/bash
set -u
mkdir -p gpurun_out
( timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "exit $?" >> gpurun_out/pytest_gpu.log )
tail -15 gpurun_out/pytest_gpu.log
for cfg in "16 8 2" "16 12 2" "16 16 2" "32 16 1" "8 8 4" "16 6 3"; do
  set -- $cfg
  echo "NW=$1 WARPS=$2 BPS=$3"
  XPBD_WORLD_NW=$1 XPBD_WORLD_WARPS=$2 XPBD_WORLD_BLOCKS_PER_SM=$3 timeout 300 python bench.py --workload c3 --steps 5 --warmup 3 --no-cpu-baseline 2>&1 | python -c "
import sys,json
for l in sys.stdin:
    try: d=json.loads(l); print('  value %.4g  ms %.2f  cons %d'%(d['value'], d['ms_per_step'], d['constraints_last_step']))
    except Exception: print(l.rstrip()[:300])
"
done
