#!/bin/bash
set -u
mkdir -p gpurun_out
( timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "exit $?" >> gpurun_out/pytest_gpu.log )
tail -15 gpurun_out/pytest_gpu.log
for path in staged fused; do
  echo "PATH=$path"
  XPBD_WORLD_PATH=$path timeout 300 python bench.py --workload c3 --steps 5 --warmup 3 --no-cpu-baseline 2>&1 | python -c "
import sys,json
for l in sys.stdin:
    try: d=json.loads(l); print('  value %.4g  ms %.2f  cons %d launches %d'%(d['value'], d['ms_per_step'], d['constraints_last_step'], d['gpu_launches']))
    except Exception: print(l.rstrip()[:300])
"
done
