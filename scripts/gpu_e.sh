#!/bin/bash
set -u
mkdir -p gpurun_out
L=constraint_solver_b200/lib
run() { XPBD_WORLD_PATH=staged timeout 300 python bench.py --workload c3 --steps 5 --warmup 3 --no-cpu-baseline 2>&1 | python -c "
import sys,json
for l in sys.stdin:
    try: d=json.loads(l); print('  value %.4g  ms %.2f  cons %d stage_ms %s'%(d['value'], d['ms_per_step'], d['constraints_last_step'], d['roofline'].get('stage_ms')))
    except Exception: print(l.rstrip()[:300])
"; }
echo "min blocks 3"; run
cp $L/libxpbd_b200.so /tmp/keep.so; cp $L/libxpbd_b200_mb4.so $L/libxpbd_b200.so
echo "min blocks 4"; run
cp /tmp/keep.so $L/libxpbd_b200.so
( timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "exit $?" >> gpurun_out/pytest_gpu.log )
tail -4 gpurun_out/pytest_gpu.log
