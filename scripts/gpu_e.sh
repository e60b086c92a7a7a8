#!/bin/bash
# gpu tests, then C3 / C4 bench lines (automatic sub-batches vs a single chain of launches)
set -u
mkdir -p gpurun_out
( timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "exit $?" >> gpurun_out/pytest_gpu.log )
tail -6 gpurun_out/pytest_gpu.log
for wl in c3 c4; do
  for k in 0 1; do
    XPBD_SUB_BATCHES=$k timeout 600 python bench.py --workload $wl --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/bench_${wl}_k$k.json 2> gpurun_out/bench_${wl}_k$k.err
    python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/bench_${wl}_k$k.json").read().strip().splitlines()[-1])
    print("$wl k=$k value %.4g e2e %.4g ms %.3f stage_ms %s launches %d" % (d["value"], d["e2e"]["value"], d["ms_per_step"], d["roofline"].get("stage_ms"), d["gpu_launches"]))
except Exception as e:
    print("$wl k=$k failed", e); print(open("gpurun_out/bench_${wl}_k$k.err").read()[-2000:])
PY
  done
done
