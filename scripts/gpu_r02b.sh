#!/bin/bash
# Round 2, call b: flattened exact pass + manifold fused into the SAT thread kernel (A/B via XPBD_SAT_FUSE).
set -u
mkdir -p gpurun_out
( timeout 900 python -m pytest tests/test_sat_fuzz.py -m gpu -x -q -k "cube_pairs and origin" > gpurun_out/pytest_fuzz.log 2>&1; echo "exit $?" >> gpurun_out/pytest_fuzz.log ); tail -3 gpurun_out/pytest_fuzz.log
( timeout 900 python -m pytest tests/test_parity_world.py -m gpu -x -q -k "not full_size" > gpurun_out/pytest_world.log 2>&1; echo "exit $?" >> gpurun_out/pytest_world.log ); tail -3 gpurun_out/pytest_world.log
for FUSE in 0 1; do
  XPBD_SAT_FUSE=$FUSE python bench.py --workload c3 --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/bench_c3_fuse$FUSE.json 2> gpurun_out/bench_c3_fuse$FUSE.err; echo "bench fuse=$FUSE exit $?"
done
python - <<'PY'
import json
for f in ("fuse0","fuse1"):
    try:
        d=json.loads(open(f"gpurun_out/bench_c3_{f}.json").read().strip().splitlines()[-1])
        print(f, "value %.4g"%d["value"], "e2e %.4g"%d["e2e"]["value"], "stage_ms", d["roofline"].get("stage_ms"))
    except Exception as e: print(f, "failed", e)
PY
CMD="python bench.py --workload c3 --steps 2 --warmup 3 --no-cpu-baseline"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:stage_sat -s 300 -c 1 -f -o gpurun_out/prof_r02b_stage_sat_thread_fused $CMD > gpurun_out/ncu_full_sat.log 2>&1; echo "ncu exit $?"
