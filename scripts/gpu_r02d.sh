#!/bin/bash
# Round 2, call d: CUDA graph per world range (A/B via XPBD_GRAPH), rigid-record e2e, comparable bench legs.
set -u
mkdir -p gpurun_out
( timeout 1200 python -m pytest tests/test_parity_world.py tests/test_parity_single.py -m gpu -x -q -k "not full_size" > gpurun_out/pytest_world.log 2>&1; echo "exit $?" >> gpurun_out/pytest_world.log ); tail -4 gpurun_out/pytest_world.log
for G in 0 1; do
  XPBD_GRAPH=$G python bench.py --steps 30 --warmup 3 --no-cpu-baseline > gpurun_out/bench_graph$G.json 2> gpurun_out/bench_graph$G.err; echo "bench graph=$G exit $?"; tail -2 gpurun_out/bench_graph$G.err
done
python - <<'PY'
import json
for g in (0,1):
    d=json.loads(open(f"gpurun_out/bench_graph{g}.json").read().strip().splitlines()[-1])
    print("graph",g,"value %.4g e2e %.4g e2e64 %.4g"%(d["value"], d["e2e"]["value"], d["e2e_rigid_f64"]["value"]), "launches", d["gpu_launches"])
    for k,v in d.get("other_workloads",{}).items(): print("   ",k, {kk:vv for kk,vv in v.items() if kk in("value","ms_per_step","failed")}, v.get("e2e",{}).get("value"))
PY
