#!/bin/bash
# Round 2, call c: new bench.py (all keys), step_host_rigid tests, reference arm.
set -u
mkdir -p gpurun_out
( timeout 900 python -m pytest tests -m gpu -x -q -k "host_rigid or nan_world or default_selector or dropins or solver_step or world_integrate" > gpurun_out/pytest_new.log 2>&1; echo "exit $?" >> gpurun_out/pytest_new.log ); tail -5 gpurun_out/pytest_new.log
python bench.py --steps 30 --warmup 3 > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err; echo "bench exit $?"; tail -3 gpurun_out/bench_default.err
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref exit $?"
python - <<'PY'
import json
d=json.loads(open("gpurun_out/bench_default.json").read().strip().splitlines()[-1])
print("value %.4g e2e %.4g e2e64 %.4g"%(d["value"], d["e2e"]["value"], d["e2e_rigid_f64"]["value"]))
print("roofline", {k:v for k,v in d["roofline"].items() if k in("bound","kernel","achieved","peak","frac","traffic","kernel_ms","stage_ms")})
print("spec", d["roofline"].get("specification",{}).get("flops_per_body_substep"), d["roofline"].get("specification",{}).get("equivalent_tflops_per_gpu"))
print("per_rank", d["per_rank"])
for k,v in d.get("other_workloads",{}).items(): print(k, {kk:vv for kk,vv in v.items() if kk in("value","ms_per_step","failed")}, v.get("e2e",{}).get("value"))
print("cpu", d.get("cpu_baseline"))
PY
