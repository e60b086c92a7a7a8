#!/bin/bash
# What the driver runs at round end: smoke(), the gpu tests, the default bench line and the reference arm.
set -u
mkdir -p gpurun_out
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke exit $?"; tail -5 gpurun_out/smoke.log
( timeout 2400 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "exit $?" >> gpurun_out/pytest_gpu.log ); tail -4 gpurun_out/pytest_gpu.log
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref exit $?"
python bench.py > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err; echo "bench exit $?"; tail -3 gpurun_out/bench_default.err
python - <<'PY'
import json
d=json.loads(open("gpurun_out/bench_default.json").read().strip().splitlines()[-1])
r=json.loads(open("gpurun_out/bench_ref.json").read().strip().splitlines()[-1])
print("value %.4g e2e %.4g e2e64 %.4g ref %.4g"%(d["value"], d["e2e"]["value"], d["e2e_rigid_f64"]["value"], r["value"]), "launches", d.get("gpu_launches"), "clocks", d.get("clocks"))
print("roofline", {k:v for k,v in d["roofline"].items() if k in("bound","kernel","achieved","peak","frac","traffic","kernel_ms","stage_ms")})
print("per_rank", d["per_rank"])
for k,v in d.get("other_workloads",{}).items(): print(k, {kk:vv for kk,vv in v.items() if kk in("value","ms_per_step","failed")}, "e2e", v.get("e2e",{}).get("value"))
print("cpu", d.get("cpu_baseline"))
PY
