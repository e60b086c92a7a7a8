#!/bin/bash
# Development check: the SAT fuzz, the multi-body parity tests (without the full-size cases), a short C3 bench and a C4 bench.
set -u
mkdir -p gpurun_out
( timeout 1200 python -m pytest tests/test_sat_fuzz.py -m gpu -x -q > gpurun_out/pytest_fuzz.log 2>&1; echo "exit $?" >> gpurun_out/pytest_fuzz.log ); tail -4 gpurun_out/pytest_fuzz.log
( timeout 1200 python -m pytest tests/test_parity_world.py tests/test_parity_blocks.py tests/test_abi_harness.py -m gpu -x -q -k "not full_size" > gpurun_out/pytest_world.log 2>&1; echo "exit $?" >> gpurun_out/pytest_world.log ); tail -4 gpurun_out/pytest_world.log
for WL in c3 c4; do
python bench.py --workload $WL --steps 30 --warmup 3 --no-cpu-baseline --no-extras > gpurun_out/bench_$WL.json 2> gpurun_out/bench_$WL.err; echo "bench $WL exit $?"; tail -2 gpurun_out/bench_$WL.err
python - $WL <<'PY'
import json,sys
d=json.loads(open(f"gpurun_out/bench_{sys.argv[1]}.json").read().strip().splitlines()[-1])
print(sys.argv[1], "value %.4g e2e %.4g"%(d["value"], d["e2e"]["value"]), "stage_ms", {k:round(v,4) for k,v in d["roofline"].get("stage_ms").items()})
PY
done
