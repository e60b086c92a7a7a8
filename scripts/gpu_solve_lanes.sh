#!/bin/bash
# The wide schedule of the solve stage (stage_solve_wide_kernel<LW>): parity tests at every lane count, then C4 at each.
set -u
mkdir -p gpurun_out
for L in 8 4 16 32; do
  ( XPBD_SOLVE_LANES=$L timeout 900 python -m pytest tests/test_parity_world.py -m gpu -x -q -k "not full_size" > gpurun_out/pytest_lanes_$L.log 2>&1; echo "exit $?" >> gpurun_out/pytest_lanes_$L.log ); echo "lanes $L: $(tail -2 gpurun_out/pytest_lanes_$L.log | tr '\n' ' ')"
done
for L in 2 4 8 16 32; do
  XPBD_SOLVE_LANES=$L python bench.py --workload c4 --steps 30 --warmup 3 --no-cpu-baseline --no-extras > gpurun_out/bench_lanes.json 2> gpurun_out/bench_lanes.err || tail -3 gpurun_out/bench_lanes.err
  python - $L <<'PY'
import json,sys
d=json.loads(open("gpurun_out/bench_lanes.json").read().strip().splitlines()[-1])
print("c4 lanes", sys.argv[1], "value %.4g e2e %.4g"%(d["value"], d["e2e"]["value"]), {k:round(v,4) for k,v in d["roofline"]["stage_ms"].items()}, "parity", d["per_rank"][0].get("parity_sample"), flush=True)
PY
done
