#!/bin/bash
# Run-time schedule sweep of the staged pipeline on C4 (environment overrides read by the library at every step call).
set -u
mkdir -p gpurun_out
run() {
  python bench.py --workload c4 --steps 30 --warmup 3 --no-cpu-baseline --no-extras > gpurun_out/bench_sweep.json 2> gpurun_out/bench_sweep.err || tail -3 gpurun_out/bench_sweep.err
  python - "$1" <<'PY'
import json,sys
d=json.loads(open("gpurun_out/bench_sweep.json").read().strip().splitlines()[-1])
print(sys.argv[1], "value %.4g e2e %.4g"%(d["value"], d["e2e"]["value"]), flush=True)
PY
}
run default
for S in 3 4; do XPBD_SAT_BLOCKS_PER_SM=$S run "sat_blocks=$S"; done
for K in 1 2 8; do XPBD_SUB_BATCHES=$K run "sub_batches=$K"; done
XPBD_SUB_BATCHES=2 XPBD_SAT_BLOCKS_PER_SM=4 run "sub_batches=2 sat_blocks=4"
XPBD_SUB_BATCHES=8 XPBD_SAT_BLOCKS_PER_SM=1 run "sub_batches=8 sat_blocks=1"
for B in 4 8; do XPBD_SOLVE_BLOCKS_PER_SM=$B run "solve_blocks=$B"; done
