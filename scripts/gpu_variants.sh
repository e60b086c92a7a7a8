#!/bin/bash
# A/B of build variants (make variant NAME=... EXTRA=...): short C3 bench per library in constraint_solver_b200/lib/variants.
set -u
mkdir -p gpurun_out
for V in default "$@"; do
  if [ "$V" = default ]; then unset XPBD_B200_LIB; else export XPBD_B200_LIB=$PWD/constraint_solver_b200/lib/variants/libxpbd_b200_$V.so; fi
  python bench.py --steps 30 --warmup 3 --no-cpu-baseline --no-extras > gpurun_out/bench_var_$V.json 2> gpurun_out/bench_var_$V.err || tail -3 gpurun_out/bench_var_$V.err
  python - "$V" <<'PY'
import json,sys
v=sys.argv[1]
d=json.loads(open(f"gpurun_out/bench_var_{v}.json").read().strip().splitlines()[-1])
print(v, "value %.4g"%d["value"], "stage_ms", {k:round(x,4) for k,x in d["roofline"]["stage_ms"].items()})
PY
done
