#!/bin/bash
# gpu tests, then C3 / C4 bench lines for every env setting in $AB_ENVS (';'-separated; first = default)
set -u
mkdir -p gpurun_out
if [ "${AB_TESTS:-1}" = 1 ]; then
( timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "exit $?" >> gpurun_out/pytest_gpu.log )
tail -6 gpurun_out/pytest_gpu.log
fi
IFS=';' read -ra ENVS <<< "${AB_ENVS:-X=1}"
for wl in ${AB_WORKLOADS:-c3 c4}; do
  i=0
  for e in "${ENVS[@]}"; do
    env $e timeout 600 python bench.py --workload $wl --steps ${AB_STEPS:-20} --warmup 3 --no-cpu-baseline > gpurun_out/bench_${wl}_v$i.json 2> gpurun_out/bench_${wl}_v$i.err
    python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/bench_${wl}_v$i.json").read().strip().splitlines()[-1])
    print("$wl [$e] value %.4g e2e %.4g ms %.3f cons %d stage_ms %s" % (d["value"], d["e2e"]["value"], d["ms_per_step"], d["constraints_last_step"], {k: round(v,4) for k,v in d["roofline"].get("stage_ms",{}).items()}))
except Exception as ex:
    print("$wl [$e] failed", ex); print(open("gpurun_out/bench_${wl}_v$i.err").read()[-2000:])
PY
    i=$((i+1))
  done
done
