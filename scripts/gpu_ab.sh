#!/bin/bash
# A/B runs on the GPU box: C3 bench (stage times) for the main library and for every variant in
# constraint_solver_b200/lib/variants/*.so and every env setting in $AB_ENVS (';'-separated), then the gpu tests.
set -u
mkdir -p gpurun_out
L=constraint_solver_b200/lib
run() { XPBD_WORLD_PATH=staged timeout 300 python bench.py --workload ${AB_WORKLOAD:-c3} --steps 5 --warmup 3 --no-cpu-baseline 2>&1 | python -c "
import sys,json
for l in sys.stdin:
    try: d=json.loads(l); print('  value %.4g  ms %.2f  cons %d stage_ms %s'%(d['value'], d['ms_per_step'], d['constraints_last_step'], d['roofline'].get('stage_ms')))
    except Exception: print(l.rstrip()[:300])
"; }
echo "== main"; run
IFS=';' read -ra ENVS <<< "${AB_ENVS:-}"
for e in "${ENVS[@]}"; do [ -n "$e" ] && { echo "== env $e"; env $e bash -c "$(declare -f run); run"; }; done
cp $L/libxpbd_b200.so /tmp/keep.so
for v in $L/variants/*.so; do
  [ -f "$v" ] || continue
  echo "== variant $(basename $v .so) ${AB_VARIANT_ENV:-}"; cp "$v" $L/libxpbd_b200.so; env ${AB_VARIANT_ENV:-} bash -c "$(declare -f run); run"
done
cp /tmp/keep.so $L/libxpbd_b200.so
if [ "${AB_TESTS:-1}" = 1 ]; then
  ( timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "exit $?" >> gpurun_out/pytest_gpu.log )
  tail -6 gpurun_out/pytest_gpu.log
fi
