#!/bin/bash
# Round 2, first GPU call: new SAT schedule (thread per candidate) — fuzz A/B, full gpu suite, C3 bench per schedule, ncu.
set -u
mkdir -p gpurun_out
( timeout 900 python -m pytest tests/test_sat_fuzz.py -m gpu -x -q > gpurun_out/pytest_fuzz.log 2>&1; echo "exit $?" >> gpurun_out/pytest_fuzz.log ); tail -5 gpurun_out/pytest_fuzz.log
for SCHED in dual thread; do
  XPBD_SAT_SCHED=$SCHED python bench.py --workload c3 --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/bench_c3_$SCHED.json 2> gpurun_out/bench_c3_$SCHED.err; echo "bench $SCHED exit $?"
done
python - <<'PY'
import json
for f in ("dual","thread"):
    try:
        d=json.loads(open(f"gpurun_out/bench_c3_{f}.json").read().strip().splitlines()[-1])
        print(f, "value %.4g"%d["value"], "e2e %.4g"%d["e2e"]["value"], "stage_ms", d["roofline"].get("stage_ms"))
    except Exception as e: print(f, "failed", e)
PY
( timeout 1800 python -m pytest tests -m gpu -x -q --deselect tests/test_sat_fuzz.py > gpurun_out/pytest_gpu.log 2>&1; echo "exit $?" >> gpurun_out/pytest_gpu.log ); tail -5 gpurun_out/pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke exit $?"; tail -4 gpurun_out/smoke.log
CMD="python bench.py --workload c3 --steps 2 --warmup 3 --no-cpu-baseline"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:stage_sat -s 300 -c 1 -f -o gpurun_out/prof_r02a_stage_sat_thread $CMD > gpurun_out/ncu_full_sat.log 2>&1; echo "ncu exit $?"
ls -la gpurun_out/
