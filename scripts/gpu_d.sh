#!/bin/bash
set -u
mkdir -p gpurun_out
( timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "exit $?" >> gpurun_out/pytest_gpu.log )
tail -15 gpurun_out/pytest_gpu.log
( timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "exit $?" >> gpurun_out/smoke.log ); tail -3 gpurun_out/smoke.log
for wl in c3 c2 c4; do
  timeout 900 python bench.py --workload $wl --steps 10 --warmup 3 > gpurun_out/bench_$wl.json 2> gpurun_out/bench_$wl.err; echo "bench $wl exit $?"
  cut -c1-1500 gpurun_out/bench_$wl.json; tail -3 gpurun_out/bench_$wl.err
done
timeout 600 python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/bench_ref_c3.json 2>&1; cut -c1-600 gpurun_out/bench_ref_c3.json
