#!/bin/bash
# First GPU pass of round 1: parity tests, bench (c2, c3), ncu launch list + full capture.
set -u
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/smi.txt 2>&1
nproc > gpurun_out/nproc.txt
( timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "exit $?" >> gpurun_out/pytest_gpu.log )
tail -5 gpurun_out/pytest_gpu.log
( timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "exit $?" >> gpurun_out/smoke.log )
tail -3 gpurun_out/smoke.log
timeout 600 python bench.py --steps 20 --warmup 3 > gpurun_out/bench_c2.json 2> gpurun_out/bench_c2.err; echo "bench c2 exit $?"
cat gpurun_out/bench_c2.json
timeout 600 python bench.py --workload c3 --steps 10 --warmup 3 > gpurun_out/bench_c3.json 2> gpurun_out/bench_c3.err; echo "bench c3 exit $?"
cat gpurun_out/bench_c3.json
timeout 300 python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/bench_ref_c2.json 2>&1
cat gpurun_out/bench_ref_c2.json
# ncu launch lists (same command exited 0 just above, without ncu)
timeout 600 python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/plain_c2.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_c2.csv \
   python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_c2.log 2>&1
timeout 600 python bench.py --workload c3 --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/plain_c3.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_c3.csv \
   python bench.py --workload c3 --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_c3.log 2>&1
# full captures of the two step kernels on reduced batches (ncu replays each launch ~40x)
timeout 300 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --worlds 262144 > gpurun_out/plain_c2s.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:step_single -s 4 -c 2 -f -o gpurun_out/prof_c2 \
   python bench.py --steps 3 --warmup 3 --no-cpu-baseline --worlds 262144 > gpurun_out/ncu_full_c2.log 2>&1
timeout 300 python bench.py --workload c3 --steps 3 --warmup 3 --no-cpu-baseline --worlds 20000 > gpurun_out/plain_c3s.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:step_world -s 4 -c 2 -f -o gpurun_out/prof_c3 \
   python bench.py --workload c3 --steps 3 --warmup 3 --no-cpu-baseline --worlds 20000 > gpurun_out/ncu_full_c3.log 2>&1
ls -la gpurun_out
