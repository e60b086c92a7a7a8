#!/bin/bash
# gpu tests, the default bench line, then the ncu launch list of the bench command (steady state: skips the pre-roll and warm-up launches)
set -u
mkdir -p gpurun_out
( timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "exit $?" >> gpurun_out/pytest_gpu.log )
tail -4 gpurun_out/pytest_gpu.log
python bench.py --steps 50 --warmup 5 > gpurun_out/bench_c3.json 2> gpurun_out/bench_c3.err
CMD="python bench.py --workload c3 --steps 2 --warmup 3 --no-cpu-baseline"
# 4 world ranges x 4 stages: pre-roll 30 x 20 x 16 = 9600 launches, warm-up 3 x 20 x 16 = 960, then 2 timed frames = 640
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -s 10560 -c 640 --csv --log-file gpurun_out/launches_c3_bench.csv $CMD > gpurun_out/ncu_launches.log 2>&1
wc -l gpurun_out/launches_c3_bench.csv
