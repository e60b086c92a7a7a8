#!/bin/bash
# Round-1j evidence: bench lines (C3 default, C2, C4, reference arm), launch list of the bench command, one full
# capture per staged kernel (single chain of launches: each kernel alone, steady state).
set -u
mkdir -p gpurun_out
python bench.py --steps 50 --warmup 5 > gpurun_out/bench_c3.json 2> gpurun_out/bench_c3.err
python bench.py --workload c4 --steps 30 --warmup 3 > gpurun_out/bench_c4.json 2> gpurun_out/bench_c4.err
python bench.py --workload c2 --steps 50 --warmup 5 > gpurun_out/bench_c2.json 2> gpurun_out/bench_c2.err
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref_c3.json 2> gpurun_out/bench_ref_c3.err
CMD="python bench.py --workload c3 --steps 2 --warmup 3 --no-cpu-baseline"
# launch list of the bench command as it runs (4 world ranges): skip the pre-roll (30 frames x 20 substeps x 4 stages x 4 ranges)
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -s 9600 -c 640 --csv --log-file gpurun_out/launches_c3_bench.csv $CMD > gpurun_out/ncu_launches.log 2>&1
for K in stage_sat stage_solve stage_prepare stage_manifold; do
  XPBD_SUB_BATCHES=1 timeout 900 ncu --set full --clock-control none --import-source on -k regex:$K -s 665 -c 1 -f -o gpurun_out/prof_$K $CMD > gpurun_out/ncu_full_$K.log 2>&1
done
ls -la gpurun_out/*.ncu-rep gpurun_out/*.json
