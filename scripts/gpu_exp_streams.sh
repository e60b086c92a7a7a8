#!/bin/bash
# Sub-batch / multi-stream experiment on C3, then the gpu tests.
set -u
mkdir -p gpurun_out
{
for occ in 4 3 2; do
  XPBD_WORLD_PATH=staged XPBD_SAT_BLOCKS_PER_SM=$occ EXP_K=${EXP_K:-1,2,3,4} timeout 600 python scripts/exp_streams.py 2>&1 | grep -v "^$"
done
} | tee gpurun_out/exp_streams.log
if [ "${AB_TESTS:-1}" = 1 ]; then
  ( timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "exit $?" >> gpurun_out/pytest_gpu.log )
  tail -6 gpurun_out/pytest_gpu.log
fi
