#!/bin/bash
set -u
mkdir -p gpurun_out
{
for occ in 1 2 3; do
  XPBD_WORLD_PATH=staged XPBD_SAT_BLOCKS_PER_SM=$occ EXP_K=4,5,6,8 timeout 600 python scripts/exp_streams.py 2>&1 | grep -v "^$"
done
for sb in 3 4 8; do
  echo "solve blocks/SM $sb"
  XPBD_WORLD_PATH=staged XPBD_SAT_BLOCKS_PER_SM=2 XPBD_SOLVE_BLOCKS_PER_SM=$sb EXP_K=4 timeout 600 python scripts/exp_streams.py 2>&1 | grep -v "^$"
done
} | tee gpurun_out/exp_streams2.log
