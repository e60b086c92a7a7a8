#!/bin/bash
# The parity tests on the bounds-checked debug build (make variant NAME=bounds EXTRA=-DXPBD_BOUNDS): every XPBD_CHECK in the
# kernels traps on a bad index (compute-sanitizer is not available on the GPU pool).  The product build compiles them out.
set -u
mkdir -p gpurun_out
export XPBD_B200_LIB=$PWD/constraint_solver_b200/lib/variants/libxpbd_b200_bounds.so
( timeout 1500 python -m pytest tests/test_parity_world.py tests/test_parity_single.py tests/test_parity_blocks.py -m gpu -x -q -k "not full_size" > gpurun_out/pytest_bounds.log 2>&1; echo "exit $?" >> gpurun_out/pytest_bounds.log )
grep -c "XPBD_CHECK failed" gpurun_out/pytest_bounds.log; tail -4 gpurun_out/pytest_bounds.log
