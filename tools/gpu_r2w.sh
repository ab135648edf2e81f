#!/bin/bash
# the whole GPU suite with the Dec kernel shape pinned either way (tests that pin it themselves keep their pin)
mkdir -p gpurun_out
MCOV_DEC_SHAPE=0 timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_shape0.log 2>&1; echo "shape0 exit $?"; tail -2 gpurun_out/pytest_gpu_shape0.log
MCOV_DEC_SHAPE=1 timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_shape1.log 2>&1; echo "shape1 exit $?"; tail -2 gpurun_out/pytest_gpu_shape1.log
