#!/bin/bash
# the static-scratch entry point: its test alone in a fresh process, then after every other GPU test
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_zz_no_alloc.py -x -q > gpurun_out/pytest_no_alloc_alone.log 2>&1; echo "alone exit $?"; tail -15 gpurun_out/pytest_no_alloc_alone.log
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "suite exit $?"; tail -3 gpurun_out/pytest_gpu.log
