#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_dec_gpu.py tests/test_random_gpu.py tests/test_sensors_gpu.py -x -q -k "lidar or random or sensor or limits" > gpurun_out/pytest_lidar.log 2>&1; echo "pytest lidar exit $?"; tail -3 gpurun_out/pytest_lidar.log
ARMS="c3_runtab:X=1|c3_cells:NVCC=-DMCOV_LIDAR_RUNTAB=0|c3_runtab_perm:NVCC=-DMCOV_LIDAR_PERM=1|c3_runtab2:X=2" WLS="c3" bash tools/gpu_ab.sh
