#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_dec_gpu.py tests/test_random_gpu.py tests/test_sensors_gpu.py tests/test_fullsize_gpu.py -x -q -k "lidar or random or sensor or limits or c3 or full" > gpurun_out/pytest_lidar.log 2>&1; echo "pytest lidar exit $?"; tail -3 gpurun_out/pytest_lidar.log
ARMS="c3_sentinel_walk:X=1|c3_sentinel_walk2:X=2" WLS="c3" bash tools/gpu_ab.sh
