#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_dec_gpu.py tests/test_random_gpu.py tests/test_fullsize_gpu.py tests/test_graph_gpu.py tests/test_zz_no_alloc.py -x -q -k "lidar or random or c3 or full or graph or no_alloc or static" > gpurun_out/pytest_lidar.log 2>&1; echo "pytest exit $?"; tail -2 gpurun_out/pytest_lidar.log
ARMS="c3_default_w2:X=1|c3_w1:MCOV_DEC_WARPS=1|c3_w8:MCOV_DEC_WARPS=8|c3_default_w2_b:X=2" WLS="c3" bash tools/gpu_ab.sh
