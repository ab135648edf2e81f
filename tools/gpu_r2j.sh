#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_dec_gpu.py -x -q -k "lidar" > gpurun_out/pytest_lidar.log 2>&1; echo "pytest lidar exit $?"; tail -3 gpurun_out/pytest_lidar.log
ARMS="c3_runs_perm:X=1|c3_noruns_noperm:NVCC=-DMCOV_LIDAR_RUNS=0,-DMCOV_LIDAR_PERM=0|c3_runs_only:NVCC=-DMCOV_LIDAR_PERM=0|c3_perm_only:NVCC=-DMCOV_LIDAR_RUNS=0" WLS="c3" bash tools/gpu_ab.sh
ARMS="c5_g8:X=1|c5_g16:MCOV_DEC_GROUP=16|c5_g32:MCOV_DEC_GROUP=32|c5_unroll8:NVCC=-DMCOV_ST_UNROLL=8|c5_unroll2:NVCC=-DMCOV_ST_UNROLL=2|c5_unroll12:NVCC=-DMCOV_ST_UNROLL=12" WLS="c5" bash tools/gpu_ab.sh
