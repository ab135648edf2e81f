#!/bin/bash
mkdir -p gpurun_out
MCOV_NVCC_EXTRA="-DMCOV_LUT8=2" python marl_coverage_b200/build.py > /dev/null 2>&1 || echo "build failed"
timeout 600 python -m pytest tests/test_dec_gpu.py tests/test_fullsize_gpu.py tests/test_random_gpu.py -x -q -k "every_group_width or batched or fullsize or full or (random_vs_oracle and not cta and not store)" > gpurun_out/pytest_v8.log 2>&1; echo "pytest v8 exit $?"; tail -3 gpurun_out/pytest_v8.log
python marl_coverage_b200/build.py > /dev/null 2>&1
ARMS="c5_base:X=1|c5_v8:NVCC=-DMCOV_LUT8=2|c5_v8_u2:NVCC=-DMCOV_LUT8=2,-DMCOV_ST_UNROLL=2|c5_v8_u6:NVCC=-DMCOV_LUT8=2,-DMCOV_ST_UNROLL=6|c5_base2:X=2" WLS="c5" bash tools/gpu_ab.sh
ARMS="v8:NVCC=-DMCOV_LUT8=2" WLS="c3 c2" bash tools/gpu_ab.sh
