#!/bin/bash
mkdir -p gpurun_out
MCOV_NVCC_EXTRA="-DMCOV_LUT8=1" python marl_coverage_b200/build.py > /dev/null 2>&1 || echo "build failed"
timeout 600 python -m pytest tests/test_dec_gpu.py tests/test_fullsize_gpu.py -x -q -k "every_group_width or batched or fullsize or full" > gpurun_out/pytest_lut8.log 2>&1; echo "pytest lut8 exit $?"; tail -3 gpurun_out/pytest_lut8.log
python marl_coverage_b200/build.py > /dev/null 2>&1
ARMS="c5_base:X=1|c5_lut8:NVCC=-DMCOV_LUT8=1|c5_nopf:NVCC=-DMCOV_NO_PF=1|c5_lut8_nopf:NVCC=-DMCOV_LUT8=1,-DMCOV_NO_PF=1|c5_base2:X=2" WLS="c5" bash tools/gpu_ab.sh
ARMS="c3_base:X=1|c3_lut8:NVCC=-DMCOV_LUT8=1|c3_nopf:NVCC=-DMCOV_NO_PF=1" WLS="c3 c2" bash tools/gpu_ab.sh
