#!/bin/bash
mkdir -p gpurun_out
MCOV_NVCC_EXTRA="-DMCOV_PF_EARLY=1" python marl_coverage_b200/build.py > /dev/null 2>&1 || echo "build failed"
timeout 300 python -m pytest tests/test_dec_gpu.py -x -q -k "every_group_width or batched" > gpurun_out/pytest_pfe.log 2>&1; echo "pytest pfe exit $?"; tail -3 gpurun_out/pytest_pfe.log
python marl_coverage_b200/build.py > /dev/null 2>&1
ARMS="c5_pfearly:NVCC=-DMCOV_PF_EARLY=1|c5_base:X=1|c5_pfearly2:NVCC=-DMCOV_PF_EARLY=1" WLS="c5" bash tools/gpu_ab.sh
ARMS="pfearly:NVCC=-DMCOV_PF_EARLY=1" WLS="c3 c2" bash tools/gpu_ab.sh
