#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_dec_gpu.py tests/test_random_gpu.py -x -q -k "cta" > gpurun_out/pytest_cta.log 2>&1; echo "pytest cta exit $?"; tail -15 gpurun_out/pytest_cta.log
ARMS="lanegroups:MCOV_DEC_SHAPE=0|cta_tma:MCOV_DEC_SHAPE=1 MCOV_DEC_TMA=1|cta_ldg:MCOV_DEC_SHAPE=1 MCOV_DEC_TMA=0" WLS="c4 c2" bash tools/gpu_ab.sh
