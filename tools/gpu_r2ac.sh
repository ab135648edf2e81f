#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_dec_gpu.py tests/test_sharding_gpu.py -x -q -k "shared_memory or variants" > gpurun_out/pytest_smt.log 2>&1; echo "pytest smt exit $?"; tail -12 gpurun_out/pytest_smt.log
ARMS="c5_base:X=1|c5_smt:MCOV_DEC_SMT=1|c5_base2:X=2|c5_smt2:MCOV_DEC_SMT=1|c5_smt_g8:MCOV_DEC_SMT=1 MCOV_DEC_GROUP=8" WLS="c5" bash tools/gpu_ab.sh
