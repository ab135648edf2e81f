#!/bin/bash
mkdir -p gpurun_out
timeout 400 python -m pytest tests/test_dec_gpu.py tests/test_sharding_gpu.py -x -q -k "store_warp" > gpurun_out/pytest_store.log 2>&1; echo "pytest store exit $?"; tail -3 gpurun_out/pytest_store.log
ARMS="direct:MCOV_DEC_STORE=0|st_sw1_b:MCOV_DEC_STORE=1 MCOV_DEC_SW=1|st_sw2_b:MCOV_DEC_STORE=1 MCOV_DEC_SW=2|st_sw1_b_nobulk:MCOV_DEC_STORE=1 MCOV_DEBUG=8|st_sw1_b_nbuf2:MCOV_DEC_STORE=1 MCOV_DEC_NBUF=2" WLS="c5" bash tools/gpu_ab.sh
ARMS="st_sw1_b:MCOV_DEC_STORE=1 MCOV_DEC_SW=1" WLS="c3" bash tools/gpu_ab.sh
