#!/bin/bash
mkdir -p gpurun_out
timeout 400 python -m pytest tests/test_dec_gpu.py tests/test_sharding_gpu.py -x -q -k "store_warp" > gpurun_out/pytest_store.log 2>&1; echo "pytest store exit $?"; tail -3 gpurun_out/pytest_store.log
ARMS="direct:MCOV_DEC_STORE=0|st_rot_sw1:MCOV_DEC_STORE=1|st_norot_sw1:MCOV_DEC_STORE=1 MCOV_DEC_ROT=0|st_rot_sw2:MCOV_DEC_STORE=1 MCOV_DEC_SW=2|st_rot_sw1_nobulk:MCOV_DEC_STORE=1 MCOV_DEBUG=8|st_rot_sw1_handover:MCOV_DEC_STORE=1 MCOV_DEBUG=4|st_rot_sw1_nbuf2:MCOV_DEC_STORE=1 MCOV_DEC_NBUF=2" WLS="c5" bash tools/gpu_ab.sh
ARMS="st_rot_sw1:MCOV_DEC_STORE=1" WLS="c3" bash tools/gpu_ab.sh
