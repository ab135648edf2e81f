#!/bin/bash
# round 2, call C: store-warp (counter hand-over) parity tests + A/B with profiling modes
mkdir -p gpurun_out
timeout 700 python -m pytest tests/test_dec_gpu.py tests/test_sharding_gpu.py tests/test_random_gpu.py -x -q -k "store_warp" > gpurun_out/pytest_store.log 2>&1; echo "pytest store exit $?"; tail -5 gpurun_out/pytest_store.log
ARMS="direct:MCOV_DEC_STORE=0|st_sw1:MCOV_DEC_STORE=1 MCOV_DEC_SW=1|st_sw2:MCOV_DEC_STORE=1 MCOV_DEC_SW=2|st_sw1_handover_only:MCOV_DEC_STORE=1 MCOV_DEBUG=4|st_sw1_nobulk:MCOV_DEC_STORE=1 MCOV_DEBUG=8|st_sw2_nobulk:MCOV_DEC_STORE=1 MCOV_DEC_SW=2 MCOV_DEBUG=8|st_sw2_handover_only:MCOV_DEC_STORE=1 MCOV_DEC_SW=2 MCOV_DEBUG=4|direct_noobs:MCOV_DEC_STORE=0 MCOV_DEBUG=1" WLS="c5" bash tools/gpu_ab.sh
ARMS="st_sw1:MCOV_DEC_STORE=1 MCOV_DEC_SW=1|st_sw2:MCOV_DEC_STORE=1 MCOV_DEC_SW=2" WLS="c3" bash tools/gpu_ab.sh
