#!/bin/bash
# round 2, call B: store-warp parity tests, then the C5 / C3 / C2 A/B of the store-warp path
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_dec_gpu.py tests/test_sharding_gpu.py tests/test_random_gpu.py -x -q -k "store_warp" > gpurun_out/pytest_store.log 2>&1; echo "pytest store exit $?"; tail -5 gpurun_out/pytest_store.log
timeout 300 python -m pytest tests/test_fullsize_gpu.py -x -q > gpurun_out/pytest_full.log 2>&1; echo "pytest fullsize exit $?"; tail -3 gpurun_out/pytest_full.log
ARMS="direct:MCOV_DEC_STORE=0|st4:MCOV_DEC_STORE=1 MCOV_DEC_NBUF=4|st3:MCOV_DEC_STORE=1 MCOV_DEC_NBUF=3|st2:MCOV_DEC_STORE=1 MCOV_DEC_NBUF=2" WLS="c5" bash tools/gpu_ab.sh
ARMS="direct:MCOV_DEC_STORE=0|st:MCOV_DEC_STORE=1" WLS="c3 c2 c4" bash tools/gpu_ab.sh
