#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_dec_gpu.py tests/test_sharding_gpu.py -x -q -k "teams or variants" > gpurun_out/pytest_teams.log 2>&1; echo "pytest teams exit $?"; tail -4 gpurun_out/pytest_teams.log
ARMS="c5_g16:X=1|c5_g8_team16:MCOV_DEC_GROUP=8 MCOV_DEC_TEAM=16|c5_g8_team32:MCOV_DEC_GROUP=8 MCOV_DEC_TEAM=32|c5_g16_team32:MCOV_DEC_TEAM=32|c5_g8:MCOV_DEC_GROUP=8" WLS="c5" bash tools/gpu_ab.sh
