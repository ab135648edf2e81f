#!/bin/bash
mkdir -p gpurun_out
MCOV_DEC_SHAPE=0 timeout 600 python -m pytest tests/test_dec_gpu.py tests/test_random_gpu.py tests/test_graph_gpu.py tests/test_sharding_gpu.py tests/test_boundary_gpu.py -x -q > gpurun_out/pytest_bal.log 2>&1; echo "pytest bal exit $?"; tail -3 gpurun_out/pytest_bal.log
ARMS="c2_balanced:X=1|c2_unbalanced:MCOV_DEC_BALANCE=0|c2_balanced2:X=2" WLS="c2" bash tools/gpu_ab.sh
for a in "MCOV_DEC_BALANCE=1" "MCOV_DEC_BALANCE=0"; do env $a timeout 300 python tools/bench_graph.py c2 6 2>&1 | tail -1 | tee -a gpurun_out/ab.jsonl; done
