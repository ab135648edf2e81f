#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_dec_gpu.py tests/test_random_gpu.py -x -q -k "cta" > gpurun_out/pytest_cta.log 2>&1; echo "pytest cta exit $?"; tail -5 gpurun_out/pytest_cta.log
ARMS="lanegroups:MCOV_DEC_SHAPE=0|cta_tma_spec:MCOV_DEC_SHAPE=1 MCOV_DEC_TMA=1|cta_ldg:MCOV_DEC_SHAPE=1 MCOV_DEC_TMA=0" WLS="c4" bash tools/gpu_ab.sh
for a in "MCOV_DEC_SHAPE=0" "MCOV_DEC_SHAPE=1 MCOV_DEC_TMA=1" "MCOV_DEC_SHAPE=1 MCOV_DEC_TMA=0"; do env $a timeout 300 python tools/bench_graph.py c4 24 2>&1 | tail -1 | tee -a gpurun_out/ab.jsonl; done
for a in "MCOV_DEC_SHAPE=0" "MCOV_DEC_SHAPE=1 MCOV_DEC_TMA=1"; do env $a timeout 300 python tools/bench_graph.py c2 6 2>&1 | tail -1 | tee -a gpurun_out/ab.jsonl; done
MCOV_NVCC_EXTRA=-DMCOV_CTA_TICKS python marl_coverage_b200/build.py > /dev/null 2>&1 || echo build failed
for t in 1 0; do MCOV_DEC_TMA=$t timeout 300 python tools/cta_ticks.py c4 2>&1 | tail -1 | tee -a gpurun_out/cta_ticks.jsonl; done
python marl_coverage_b200/build.py > /dev/null 2>&1
