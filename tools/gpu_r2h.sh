#!/bin/bash
mkdir -p gpurun_out
MCOV_NVCC_EXTRA=-DMCOV_CTA_TICKS python marl_coverage_b200/build.py > /dev/null 2>&1 || echo build failed
for t in 1 0; do MCOV_DEC_TMA=$t timeout 300 python tools/cta_ticks.py c4 2>&1 | tail -1 | tee -a gpurun_out/cta_ticks.jsonl; done
python marl_coverage_b200/build.py > /dev/null 2>&1
