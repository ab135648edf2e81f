#!/bin/bash
mkdir -p gpurun_out
ARMS="c5_base:X=1|c5_pftrim1:NVCC=-DMCOV_PF_TRIM=1|c5_pftrim2:NVCC=-DMCOV_PF_TRIM=2|c5_base2:X=2" WLS="c5" bash tools/gpu_ab.sh
ARMS="pftrim1:NVCC=-DMCOV_PF_TRIM=1" WLS="c2 c3" bash tools/gpu_ab.sh
