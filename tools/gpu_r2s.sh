#!/bin/bash
mkdir -p gpurun_out
ARMS="c3_cta_tma:MCOV_DEC_SHAPE=1|c3_cta_ldg:MCOV_DEC_SHAPE=1 MCOV_DEC_TMA=0|c3_lane:MCOV_DEC_SHAPE=0" WLS="c3" bash tools/gpu_ab.sh
ARMS="c5_cta_tma:MCOV_DEC_SHAPE=1|c5_cta_ldg:MCOV_DEC_SHAPE=1 MCOV_DEC_TMA=0" WLS="c5" bash tools/gpu_ab.sh
