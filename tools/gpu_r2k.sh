#!/bin/bash
mkdir -p gpurun_out
ARMS="c5_g16_default:X=1|c5_g16_st:MCOV_DEC_STORE=1 MCOV_DEC_ROT=0|c5_g16_st_rot:MCOV_DEC_STORE=1|c5_g16_minb3:NVCC=-DMCOV_DEC_MINB=3|c5_g16_minb5:NVCC=-DMCOV_DEC_MINB=5|c5_g16_unroll8:NVCC=-DMCOV_ST_UNROLL=8|c5_g16_noobs:MCOV_DEBUG=1" WLS="c5" bash tools/gpu_ab.sh
ARMS="default:X=1" WLS="c2 c3 c4" bash tools/gpu_ab.sh
