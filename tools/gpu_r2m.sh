#!/bin/bash
mkdir -p gpurun_out
ARMS="c5_g16_a:X=1|c5_g16_minb3:NVCC=-DMCOV_DEC_MINB=3|c5_g16_minb3_u8:NVCC=-DMCOV_DEC_MINB=3,-DMCOV_ST_UNROLL=8|c5_g16_pfnext:NVCC=-DMCOV_PF_NEXT=1|c5_g16_stdefault:NVCC=-DMCOV_ST_MODE=1|c5_g16_stcg:NVCC=-DMCOV_ST_MODE=2|c5_g16_b:X=2|c5_g16_minb3_b:NVCC=-DMCOV_DEC_MINB=3" WLS="c5" bash tools/gpu_ab.sh
