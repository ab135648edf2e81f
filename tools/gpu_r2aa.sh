#!/bin/bash
mkdir -p gpurun_out
ARMS="c3_g32:X=1|c3_g16:MCOV_DEC_GROUP=16|c3_g32_w4:MCOV_DEC_WARPS=4|c3_g32_w2:MCOV_DEC_WARPS=2" WLS="c3" bash tools/gpu_ab.sh
ARMS="c5_g16_w8:X=1|c5_g16_w4:MCOV_DEC_WARPS=4|c5_g16_w2:MCOV_DEC_WARPS=2" WLS="c5" bash tools/gpu_ab.sh
