#!/bin/bash
mkdir -p gpurun_out
ARMS="c2_g32_bal:X=1|c2_g16_bal:MCOV_DEC_GROUP=16|c2_g8_bal:MCOV_DEC_GROUP=8|c2_g32_bal2:X=2" WLS="c2" bash tools/gpu_ab.sh
