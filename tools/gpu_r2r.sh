#!/bin/bash
mkdir -p gpurun_out
ARMS="c2_w8:MCOV_DEC_WARPS=8|c2_w4:MCOV_DEC_WARPS=4|c2_w2:MCOV_DEC_WARPS=2|c2_w1:MCOV_DEC_WARPS=1|c2_g16_w4:MCOV_DEC_GROUP=16 MCOV_DEC_WARPS=4|c2_g16_w2:MCOV_DEC_GROUP=16 MCOV_DEC_WARPS=2" WLS="c2" bash tools/gpu_ab.sh
for a in "MCOV_DEC_WARPS=4" "MCOV_DEC_WARPS=2" "MCOV_DEC_WARPS=1" "MCOV_DEC_GROUP=16 MCOV_DEC_WARPS=4"; do env $a timeout 300 python tools/bench_graph.py c2 6 2>&1 | tail -1 | tee -a gpurun_out/ab.jsonl; done
