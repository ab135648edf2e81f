#!/bin/bash
mkdir -p gpurun_out
ARMS="c2_g32:MCOV_DEC_GROUP=32|c2_g16:MCOV_DEC_GROUP=16|c2_g8:MCOV_DEC_GROUP=8" WLS="c2" bash tools/gpu_ab.sh
for a in "MCOV_DEC_GROUP=32" "MCOV_DEC_GROUP=16" "MCOV_DEC_GROUP=8"; do env $a timeout 300 python tools/bench_graph.py c2 6 2>&1 | tail -1 | tee -a gpurun_out/ab.jsonl; done
