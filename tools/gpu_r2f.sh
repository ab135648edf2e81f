#!/bin/bash
# ncu --set full of one warm C5 step launch of the store-warp kernel
mkdir -p gpurun_out
export MCOV_DEC_STORE=1 MCOV_DEC_ROT=${ROT:-0}
SHORT="python bench.py --workload c5 --steps 4 --warmup 3 --no-cpu --no-extras"
$SHORT > gpurun_out/plain_c5st.log 2>&1 &&
timeout 600 ncu --set full --clock-control none --import-source on -k regex:dec_env_kernel -s 10 -c 1 -f -o gpurun_out/prof_c5st $SHORT > gpurun_out/ncu_full_c5st.log 2>&1
echo "ncu exit $?"; ls -la gpurun_out/*.ncu-rep
