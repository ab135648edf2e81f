#!/bin/bash
# ncu --set full of one WARM step launch of a workload's dominant kernel.
# usage: WL=c5 KREGEX=dec_env_kernel SKIP=10 bash tools/gpu_prof.sh
mkdir -p gpurun_out
WL=${WL:-c5}; KREGEX=${KREGEX:-dec_env_kernel}; SKIP=${SKIP:-10}
SHORT="python bench.py --workload $WL --steps 4 --warmup 3 --no-cpu --no-extras"
$SHORT > gpurun_out/plain_$WL.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:$KREGEX -s $SKIP -c 1 -f -o gpurun_out/prof_$WL $SHORT > gpurun_out/ncu_full_$WL.log 2>&1
echo "ncu $WL exit $?"
