#!/bin/bash
# One GPU-box visit: parity tests, the bench line, the ncu launch list and one full capture of the
# dominant kernel.  Usage (from the build container):  gpurun --timeout 1800 -- 'bash tools/gpu_round.sh'
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/smi.txt 2>&1
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1
echo "pytest exit $?" | tee -a gpurun_out/pytest_gpu.log
tail -5 gpurun_out/pytest_gpu.log
timeout 600 python bench.py --steps ${STEPS:-300} --warmup 30 > gpurun_out/bench.json 2> gpurun_out/bench.err
echo "bench exit $?"; tail -c 3000 gpurun_out/bench.json
if [ -z "$NO_NCU" ]; then
SHORT="python bench.py --steps 4 --warmup 3 --no-cpu --no-extras"
$SHORT > gpurun_out/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv $SHORT > gpurun_out/ncu_launches.log 2>&1
echo "ncu launches exit $?"
$SHORT > gpurun_out/plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:dec_env_kernel -s 8 -c 2 -f -o gpurun_out/prof_dec_c5 $SHORT > gpurun_out/ncu_full.log 2>&1
echo "ncu full exit $?"
fi
