#!/bin/bash
# round 2 validation: full GPU parity suite, driver-flag bench, reference arm, ncu captures of every kernel of the path
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/smi.txt 2>&1
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke exit $?"; tail -2 gpurun_out/smoke.log
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?"; tail -4 gpurun_out/pytest_gpu.log
timeout 600 python bench.py --steps 20 --warmup 5 > gpurun_out/bench_driver.json 2> gpurun_out/bench_driver.err; echo "bench exit $?"; tail -2 gpurun_out/bench_driver.err
timeout 300 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_reference.json 2> gpurun_out/bench_reference.err; echo "ref exit $?"
prof() { # WL KREGEX SKIP NAME
  SHORT="python bench.py --workload $1 --steps 4 --warmup 3 --no-cpu --no-extras"
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:$2 -s $3 -c 1 -f -o gpurun_out/prof_$4 $SHORT > gpurun_out/ncu_full_$4.log 2>&1; echo "ncu $4 exit $?"
}
prof c5 dec_env_kernel 10 c5
prof c3 dec_env_kernel 10 c3
prof c4 dec_cta_kernel 10 c4
prof c2 dec_env_kernel 10 c2
prof c2_share dec_share_kernel 10 c2_share
prof c2_dist dec_extras_kernel 10 c2_dist
prof c2_dijkstra dec_extras_kernel 10 c2_dijkstra
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_c5.csv python bench.py --steps 4 --warmup 3 --no-cpu --no-extras > gpurun_out/ncu_launches.log 2>&1; echo "launch list exit $?"
