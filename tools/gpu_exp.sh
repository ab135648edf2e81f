#!/bin/bash
# Experiment round: A/B build variants on one box (same clocks), then the aux bench.
# usage: VARIANTS="-DMCOV_PF_NEXT=0|-DMCOV_PF_NEXT=1" WLS="c5 c3" bash tools/gpu_exp.sh
mkdir -p gpurun_out
IFS="|" read -ra VS <<< "${VARIANTS:- }"
for v in "${VS[@]}"; do
  MCOV_NVCC_EXTRA="$v" python marl_coverage_b200/build.py > /dev/null 2>&1 || { echo "build failed: $v"; continue; }
  echo "== variant: $v"
  WORKLOADS="${WLS:-c5}" STEPS=${STEPS:-200} bash tools/gpu_workloads.sh | grep value
done
python marl_coverage_b200/build.py > /dev/null
if [ -n "$TESTS" ]; then timeout 900 python -m pytest $TESTS -x -q 2>&1 | tail -5; fi
if [ -n "$AUX" ]; then python tools/bench_aux.py > gpurun_out/bench_aux.jsonl 2> gpurun_out/bench_aux.err; echo "aux exit $?"; cat gpurun_out/bench_aux.jsonl | python -c "
import sys, json
for l in sys.stdin:
    if not l.startswith('{'): continue
    d = json.loads(l); print('%-58s %8.4f ms  %7.0f GB/s  frac %.3f' % (d['kernel'], d['ms'], d['roofline']['achieved'], d['roofline']['frac']))"; tail -3 gpurun_out/bench_aux.err; fi
