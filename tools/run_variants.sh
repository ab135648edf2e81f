#!/bin/bash
# rebuild on the GPU box with different -D flags and bench one workload per variant
WL=${WL:-super}
for v in "$@"; do
  MCOV_NVCC_EXTRA="$v" python marl_coverage_b200/build.py > /dev/null 2>&1 || { echo "build failed: $v"; continue; }
  echo "== $v"
  WORKLOADS=$WL STEPS=100 bash tools/gpu_workloads.sh | grep value
done
python marl_coverage_b200/build.py > /dev/null
