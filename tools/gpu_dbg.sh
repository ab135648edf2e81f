#!/bin/bash
# A/B of build variants x MCOV_DEBUG modes on one workload.  VARIANTS="a|b" DBG="0 2 4" WL=c5
IFS="|" read -ra VS <<< "${VARIANTS:- }"
for v in "${VS[@]}"; do
  MCOV_NVCC_EXTRA="$v" python marl_coverage_b200/build.py > /dev/null 2>&1 || { echo "build failed: $v"; continue; }
  for d in ${DBG:-0}; do
    echo "== variant: $v  MCOV_DEBUG=$d"
    MCOV_DEBUG=$d WORKLOADS="${WL:-c5}" STEPS=${STEPS:-100} bash tools/gpu_workloads.sh | grep value
  done
done
python marl_coverage_b200/build.py > /dev/null
