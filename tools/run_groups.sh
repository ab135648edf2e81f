#!/bin/bash
# bench one workload with each lane-group width pinned
WL=${WL:-c3}
for g in 8 16 32; do
  echo "== MCOV_DEC_GROUP=$g"
  MCOV_DEC_GROUP=$g WORKLOADS=$WL STEPS=100 bash tools/gpu_workloads.sh | grep value
done
