#!/bin/bash
# parity soak over the whole accepted parameter range (tests/test_random_gpu.py::test_whole_parameter_range_vs_oracle),
# through the auto-selected kernel shape and with each Dec shape pinned
mkdir -p gpurun_out
for sh in auto 0 1; do
  if [ "$sh" = auto ]; then unset MCOV_DEC_SHAPE; else export MCOV_DEC_SHAPE=$sh; fi
  MCOV_SOAK=${SOAK:-30} timeout 600 python -m pytest tests/test_random_gpu.py -x -q -k "whole_parameter_range" > gpurun_out/soak_$sh.log 2>&1
  echo "soak shape=$sh exit $?: $(tail -1 gpurun_out/soak_$sh.log)"
done
