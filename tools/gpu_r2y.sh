#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_graph_gpu.py -x -q > gpurun_out/pytest_graph.log 2>&1; echo "pytest graph exit $?"; tail -15 gpurun_out/pytest_graph.log
