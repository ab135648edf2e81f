timeout 600 python -m pytest tests/test_super_gpu.py "tests/test_random_gpu.py::test_super_random_vs_oracle" -x -q 2>&1 | tail -3
WORKLOADS=super bash tools/gpu_workloads.sh
