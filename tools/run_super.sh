timeout 600 python -m pytest tests/test_super_gpu.py tests/test_random_gpu.py "tests/test_dec_gpu.py::test_dec_golden_single_env" -x -q 2>&1 | tail -3
WORKLOADS=super bash tools/gpu_workloads.sh
