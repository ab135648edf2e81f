timeout 900 python -m pytest tests/test_dec_gpu.py "tests/test_random_gpu.py::test_dec_random_vs_oracle" -x -q 2>&1 | tail -3
WORKLOADS="${WORKLOADS:-c3 c2 c4 c5}" bash tools/gpu_workloads.sh
