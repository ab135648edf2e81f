#!/bin/bash
# compute-sanitizer over the small parity tests: memcheck, then racecheck / synccheck for the kernels that
# hand data between lanes through shared memory (bit stream, position patches, rowbuf).  10-100x slower than a
# plain run: pick tests with K (pytest -k expression) and give gpurun a generous --timeout.
# usage: K="square_n1 or c1_100" TOOLS="memcheck racecheck" bash tools/gpu_sanitize.sh
mkdir -p gpurun_out
T=${T:-"tests/test_dec_gpu.py tests/test_super_gpu.py tests/test_replay_gpu.py tests/test_sim_render_gpu.py"}
for tool in ${TOOLS:-memcheck racecheck synccheck}; do
    timeout ${LIMIT:-900} compute-sanitizer --tool $tool --error-exitcode 9 --target-processes all \
        python -m pytest $T -x -q ${K:+-k "$K"} > gpurun_out/sanitize_$tool.log 2>&1
    echo "$tool exit $?"; tail -3 gpurun_out/sanitize_$tool.log
done
