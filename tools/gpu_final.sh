#!/bin/bash
# last check of the round on the committed library: smoke, whole GPU suite, driver-flag bench
mkdir -p gpurun_out
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke exit $?"; tail -1 gpurun_out/smoke.log
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?"; tail -2 gpurun_out/pytest_gpu.log
timeout 600 python bench.py --steps 20 --warmup 5 > gpurun_out/bench_driver.json 2> gpurun_out/bench_driver.err; echo "bench exit $?"
python - <<'PY'
import json
d = json.loads([l for l in open('gpurun_out/bench_driver.json') if l.startswith('{')][-1])
print('value %.4g ms/step %.4f e2e %.4g kernel %.4f frac %.3f ok %s' % (d['value'], d['ms_per_step'], d['e2e']['value'], d['roofline']['kernel_ms'], d['roofline']['frac'], d['consistency']['ok']))
PY
