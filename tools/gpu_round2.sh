#!/bin/bash
# Round-2 validation call: parity tests, default bench (driver flags), reference arm bookkeeping.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/smi.txt 2>&1
timeout ${TEST_LIMIT:-900} python -m pytest tests -m gpu -x -q ${K:+-k "$K"} > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?"; tail -4 gpurun_out/pytest_gpu.log
python bench.py --steps 20 --warmup 5 > gpurun_out/bench_driver.json 2> gpurun_out/bench_driver.err; echo "bench exit $?"; tail -2 gpurun_out/bench_driver.err
python - <<'PY'
import json
try:
    d = json.loads([l for l in open('gpurun_out/bench_driver.json') if l.startswith('{')][-1])
    print('value %.4g  ms/step %.4f  median %.4f  e2e %.4g (%.4f ms)  kernel %.4f ms frac %.3f  consistency %s' % (
        d['value'], d['ms_per_step'], d['ms_per_step_median'], d['e2e']['value'], d['e2e']['ms_per_step'],
        d['roofline']['kernel_ms'], d['roofline']['frac'], d['consistency']))
    print('obs_to_host', d['e2e'].get('obs_to_host'))
    for k, v in (d.get('workloads') or {}).items():
        print(k, {a: b for a, b in v.items() if a != 'workload'})
    print('cpu', d['cpu_baseline'])
except Exception as ex:
    print('parse failed', ex)
PY
