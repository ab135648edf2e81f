#!/bin/bash
# GPU-box check: parity tests, then a short bench (use under gpurun)
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -8
timeout 400 python bench.py --steps ${STEPS:-300} --warmup 20 --no-cpu ${BENCH_ARGS} > gpurun_out/bench_quick.log 2>gpurun_out/bench_quick.err
python - <<'PY'
import json
try:
    d=json.loads(open('gpurun_out/bench_quick.log').read().strip().splitlines()[-1])
    print('value %.4g  ms/step %.4f  kernel_ms %.4f frac %.3f  e2e %.4g  c2_4096 %s' % (d['value'], d['ms_per_step'], d['roofline']['kernel_ms'], d['roofline']['frac'], d['e2e']['value'], (d.get('c2_4096') or {}).get('value')))
    print('stats', d['episode_stats'], 'clocks', d['clocks'])
except Exception as ex:
    print('bench parse failed', ex); print(open('gpurun_out/bench_quick.err').read()[-2000:])
PY
