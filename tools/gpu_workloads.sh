#!/bin/bash
# Bench every named workload once (kernel-only roofline + whole-step numbers), no CPU arm.
mkdir -p gpurun_out
for w in ${WORKLOADS:-c2 c3 c4 super c5}; do
  timeout 600 python bench.py --workload $w --steps ${STEPS:-200} --warmup 20 --no-cpu --no-extras > gpurun_out/bench_$w.json 2> gpurun_out/bench_$w.err
  echo "$w exit $?"
  python - "$w" <<'PY'
import json, sys
w = sys.argv[1]
try:
    d = json.loads(open('gpurun_out/bench_%s.json' % w).read().strip().splitlines()[-1])
    r = d['roofline']
    print('%s value %.4g ms/step %.4f kernel_ms %.4f achieved %.0f GB/s frac %.3f e2e %.4g' % (
        w, d['value'], d['ms_per_step'], r['kernel_ms'], r['achieved'], r['frac'], d['e2e']['value']))
except Exception as ex:
    print(w, 'parse failed', ex)
    print(open('gpurun_out/bench_%s.err' % w).read()[-1500:])
PY
done
