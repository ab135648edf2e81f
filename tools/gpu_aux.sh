#!/bin/bash
mkdir -p gpurun_out
timeout 600 python tools/bench_aux.py > gpurun_out/bench_aux.jsonl 2> gpurun_out/bench_aux.err; echo "aux exit $?"
python - <<'PY'
import json
for l in open('gpurun_out/bench_aux.jsonl'):
    if l.startswith('{'):
        d = json.loads(l); print('%-58s %8.4f ms  %7.0f GB/s  frac %.3f' % (d['kernel'], d['ms'], d['roofline']['achieved'], d['roofline']['frac']))
PY
for w in super; do timeout 300 python bench.py --workload $w --steps 100 --warmup 20 --no-cpu --no-extras > gpurun_out/bench_$w.json 2> gpurun_out/bench_$w.err; echo "$w exit $?"; done
