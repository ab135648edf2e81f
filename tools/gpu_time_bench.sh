#!/bin/bash
mkdir -p gpurun_out
s=$(date +%s); python bench.py --steps 20 --warmup 5 > gpurun_out/bench_t.json 2> gpurun_out/bench_t.err; e=$(date +%s); echo "default bench wall $((e-s)) s exit $?"
s=$(date +%s); python bench.py --impl reference --steps 20 --warmup 5 > gpurun_out/bench_tr.json 2> gpurun_out/bench_tr.err; e=$(date +%s); echo "reference arm wall $((e-s)) s"; python -c "
import json; d=json.loads([l for l in open('gpurun_out/bench_tr.json') if l.startswith('{')][-1]); print({k:d.get(k) for k in ('impl','value','steps','warmup','ms_per_step')}, d['cpu_baseline']['sample'])"
