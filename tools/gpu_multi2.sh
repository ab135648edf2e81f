#!/bin/bash
# N-GPU bench exactly as the driver launches it (torchrun, one rank per GPU): driver flags, then a longer run
N=${N:-2}
mkdir -p gpurun_out
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 \
  bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/bench_n${N}_driver.json 2> gpurun_out/bench_n${N}_driver.err
echo "bench n=$N (20/5) exit $?"; tail -3 gpurun_out/bench_n${N}_driver.err
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29519 \
  bench.py --gpus $N --steps 200 --warmup 20 > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err
echo "bench n=$N (200/20) exit $?"; tail -3 gpurun_out/bench_n$N.err
python - <<PY
import json
for f in ('gpurun_out/bench_n${N}_driver.json', 'gpurun_out/bench_n$N.json'):
    try:
        d = json.loads([l for l in open(f) if l.startswith('{')][-1])
        print(f, 'value %.4g ms/step %.4f med %.4f e2e %.4g (%.4f ms, allreduce in loop %s) kernel %.4f frac %.3f' % (
            d['value'], d['ms_per_step'], d['ms_per_step_median'], d['e2e']['value'], d['e2e']['ms_per_step'],
            d['e2e'].get('stats_allreduce_in_loop'), d['roofline']['kernel_ms'], d['roofline']['frac']))
        print('  consistency', d['consistency'], 'block_ab', d.get('stats_block_ab'))
    except Exception as ex:
        print(f, 'parse failed', ex)
PY
