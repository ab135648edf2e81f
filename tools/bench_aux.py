#!/usr/bin/env python
"""Kernel-level measurement of the §8f "next" rows on one B200: SuperGrid_Sim lookahead batch, render()
frames, replay-ring push.  Each line: algorithmic bytes per launch / CUDA-event time vs the measured HBM
peak (MEASURED_PEAKS.json).  Inputs larger than the 126 MB L2, 5 warm-up launches, 30 timed.
usage: python tools/bench_aux.py > gpurun_out/bench_aux.jsonl"""
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as ge                     # noqa: E402

ge.build()
from marl_coverage_b200 import ReplayBuffer, VecDecGridRL, VecSuperGridRL, VecSuperGridSim, load_grid_set   # noqa: E402
from bench import DEC_BASE, measured_peak        # noqa: E402


def timed(fn, warm=5, reps=30):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(reps)]
    for a, b in ev:
        a.record()
        fn()
        b.record()
    torch.cuda.synchronize()
    return float(np.median([a.elapsed_time(b) for a, b in ev]))


def line(name, unit_name, units, alg_bytes, ms, peak, src, extra=None):
    d = {"kernel": name, "units_per_launch": units, "unit": unit_name, "ms": ms,
         "value": units / (ms * 1e-3), "alg_bytes_per_unit": alg_bytes / units,
         "roofline": {"bound": "hbm", "achieved": alg_bytes / (ms * 1e-3) / 1e9, "peak": peak, "unit": "GB/s",
                      "frac": alg_bytes / (ms * 1e-3) / 1e9 / peak, "peak_source": src}}
    d.update(extra or {})
    print(json.dumps(d), flush=True)


def main():
    peak, src = measured_peak()
    grids = load_grid_set("bg2_100x100", 37)
    scfg = dict(numrobot=1, train_maxsteps=1000, test_maxsteps=1000, collision_penalty=5, senseradius=2,
                free_penalty=0.2, done_thresh=1, done_incr=0, terminal_reward=30, dist_reward=0, use_scanning=0)
    # ---- lookahead: 8192 (state, action) pairs of [4,100,100] f32 = 1.3 GB in, 1.3 GB out ---------------
    B, W, L = 8192, 100, 100
    env = VecSuperGridRL(grids, scfg, num_envs=B)
    mi = (torch.arange(B, dtype=torch.int32, device="cuda") % len(grids))
    env.reset(map_idx=mi)
    acts = torch.randint(0, 4, (B, 1), dtype=torch.uint8, device="cuda")
    for _ in range(30):
        env.step(acts)
    states = env.obs.clone()                                   # [B, 4, W, L]: numrobot = 1
    pos = env.positions()[:, 0].contiguous()
    sim = VecSuperGridSim(grids, scfg)
    out = torch.empty_like(states)
    a1 = acts[:, 0].contiguous()
    plane = W * L * 4
    ms = timed(lambda: sim.step(states, a1, map_idx=mi, out=out))
    line("sim_lookahead_kernel (pose from the position plane)", "lookaheads", B, B * 8 * plane, ms, peak, src,
         {"note": "reads 4 planes, writes 4 planes"})
    ms = timed(lambda: sim.step(states, a1, map_idx=mi, pos_xy=pos, out=out))
    line("sim_lookahead_kernel (pose supplied)", "lookaheads", B, B * 7 * plane, ms, peak, src,
         {"note": "reads 3 planes, writes 4 planes"})
    ms = timed(lambda: sim.step(states, a1, map_idx=mi, pos_xy=pos, inplace=True))
    line("sim_lookahead_kernel (in place, pose supplied)", "lookaheads", B, B * 1 * plane, ms, peak, src,
         {"note": "reads the not-yet-covered plane for percent_covered, patches <= 26 cells"})
    del env, sim, states, out
    # ---- render: 64 DecGridRL frames of 1075x1075x3 = 222 MB -------------------------------------------
    dcfg = dict(DEC_BASE, numrobot=4, sensor_type="square_sensor", sensor_config={"range": 2})
    E = 64
    dec = VecDecGridRL(grids, dcfg, num_envs=E)
    dec.reset(map_idx=(torch.arange(E, dtype=torch.int32, device="cuda") % len(grids)))
    da = torch.randint(0, 4, (E, 4), dtype=torch.uint8, device="cuda")
    for _ in range(50):
        dec.step(da)
    frames = dec.render()
    ms = timed(lambda: dec.render(out=frames))
    line("render_kernel (DecGridRL 102x102 -> 1075x1075x3 u8)", "frames", E, frames.numel(), ms, peak, src)
    del dec, frames
    # ---- replay push: 131072 rows of one robot's [3,11,11] obs -> ring (state + next) ---------------------
    E = 131072
    buf = ReplayBuffer((3, 11, 11), 1, 4 * E)
    st = torch.rand((E, 4, 3, 11, 11), device="cuda")
    nx = torch.rand((E, 4, 3, 11, 11), device="cuda")
    ac = torch.randint(0, 4, (E, 1), dtype=torch.uint8, device="cuda")
    rw = torch.rand((E,), device="cuda")
    dn = torch.zeros((E,), dtype=torch.uint8, device="cuda")
    ms = timed(lambda: buf.addbatch(st, ac, rw, nx, dn, src_stride=4 * 363))
    line("replay_push_kernel (obs[:,0] rows, 363 floats)", "transitions", E, E * (363 * 4 * 4 + 1 + 4 + 1 + 12), ms,
         peak, src)


if __name__ == "__main__":
    main()
