#!/usr/bin/env python
"""Phase timing of the CTA-per-env DecGridRL kernel (profiling build -DMCOV_CTA_TICKS): runs C4 steps with the
L2 flushed in between and prints the mean clock64() deltas between the phase boundaries thread 0 of every CTA
recorded (see CTA_TICK in mcov_dec_cta.cu).  usage (GPU box): MCOV_NVCC_EXTRA=-DMCOV_CTA_TICKS python
marl_coverage_b200/build.py && python tools/cta_ticks.py [workload]"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

import bench

NAMES = ["pass start", "poses+actions arrived", "wall bits arrived", "moves resolved", "staged runs landed (after barrier)",
         "rows done (barrier)", "bookkeeping done (barrier)", "obs stores issued"]


def main():
    wl = sys.argv[1] if len(sys.argv) > 1 else "c4"
    kind, grid_name, E, cfg, _ = bench.WORKLOADS[wl]
    dev = torch.device("cuda:0")
    env, n_maps = bench.make_env(kind, grid_name, cfg, E, dev, 0, auto_reset=False)
    acts = bench.gen_actions(torch, dev, 32, E, cfg["numrobot"], 7)
    env.reset(map_idx=torch.arange(E, device=dev, dtype=torch.int32) % n_maps)
    flush = torch.zeros(64 << 20, dtype=torch.float32, device=dev)
    rows = []
    for k in range(24):
        flush.add_(1)
        env.step(acts[k % 32])
        torch.cuda.synchronize()
        off = env.layout.off_scratch
        t = env._arena[off:off + E * 64].view(torch.int64).view(E, 8).cpu().numpy()
        if k >= 4:
            rows.append(t)
    t = np.stack(rows).astype(np.float64)                          # [steps, CTAs, 8]
    d = np.diff(t, axis=2)
    out = {"workload": wl, "ctas": int(E), "unit": "SM clock cycles (1965 MHz)",
           "phases": {NAMES[k + 1]: {"mean": float(d[..., k].mean()), "p90": float(np.percentile(d[..., k], 90))} for k in range(7)},
           "pass_total_mean": float((t[..., 7] - t[..., 0]).mean())}
    print(json.dumps(out))


if __name__ == "__main__":
    main()
