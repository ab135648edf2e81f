#!/usr/bin/env python
"""Times the UNMODIFIED reference env and the numpy oracle port side by side (build container only; the reference
cannot travel to the GPU box, so bench.py's CPU arm runs the port — this record says how the two relate).

Same settings, maps, action distribution and reset rule for both (random base-4 actions, reset on done /
train_maxsteps, stdout silenced): C1 = SuperGridRL 4 robots bg2_100x100, C2 = DecGridRL 4 robots square r=2,
C3 = DecGridRL 8 robots lidar 21x10 on bg2_256x256; one process, then one process per core.

  python tools/cpu_reference_vs_port.py [--seconds 6]   ->  profiles/cpu_reference_vs_port.json
"""
import argparse
import json
import multiprocessing as mp
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))

import bench                                                # noqa: E402

CASES = {"c1_super": "super", "c2_dec_square": "c2", "c3_dec_lidar": "c3"}


def worker(args):
    impl, workload, seconds, seed = args
    import contextlib
    import io
    import numpy as np
    kind, grid_name, _, cfg, _ = bench.WORKLOADS[workload]
    from marl_coverage_b200.maps import load_grid_set
    grids = load_grid_set(grid_name)
    train = grids[:len(grids) // 2]
    np.random.seed(seed)
    rng = np.random.default_rng(seed)
    N = cfg["numrobot"]
    with contextlib.redirect_stdout(io.StringIO()):
        if impl == "reference":
            import ref_harness as rh
            ns = rh.load_reference()
            env = ns.DecGridRL(train, cfg) if kind == "dec" else ns.SuperGridRL(train, cfg)
        else:
            from oracle import coverage_oracle as oc
            env = oc.OracleDecGridRL(train, cfg) if kind == "dec" else oc.OracleSuperGridRL(train, cfg)
        t_warm = time.perf_counter() + 1.0
        steps, t0 = 0, None
        while True:
            _, _, done = env.step(int(rng.integers(0, 4 ** N)))
            if done or env._currstep >= cfg["train_maxsteps"]:
                env.reset(False, None)
            now = time.perf_counter()
            if t0 is None:
                if now >= t_warm:
                    t0 = now
                continue
            steps += 1
            if now - t0 >= seconds:
                return steps * N / (now - t0)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--seconds", type=float, default=6.0)
    args = ap.parse_args()
    cores = len(os.sched_getaffinity(0))
    ctx = mp.get_context("spawn")
    out = {"cores": cores, "seconds_per_sample": args.seconds, "unit": "agent-steps/s", "cases": {},
           "how": "tools/cpu_reference_vs_port.py: unmodified reference env (imported from /root/reference with pygame / "
                  "matplotlib stubs) vs oracle/coverage_oracle.py, same maps / settings / random base-4 actions / reset rule"}
    ratios = []
    for name, wl in CASES.items():
        row = {}
        for procs in (1, cores):
            for impl in ("reference", "port"):
                with ctx.Pool(procs) as pool:
                    rates = pool.map(worker, [(impl, wl, args.seconds, 100 + i) for i in range(procs)])
                row["%s_%dproc" % (impl, procs)] = sum(rates)
            row["port_over_reference_%dproc" % procs] = row["port_%dproc" % procs] / row["reference_%dproc" % procs]
        ratios.append(row["port_over_reference_%dproc" % cores])
        out["cases"][name] = row
        print(name, {k: round(v, 2) for k, v in row.items()})
    out["summary"] = {"port_over_reference": out["cases"]["c2_dec_square"]["port_over_reference_%dproc" % cores],
                      "port_over_reference_by_case": {n: out["cases"][n]["port_over_reference_%dproc" % cores] for n in CASES},
                      "note": "bench.py's CPU arm runs the port on the GPU box; the unmodified reference is slower by this "
                              "factor, so every GPU / CPU ratio built on the port is a lower bound by the same factor "
                              "(headline workload c5 uses the c2 settings)"}
    with open(os.path.join(ROOT, "profiles", "cpu_reference_vs_port.json"), "w") as f:
        json.dump(out, f, indent=1)
    return 0


if __name__ == "__main__":
    sys.exit(main())
