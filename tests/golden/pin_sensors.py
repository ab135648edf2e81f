"""Pin the sensor plugin point (Environments/Sensors/Sensor.py:8) against the live reference (build container only).

Runs the unmodified reference SquareSensor / LidarSensor ``getMeasurement(x, y, oc, free_map, obst_map, pad)`` on
random bordered occupancy grids, poses (walkable cells and, for a few, obstacle cells), paddings and STALE input maps
(random 0/1 — the square sensor overwrites its window, the lidar only sets cells), checks that the oracle sensors
(oracle.OracleSquareSensor / OracleLidarSensor ``footprint``) give the same maps, and records inputs + the reference's
outputs into tests/golden/sensor_probe.npz for the GPU test of the product sensors (tests/test_sensors_gpu.py).

Usage:  python tests/golden/pin_sensors.py [--check]
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, HERE)
sys.path.insert(0, ROOT)

import ref_harness as rh                                   # noqa: E402
from oracle import coverage_oracle as oc                   # noqa: E402

OUT = os.path.join(HERE, "sensor_probe.npz")
MAXW = 40                                                   # cases are stored in fixed [MAXW + 2*MAXPAD]^2 frames
MAXPAD = 6


def make_cases(rng, n):
    cases = []
    for k in range(n):
        W, L = int(rng.integers(5, MAXW + 1)), int(rng.integers(5, MAXW + 1))
        grid = np.where(rng.random((W, L)) < rng.choice([0.0, 0.15, 0.35]), -1.0, 1.0)
        if k % 5 == 0:
            grid = np.where((grid > 0) & (rng.random((W, L)) < 0.2), 0.0, grid)      # walkable, never known-free
        if k % 7 == 0:
            grid = np.where(grid > 0, rng.choice([1.0, 2.0, 3.5], size=(W, L)), grid * rng.choice([1.0, 4.0], size=(W, L)))
        grid[0, :] = grid[-1, :] = grid[:, 0] = grid[:, -1] = -1.0                   # DecGridRL.reset's ring (:471-472)
        lidar = bool(k % 2)
        if lidar:
            sc = {"num_lasers": int(rng.choice([1, 5, 9, 21, 33])), "range": int(rng.choice([0, 1, 3, 6, 10]))}
        else:
            sc = {"range": int(rng.choice([0, 1, 2, 4, 7]))}
        walk = np.argwhere(grid >= 0)
        if k % 9 == 4 or len(walk) == 0:
            x, y = int(rng.integers(0, W)), int(rng.integers(0, L))                  # any cell, obstacles and ring included
        else:
            x, y = (int(v) for v in walk[rng.integers(len(walk))])
        pad = int(rng.choice([0, 1, 3, MAXPAD]))
        stale = rng.random() < 0.6
        fm = (rng.random((W + 2 * pad, L + 2 * pad)) < 0.3).astype(np.float64) if stale else np.zeros((W + 2 * pad, L + 2 * pad))
        om = (rng.random((W + 2 * pad, L + 2 * pad)) < 0.3).astype(np.float64) if stale else np.zeros((W + 2 * pad, L + 2 * pad))
        cases.append(dict(grid=grid, lidar=lidar, sc=sc, x=x, y=y, pad=pad, free=fm, obst=om))
    return cases


def oracle_measure(c):
    """The oracle's footprint applied the way each reference sensor applies it."""
    g = c["grid"]
    W, L = g.shape
    pad = c["pad"]
    nf, no = c["free"].copy(), c["obst"].copy()
    if c["lidar"]:
        fr, ob = oc.OracleLidarSensor(c["sc"]).footprint(c["x"], c["y"], g)
        nf[pad:pad + W, pad:pad + L][fr] = 1
        no[pad:pad + W, pad:pad + L][ob] = 1
    else:
        r = c["sc"]["range"]
        fr, ob = oc.OracleSquareSensor(c["sc"]).footprint(c["x"], c["y"], g)
        xa, xb = max(c["x"] - r, 0), min(c["x"] + r + 1, W)
        ya, yb = max(c["y"] - r, 0), min(c["y"] + r + 1, L)
        nf[xa + pad:xb + pad, ya + pad:yb + pad] = fr[xa:xb, ya:yb]
        no[xa + pad:xb + pad, ya + pad:yb + pad] = ob[xa:xb, ya:yb]
    return nf, no


def main():
    ns = rh.load_reference()
    cases = make_cases(np.random.default_rng(20251105), 120)
    F = MAXW + 2 * MAXPAD
    n = len(cases)
    rec = dict(grid=np.zeros((n, MAXW, MAXW)), shape=np.zeros((n, 2), np.int32), lidar=np.zeros(n, np.uint8),
               rng_=np.zeros(n, np.int32), lasers=np.zeros(n, np.int32), xy=np.zeros((n, 2), np.int32),
               pad=np.zeros(n, np.int32), free_in=np.zeros((n, F, F), np.uint8), obst_in=np.zeros((n, F, F), np.uint8),
               free_out=np.zeros((n, F, F), np.uint8), obst_out=np.zeros((n, F, F), np.uint8))
    bad = 0
    for k, c in enumerate(cases):
        sensor = ns.LidarSensor(c["sc"]) if c["lidar"] else ns.SquareSensor(c["sc"])
        rf, ro = sensor.getMeasurement(c["x"], c["y"], c["grid"], c["free"], c["obst"], c["pad"])
        of, oo = oracle_measure(c)
        # the reference's square sensor stores clip(cell, 0, 1): all cells here have magnitude 0 or >= 1 -> 0 / 1
        assert set(np.unique(rf)) <= {0.0, 1.0} and set(np.unique(ro)) <= {0.0, 1.0}
        if not (np.array_equal(rf, of) and np.array_equal(ro, oo)):
            bad += 1
            print("case %d MISMATCH (%s %s at %s pad %d)" % (k, "lidar" if c["lidar"] else "square", c["sc"],
                                                            (c["x"], c["y"]), c["pad"]))
        W, L = c["grid"].shape
        P = c["pad"]
        rec["grid"][k, :W, :L] = c["grid"]
        rec["shape"][k] = (W, L)
        rec["lidar"][k] = c["lidar"]
        rec["rng_"][k] = c["sc"]["range"]
        rec["lasers"][k] = c["sc"].get("num_lasers", 0)
        rec["xy"][k] = (c["x"], c["y"])
        rec["pad"][k] = P
        rec["free_in"][k, :W + 2 * P, :L + 2 * P] = c["free"]
        rec["obst_in"][k, :W + 2 * P, :L + 2 * P] = c["obst"]
        rec["free_out"][k, :W + 2 * P, :L + 2 * P] = rf
        rec["obst_out"][k, :W + 2 * P, :L + 2 * P] = ro
    print("sensor pin: %d cases, %d mismatches against the oracle sensors" % (n, bad))
    if "--check" not in sys.argv[1:] and bad == 0:
        np.savez_compressed(OUT, **rec)
    return 0 if bad == 0 else 1


if __name__ == "__main__":
    sys.exit(main())
