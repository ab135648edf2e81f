"""Sensor plugin point (Environments/Sensors/Sensor.py:8), CPU side: the oracle sensors against the outputs the live
reference's SquareSensor / LidarSensor.getMeasurement produced (tests/golden/sensor_probe.npz, recorded by
tests/golden/pin_sensors.py), and the pin itself re-run wherever /root/reference is present."""
import os
import subprocess
import sys

import numpy as np
import pytest

from conftest import GOLDEN_DIR

sys.path.insert(0, GOLDEN_DIR)


def load_probe_cases():
    z = np.load(os.path.join(GOLDEN_DIR, "sensor_probe.npz"))
    cases = []
    for k in range(z["shape"].shape[0]):
        W, L = (int(v) for v in z["shape"][k])
        P = int(z["pad"][k])
        lidar = bool(z["lidar"][k])
        sc = {"range": int(z["rng_"][k])}
        if lidar:
            sc["num_lasers"] = int(z["lasers"][k])
        cases.append(dict(grid=z["grid"][k, :W, :L], lidar=lidar, sc=sc, x=int(z["xy"][k, 0]), y=int(z["xy"][k, 1]),
                          pad=P, free=z["free_in"][k, :W + 2 * P, :L + 2 * P].astype(np.float64),
                          obst=z["obst_in"][k, :W + 2 * P, :L + 2 * P].astype(np.float64),
                          free_out=z["free_out"][k, :W + 2 * P, :L + 2 * P].astype(np.float64),
                          obst_out=z["obst_out"][k, :W + 2 * P, :L + 2 * P].astype(np.float64)))
    return cases


def test_oracle_sensors_match_reference_recordings():
    import pin_sensors
    cases = load_probe_cases()
    assert len(cases) >= 100
    assert any(c["lidar"] for c in cases) and any(not c["lidar"] for c in cases)
    assert any(c["grid"][c["x"], c["y"]] < 0 for c in cases)            # probes standing on an obstacle
    assert any(c["free"].any() for c in cases)                          # stale input maps
    for k, c in enumerate(cases):
        nf, no = pin_sensors.oracle_measure(c)
        assert np.array_equal(nf, c["free_out"]) and np.array_equal(no, c["obst_out"]), "case %d" % k


def test_sensor_pin_live():
    ref = os.environ.get("MCOV_REFERENCE_ROOT", "/root/reference")
    if not os.path.isdir(os.path.join(ref, "Environments")):
        pytest.skip("reference not present")
    res = subprocess.run([sys.executable, os.path.join(GOLDEN_DIR, "pin_sensors.py"), "--check"],
                         capture_output=True, text=True, timeout=300)
    assert res.returncode == 0 and " 0 mismatches" in res.stdout, res.stdout[-1500:] + res.stderr[-1500:]
