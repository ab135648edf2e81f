"""Sensor plugin point on the device: SquareSensor / LidarSensor.getMeasurement of the product (the fused sensing code
of dec_env_kernel behind the reference's interface, Sensor.py:8) against the maps the live reference's sensors
returned (tests/golden/sensor_probe.npz) and against the oracle sensors on further random probes."""
import numpy as np
import pytest

from oracle import coverage_oracle as oc
from test_sensors_cpu import load_probe_cases

pytestmark = pytest.mark.gpu


def _sensor(c):
    from marl_coverage_b200 import LidarSensor, SquareSensor
    return LidarSensor(c["sc"]) if c["lidar"] else SquareSensor(c["sc"])


@pytest.mark.parametrize("fp64_march", [False, True])
def test_get_measurement_matches_reference_recordings(fp64_march, monkeypatch):
    if fp64_march:
        monkeypatch.setenv("MCOV_LIDAR_FP64", "1")           # the float64 march instead of the integer beam table
    cases = load_probe_cases()
    for k, c in enumerate(cases):
        if fp64_march and not c["lidar"]:
            continue
        f0, o0 = c["free"].copy(), c["obst"].copy()
        nf, no = _sensor(c).getMeasurement(c["x"], c["y"], c["grid"], c["free"], c["obst"], c["pad"])
        assert np.array_equal(c["free"], f0) and np.array_equal(c["obst"], o0)       # inputs are copied, not mutated
        assert nf.dtype == c["free"].dtype and nf.shape == c["free"].shape
        assert np.array_equal(nf, c["free_out"]), "free map, case %d (%s %s)" % (k, "lidar" if c["lidar"] else "square", c["sc"])
        assert np.array_equal(no, c["obst_out"]), "obstacle map, case %d" % k


def test_footprints_vs_oracle_one_grid_many_poses():
    """One sensor object probing the same grid from every walkable cell (the device env is reused)."""
    from marl_coverage_b200 import LidarSensor, SquareSensor
    rng = np.random.default_rng(4)
    g = np.where(rng.random((19, 23)) < 0.2, -1.0, 1.0)
    g[0, :] = g[-1, :] = g[:, 0] = g[:, -1] = -1.0
    for mine, theirs in ((SquareSensor({"range": 3}), oc.OracleSquareSensor({"range": 3})),
                         (LidarSensor({"range": 5, "num_lasers": 21}), oc.OracleLidarSensor({"range": 5, "num_lasers": 21}))):
        for x, y in np.argwhere(g >= 0)[::7]:
            fr, ob = mine.footprint(int(x), int(y), g)
            wf, wo = theirs.footprint(int(x), int(y), g)
            assert np.array_equal(fr, wf) and np.array_equal(ob, wo), (type(mine).__name__, x, y)


def test_unsupported_probes_are_refused():
    from marl_coverage_b200 import LidarSensor, SquareSensor
    g = np.ones((8, 8))                                       # no obstacle border
    with pytest.raises(ValueError, match="border"):
        LidarSensor({"range": 3, "num_lasers": 5}).getMeasurement(3, 3, g, np.zeros((8, 8)), np.zeros((8, 8)), 0)
    with pytest.raises(ValueError, match="outside"):
        SquareSensor({"range": 1}).getMeasurement(8, 3, g, np.zeros((8, 8)), np.zeros((8, 8)), 0)
    # the square sensor sees through everything and needs no border (squaresensor.py:22-31 clamps the window)
    nf, no = SquareSensor({"range": 2}).getMeasurement(0, 7, g, np.zeros((12, 12)), np.ones((12, 12)), 2)
    assert nf[2:5, 7:10].all() and nf.sum() == 9 and no[2:5, 7:10].sum() == 0 and no.sum() == 144 - 9
