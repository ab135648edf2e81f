"""Sensor plugin point of the reference (Environments/Sensors/Sensor.py:4-9) on the B200 build.

In the reference a sensor object computes ``getMeasurement(x, y, oc, free_map, obst_map, pad)`` in
Python for one robot.  Here the two shipped sensors are *descriptors*: they carry the reference's
``sensor_config`` and select the fused in-kernel implementation (square footprint copy / float64
DDA ray march); the arithmetic runs inside ``dec_env_kernel``.  ``getMeasurement`` is kept so code
that probes a sensor directly still works — it drives a one-env device kernel, never a CPU path.
"""
import numpy as np


class Sensor(object):
    """Interface — Environments/Sensors/Sensor.py:4-9."""

    sensor_type = None

    def __init__(self):
        super().__init__()

    def getMeasurement(self, x, y, oc, free_map, obst_map, pad):
        raise NotImplementedError()


def lidar_beam_table(num_lasers):
    """float64 [num_lasers, 3] = (xinc, yinc, distinc) per laser, built with the same numpy calls
    as lidar.py:14 (linspace) and lidar.py:39-48 (cos/sin, normalise by the larger component,
    sqrt of the squares) so the device marches with bit-identical increments."""
    thetas = np.linspace(0, 2 * np.pi, num=num_lasers, endpoint=False)
    table = np.empty((num_lasers, 3), dtype=np.float64)
    for k, theta in enumerate(thetas):
        xinc = np.cos(theta)
        yinc = np.sin(theta)
        larger = max(abs(xinc), abs(yinc))
        xinc /= larger
        yinc /= larger
        table[k, 0] = xinc
        table[k, 1] = yinc
        table[k, 2] = np.sqrt(xinc ** 2 + yinc ** 2)
    return table


class _DeviceSensor(Sensor):
    def env_keys(self):
        raise NotImplementedError

    def getMeasurement(self, x, y, oc, free_map, obst_map, pad):
        """Single measurement on the device: builds a one-robot, one-env DecGridRL over ``oc``
        (which the caller has already bordered, as DecGridRL.reset does), injects the pose, and
        ORs the sensed cells into copies of the caller's padded maps."""
        from .vec_env import VecDecGridRL
        oc = np.asarray(oc)
        inner = oc[1:-1, 1:-1]
        cfg = dict(numrobot=1, maxsteps=0, collision_penalty=0, done_thresh=2, done_incr=0,
                   terminal_reward=0, dist_reward=0, train_maxsteps=0, test_maxsteps=0, egoradius=0,
                   mini_map_rad=0, comm_radius=0, allow_comm=0, map_sharing=0, single_square_tool=0,
                   dijkstra_input=0)
        cfg.update(self.env_keys())
        env = VecDecGridRL([inner], cfg, num_envs=1)
        env.reset(map_idx=np.zeros(1, np.int32), start_xy=np.array([[[x, y]]], np.int16))
        free, obst, _ = env.planes()
        new_free = np.copy(free_map)
        new_obst = np.copy(obst_map)
        W, L = oc.shape
        new_free[pad:pad + W, pad:pad + L] = np.maximum(new_free[pad:pad + W, pad:pad + L], free[0, 0])
        new_obst[pad:pad + W, pad:pad + L] = np.maximum(new_obst[pad:pad + W, pad:pad + L], obst[0, 0])
        return new_free, new_obst


class SquareSensor(_DeviceSensor):
    """Chebyshev square footprint — Environments/Sensors/squaresensor.py:4-37."""
    sensor_type = "square_sensor"

    def __init__(self, sensor_config):
        super().__init__()
        self._radius = sensor_config["range"]

    def env_keys(self):
        return {"sensor_type": "square_sensor", "sensor_config": {"range": self._radius}}


class LidarSensor(_DeviceSensor):
    """Equally spaced DDA rays — Environments/Sensors/lidar.py:5-68."""
    sensor_type = "lidar"

    def __init__(self, sensor_config):
        super().__init__()
        self._num_lasers = sensor_config["num_lasers"]
        self._max_range = sensor_config["range"]
        assert self._num_lasers % 2 == 1, "odd number of lasers needed"
        self._thetalist = np.linspace(0, 2 * np.pi, num=self._num_lasers, endpoint=False)

    def env_keys(self):
        return {"sensor_type": "lidar",
                "sensor_config": {"range": self._max_range, "num_lasers": self._num_lasers}}
