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
    # squaresensor.py:34-35 ASSIGNS the window (stale cells inside it are cleared), lidar.py:52-63 only sets cells
    OVERWRITES_WINDOW = False

    def env_keys(self):
        raise NotImplementedError

    def _device_env(self, oc):
        """One-robot, one-env DecGridRL over ``oc`` on the device, kept while the caller keeps probing the same grid."""
        from .vec_env import VecDecGridRL
        key = (oc.shape, oc.tobytes())
        cached = getattr(self, "_probe", None)
        if cached is not None and cached[0] == key:
            return cached[1]
        cfg = dict(numrobot=1, maxsteps=0, collision_penalty=0, done_thresh=2, done_incr=0,
                   terminal_reward=0, dist_reward=0, train_maxsteps=0, test_maxsteps=0, egoradius=0,
                   mini_map_rad=0, comm_radius=0, allow_comm=0, map_sharing=0, single_square_tool=0,
                   dijkstra_input=0)
        cfg.update(self.env_keys())
        env = VecDecGridRL([oc], cfg, num_envs=1)
        self._probe = (key, env)
        return env

    def footprint(self, x, y, oc):
        """(free_hits, obst_hits) boolean planes the size of ``oc`` for one measurement from cell (x, y), computed
        by the fused sensing code of ``dec_env_kernel`` on the device (never on the CPU).  The device env surrounds
        ``oc`` with the obstacle ring DecGridRL.reset adds, so the pose shifts by one and the ring is cut off again."""
        oc = np.ascontiguousarray(np.asarray(oc, dtype=np.float64))
        W, L = oc.shape
        if not (0 <= int(x) < W and 0 <= int(y) < L):
            raise ValueError("sensor pose (%d, %d) outside the %d x %d grid" % (x, y, W, L))
        env = self._device_env(oc)
        env.reset(map_idx=np.zeros(1, np.int32), start_xy=np.array([[[int(x) + 1, int(y) + 1]]], np.int16))
        free, obst, _ = env.planes()
        return free[0, 0, 1:-1, 1:-1], obst[0, 0, 1:-1, 1:-1]

    def getMeasurement(self, x, y, oc, free_map, obst_map, pad):
        """Sensor.getMeasurement (Sensor.py:8): copies of the caller's padded maps with one measurement from
        (x, y) on occupancy grid ``oc`` applied — the square sensor assigns its window, the lidar sets the cells
        its rays touch — values and dtype as the reference writes them (1 / 0 floats)."""
        oc = np.asarray(oc)
        W, L = oc.shape
        fr, ob = self.footprint(x, y, oc)
        new_free = np.copy(free_map)
        new_obst = np.copy(obst_map)
        if self.OVERWRITES_WINDOW:
            r = self._radius
            xa, xb = max(x - r, 0), min(x + r + 1, W)
            ya, yb = max(y - r, 0), min(y + r + 1, L)
            new_free[xa + pad:xb + pad, ya + pad:yb + pad] = fr[xa:xb, ya:yb]
            new_obst[xa + pad:xb + pad, ya + pad:yb + pad] = ob[xa:xb, ya:yb]
        else:
            new_free[pad:pad + W, pad:pad + L][fr] = 1
            new_obst[pad:pad + W, pad:pad + L][ob] = 1
        return new_free, new_obst


class SquareSensor(_DeviceSensor):
    """Chebyshev square footprint — Environments/Sensors/squaresensor.py:4-37."""
    sensor_type = "square_sensor"
    OVERWRITES_WINDOW = True

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

    def footprint(self, x, y, oc):
        oc = np.asarray(oc)
        ring = np.concatenate([oc[0], oc[-1], oc[:, 0], oc[:, -1]])
        if np.any(ring >= 0):
            # lidar.py:58-63 indexes with int() of a coordinate that has LEFT the grid (truncation toward zero,
            # negative indices wrapping around); DecGridRL never gets there because reset() borders the map
            raise ValueError("LidarSensor needs an occupancy grid closed by an obstacle border (DecGridRL.reset adds "
                             "one, dec_grid_rl.py:471-472): a ray leaving the grid is outside the supported path")
        return super().footprint(x, y, oc)

    def env_keys(self):
        return {"sensor_type": "lidar",
                "sensor_config": {"range": self._max_range, "num_lasers": self._num_lasers}}
