"""CPU oracle for the MARL-Coverage env-step hot path (TEST INFRASTRUCTURE — not product code).

A from-scratch numpy restatement of what the reference computes in
``Environments/dec_grid_rl.py``, ``Environments/super_grid_rl.py``,
``Environments/Sensors/{squaresensor,lidar}.py`` and ``Environments/dijkstra.py`` of
ExistentialRobotics/MARL-Coverage.  Only ``tests/``, ``__graft_entry__.smoke()`` and the
``cpu_baseline`` / ``--impl reference`` legs of ``bench.py`` may import it; the product path
(``marl_coverage_b200``) never does and fails loudly without its CUDA library.

Parity status: PINNED.  ``tests/golden/pin_oracle.py`` runs this file step-for-step against the live,
unmodified reference (imported from /root/reference in the build container) over the config matrix
listed in ``tests/golden/PIN_REPORT.json``; ``tests/golden/make_golden.py`` records reference traces into
``tests/golden/*.npz`` and ``tests/test_oracle_golden.py`` replays them against this file on every
CPU test run.  The reference itself ships no tests for this path.

State representation differs from the reference on purpose (it is a restatement, not a copy):
planes are boolean and un-padded (bordered-map sized); windows are cut with explicit zero fill
instead of pre-padding.  ``pad`` only survives where the reference's padding is observable
(distance-map quirk, dijkstra domain).

The global ``np.random`` stream is consumed in exactly the reference's order
(map draw, then x, y rejection sampling) so a seeded oracle env and a seeded reference env
walk the same episodes.
"""
import heapq

import numpy as np

# --------------------------------------------------------------------------------------
# shared helpers
# --------------------------------------------------------------------------------------

_FLT_MAX = np.float32(3.4028235e+38)


def l1_distance_to_zero(mask_nonzero):
    """Exact L1 (taxicab) distance from every True cell to the nearest False cell, float32.

    Restates ``cv2.distanceTransform(img, cv2.DIST_L1, cv2.DIST_MASK_PRECISE)`` as called at
    super_grid_rl.py:295 / dec_grid_rl.py:274 (OpenCV is an un-vendored, un-pinned dependency of
    the reference; 4.13.0 in this image).  Two-pass chamfer with unit steps is exact for L1.
    When the image has no False cell OpenCV returns FLT_MAX everywhere; so do we.
    """
    m = np.asarray(mask_nonzero, dtype=bool)
    if m.all():
        return np.full(m.shape, _FLT_MAX, dtype=np.float32)
    big = np.int64(m.shape[0] + m.shape[1] + 2)
    d = np.where(m, big, 0).astype(np.int64)
    # along axis 1 (forward / backward), vectorised over axis 0
    for k in range(1, d.shape[1]):
        np.minimum(d[:, k], d[:, k - 1] + 1, out=d[:, k])
    for k in range(d.shape[1] - 2, -1, -1):
        np.minimum(d[:, k], d[:, k + 1] + 1, out=d[:, k])
    # along axis 0
    for k in range(1, d.shape[0]):
        np.minimum(d[k], d[k - 1] + 1, out=d[k])
    for k in range(d.shape[0] - 2, -1, -1):
        np.minimum(d[k], d[k + 1] + 1, out=d[k])
    return d.astype(np.float32)


def inverted_distance_channel(known_mask_is_zero_pixel):
    """``1 - d/max(d)`` in float32 — dec_grid_rl.py:273-282 / super_grid_rl.py:294-303.

    ``known_mask_is_zero_pixel``: boolean plane, True where the reference's ``free`` plane is
    non-zero (those become the zero pixels of the transform).
    """
    d = l1_distance_to_zero(~np.asarray(known_mask_is_zero_pixel, dtype=bool))
    mx = d.max()
    if mx > 0:
        d = d / mx          # float32 / float32
    return np.float32(1) - d


def window(plane, cx, cy, rad, dtype=np.float64):
    """(2*rad+1)^2 window of ``plane`` centred on (cx, cy), zeros outside the plane.

    Equivalent of DecGridRL.arraySubset (dec_grid_rl.py:393-421) on a zero-padded plane."""
    n = 2 * rad + 1
    out = np.zeros((n, n), dtype=dtype)
    W, L = plane.shape
    x0, y0 = cx - rad, cy - rad
    xa, xb = max(x0, 0), min(x0 + n, W)
    ya, yb = max(y0, 0), min(y0 + n, L)
    if xa < xb and ya < yb:
        out[xa - x0:xb - x0, ya - y0:yb - y0] = plane[xa:xb, ya:yb]
    return out


def bilinear_resize(src, out_n):
    """Square bilinear resize, OpenCV INTER_LINEAR conventions (pixel centres, edge clamp).

    Restates ``cv2.resize(..., interpolation=cv2.INTER_LINEAR)`` at dec_grid_rl.py:365-370.
    OpenCV's coefficient precision is version dependent, hence parity 1e-6.
    """
    src = np.asarray(src, dtype=np.float64)
    n_in = src.shape[0]
    scale = n_in / out_n
    idx = np.empty(out_n, dtype=np.int64)
    frac = np.empty(out_n, dtype=np.float64)
    for d in range(out_n):
        f = (d + 0.5) * scale - 0.5
        s = int(np.floor(f))
        f -= s
        if s < 0:
            s, f = 0, 0.0
        if s >= n_in - 1:
            s, f = n_in - 1, 0.0
        idx[d], frac[d] = s, np.float32(f)
    idx1 = np.minimum(idx + 1, n_in - 1)
    # cv2 dsize/axes: output[r, c] interpolates rows with the r-coefficients, cols with c's
    rows = src[idx, :] * (1 - frac)[:, None] + src[idx1, :] * frac[:, None]
    return rows[:, idx] * (1 - frac)[None, :] + rows[:, idx1] * frac[None, :]


def dijkstra_path_plane(code, sx, sy):
    """0/1 plane holding a shortest path from (sx, sy) to the nearest unknown cell.

    ``code``: int plane, 1 explored-free, -1 obstacle, 0 unknown (Environments/dijkstra.py:112-187).
    Unit edge costs, so the reference's PriorityQueue of ``(cost, (x, y))`` pops cells in
    (cost, x, y) order; the first popped unknown cell is the goal.  Back-trace takes the first
    neighbour in order x+1, x-1, y+1, y-1 whose cost is exactly one less.
    """
    W, L = code.shape
    cost = -np.ones((W, L), dtype=np.int64)
    seen = np.zeros((W, L), dtype=bool)
    heap = [(0, (sx, sy))]
    goal = None
    nbrs = ((1, 0), (-1, 0), (0, 1), (0, -1))
    while heap:
        c, (x, y) = heapq.heappop(heap)
        if seen[x, y]:
            continue
        seen[x, y] = True
        cost[x, y] = c
        if code[x, y] == 0:
            goal = (x, y)
            break
        for dx, dy in nbrs:
            u, v = x + dx, y + dy
            if 0 <= u < W and 0 <= v < L and not seen[u, v] and code[u, v] != -1:
                heapq.heappush(heap, (c + 1, (u, v)))
    path = np.zeros((W, L), dtype=np.float64)
    if goal is None:
        return path
    x, y = goal
    c = cost[x, y]
    path[x, y] = 1
    while (x, y) != (sx, sy):
        for dx, dy in nbrs:
            u, v = x + dx, y + dy
            # the reference reuses get_valid_neighbors with visited inverted: only finalised cells
            if 0 <= u < W and 0 <= v < L and seen[u, v] and code[u, v] != -1 \
                    and cost[u, v] == c - 1:
                x, y, c = u, v, c - 1
                break
        path[x, y] = 1
    return path


def decode_base4(action, n):
    """Joint integer -> per-robot digits, least significant digit = robot 0
    (dec_grid_rl.py:110-115, super_grid_rl.py:93-98)."""
    a = int(action)
    out = np.zeros(n, dtype=np.int64)
    for i in range(n):
        out[i] = a % 4
        a //= 4
    return out


def _is_sentinel(action):
    if action is None:
        return True
    if isinstance(action, (int, np.integer)):
        return int(action) == -1
    if hasattr(action, "ndim") and getattr(action, "ndim") == 0:
        return int(action) == -1
    return False


# --------------------------------------------------------------------------------------
# sensors (plugin interface of Environments/Sensors/Sensor.py:8)
# --------------------------------------------------------------------------------------

class OracleSquareSensor:
    """Chebyshev square footprint that sees through walls — squaresensor.py:15-37."""

    def __init__(self, sensor_config):
        self._radius = sensor_config["range"]

    def footprint(self, x, y, grid):
        """Returns (free_hits, obst_hits) boolean planes the size of ``grid``."""
        W, L = grid.shape
        r = self._radius
        xa, xb = max(x - r, 0), min(x + r + 1, W)
        ya, yb = max(y - r, 0), min(y + r + 1, L)
        fr = np.zeros((W, L), dtype=bool)
        ob = np.zeros((W, L), dtype=bool)
        fr[xa:xb, ya:yb] = grid[xa:xb, ya:yb] > 0
        ob[xa:xb, ya:yb] = grid[xa:xb, ya:yb] < 0
        return fr, ob


class OracleLidarSensor:
    """DDA ray march in float64 — lidar.py:6-65."""

    def __init__(self, sensor_config):
        self._num_lasers = sensor_config["num_lasers"]
        self._max_range = sensor_config["range"]
        assert self._num_lasers % 2 == 1, "odd number of lasers needed"
        self.beams = lidar_beam_table(self._num_lasers)

    def footprint(self, x, y, grid):
        W, L = grid.shape
        fr = np.zeros((W, L), dtype=bool)
        ob = np.zeros((W, L), dtype=bool)
        rng = self._max_range
        for xinc, yinc, dinc in self.beams:
            cx, cy, dist = np.float64(x), np.float64(y), np.float64(0)
            while True:
                inside = cx >= 0 and cy >= 0 and cx < W and cy < L
                if not (inside and grid[int(cx), int(cy)] >= 0 and dist < rng):
                    break
                fr[int(cx), int(cy)] = True
                cx += xinc
                cy += yinc
                dist += dinc
            inside = cx >= 0 and cy >= 0 and cx < W and cy < L
            if inside and grid[int(cx), int(cy)] >= 0:
                fr[int(cx), int(cy)] = True
            else:
                # NB the reference indexes with int() even when outside; with the obstacle
                # border DecGridRL adds (dec_grid_rl.py:471) a ray always ends inside the map.
                ob[int(cx), int(cy)] = True
        return fr, ob


def lidar_beam_table(num_lasers):
    """Per-laser (xinc, yinc, distinc) float64 triples exactly as lidar.py:14,39-48 builds them."""
    thetas = np.linspace(0, 2 * np.pi, num=num_lasers, endpoint=False)
    out = np.empty((num_lasers, 3), dtype=np.float64)
    for k, th in enumerate(thetas):
        xinc = np.cos(th)
        yinc = np.sin(th)
        larger = max(abs(xinc), abs(yinc))
        xinc /= larger
        yinc /= larger
        out[k] = (xinc, yinc, np.sqrt(xinc ** 2 + yinc ** 2))
    return out


# --------------------------------------------------------------------------------------
# DecGridRL
# --------------------------------------------------------------------------------------

class OracleDecGridRL:
    """Decentralised partially observed coverage env — restates dec_grid_rl.py:21-552."""

    # Dec move table (dec_grid_rl.py:131-145): 0:x+1 1:y+1 2:x-1 3:y-1
    MOVES = ((1, 0), (0, 1), (-1, 0), (0, -1))

    def __init__(self, train_set, env_config, use_graph=False, test_set=None):
        c = env_config
        self._train_gridlis = train_set
        self._test_gridlis = test_set
        self._numrobot = c["numrobot"]
        self._maxsteps = c["maxsteps"]
        self._collision_penalty = c["collision_penalty"]
        self._done_thresh = c["done_thresh"]
        self._done_incr = c["done_incr"]
        self._terminal_reward = c["terminal_reward"]
        self._dist_r = c["dist_reward"]
        self._train_maxsteps = c["train_maxsteps"]
        self._test_maxsteps = c["test_maxsteps"]
        self._egoradius = c["egoradius"]
        self._mini_map_rad = c["mini_map_rad"]
        self._comm_radius = c["comm_radius"]
        self._allow_comm = c["allow_comm"]
        self._map_sharing = c["map_sharing"]
        self._use_graph = use_graph
        self._single_square_tool = c["single_square_tool"]
        self._dijkstra_input = c["dijkstra_input"]
        if c["sensor_type"] == "lidar":
            self._sensor = OracleLidarSensor(c["sensor_config"])
        elif c["sensor_type"] == "square_sensor":
            self._sensor = OracleSquareSensor(c["sensor_config"])
        self._pad = max(self._egoradius, self._mini_map_rad)
        self.reset(False, None)
        self._obs_dim = self.get_egocentric_observations()[0].shape
        self._num_actions = 4

    # ---- episode start (dec_grid_rl.py:449-531) -------------------------------------
    def reset(self, testing, ind, start_xy=None, map_index=None):
        """``start_xy`` ((N,2) ints, bordered coordinates) / ``map_index`` bypass the random draws
        (harness extension used to inject reference start poses)."""
        if map_index is not None:
            base = self._train_gridlis[map_index]
        elif testing and self._test_gridlis is not None:
            base = self._test_gridlis[ind]
        else:
            base = self._train_gridlis[np.random.randint(len(self._train_gridlis))]
        self._grid = np.pad(base, (1,), "constant", constant_values=(-1,))
        self._gridwidth, self._gridlen = self._grid.shape
        W, L, n = self._gridwidth, self._gridlen, self._numrobot
        self._currstep = 0
        self._xinds = np.zeros(n, dtype=int)
        self._yinds = np.zeros(n, dtype=int)
        if start_xy is not None:
            self._xinds[:] = np.asarray(start_xy)[:, 0]
            self._yinds[:] = np.asarray(start_xy)[:, 1]
        else:
            taken = set()
            k = 0
            while k != n:
                x = np.random.randint(W)
                y = np.random.randint(L)
                if self._grid[x][y] >= 0 and (x, y) not in taken:
                    taken.add((x, y))
                    self._xinds[k], self._yinds[k] = x, y
                    k += 1
        # layer-0 source: a robot only shows up after its first successful move (quirk A.3.4)
        self._shown = np.zeros(n, dtype=bool)
        self._free = np.zeros((n, W, L), dtype=bool)      # per-robot known-free memory
        self._obst = np.zeros((n, W, L), dtype=bool)      # per-robot known-obstacle memory
        self._visited = np.zeros((W, L), dtype=bool)
        self._observed_obstacles = np.zeros((W, L), dtype=bool)
        self._numfree = int(np.count_nonzero(self._grid > 0))
        self._update_graph()
        self._observe()
        obs = self.get_egocentric_observations()
        if self._allow_comm and self._use_graph:
            return obs, self._grid, self._adjacency_matrix
        return obs, self._grid

    # ---- one transition (dec_grid_rl.py:91-169) --------------------------------------
    def step(self, action):
        done = False
        if _is_sentinel(action):
            done = True
            reward = 0
        else:
            if isinstance(action, np.ndarray) and action.ndim >= 1:
                moves = action.astype(np.int64).ravel()
            elif isinstance(action, (list, tuple)):
                moves = np.asarray(action, dtype=np.int64)
            else:
                moves = decode_base4(action, self._numrobot)
            reward = 0
            if self._map_sharing:
                self._share_maps()
            for i in range(self._numrobot):
                dx, dy = self.MOVES[int(moves[i])]
                reward += self._try_move(i, self._xinds[i] + dx, self._yinds[i] + dy)
            self._update_graph()
            reward += self._observe()
            self._currstep += 1
            if min(self._done_thresh, 1) <= self.percent_covered():
                reward += self._terminal_reward
        obs = self.get_egocentric_observations()
        if done is False:
            done = self.done()
        if self._allow_comm and self._use_graph:
            return [obs, self._adjacency_matrix], reward, done
        return obs, reward, done

    def isInBounds(self, x, y):
        return 0 <= x < self._gridwidth and 0 <= y < self._gridlen

    def isOccupied(self, x, y):
        if self._grid[x][y] < 0:
            return True
        return bool(np.any((self._xinds == x) & (self._yinds == y)))

    def _try_move(self, i, x, y):
        # dec_grid_rl.py:171-204
        if self.isInBounds(x, y) and not self.isOccupied(x, y):
            self._xinds[i], self._yinds[i] = x, y
            self._shown[i] = True
            return 0
        return -self._collision_penalty

    def _update_graph(self):
        # dec_grid_rl.py:374-391 — Chebyshev ball, diagonal included
        dx = np.abs(self._xinds[:, None] - self._xinds[None, :])
        dy = np.abs(self._yinds[:, None] - self._yinds[None, :])
        self._adjacency_matrix = (np.maximum(dx, dy) <= self._comm_radius).astype(np.float64)

    def _share_maps(self):
        # dec_grid_rl.py:423-447 — simultaneous OR over neighbours (self included)
        link = (self._adjacency_matrix != 0) | np.eye(self._numrobot, dtype=bool)
        free_new = np.zeros_like(self._free)
        obst_new = np.zeros_like(self._obst)
        for i in range(self._numrobot):
            free_new[i] = self._free[link[i]].any(axis=0)
            obst_new[i] = self._obst[link[i]].any(axis=0)
        self._free, self._obst = free_new, obst_new

    def _padded(self, plane):
        return np.pad(plane, self._pad, "constant", constant_values=0)

    def _distance_channel(self, i):
        """dec_grid_rl.py:260-282 on the robot's PADDED free plane (padding counts as unknown)."""
        return inverted_distance_channel(self._padded(self._free[i]))

    def _observe(self):
        # dec_grid_rl.py:206-258
        obs_reward = 0
        for i in range(self._numrobot):
            x, y = int(self._xinds[i]), int(self._yinds[i])
            if self._dist_r:
                dmap = self._distance_channel(i)
            fr, ob = self._sensor.footprint(x, y, self._grid)
            self._obst[i] |= ob
            if self._single_square_tool:
                self._free[i, x, y] = True
            else:
                self._free[i] |= fr
            if self._dist_r:
                # quirk A.3.11: padded map indexed with UNpadded coordinates, float32 accumulate
                obs_reward += dmap[x, y]
        before = int(np.count_nonzero(self._visited))
        self._visited = self._free.any(axis=0)
        self._observed_obstacles = self._obst.any(axis=0)
        obs_reward += np.float64(int(np.count_nonzero(self._visited)) - before)
        return obs_reward

    def get_egocentric_observations(self):
        # dec_grid_rl.py:312-372
        n, e = self._numrobot, self._egoradius
        layers = 5 if self._mini_map_rad > 0 else 3
        if self._dist_r:
            layers += 1
        if self._dijkstra_input:
            layers += 1
        z = np.zeros((n, layers, 2 * e + 1, 2 * e + 1))
        robots = np.zeros((self._gridwidth, self._gridlen), dtype=bool)
        robots[self._xinds[self._shown], self._yinds[self._shown]] = True
        for i in range(n):
            x, y = int(self._xinds[i]), int(self._yinds[i])
            z[i, 0] = window(robots, x, y, e)
            z[i, 1] = window(self._free[i], x, y, e)
            z[i, 2] = window(self._obst[i], x, y, e)
            if self._dist_r:
                dmap = self._distance_channel(i)
                z[i, 3] = window(dmap, x + self._pad, y + self._pad, e)
            if self._dijkstra_input:
                code = self._padded(self._free[i]).astype(np.int64) \
                    - self._padded(self._obst[i]).astype(np.int64)
                path = dijkstra_path_plane(code, x + self._pad, y + self._pad)
                z[i, 3] = window(path, x + self._pad, y + self._pad, e)
            if self._mini_map_rad > 0:
                m = self._mini_map_rad
                z[i, 3] = bilinear_resize(window(self._free[i], x, y, m), 2 * e + 1)
                z[i, 4] = bilinear_resize(window(self._obst[i], x, y, m), 2 * e + 1)
        return z

    def done(self):
        # dec_grid_rl.py:533-546
        if min(self._done_thresh, 1) <= self.percent_covered():
            self._done_thresh += self._done_incr
            return True
        return self._currstep == self._maxsteps

    def percent_covered(self):
        # dec_grid_rl.py:548-552 — sums over robots, not the union (quirk A.3.3)
        return int(np.count_nonzero(self._free)) / self._numfree

    def render_rgb(self):
        """Un-scaled RGB composition of dec_grid_rl.py:562-580 (window/scale/flip left out)."""
        img = np.zeros((self._gridwidth, self._gridlen, 3))
        img[self._observed_obstacles] += (200, 0, 255)
        img[self._visited] += (0, 225, 255)
        for x, y in zip(self._xinds, self._yinds):
            img[x, y] += (255, 0, 0)
        return img


# --------------------------------------------------------------------------------------
# SuperGridRL
# --------------------------------------------------------------------------------------

class OracleSuperGridRL:
    """Centralised fully observed coverage env — restates super_grid_rl.py:18-421."""

    # Super move table (super_grid_rl.py:132-174): 0:x-1 1:x+1 2:y+1 3:y-1
    MOVES = ((-1, 0), (1, 0), (0, 1), (0, -1))

    def __init__(self, train_set, env_config, test_set=None):
        c = env_config
        self._train_gridlis = train_set
        self._test_gridlis = test_set
        self._numrobot = c["numrobot"]
        self._train_maxsteps = c["train_maxsteps"]
        self._test_maxsteps = c["test_maxsteps"]
        self._collision_penalty = c["collision_penalty"]
        self._senseradius = c["senseradius"]
        self._free_penalty = c["free_penalty"]
        self._done_thresh = c["done_thresh"]
        self._done_incr = c["done_incr"]
        self._terminal_reward = c["terminal_reward"]
        self._dist_r = c["dist_reward"]
        self._use_scanning = c["use_scanning"]
        self._never_stepped = True            # stands in for ``a_prev is None`` (quirk A.2.6)
        self.reset(False, False)
        self._obs_dim = self.get_state()[0].shape
        self._num_actions = 4 ** self._numrobot

    def reset(self, testing, ind, start_xy=None, map_index=None):
        # super_grid_rl.py:343-399
        if map_index is not None:
            self._grid = self._train_gridlis[map_index]
        elif testing and self._test_gridlis is not None:
            self._grid = self._test_gridlis[ind]
        else:
            self._grid = self._train_gridlis[np.random.randint(len(self._train_gridlis))]
        self._gridwidth, self._gridlen = self._grid.shape
        n = self._numrobot
        self._currstep = 0
        self._xinds = np.zeros(n, dtype=int)
        self._yinds = np.zeros(n, dtype=int)
        if start_xy is not None:
            self._xinds[:] = np.asarray(start_xy)[:, 0]
            self._yinds[:] = np.asarray(start_xy)[:, 1]
        else:
            taken = set()
            k = 0
            while k != n:
                x = np.random.randint(self._gridwidth)
                y = np.random.randint(self._gridlen)
                if self._grid[x][y] >= 0 and (x, y) not in taken:
                    taken.add((x, y))
                    self._xinds[k], self._yinds[k] = x, y
                    k += 1
        self._covered = np.zeros(self._grid.shape, dtype=bool)     # reference: _free == 0
        self._observed_obstacles = np.zeros(self._grid.shape, dtype=bool)
        return self.get_state(), self._grid

    def isInBounds(self, x, y):
        return 0 <= x < self._gridwidth and 0 <= y < self._gridlen

    def isOccupied(self, x, y):
        if self._grid[x][y] < 0:
            return True
        return bool(np.any((self._xinds == x) & (self._yinds == y)))

    def _distance_channel(self):
        # super_grid_rl.py:281-303: zero pixels are the NOT-covered cells (obstacles included)
        return inverted_distance_channel(~self._covered)

    def step(self, action):
        # super_grid_rl.py:74-225
        done = False
        if _is_sentinel(action):
            done = True
            reward = 0
        else:
            n = self._numrobot
            if isinstance(action, (list, tuple)) or \
                    (isinstance(action, np.ndarray) and action.ndim >= 1):
                moves = np.asarray(action, dtype=np.int64).ravel()
            else:
                moves = decode_base4(action, n)
            reward = np.zeros(n)
            # raster-scan order of the start-of-step poses, ties by robot index (:108-111)
            order = sorted(range(n), key=lambda r: (self._xinds[r]
                                                    + self._yinds[r] * self._gridwidth, r))
            slot_of_robot = np.zeros(n, dtype=int)
            if self._dist_r:
                dmap = self._distance_channel()
            for i in range(n):
                z = order[i] if self._use_scanning else i
                slot_of_robot[z] = i
                dx, dy = self.MOVES[int(moves[i])]
                x, y = self._xinds[z] + dx, self._yinds[z] + dy
                if self.isInBounds(x, y) and not self.isOccupied(x, y):
                    self._xinds[z], self._yinds[z] = x, y
                    if self._dist_r:
                        reward[i] += dmap[x, y]
                else:
                    reward[i] -= self._collision_penalty
            r = self._senseradius
            for i in range(n):
                x, y = int(self._xinds[i]), int(self._yinds[i])
                s = slot_of_robot[i]
                for j in range(max(x - r, 0), min(x + r + 1, self._gridwidth)):
                    for k in range(max(y - r, 0), min(y + r + 1, self._gridlen)):
                        g = self._grid[j][k]
                        if g >= 0:
                            if not self._covered[j, k]:
                                reward[s] += g
                                self._covered[j, k] = True
                            else:
                                reward[s] -= self._free_penalty
                        else:
                            self._observed_obstacles[j, k] = True
            # quirk A.2.6 / row S7: the decoded-away action is always 0, so the motion penalty is
            # -1 (broadcast over the N slots) on the first step of the object's life, then 0.
            if self._never_stepped:
                reward += -1
                self._never_stepped = False
            self._currstep += 1
            reward = np.sum(reward)
            if min(self._done_thresh, 1) <= self.percent_covered():
                reward += self._terminal_reward
        state = self.get_state()
        if done is False:
            done = self.done()
        return state, reward, done

    def get_state(self):
        # super_grid_rl.py:305-341
        W, L, n = self._gridwidth, self._gridlen, self._numrobot
        if self._use_scanning:
            pos = np.zeros((1, W, L))
            pos[0, self._xinds, self._yinds] = 1
        else:
            pos = np.zeros((n, W, L))
            pos[np.arange(n), self._xinds, self._yinds] = 1
        tail = np.stack([self._observed_obstacles.astype(np.float64),
                         (~self._covered).astype(np.float64),
                         self._distance_channel().astype(np.float64)], axis=0)
        return (np.concatenate([pos, tail], axis=0), self._currstep)

    def done(self):
        if min(self._done_thresh, 1) <= self.percent_covered():
            self._done_thresh += self._done_incr
            return True
        return False

    def percent_covered(self):
        return int(np.count_nonzero(self._covered)) / int(np.count_nonzero(self._grid > 0))


# --------------------------------------------------------------------------------------
# action space (Action_Spaces/discrete.py:4-14)
# --------------------------------------------------------------------------------------

class OracleDiscrete:
    def __init__(self, num_actions):
        self.num_actions = num_actions

    def sample(self, s=None):
        return np.random.randint(self.num_actions, size=s)


# --------------------------------------------------------------------------------------
# replay buffer rows (Policies/replaybuffer.py:3-72) — the consumer right after the env step
# --------------------------------------------------------------------------------------

class OracleReplayBuffer:
    """float32 ring arrays with the reference's write pointer arithmetic (replaybuffer.py:33-48)
    and index gather (replaybuffer.py:66-72)."""

    def __init__(self, obs_dim, act_dim, size):
        def shp(d):
            return (size,) if d is None else ((size, d) if np.isscalar(d) else (size, *d))
        self._maxsize = size
        self._size = 0
        self._ptr = 0
        self._state = np.zeros(shp(obs_dim), dtype=np.float32)
        self._action = np.zeros(shp(act_dim), dtype=np.float32)
        self._reward = np.zeros(size, dtype=np.float32)
        self._nextstate = np.zeros(shp(obs_dim), dtype=np.float32)
        self._done = np.zeros(size, dtype=np.float32)

    def addtransition(self, state, action, reward, next_state, done):
        k = self._ptr
        self._state[k] = state
        self._action[k] = action
        self._reward[k] = reward
        self._nextstate[k] = next_state
        self._done[k] = done
        self._ptr = (self._ptr + 1) % self._maxsize
        self._size = min(self._size + 1, self._maxsize)

    def gather(self, indices):
        i = np.asarray(indices)
        return (self._state[i], self._action[i], self._reward[i], self._nextstate[i], self._done[i])


# --------------------------------------------------------------------------------------
# SuperGrid_Sim one-step lookahead (Environments/sim_env.py:57-190) — SURVEY.md §8f rank 3
# --------------------------------------------------------------------------------------

def sim_lookahead(state4, action, grid, env_config, done_thresh=None):
    """Stateless one-robot lookahead: (state (4,W,L), action 0..3, grid) -> (next_state, reward).

    Restates ``SuperGrid_Sim.step`` (sim_env.py:57-190).  The reference function is broken at HEAD in
    two places that the restatement reads the way the surrounding code intends:
      * ``np.asscalar`` (sim_env.py:146-147) was removed in numpy 1.23 -> ``.item()``;
      * ``self.percent_covered(state)`` (sim_env.py:187) omits the ``grid`` argument of
        ``percent_covered(self, state, grid)`` (sim_env.py:248) -> the step's own ``grid``.
    ``tests/golden/pin_sim_render.py`` runs the reference with exactly those two shims to pin this function.
    Moves use SuperGridRL's mapping 0:x-1 1:x+1 2:y+1 3:y-1; only obstacles block (isOccupied,
    sim_env.py:210-228); the distance plane is carried through unchanged (not recomputed); the float64
    reward is accumulated in the reference's order: move term, then footprint cells in raster order.
    Unlike the reference this does not mutate ``state4`` (the compat class re-applies that side effect).
    """
    c = env_config
    st = np.array(state4, dtype=np.float64)
    pos_img, seen, free, dist = st[0], st[1].copy(), st[2].copy(), st[3]
    xs, ys = np.nonzero(pos_img)
    if xs.size != 1:
        raise ValueError("sim lookahead needs exactly one robot in the position plane")
    x0, y0 = int(xs[0]), int(ys[0])
    W, L = grid.shape
    reward = 0.0
    dx, dy = OracleSuperGridRL.MOVES[int(action)]
    x, y = x0 + dx, y0 + dy
    if 0 <= x < W and 0 <= y < L and not grid[x, y] < 0:
        x0, y0 = x, y
        if c["dist_reward"]:
            reward += dist[x, y]
    else:
        reward -= c["collision_penalty"]
    r = c["senseradius"]
    for j in range(x0 - r, x0 + r + 1):
        for k in range(y0 - r, y0 + r + 1):
            if not (0 <= j < W and 0 <= k < L):
                continue
            if grid[j][k] >= 0 and free[j][k] == 1:
                reward += grid[j][k]
                free[j][k] = 0
            elif grid[j][k] >= 0 and free[j][k] == 0:
                reward -= c["free_penalty"]
            elif grid[j][k] < 0 and seen[j][k] == 0:
                seen[j][k] = 1
    new_pos = np.zeros((W, L))
    new_pos[x0, y0] = 1
    out = np.stack([new_pos, seen, free, dist], axis=0)
    thresh = c["done_thresh"] if done_thresh is None else done_thresh
    if min(thresh, 1) <= np.count_nonzero(free < 1) / np.count_nonzero(grid > 0):
        reward += c["terminal_reward"]
    return out, float(reward)


# --------------------------------------------------------------------------------------
# render() frames (super_grid_rl.py:423-463, dec_grid_rl.py:554-616) — SURVEY.md §8f rank 4
# --------------------------------------------------------------------------------------

def render_frame(seen_obst, covered, robot_map, scale_base, flip):
    """uint8 frame of the reference's render(): colour sum of the three layers (observed obstacle
    (200,0,255), covered (0,225,255), robot (255,0,0)), nearest-neighbour upscale by
    s = max(min(scale_base//W, scale_base//L), 1) (cv2.resize INTER_NEAREST with integer factors =
    pixel replication), zero border of 1075 - L*s rows at the bottom of axis 0 and 1075 - W*s columns on
    the right of axis 1 (cv2.copyMakeBorder argument order as written: the two are swapped for
    non-square maps), DecGridRL additionally flips axis 1 (cv2.flip(image, 1)).  scale_base is 1075
    for SuperGridRL and 1024 for DecGridRL.  The window (pygame) is left out."""
    W, L = covered.shape
    img = np.zeros((W, L, 3))
    img += np.asarray(seen_obst, dtype=np.float64)[:, :, None] * np.array([200., 0., 255.])
    img += np.asarray(covered, dtype=np.float64)[:, :, None] * np.array([0., 225., 255.])
    img += np.asarray(robot_map, dtype=np.float64)[:, :, None] * np.array([255., 0., 0.])
    s = max(min(scale_base // W, scale_base // L), 1)
    img = np.repeat(np.repeat(img, s, axis=0), s, axis=1)
    bottom, right = 1075 - img.shape[1], 1075 - img.shape[0]
    if bottom < 0 or right < 0:
        raise ValueError("map too large for the 1075 canvas (cv2.copyMakeBorder would fail)")
    out = np.zeros((img.shape[0] + bottom, img.shape[1] + right, 3))
    out[:img.shape[0], :img.shape[1]] = img
    if flip:
        out = out[:, ::-1]
    return np.ascontiguousarray(out).astype(np.uint8)
