"""Size-independent properties of the coverage path, checked on the oracle over random small worlds.
The same invariants are checked on the GPU batch at BASELINE.json's full sizes in tests/test_fullsize_gpu.py;
here they guard the checker itself (oracle/ is test infrastructure)."""
import numpy as np
import pytest

from oracle import coverage_oracle as oc


def _world(rng, W, L, p_obst):
    g = np.where(rng.random_sample((W, L)) < p_obst, -1, 1).astype(np.int8)
    g[rng.randint(W), rng.randint(L)] = 1                       # at least one free cell
    return g


def _cfg(**kw):
    c = dict(numrobot=3, maxsteps=1000, collision_penalty=5, done_thresh=1, done_incr=0, terminal_reward=30,
             dist_reward=0, train_maxsteps=1000, test_maxsteps=1000, egoradius=3, mini_map_rad=0, comm_radius=0,
             allow_comm=0, map_sharing=0, single_square_tool=0, dijkstra_input=0, sensor_type="square_sensor",
             sensor_config={"range": 2})
    c.update(kw)
    return c


@pytest.mark.parametrize("seed", range(6))
@pytest.mark.parametrize("sensor", ["square_sensor", "lidar"])
def test_dec_step_invariants(seed, sensor):
    rng = np.random.RandomState(100 + seed)
    np.random.seed(seed)
    W, L = rng.randint(6, 20), rng.randint(6, 20)
    n = int(rng.randint(1, 5))
    scfg = {"range": int(rng.randint(1, 4))}
    if sensor == "lidar":
        scfg["num_lasers"] = int(rng.choice([5, 9, 21]))
    cfg = _cfg(numrobot=n, sensor_type=sensor, sensor_config=scfg, egoradius=int(rng.randint(1, 5)))
    env = oc.OracleDecGridRL([_world(rng, W, L, 0.2) for _ in range(2)], cfg)
    env.reset(False, None)
    grid = env._grid                                            # bordered: (W+2, L+2)
    assert (grid[0] < 0).all() and (grid[-1] < 0).all() and (grid[:, 0] < 0).all() and (grid[:, -1] < 0).all()
    for t in range(80):
        pos0 = np.stack([env._xinds, env._yinds], 1).copy()
        free0, obst0, vis0 = env._free.copy(), env._obst.copy(), env._visited.copy()
        moves = rng.randint(0, 4, size=n)
        obs, reward, done = env.step(moves)
        pos1 = np.stack([env._xinds, env._yinds], 1)
        # robots stand on free cells, never two on one cell, move by at most one cell
        assert (grid[pos1[:, 0], pos1[:, 1]] >= 0).all()
        assert len({tuple(p) for p in pos1}) == n
        assert (np.abs(pos1 - pos0).sum(1) <= 1).all()
        # memory only grows and never contradicts the map
        assert (env._free >= free0).all() and (env._obst >= obst0).all()
        assert not (env._free & (grid < 0)).any() and not (env._obst & (grid >= 0)).any()
        assert np.array_equal(env._visited, env._free.any(0))
        # reward = newly visited cells of the union - penalty per blocked robot (+ terminal bonus)
        blocked = int((np.abs(pos1 - pos0).sum(1) == 0).sum())
        want = int(env._visited.sum()) - int(vis0.sum()) - cfg["collision_penalty"] * blocked
        if done:
            want += cfg["terminal_reward"]
        assert reward == want
        # observation layers are 0/1 and centred on the robot's own (free) cell
        e = cfg["egoradius"]
        assert set(np.unique(obs)) <= {0.0, 1.0}
        assert obs.shape == (n, 3, 2 * e + 1, 2 * e + 1)
        assert (obs[:, 1, e, e] == 1).all() and (obs[:, 2, e, e] == 0).all()
        if done:
            break


@pytest.mark.parametrize("seed", range(4))
def test_lidar_footprint_inside_square_of_same_range(seed):
    """A beam never reaches further than its range in Chebyshev distance + the end cell, so the lidar's hits
    lie in the square of radius range + 1 around the robot, and its own cell is always seen."""
    rng = np.random.RandomState(seed)
    g = np.pad(_world(rng, 15, 17, 0.15), 1, constant_values=-1)
    r = int(rng.randint(1, 5))
    lidar = oc.OracleLidarSensor({"range": r, "num_lasers": 21})
    square = oc.OracleSquareSensor({"range": r + 1})
    xs, ys = np.nonzero(g >= 0)
    for k in rng.choice(len(xs), size=12, replace=False):
        x, y = int(xs[k]), int(ys[k])
        fr, ob = lidar.footprint(x, y, g)
        sq_fr, sq_ob = square.footprint(x, y, g)
        assert fr[x, y]
        assert not (fr & ~sq_fr).any() and not (ob & ~sq_ob).any()


@pytest.mark.parametrize("seed", range(4))
def test_super_step_invariants(seed):
    rng = np.random.RandomState(seed)
    np.random.seed(seed)
    n = int(rng.randint(1, 5))
    cfg = dict(numrobot=n, train_maxsteps=1000, test_maxsteps=1000, collision_penalty=5,
               senseradius=int(rng.randint(1, 4)), free_penalty=0, done_thresh=1, done_incr=0, terminal_reward=30,
               dist_reward=0, use_scanning=0)
    env = oc.OracleSuperGridRL([_world(rng, 12, 14, 0.2)], cfg)
    (state, _), grid = env.reset(False, None)
    assert state.shape == (n + 3, 12, 14)
    for t in range(60):
        cov0 = env._covered.copy()
        state, reward, done = env.step(int(rng.randint(0, 4 ** n)))
        planes = state[0]
        assert (planes[:n].sum(axis=(1, 2)) == 1).all()                       # one-hot pose per robot
        assert np.array_equal(planes[n] > 0, env._observed_obstacles)
        assert np.array_equal(planes[n + 1] < 1, env._covered)
        assert (env._covered >= cov0).all() and not (env._covered & (grid < 0)).any()
        # inverted, normalised L1 distance plane (super_grid_rl.py:327-339) stays in [0, 1]
        assert planes[n + 2].min() >= 0 and planes[n + 2].max() <= 1.0 + 1e-12
        if done:
            break
