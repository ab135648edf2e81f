"""The drop-in classes (marl_coverage_b200.DecGridRL / SuperGridRL: the reference's constructor, methods,
attributes and return types) against the oracle, driven the way grid_rl_main.py / Utils/utils.py
drive the reference: seeded global np.random, reset(testing, ind), integer base-4 actions, done() / maxsteps
cut-off, several episodes on one env object."""
import contextlib
import io

import numpy as np
import pytest

from oracle import coverage_oracle as oc

pytestmark = pytest.mark.gpu

DEC = dict(numrobot=3, maxsteps=40, collision_penalty=5, done_thresh=0.5, done_incr=0.1, terminal_reward=30,
           dist_reward=0, train_maxsteps=40, test_maxsteps=25, egoradius=4, mini_map_rad=0, comm_radius=6,
           allow_comm=1, map_sharing=1, single_square_tool=0, dijkstra_input=0, sensor_type="square_sensor",
           sensor_config={"range": 2})
SUPER = dict(numrobot=3, train_maxsteps=40, test_maxsteps=25, collision_penalty=5, senseradius=2, free_penalty=0.2,
             done_thresh=0.3, done_incr=0.05, terminal_reward=30, dist_reward=1, use_scanning=0)


def _session(make, kind, cfg, seed, episodes=4):
    """Utils/utils.py:generate_episode's loop (reset, step until done or maxsteps), recorded."""
    log = []
    np.random.seed(seed)
    with contextlib.redirect_stdout(io.StringIO()):
        env = make()
        log.append(("ctor", env._num_actions, tuple(env._obs_dim)))
        arng = np.random.RandomState(seed + 1)
        for ep in range(episodes):
            testing = ep % 2 == 1
            out = env.reset(testing, ep % 2 if testing else None)
            first = out[0] if kind == "dec" else out[0][0]
            log.append(("reset", np.array(first, dtype=np.float64), np.array(out[1]),
                        np.array(env._xinds), np.array(env._yinds), env._currstep))
            steps = 0
            maxsteps = env._test_maxsteps if testing else env._train_maxsteps
            while True:
                a = int(arng.randint(env._num_actions if kind == "super" else 4 ** cfg["numrobot"]))
                if steps == 3 and ep == 0:
                    a = -1 if kind == "dec" else None                  # the sentinel forms
                state, reward, done = env.step(a)
                obs = state[0] if (kind == "super" or isinstance(state, list)) else state
                adj = state[1] if (kind == "dec" and isinstance(state, list)) else None
                log.append(("step", np.array(obs, dtype=np.float64), float(reward), bool(done),
                            np.array(env._xinds), np.array(env._yinds), env._currstep,
                            float(env.percent_covered()), float(env._done_thresh),
                            None if adj is None else np.array(adj),
                            bool(env.isInBounds(3, 4)), bool(env.isOccupied(int(env._xinds[0]), int(env._yinds[0])))))
                steps += 1
                if a in (-1, None):
                    continue
                if done or steps >= maxsteps:
                    break
    return log


def _same(a, b, path="log"):
    assert type(a) is type(b) or (np.isscalar(a) and np.isscalar(b)), (path, type(a), type(b))
    if isinstance(a, (tuple, list)):
        assert len(a) == len(b), path
        for k, (x, y) in enumerate(zip(a, b)):
            _same(x, y, "%s[%d]" % (path, k))
    elif isinstance(a, np.ndarray):
        assert a.shape == b.shape and a.dtype == b.dtype and np.array_equal(a, b), path
    else:
        assert a == b, (path, a, b)


@pytest.mark.parametrize("kind,use_graph", [("dec", False), ("dec", True), ("super", False)])
def test_drop_in_class_walks_the_same_episodes_as_the_oracle(kind, use_graph):
    from marl_coverage_b200 import DecGridRL, SuperGridRL, gridload
    train, test = gridload(None)                                       # the built-in 15x15 maps
    cfg = dict(DEC if kind == "dec" else SUPER)
    if kind == "dec":
        ours = lambda: DecGridRL(train, cfg, use_graph=use_graph, test_set=test)
        ref = lambda: oc.OracleDecGridRL(train, cfg, use_graph=use_graph, test_set=test)
    else:
        ours = lambda: SuperGridRL(train, cfg, test_set=test)
        ref = lambda: oc.OracleSuperGridRL(train, cfg, test_set=test)
    got = _session(ours, kind, cfg, seed=31)
    want = _session(ref, kind, cfg, seed=31)
    assert len(got) == len(want) and sum(1 for r in want if r[0] == "step" and r[3]) >= 2
    for k, (g, w) in enumerate(zip(got, want)):
        _same(g, w, "op %d (%s)" % (k, w[0]))
