"""CPU checks for SURVEY §8f ranks 3-4: the oracle's SuperGrid_Sim lookahead and render() frames against the
golden vectors tests/golden/pin_sim_render.py recorded from the live reference (sim_env.py with the two
documented shims; render() with the pygame stub)."""
import json
import os
import zlib

import numpy as np
import pytest

from conftest import GOLDEN_DIR
from oracle import coverage_oracle as oc


def _meta(g):
    return json.loads(bytes(g["meta"]).decode())


def test_oracle_sim_lookahead_matches_reference_golden():
    g = np.load(os.path.join(GOLDEN_DIR, "sim_lookahead.npz"))
    n = 0
    for m in _meta(g):
        name, cfg = m["name"], m["cfg"]
        maps = g["maps_" + name].astype(np.float64)
        full = {int(i): k for k, i in enumerate(g["fulli_" + name])}
        for k in range(g["st_" + name].shape[0]):
            grid = maps[int(g["mi_" + name][k])]
            nxt, r = oc.sim_lookahead(g["st_" + name][k], int(g["act_" + name][k]), grid, cfg)
            assert r == float(g["rew_" + name][k]), (name, k)
            assert zlib.crc32(nxt.astype(np.float32).tobytes()) == int(g["crc_" + name][k]), (name, k)
            if k in full:
                assert np.array_equal(nxt.astype(np.float32), g["full_" + name][full[k]])
            n += 1
    assert n > 200


def test_oracle_render_matches_reference_golden():
    g = np.load(os.path.join(GOLDEN_DIR, "render_frames.npz"))
    for m in _meta(g):
        name, kind, cfg = m["name"], m["kind"], m["cfg"]
        maps = [a.astype(np.float64) for a in g["maps_" + name]]
        cls = oc.OracleDecGridRL if kind == "dec" else oc.OracleSuperGridRL
        np.random.seed(0)
        env = cls(maps, cfg)
        env.reset(False, None, start_xy=g["start_" + name], map_index=int(m["map_index"]))
        for a in g["acts_" + name]:
            env.step(int(a))
        if kind == "dec":
            robots = np.zeros((env._gridwidth, env._gridlen))
            robots[env._xinds, env._yinds] = 1
            frame = oc.render_frame(env._observed_obstacles, env._visited, robots, 1024, True)
        else:
            robots = np.zeros((env._gridwidth, env._gridlen))
            k = slice(None) if cfg["use_scanning"] else slice(0, 1)
            robots[env._xinds[k], env._yinds[k]] = 1
            frame = oc.render_frame(env._observed_obstacles, env._covered, robots, 1075, False)
        assert frame.shape == g["frame_" + name].shape, name
        assert np.array_equal(frame, g["frame_" + name]), name


def test_sim_lookahead_does_not_mutate_and_terminal():
    grid = np.ones((6, 7))
    grid[2, 3] = -1
    cfg = dict(numrobot=1, train_maxsteps=10, test_maxsteps=10, collision_penalty=5, senseradius=1,
               free_penalty=0.25, done_thresh=0.19, done_incr=0, terminal_reward=30, dist_reward=0,
               use_scanning=0)
    st = np.zeros((4, 6, 7))
    st[0, 2, 2] = 1
    st[2] = 1
    keep = st.copy()
    nxt, r = oc.sim_lookahead(st, 2, grid, cfg)        # y+1 is the obstacle: blocked
    assert np.array_equal(st, keep)
    assert nxt[0, 2, 2] == 1 and nxt[1, 2, 3] == 1
    assert r == -5 + 8 + 30                             # 8 free cells of the 3x3; 8/41 >= 0.19: terminal bonus
    cfg["done_thresh"] = 0.2
    assert oc.sim_lookahead(st, 2, grid, cfg)[1] == -5 + 8


def test_oracle_matches_live_reference():
    """Re-runs the pin against the live reference where it is present (never on the GPU box)."""
    import subprocess
    import sys
    ref = os.environ.get("MCOV_REFERENCE_ROOT", "/root/reference")
    if not os.path.isdir(os.path.join(ref, "Environments")):
        pytest.skip("reference not present")
    res = subprocess.run([sys.executable, os.path.join(GOLDEN_DIR, "pin_sim_render.py"), "--check"],
                         capture_output=True, text=True, timeout=600)
    assert res.returncode == 0, res.stdout[-2000:] + res.stderr[-2000:]
    assert res.stdout.count("matches the oracle") >= 6


def test_lookahead_agrees_with_env_step():
    """Two independently pinned restatements against each other: for one robot, SuperGrid_Sim's lookahead of
    the env's state is the env's own next state (position / seen-obstacle / not-covered planes) and reward —
    except the env's first-step motion penalty (quirk A.2.6), which the simulator does not have."""
    from marl_coverage_b200 import load_grid_set
    grids = load_grid_set("bg2_100x100", 3)
    cfg = dict(numrobot=1, train_maxsteps=1000, test_maxsteps=1000, collision_penalty=5, senseradius=2,
               free_penalty=0.25, done_thresh=1, done_incr=0, terminal_reward=30, dist_reward=0, use_scanning=0)
    rng = np.random.RandomState(4)
    np.random.seed(4)
    env = oc.OracleSuperGridRL(grids, cfg)
    for mi in range(len(grids)):
        state = env.reset(False, None, map_index=mi)[0][0]      # ((planes, step), grid)
        first = env._never_stepped
        for t in range(120):
            a = int(rng.randint(0, 4))
            want, want_r = oc.sim_lookahead(state, a, env._grid, cfg)
            nxt, r, done = env.step(a)
            nxt = nxt[0]
            assert np.array_equal(nxt[:3], want[:3]), (mi, t)
            assert r == want_r + (-1 if first else 0), (mi, t, r, want_r)
            first = False
            state = nxt
            assert not done
