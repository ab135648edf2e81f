"""GPU parity for SURVEY §8f ranks 3-4 through the C ABI: mcov_sim_step (SuperGrid_Sim.step lookahead) and
mcov_render (render() frames) against the reference-recorded golden vectors and the oracle."""
import json
import os
import zlib

import numpy as np
import pytest

from conftest import GOLDEN_DIR
from oracle import coverage_oracle as oc

pytestmark = pytest.mark.gpu


def _meta(g):
    return json.loads(bytes(g["meta"]).decode())


def test_sim_lookahead_matches_reference_golden():
    import torch
    from marl_coverage_b200 import VecSuperGridSim
    g = np.load(os.path.join(GOLDEN_DIR, "sim_lookahead.npz"))
    for m in _meta(g):
        name, cfg = m["name"], m["cfg"]
        maps = [a.astype(np.float64) for a in g["maps_" + name]]
        sim = VecSuperGridSim(maps, cfg)
        st = torch.from_numpy(g["st_" + name]).cuda()
        keep = st.clone()
        nxt, r64, _ = sim.step(st, g["act_" + name], map_idx=g["mi_" + name])
        assert torch.equal(st, keep), "out-of-place lookahead must not touch its input"
        assert np.array_equal(r64.cpu().numpy(), g["rew_" + name]), name       # float64, bit for bit
        out = nxt.cpu().numpy()
        crc = np.array([zlib.crc32(out[k].tobytes()) for k in range(out.shape[0])], dtype=np.uint32)
        assert np.array_equal(crc, g["crc_" + name]), name
        for k, i in enumerate(g["fulli_" + name]):
            assert np.array_equal(out[int(i)], g["full_" + name][k])
        # in place (the reference's side effect on the caller's planes) gives the same planes
        nxt2, r2, _ = sim.step(st, g["act_" + name], map_idx=g["mi_" + name], inplace=True)
        assert nxt2.data_ptr() == st.data_ptr()
        assert torch.equal(nxt2, nxt) and torch.equal(r2, r64)


@pytest.mark.parametrize("W,L,R", [(100, 100, 2), (37, 53, 3), (15, 15, 15), (64, 31, 0)])
def test_sim_lookahead_random_vs_oracle(W, L, R):
    import torch
    from marl_coverage_b200 import VecSuperGridSim
    rng = np.random.RandomState(W * 1000 + L)
    maps = [np.where(rng.rand(W, L) < 0.2, -1.0, 1.0) for _ in range(3)]
    cfg = dict(numrobot=1, train_maxsteps=10, test_maxsteps=10, collision_penalty=0.3, senseradius=R,
               free_penalty=0.7, done_thresh=0.5, done_incr=0, terminal_reward=2.5, dist_reward=1, use_scanning=0)
    B = 97
    st = np.zeros((B, 4, W, L), np.float32)
    mi = rng.randint(0, 3, B).astype(np.int32)
    act = rng.randint(0, 4, B).astype(np.uint8)
    pos = np.zeros((B, 2), np.int16)
    for b in range(B):
        grid = maps[mi[b]]
        fx, fy = np.nonzero(grid >= 0)
        k = rng.randint(fx.size) if b % 5 else int(np.argmax((fx == 0) | (fy == 0) | (fx == W - 1) | (fy == L - 1)))
        pos[b] = (fx[k], fy[k])
        st[b, 0, fx[k], fy[k]] = 1
        cov = (rng.rand(W, L) < rng.rand()) & (grid >= 0)
        st[b, 2] = 1.0 - cov
        st[b, 1] = (rng.rand(W, L) < 0.3) & (grid < 0)
        st[b, 3] = rng.rand(W, L).astype(np.float32)
    sim = VecSuperGridSim(maps, cfg)
    for use_pos in (False, True):
        nxt, r64, term = sim.step(torch.from_numpy(st).cuda(), act, map_idx=mi, pos_xy=pos if use_pos else None)
        nxt, r64, term = nxt.cpu().numpy(), r64.cpu().numpy(), term.cpu().numpy()
        for b in range(B):
            want, wr = oc.sim_lookahead(st[b], int(act[b]), maps[mi[b]], cfg)
            assert np.array_equal(nxt[b], want.astype(np.float32)), (b, use_pos)
            assert r64[b] == wr, (b, r64[b], wr)
            pct = np.count_nonzero(want[2] < 1) / np.count_nonzero(maps[mi[b]] > 0)
            assert bool(term[b]) == (min(cfg["done_thresh"], 1) <= pct)


def test_compat_sim_class_mutates_like_the_reference():
    from marl_coverage_b200 import SuperGrid_Sim
    g = np.load(os.path.join(GOLDEN_DIR, "sim_lookahead.npz"))
    m = _meta(g)[3]
    name, cfg = m["name"], m["cfg"]
    maps = g["maps_" + name].astype(np.float64)
    sim = SuperGrid_Sim((4, 15, 15), cfg)
    for k in range(0, g["st_" + name].shape[0], 7):
        grid = maps[int(g["mi_" + name][k])]
        planes = g["st_" + name][k].astype(np.float64)
        state, r = sim.step((planes, 0), int(g["act_" + name][k]), grid)
        assert state.dtype == np.float64 and state.shape == (4, 15, 15)
        assert r == float(g["rew_" + name][k])
        assert zlib.crc32(state.astype(np.float32).tobytes()) == int(g["crc_" + name][k])
        assert np.array_equal(planes[1], state[1]) and np.array_equal(planes[2], state[2])
        assert sim.percent_covered(state, grid) == np.count_nonzero(state[2] < 1) / np.count_nonzero(grid > 0)


def test_render_matches_reference_golden():
    import torch
    from marl_coverage_b200 import VecDecGridRL, VecSuperGridRL
    g = np.load(os.path.join(GOLDEN_DIR, "render_frames.npz"))
    for m in _meta(g):
        name, kind, cfg = m["name"], m["kind"], m["cfg"]
        maps = [a.astype(np.float64) for a in g["maps_" + name]]
        E = 3
        vec = (VecDecGridRL if kind == "dec" else VecSuperGridRL)(maps, cfg, num_envs=E)
        n = cfg["numrobot"]
        start = np.broadcast_to(g["start_" + name][None], (E, n, 2)).astype(np.int16)
        vec.reset(map_idx=np.full((E,), int(m["map_index"]), np.int32), start_xy=start)
        acts = g["acts_" + name]
        for t, a in enumerate(acts):
            joint = torch.full((E,), int(a), dtype=torch.int64)
            if t >= len(acts) - 5:
                joint[1] = -1                    # env 1 stops early: its frame must differ
            vec.step(joint)
        frames = vec.render().cpu().numpy()
        want = g["frame_" + name]
        assert frames.shape == (E,) + want.shape, name
        assert np.array_equal(frames[0], want), name
        assert np.array_equal(frames[2], want), name
        sel = vec.render(env_ids=[2, 1]).cpu().numpy()
        assert np.array_equal(sel[0], want) and np.array_equal(sel[1], frames[1])


def test_compat_render_types():
    from marl_coverage_b200 import DecGridRL, SuperGridRL
    g = np.load(os.path.join(GOLDEN_DIR, "render_frames.npz"))
    metas = {m["name"]: m for m in _meta(g)}
    maps = [a.astype(np.float64) for a in g["maps_dec_rect"]]
    np.random.seed(3)
    dec = DecGridRL(maps, metas["dec_rect"]["cfg"])
    f = dec.render()
    assert f.dtype == np.uint8 and f.shape == g["frame_dec_rect"].shape
    robots = np.zeros(dec._grid.shape)
    robots[dec._xinds, dec._yinds] = 1
    assert np.array_equal(f, oc.render_frame(dec._observed_obstacles, dec._visited, robots, 1024, True))
    sup = SuperGridRL(maps, metas["super_rect"]["cfg"])
    sup.step(5)
    f = sup.render()
    assert f.dtype == np.float64 and f.shape == g["frame_super_rect"].shape
    assert np.array_equal(f, oc.render_frame(sup._observed_obstacles, sup._free < 1,
                                             sup.get_pos_image()[0], 1075, False).astype(np.float64))


@pytest.mark.parametrize("kind", ["dec", "super"])
def test_render_large_map_scale_one(kind):
    """1073 x 1073 maps: scale factor 1, DecGridRL's bordered 1075 fills the canvas exactly (no border rows)."""
    import torch
    from marl_coverage_b200 import VecDecGridRL, VecSuperGridRL, load_grid_set
    grids = load_grid_set("bg2_1073x1073", 1)
    N = 5
    if kind == "dec":
        cfg = dict(numrobot=N, maxsteps=100, collision_penalty=5, done_thresh=1, done_incr=0, terminal_reward=30,
                   dist_reward=0, train_maxsteps=100, test_maxsteps=100, egoradius=5, mini_map_rad=0, comm_radius=0,
                   allow_comm=0, map_sharing=0, single_square_tool=0, dijkstra_input=0, sensor_type="lidar",
                   sensor_config={"range": 9, "num_lasers": 21})
        vec, orc = VecDecGridRL(grids, cfg, num_envs=2, seed=4), oc.OracleDecGridRL(grids, cfg)
    else:
        cfg = dict(numrobot=N, train_maxsteps=100, test_maxsteps=100, collision_penalty=5, senseradius=3,
                   free_penalty=0.2, done_thresh=1, done_incr=0, terminal_reward=30, dist_reward=0, use_scanning=1)
        vec, orc = VecSuperGridRL(grids, cfg, num_envs=2, seed=4), oc.OracleSuperGridRL(grids, cfg)
    vec.reset(map_idx=np.zeros(2, np.int32))
    start = vec.positions()[1].cpu().numpy()
    np.random.seed(0)
    orc.reset(False, None, start_xy=start, map_index=0)
    rng = np.random.RandomState(8)
    for _ in range(15):
        a = rng.randint(0, 4, size=(2, N)).astype(np.uint8)
        vec.step(torch.from_numpy(a))
        orc.step(a[1].astype(np.int64))
    frame = vec.render(env_ids=[1]).cpu().numpy()[0]
    robots = np.zeros((orc._gridwidth, orc._gridlen))
    robots[orc._xinds, orc._yinds] = 1
    if kind == "dec":
        want = oc.render_frame(orc._observed_obstacles, orc._visited, robots, 1024, True)
    else:
        want = oc.render_frame(orc._observed_obstacles, orc._covered, robots, 1075, False)
    assert frame.shape == want.shape == (1075, 1075, 3) if kind == "dec" else frame.shape == want.shape
    assert np.array_equal(frame, want)
