"""Parity at BASELINE.json's full sizes (configs[1..4] and the SuperGridRL batch the bench runs), through
properties that do not need the oracle to step the whole batch:

* replication: every env of the full batch replays one reference-recorded golden session — env 0 / middle /
  last must match the golden snapshots bit for bit and every other env must equal env 0 (one device-side
  comparison per step), which exercises the persistent loops, 32-bit arena offsets and grid sizing at size;
* sampling: a full-size batch with device-RNG resets and random actions is stepped, a few dozen envs spread
  over the batch are re-stepped by the oracle from the poses read back after reset;
* invariants over the whole batch: known-free counts never shrink within an episode, percent_covered matches the
  counts, reward64 - float32 reward agree, observe() after step() is idempotent, needs_reset == done.
"""
import os

import numpy as np
import pytest

from conftest import GOLDEN_DIR
from oracle import coverage_oracle as oc
from oracle.trace import OP_RESET, Trace, compare_snap

pytestmark = pytest.mark.gpu

DEC = dict(maxsteps=1000, collision_penalty=5, done_thresh=1, done_incr=0, terminal_reward=30, dist_reward=0,
           train_maxsteps=1000, test_maxsteps=1000, egoradius=5, mini_map_rad=0, comm_radius=0, allow_comm=0,
           map_sharing=0, single_square_tool=0, dijkstra_input=0)
SUPER = dict(numrobot=4, train_maxsteps=1000, test_maxsteps=1000, collision_penalty=5, senseradius=2,
             free_penalty=0.2, done_thresh=1, done_incr=0, terminal_reward=30, dist_reward=0, use_scanning=0)

FULL = {
    # name: (kind, grid set, maps, envs, cfg, oracle-sampled envs, steps)
    "c2": ("dec", "bg2_100x100", 37, 4096, dict(DEC, numrobot=4, sensor_type="square_sensor",
                                                 sensor_config={"range": 2}), 24, 40),
    "c3": ("dec", "bg2_256x256", 8, 16384, dict(DEC, numrobot=8, sensor_type="lidar",
                                                sensor_config={"range": 10, "num_lasers": 21}), 12, 25),
    "c4": ("dec", "bg2_1073x1073", 2, 256, dict(DEC, numrobot=16, sensor_type="square_sensor",
                                                sensor_config={"range": 2}), 3, 12),
    "c5": ("dec", "bg2_100x100", 37, 131072, dict(DEC, numrobot=4, sensor_type="square_sensor",
                                                  sensor_config={"range": 2}), 32, 40),
    "super": ("super", "bg2_100x100", 37, 8192, SUPER, 16, 25),
}


def _snap_env(vec, kind, e, reward, done, obs):
    """gpu_helpers.vec_snapshot for ONE env of a large batch: slices the arena instead of unpacking all E envs."""
    from oracle.trace import crc_f32
    head = vec.head()
    lay, N, PW = vec.layout, vec.num_robots, vec.layout.plane_words
    s = dict(pos=vec.positions()[e].cpu().numpy(), reward=float(reward), done=int(done),
             pct=float(vec.percent_covered()[e].item()), done_thresh=float(head["done_thresh"][e].item()),
             currstep=int(head["currstep"][e].item()))
    if kind == "dec":
        fo = vec._plane_words(lay.off_fo, N * PW * 2)[e:e + 1].view(1, N, PW, 2)
        s["obs"] = obs[e].cpu().numpy()
        s["adj"] = vec.adjacency()[e].astype(np.uint8)
        s["ckpt_a"] = vec._unpack(fo[..., 0].contiguous())[0].astype(np.uint8)
        s["ckpt_b"] = vec._unpack(fo[..., 1].contiguous())[0].astype(np.uint8)
        s["ckpt_visited"] = vec._unpack(vec._plane_words(lay.off_visited, PW)[e:e + 1])[0].astype(np.uint8)
    else:
        st = obs[e].cpu().numpy()
        s["state_crc"] = crc_f32(st)
        s["ckpt_a"] = vec._unpack(vec._plane_words(lay.off_covered, PW)[e:e + 1])[0].astype(np.uint8)
        s["ckpt_b"] = vec._unpack(vec._plane_words(lay.off_seen_obst, PW)[e:e + 1])[0].astype(np.uint8)
        s["ckpt_state"] = st
    return s


def _make(kind, grids, cfg, E, **kw):
    from marl_coverage_b200 import VecDecGridRL, VecSuperGridRL
    return (VecDecGridRL if kind == "dec" else VecSuperGridRL)(grids, cfg, num_envs=E, **kw)


@pytest.mark.parametrize("name,E,steps", [("dec_c2_square_100", 4096, 300), ("dec_c2_square_100", 131072, 120),
                                          ("dec_c3_lidar_256_n8", 16384, 80), ("dec_c4_square_1073_n16", 256, 40),
                                          ("super_c1_100", 8192, 60)])
def test_golden_session_replicated_at_full_size(name, E, steps):
    import torch
    tr = Trace.load(os.path.join(GOLDEN_DIR, name + ".npz"))
    vec = _make(tr.kind, tr.float_maps(), tr.config, E)
    N = tr.config["numrobot"]
    check = (0, E // 2 + 1, E - 1)
    nsteps = 0
    for k in range(tr.num_ops):
        if tr.op_kind[k] == OP_RESET:
            if nsteps >= steps:
                break
            mi = torch.full((E,), int(tr.op_map[k]), dtype=torch.int32, device="cuda")
            sx = torch.from_numpy(np.ascontiguousarray(tr.op_start[k])).cuda().expand(E, N, 2).contiguous()
            obs = vec.reset(map_idx=mi, start_xy=sx)
            rew, done = np.zeros(E), np.zeros(E, dtype=np.uint8)
        else:
            if nsteps >= steps:
                continue
            nsteps += 1
            act = torch.full((E,), int(tr.op_action[k]), dtype=torch.int64, device="cuda")
            obs, _, d = vec.step(act)
            rew, done = vec.reward64.cpu().numpy(), d.cpu().numpy()
            assert (rew == rew[0]).all() and (done == done[0]).all()
        # every env equals env 0 (device-side), a few are compared with the reference's snapshot
        assert bool((obs == obs[0:1]).all().item()), "op %d: envs diverged" % k
        assert bool((vec.positions() == vec.positions()[0:1]).all().item())
        if k % 10 == 0 or tr.op_kind[k] == OP_RESET:
            for e in check:
                compare_snap(tr, k, _snap_env(vec, tr.kind, e, rew[e], done[e], obs), where="env %d of %d" % (e, E))
    assert nsteps > 0


@pytest.mark.parametrize("name", sorted(FULL))
def test_full_size_random_rollout_sampled_against_oracle(name):
    import torch
    from marl_coverage_b200 import load_grid_set
    kind, gset, nmaps, E, cfg, nsample, T = FULL[name]
    grids = load_grid_set(gset, nmaps)
    N = cfg["numrobot"]
    vec = _make(kind, grids, cfg, E, seed=11)
    mi = (torch.arange(E, dtype=torch.int32, device="cuda") * 7 + 3) % len(grids)
    obs = vec.reset(map_idx=mi)                                      # poses: device Philox
    pos0 = vec.positions().cpu().numpy()
    sample = np.unique(np.linspace(0, E - 1, nsample).astype(np.int64))
    np.random.seed(0)
    oracles = {}
    for e in sample:
        o = (oc.OracleDecGridRL if kind == "dec" else oc.OracleSuperGridRL)(grids, cfg)
        first = o.reset(False, None, start_xy=pos0[e], map_index=int(mi[e].item()))[0]
        first = first if kind == "dec" else first[0]
        assert np.array_equal(obs[e].cpu().numpy(), np.asarray(first, dtype=np.float32)), "reset obs env %d" % e
        oracles[int(e)] = o
    # start poses drawn on the device: distinct walkable cells of the env's own map
    walls = np.stack([np.pad(g, 1, constant_values=-1) if kind == "dec" else g for g in grids]) < 0
    mih = mi.cpu().numpy()
    for e in sample:
        p = pos0[e]
        assert not walls[mih[e], p[:, 0], p[:, 1]].any() and len({(int(a), int(b)) for a, b in p}) == N
    gen = torch.Generator(device="cuda")
    gen.manual_seed(99)
    cov_prev = vec.coverage_counts().clone() if kind == "dec" else None
    sidx = torch.from_numpy(sample).cuda()
    for t in range(T):
        act = torch.randint(0, 4, (E, N), dtype=torch.uint8, device="cuda", generator=gen)
        obs, rew, done = vec.step(act)
        # whole-batch invariants
        assert torch.equal(vec.reward64.float(), rew)
        assert torch.equal(vec.needs_reset(), done)
        again = vec.observe(out=torch.empty_like(obs))
        assert torch.equal(again, obs), "observe() after step() must be idempotent"
        if kind == "dec":
            cov = vec.coverage_counts()
            assert bool((cov >= cov_prev).all().item())
            cov_prev = cov.clone()
            fc = vec.maps.free_counts[mi.long()].double()
            assert torch.equal(vec.percent_covered(), cov.sum(dim=1).double() / fc)
        # sampled envs against the oracle
        a_h = act[sidx].cpu().numpy()
        o_h = obs[sidx].cpu().numpy()
        r_h = vec.reward64[sidx].cpu().numpy()
        d_h = done[sidx].cpu().numpy()
        p_h = vec.positions()[sidx].cpu().numpy()
        for j, e in enumerate(sample):
            o = oracles[int(e)]
            if o is None:
                continue
            oo, orr, od = o.step(a_h[j].astype(np.int64))
            oo = oo if kind == "dec" else oo[0]
            tag = "%s env %d step %d" % (name, e, t)
            assert np.array_equal(o_h[j], np.asarray(oo, dtype=np.float32)), "obs " + tag
            assert float(r_h[j]) == float(orr), "reward " + tag
            assert bool(d_h[j]) == bool(od), "done " + tag
            assert np.array_equal(p_h[j, :, 0], o._xinds) and np.array_equal(p_h[j, :, 1], o._yinds), "pos " + tag
            if od:
                oracles[int(e)] = None                                # finished: stop following it
