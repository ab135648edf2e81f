"""The drop-in boundary under stress (GPU): the truncation bit of the done byte, caller-supplied device values that
would be addresses (map indices, start poses, replay sample indices), free cells worth something other than 1
(super_grid_rl.py:189), and the one-handle-one-device rule."""
import numpy as np
import pytest

from oracle import coverage_oracle as oc

pytestmark = pytest.mark.gpu

DEC = dict(numrobot=3, maxsteps=1000, collision_penalty=5, done_thresh=1, done_incr=0, terminal_reward=30,
           dist_reward=0, train_maxsteps=1000, test_maxsteps=1000, egoradius=3, mini_map_rad=0, comm_radius=0,
           allow_comm=0, map_sharing=0, single_square_tool=0, dijkstra_input=0, sensor_type="square_sensor",
           sensor_config={"range": 1})
SUPER = dict(numrobot=3, train_maxsteps=1000, test_maxsteps=1000, collision_penalty=5, senseradius=1,
             free_penalty=0.2, done_thresh=1, done_incr=0, terminal_reward=30, dist_reward=0, use_scanning=0)


def _grids(rng, n=3, W=17, L=22, valued=False, worths=(0.0, 0.25, 1.0, 2.5, 3.0, 0.1)):
    out = []
    for _ in range(n):
        g = np.where(rng.random((W, L)) < 0.15, -1.0, 1.0)
        if valued:
            worth = rng.choice(list(worths), size=(W, L))
            g = np.where(g > 0, worth, g * rng.choice([1.0, 2.0], size=(W, L)))
        out.append(g)
    return out


@pytest.mark.parametrize("kind", ["dec", "super"])
@pytest.mark.parametrize("auto_reset", [False, True])
def test_truncation_sets_bit_1_of_done(kind, auto_reset):
    """with auto_reset the env restarts in place at truncate_steps and the caller must still see the
    episode boundary — generate_episode stores done = True at the cut (Utils/utils.py:25-31)."""
    import torch
    from marl_coverage_b200 import VecDecGridRL, VecSuperGridRL, _lib
    rng = np.random.default_rng(5)
    grids = _grids(rng)
    cls, cfg = (VecDecGridRL, DEC) if kind == "dec" else (VecSuperGridRL, SUPER)
    E, N, cut = 6, cfg["numrobot"], 7
    vec = cls(grids, cfg, num_envs=E, auto_reset=auto_reset, truncate_steps=cut, seed=11)
    vec.reset()
    for t in range(1, 2 * cut + 2):
        act = torch.from_numpy(rng.integers(0, 4, size=(E, N)).astype(np.uint8))
        _, _, done = vec.step(act)
        d = done.cpu().numpy()
        head = {k: v.cpu().numpy() for k, v in vec.head().items()}
        nr = vec.needs_reset().cpu().numpy()
        assert not (d & _lib.DONE_TERMINAL).any()                # nobody covers a 17 x 22 map in 15 steps
        if auto_reset:
            at_cut = (t % cut == 0)
            assert ((d & _lib.DONE_TRUNCATED) != 0).all() == at_cut and ((d != 0).any() == at_cut)
            assert (head["currstep"] == (t % cut)).all() and (head["episode"] == 1 + t // cut).all()
            assert not nr.any()                                  # the fused re-seed consumed the flag
        else:
            assert (((d & _lib.DONE_TRUNCATED) != 0) == (t >= cut)).all()
            assert (nr == (1 if t >= cut else 0)).all() and (head["currstep"] == t).all()


def test_terminal_and_truncated_together():
    """An env that reaches its coverage threshold on the very step of the cut reports both bits."""
    import torch
    from marl_coverage_b200 import VecDecGridRL, _lib
    g = [np.ones((3, 3))]
    cfg = dict(DEC, numrobot=1, done_thresh=0.5, sensor_config={"range": 1}, egoradius=1)
    vec = VecDecGridRL(g, cfg, num_envs=2, auto_reset=False, truncate_steps=1)
    vec.reset(map_idx=np.zeros(2, np.int32), start_xy=np.full((2, 1, 2), 2, np.int16))
    _, _, done = vec.step(torch.zeros((2, 1), dtype=torch.uint8))
    assert (done.cpu().numpy() == (_lib.DONE_TERMINAL | _lib.DONE_TRUNCATED)).all()


@pytest.mark.parametrize("kind", ["dec", "super"])
def test_injected_draws_host_checked_device_clamped(kind):
    import torch
    from marl_coverage_b200 import VecDecGridRL, VecSuperGridRL
    rng = np.random.default_rng(6)
    grids = _grids(rng)
    cls, cfg = (VecDecGridRL, DEC) if kind == "dec" else (VecSuperGridRL, SUPER)
    E, N = 5, cfg["numrobot"]
    vec = cls(grids, cfg, num_envs=E)
    ok_xy = np.zeros((E, N, 2), np.int16)
    ok_xy[:, :, 0] = 1 + np.arange(N)[None]
    ok_xy[:, :, 1] = 2
    # host arrays: refused before a pointer reaches a kernel
    with pytest.raises(ValueError):
        vec.reset(map_idx=np.full(E, len(grids), np.int32), start_xy=ok_xy)
    with pytest.raises(ValueError):
        vec.reset(map_idx=np.full(E, -1, np.int32), start_xy=ok_xy)
    bad = ok_xy.copy()
    bad[2, 1, 0] = vec.W
    with pytest.raises(ValueError):
        vec.reset(map_idx=np.zeros(E, np.int32), start_xy=bad)
    # device tensors: clamped into the table / the plane by the kernel, the batch stays usable
    mi = torch.tensor([0, 99, -7, 1, 2 ** 30], dtype=torch.int32, device="cuda:0")
    xy = torch.from_numpy(ok_xy).cuda()
    xy[1, 0, 0] = 30000
    xy[3, 2, 1] = -30000
    vec.reset(map_idx=mi, start_xy=xy)
    torch.cuda.synchronize()
    assert vec.head()["map_idx"].cpu().tolist() == [0, len(grids) - 1, 0, 1, len(grids) - 1]
    pos = vec.positions().cpu().numpy()
    assert pos[1, 0, 0] == vec.W - 1 and pos[3, 2, 1] == 0
    assert (pos[..., 0] >= 0).all() and (pos[..., 0] < vec.W).all() and (pos[..., 1] >= 0).all() and (pos[..., 1] < vec.L).all()
    for _ in range(5):
        vec.step(torch.from_numpy(rng.integers(0, 4, size=(E, N)).astype(np.uint8)))
    torch.cuda.synchronize()


def test_replay_gather_indices_host_checked_device_clamped():
    import torch
    from marl_coverage_b200 import ReplayBuffer
    buf = ReplayBuffer((3, 5, 5), 1, 16, device="cuda:0")
    E = 16
    st = torch.arange(E * 75, dtype=torch.float32, device="cuda:0").view(E, 3, 5, 5)
    buf.addbatch(st, torch.zeros((E, 1), dtype=torch.uint8, device="cuda:0"), torch.arange(E, dtype=torch.float32, device="cuda:0"),
                 st + 0.5, torch.zeros(E, dtype=torch.uint8, device="cuda:0"))
    with pytest.raises(IndexError):
        buf.gather(np.array([0, 16]))
    with pytest.raises(IndexError):
        buf.gather(np.array([-1]))
    s, a, r, n, d = buf.gather(torch.tensor([3, 10 ** 12, -5, 15], dtype=torch.int64, device="cuda:0"))
    assert r.cpu().tolist() == [3.0, 15.0, 0.0, 15.0]
    assert torch.equal(s[1], st[15]) and torch.equal(s[2], st[0]) and torch.equal(n[3], st[15] + 0.5)


@pytest.mark.parametrize("seed", range(6))
def test_super_cell_values_vs_oracle(seed):
    """Free cells worth 0 / 0.25 / 1 / 2.5 / 3 / 0.1: the covered cell's value enters the float64 reward in the
    reference's raster order (super_grid_rl.py:189) — bit-exact against the oracle."""
    import torch
    from marl_coverage_b200 import VecSuperGridRL
    rng = np.random.default_rng(300 + seed)
    grids = _grids(rng, W=int(rng.integers(6, 30)), L=int(rng.integers(6, 40)), valued=True)
    cfg = dict(SUPER, numrobot=int(rng.integers(1, 6)), senseradius=int(rng.integers(0, 4)),
               free_penalty=float(rng.choice([0.0, 0.2, 0.37])), use_scanning=int(rng.random() < 0.4),
               done_thresh=float(rng.choice([0.5, 1.0])))
    E, N = 7, cfg["numrobot"]
    np.random.seed(seed)
    oracles = [oc.OracleSuperGridRL(grids, cfg) for _ in range(E)]
    vec = VecSuperGridRL(grids, cfg, num_envs=E)
    assert vec.maps.values is not None
    maps = np.arange(E, dtype=np.int32) % len(grids)
    start = np.zeros((E, N, 2), np.int16)
    for e, o in enumerate(oracles):
        o.reset(False, None, map_index=int(maps[e]))
        start[e, :, 0], start[e, :, 1] = o._xinds, o._yinds
        o.reset(False, None, start_xy=start[e], map_index=int(maps[e]))
    vec.reset(map_idx=maps, start_xy=start)
    for t in range(40):
        act = rng.integers(0, 4, size=(E, N)).astype(np.uint8)
        obs, _, done = vec.step(torch.from_numpy(act))
        r64, dn, ob = vec.reward64.cpu().numpy(), done.cpu().numpy(), obs.cpu().numpy()
        pct = vec.percent_covered().cpu().numpy()
        for e, o in enumerate(oracles):
            if getattr(o, "_finished", False):
                continue
            oo, orr, od = o.step(act[e].astype(np.int64))
            assert float(r64[e]) == float(orr), "reward %r vs %r (seed %d step %d env %d)" % (r64[e], orr, seed, t, e)
            assert bool(dn[e]) == bool(od) and float(pct[e]) == float(o.percent_covered())
            assert np.array_equal(ob[e], oo[0].astype(np.float32))
            o._finished = bool(od)


def test_dec_ignores_cell_values_like_the_reference():
    """DecGridRL's reward counts newly visited cells (dec_grid_rl.py:243-256) and its sensor stores clip(cell, 0, 1)
    (squaresensor.py:34-35): maps with magnitudes >= 1 behave like their sign maps; magnitudes inside (0, 1) are refused."""
    import torch
    from marl_coverage_b200 import VecDecGridRL
    rng = np.random.default_rng(9)
    with pytest.raises(ValueError):
        VecDecGridRL(_grids(rng, valued=True), DEC, num_envs=2)
    grids = _grids(rng, valued=True, worths=(0.0, 1.0, 2.5, 3.0))
    signs = [np.sign(g) for g in grids]
    E, N = 4, DEC["numrobot"]
    a = VecDecGridRL(grids, DEC, num_envs=E, seed=3)
    b = VecDecGridRL(signs, DEC, num_envs=E, seed=3)
    np.random.seed(1)
    o = oc.OracleDecGridRL(grids, DEC)
    o.reset(False, None, map_index=0)
    start = np.stack([o._xinds, o._yinds], axis=1).astype(np.int16)
    o.reset(False, None, start_xy=start, map_index=0)
    sx = np.broadcast_to(start[None], (E, N, 2)).copy()
    a.reset(map_idx=np.zeros(E, np.int32), start_xy=sx)
    b.reset(map_idx=np.zeros(E, np.int32), start_xy=sx)
    for t in range(25):
        act = np.broadcast_to(rng.integers(0, 4, size=(1, N)).astype(np.uint8), (E, N)).copy()
        oa, _, _ = a.step(torch.from_numpy(act))
        ob, _, _ = b.step(torch.from_numpy(act))
        oo, orr, od = o.step(act[0].astype(np.int64))
        assert torch.equal(oa, ob) and torch.equal(a.reward64, b.reward64)
        assert float(a.reward64[0]) == float(orr) and np.array_equal(oa[0].cpu().numpy(), oo.astype(np.float32))


def test_sim_lookahead_cell_values_vs_oracle():
    """sim_env.py:159 `reward += grid[j][k]` with valued free cells."""
    import torch
    from marl_coverage_b200 import VecSuperGridSim
    rng = np.random.default_rng(12)
    grids = _grids(rng, n=2, W=12, L=16, valued=True)
    cfg = dict(SUPER, numrobot=1, senseradius=2, use_scanning=1, done_thresh=0.5)
    B, W, L = 24, 12, 16
    state = np.zeros((B, 4, W, L), np.float32)
    acts = rng.integers(0, 4, size=B).astype(np.uint8)
    mi = rng.integers(0, 2, size=B).astype(np.int32)
    for b in range(B):
        grid = grids[mi[b]]
        free = np.argwhere(grid >= 0)
        x, y = free[rng.integers(len(free))]
        state[b, 0, x, y] = 1.0
        state[b, 1] = (rng.random((W, L)) < 0.1) & (grid < 0)
        state[b, 2] = 1.0 - ((rng.random((W, L)) < 0.5) & (grid >= 0))
        state[b, 3] = rng.random((W, L)).astype(np.float32)
    vsim = VecSuperGridSim(grids, cfg)
    nxt, r64, term = vsim.step(torch.from_numpy(state).cuda(), acts, map_idx=mi)
    nxt, r64, term = nxt.cpu().numpy(), r64.cpu().numpy(), term.cpu().numpy()
    for b in range(B):
        want, wr = oc.sim_lookahead(state[b], int(acts[b]), grids[mi[b]], cfg)
        assert r64[b] == wr, "pair %d: %r vs %r" % (b, r64[b], wr)
        assert np.array_equal(nxt[b], want.astype(np.float32))
        pct = np.count_nonzero(want[2] < 1) / np.count_nonzero(grids[mi[b]] > 0)
        assert bool(term[b]) == (min(cfg["done_thresh"], 1) <= pct)


def test_handle_is_bound_to_its_device():
    """One handle <-> one device: a call with another device current is refused (MCOV_ERR_STATE), and a second
    device in the same process gets its own kernel attributes (1073^2 SuperGridRL needs the opt-in shared memory)."""
    import torch
    from marl_coverage_b200 import VecSuperGridRL, _lib
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    g = [np.ones((40, 40))]
    envs = [VecSuperGridRL(g, SUPER, num_envs=3, device="cuda:%d" % d) for d in (0, 1)]
    for env in envs:
        env.reset()
        env.step(torch.zeros((3, 3), dtype=torch.uint8, device=env.device))
        torch.cuda.synchronize(env.device)
    with torch.cuda.device(1):
        rc = envs[0]._lib.mcov_observe(envs[0]._handle, envs[0].obs.data_ptr(), None)
    assert rc == -5 and b"device" in envs[0]._lib.mcov_last_error()
