"""Randomised parity: small random maps / configs, oracle vs the CUDA env, bit-exact.

Covers what the golden fixtures cannot enumerate: map sizes that are not multiples of 4 or 32
(generic store path, strip tails), all-free and nearly-blocked maps, N up to 16, radii up to the
window limit, per-robot uint8 actions, many envs with different maps/poses in one launch."""
import numpy as np
import pytest

from oracle import coverage_oracle as oc

pytestmark = pytest.mark.gpu


def _rand_map(rng, W, L, p_obst):
    g = np.where(rng.random((W, L)) < p_obst, -1.0, 1.0)
    free = np.argwhere(g > 0)
    if len(free) < 20:                       # keep enough room for the robots
        g[:] = 1.0
    return g


def _dec_cfg(rng, N):
    lidar = rng.random() < 0.4
    R = int(rng.integers(1, 7 if lidar else 5))
    cfg = dict(numrobot=N, maxsteps=int(rng.integers(20, 60)), collision_penalty=float(rng.integers(0, 6)),
               done_thresh=float(rng.choice([0.2, 0.6, 1.0, 1.5])), done_incr=float(rng.choice([0.0, 0.1])),
               terminal_reward=float(rng.integers(0, 40)), dist_reward=int(rng.random() < 0.3),
               train_maxsteps=1000, test_maxsteps=1000, egoradius=int(rng.integers(1, 8)),
               mini_map_rad=int(rng.integers(1, 11)) if rng.random() < 0.2 else 0,
               comm_radius=int(rng.integers(0, 6)), allow_comm=1, map_sharing=int(rng.random() < 0.3),
               single_square_tool=int(rng.random() < 0.2), dijkstra_input=int(rng.random() < 0.25))
    if lidar:
        cfg.update(sensor_type="lidar", sensor_config={"range": R, "num_lasers": int(rng.choice([5, 9, 21, 33]))})
    else:
        cfg.update(sensor_type="square_sensor", sensor_config={"range": R})
    return cfg


def _super_cfg(rng, N):
    return dict(numrobot=N, train_maxsteps=1000, test_maxsteps=1000,
                collision_penalty=float(rng.integers(0, 6)), senseradius=int(rng.integers(0, 5)),
                free_penalty=float(rng.choice([0.0, 0.2, 0.37, 1.0])),
                done_thresh=float(rng.choice([0.3, 0.7, 1.0])), done_incr=float(rng.choice([0.0, 0.05])),
                terminal_reward=float(rng.integers(0, 40)), dist_reward=int(rng.random() < 0.4),
                use_scanning=int(rng.random() < 0.4))


def _wide(rng, kind):
    """Forced values drawn over the whole range include/mcov.h accepts (soak runs: MCOV_SOAK=n)."""
    f = dict(W=int(rng.integers(3, 70)), L=int(rng.integers(3, 70)), N=int(rng.integers(1, 33)))
    while f["W"] * f["L"] < 3 * f["N"] + 20:
        f["W"] += 3
        f["L"] += 2
    if kind == "dec":
        f["egoradius"] = int(rng.integers(0, 16))
        f["mini_map_rad"] = int(rng.integers(1, 16)) if rng.random() < 0.2 else 0
        f["comm_radius"] = int(rng.integers(0, 40))
        if rng.random() < 0.5:
            f.update(sensor_type="lidar", sensor_config={"range": int(rng.integers(0, 16)),
                                                         "num_lasers": int(rng.choice([1, 3, 7, 21, 45, 201]))})
        else:
            f.update(sensor_type="square_sensor", sensor_config={"range": int(rng.integers(0, 16))})
    else:
        f["senseradius"] = int(rng.integers(0, 16))
    return f


def _run_case(kind, seed, E=9, T=40, force=None):
    import torch
    from marl_coverage_b200 import VecDecGridRL, VecSuperGridRL
    rng = np.random.default_rng(seed)
    W, L = int(rng.integers(6, 45)), int(rng.integers(6, 45))
    if seed % 3 == 0:
        L = (L // 4) * 4 + 4                 # exercise the vectorised store path too
    N = int(rng.integers(1, 9 if kind == "dec" else 17))
    force = dict(force or {})
    W, L, N = force.pop("W", W), force.pop("L", L), force.pop("N", N)
    grids = [_rand_map(rng, W, L, float(rng.choice([0.0, 0.1, 0.3]))) for _ in range(3)]
    cfg = _dec_cfg(rng, N) if kind == "dec" else _super_cfg(rng, N)
    cfg.update(force)
    np.random.seed(seed)
    mk = (lambda: oc.OracleDecGridRL(grids, cfg, use_graph=True)) if kind == "dec" else \
         (lambda: oc.OracleSuperGridRL(grids, cfg))
    oracles = [mk() for _ in range(E)]
    cls = VecDecGridRL if kind == "dec" else VecSuperGridRL
    vec = cls(grids, cfg, num_envs=E, device="cuda:0")

    def reset_all(which):
        maps = np.zeros((E,), np.int32)
        start = np.zeros((E, N, 2), np.int16)
        mask = np.zeros((E,), np.uint8)
        outs = {}
        for e in which:
            o = oracles[e]
            mi = int(rng.integers(len(grids)))
            o.reset(False, None, map_index=mi)               # draws poses from np.random
            maps[e], mask[e] = mi, 1
            start[e, :, 0], start[e, :, 1] = o._xinds, o._yinds
            outs[e] = o.reset(False, None, start_xy=start[e], map_index=mi)[0]
        obs = vec.reset(map_idx=maps, start_xy=start, mask=mask)
        for e in which:
            ref = outs[e] if kind == "dec" else outs[e][0]
            got = obs[e].cpu().numpy()
            if kind == "dec" and cfg["mini_map_rad"] > 0:
                assert np.abs(got - np.asarray(ref, dtype=np.float32)).max() <= 1e-6
            else:
                assert np.array_equal(got, np.asarray(ref, dtype=np.float32)), \
                    "seed %d reset obs env %d" % (seed, e)

    reset_all(range(E))
    for t in range(T):
        act = rng.integers(0, 4, size=(E, N)).astype(np.uint8)
        obs, rew, done = vec.step(torch.from_numpy(act))
        r64 = vec.reward64.cpu().numpy()
        dn = done.cpu().numpy()
        ob = obs.cpu().numpy()
        pos = vec.positions().cpu().numpy()
        pct = vec.percent_covered().cpu().numpy()
        finished = []
        for e, o in enumerate(oracles):
            oo, orr, od = o.step(act[e].astype(np.int64))
            tag = "seed %d step %d env %d (%s W=%d L=%d N=%d cfg=%s)" % (seed, t, e, kind, W, L, N, cfg)
            if kind == "dec":
                oo = oo[0] if isinstance(oo, list) else oo
                if cfg["mini_map_rad"] > 0:      # bilinear layers: 1e-6
                    assert np.abs(ob[e] - oo.astype(np.float32)).max() <= 1e-6, "obs " + tag
                else:
                    assert np.array_equal(ob[e], oo.astype(np.float32)), "obs " + tag
            else:
                assert np.array_equal(ob[e], oo[0].astype(np.float32)), "state " + tag
            assert np.array_equal(pos[e, :, 0], o._xinds) and np.array_equal(pos[e, :, 1], o._yinds), "pos " + tag
            assert float(r64[e]) == float(orr), "reward %r vs %r %s" % (r64[e], orr, tag)
            assert bool(dn[e]) == bool(od), "done " + tag
            assert float(pct[e]) == float(o.percent_covered()), "pct " + tag
            if od:
                finished.append(e)
        if kind == "dec":
            adj = vec.adjacency()
            for e, o in enumerate(oracles):
                assert np.array_equal(adj[e], o._adjacency_matrix), "adjacency seed %d" % seed
        if finished:
            reset_all(finished)


@pytest.mark.parametrize("seed", range(40))
def test_dec_random_vs_oracle(seed):
    _run_case("dec", 1000 + seed)


@pytest.mark.parametrize("seed", range(16))
def test_dec_random_vs_oracle_store_warp(seed, monkeypatch):
    """Same cases through the store-warp path (slot hand-over + TMA bulk stores), one CTA looping over the
    batch: masked resets publish empty slots, generic shapes fall back to direct stores when the block is
    not a 16-byte multiple."""
    monkeypatch.setenv("MCOV_DEC_SHAPE", "0")
    monkeypatch.setenv("MCOV_DEC_STORE", "1")
    monkeypatch.setenv("MCOV_DEC_GRID", "1")
    _run_case("dec", 1000 + seed, E=70)


@pytest.mark.parametrize("tma", [0, 1])
@pytest.mark.parametrize("seed", range(20))
def test_dec_random_vs_oracle_cta_shape(seed, tma, monkeypatch):
    """Same cases through the CTA-per-env kernel (auto-selected for small batches; pinned here), with and
    without bulk-copy staging of the map windows: odd map sizes put the staged runs at every alignment."""
    monkeypatch.setenv("MCOV_DEC_SHAPE", "1")
    monkeypatch.setenv("MCOV_DEC_TMA", str(tma))
    monkeypatch.setenv("MCOV_DEC_GRID", "4")
    _run_case("dec", 1000 + seed)


@pytest.mark.parametrize("seed", range(24))
def test_super_random_vs_oracle(seed):
    _run_case("super", 2000 + seed)


# ---- limits of the ABI (include/mcov.h: MCOV_MAX_ROBOTS 32, MCOV_MAX_RADIUS 15, odd num_lasers) and degenerate shapes
EDGE = [
    ("dec", dict(N=32, W=40, L=33, egoradius=2, sensor_type="square_sensor", sensor_config={"range": 1},
                 mini_map_rad=0, dijkstra_input=0, dist_reward=0)),
    ("dec", dict(N=17, W=35, L=70, egoradius=3, sensor_type="lidar", sensor_config={"range": 4, "num_lasers": 9},
                 mini_map_rad=0)),
    ("dec", dict(N=2, W=50, L=50, egoradius=15, sensor_type="square_sensor", sensor_config={"range": 15},
                 mini_map_rad=0)),
    ("dec", dict(N=3, W=40, L=40, egoradius=4, sensor_type="lidar", sensor_config={"range": 15, "num_lasers": 1},
                 mini_map_rad=0)),
    ("dec", dict(N=2, W=30, L=30, egoradius=5, sensor_type="lidar", sensor_config={"range": 6, "num_lasers": 1023},
                 mini_map_rad=0)),
    ("dec", dict(N=1, W=5, L=6, egoradius=1, sensor_type="square_sensor", sensor_config={"range": 0},
                 mini_map_rad=0)),
    ("dec", dict(N=4, W=33, L=31, egoradius=7, mini_map_rad=15, sensor_type="square_sensor",
                 sensor_config={"range": 3})),
    ("super", dict(N=31, W=30, L=36, senseradius=1)),
    ("super", dict(N=2, W=5, L=4, senseradius=4)),
    ("super", dict(N=3, W=64, L=32, senseradius=15)),
    ("super", dict(N=1, W=7, L=129, senseradius=0, use_scanning=1)),
]


@pytest.mark.parametrize("k", range(int(__import__("os").environ.get("MCOV_SOAK", "6"))))
@pytest.mark.parametrize("kind", ["dec", "super"])
def test_whole_parameter_range_vs_oracle(kind, k):
    seed = 9000 + k
    _run_case(kind, seed, E=4, T=20, force=_wide(np.random.default_rng(seed), kind))


@pytest.mark.parametrize("case", range(len(EDGE)))
def test_limits_and_degenerate_shapes_vs_oracle(case):
    kind, force = EDGE[case]
    _run_case(kind, 7000 + case, E=3, T=25, force=force)


@pytest.mark.parametrize("case", [k for k in range(len(EDGE)) if EDGE[k][0] == "dec"])
def test_limits_and_degenerate_shapes_cta_shape(case, monkeypatch):
    """The same limits (32 robots = 32 warps, radius 15 = 31-row bands, 1023 lasers, 5 x 6 maps) through the
    CTA-per-env kernel."""
    monkeypatch.setenv("MCOV_DEC_SHAPE", "1")
    kind, force = EDGE[case]
    _run_case(kind, 7000 + case, E=3, T=25, force=force)


def test_bad_configs_are_refused():
    from marl_coverage_b200 import VecDecGridRL, VecSuperGridRL, _lib
    g = [np.ones((10, 10))]
    base = _dec_cfg(np.random.default_rng(0), 2)
    base.update(sensor_type="square_sensor", sensor_config={"range": 2}, egoradius=3, mini_map_rad=0)
    for bad in (dict(numrobot=33), dict(egoradius=16), dict(sensor_config={"range": 16}),
                dict(sensor_type="lidar", sensor_config={"range": 3, "num_lasers": 8})):
        with pytest.raises((_lib.McovError, AssertionError, ValueError)):
            VecDecGridRL(g, dict(base, **bad), num_envs=2)
    with pytest.raises(_lib.McovError):
        VecSuperGridRL(g, dict(_super_cfg(np.random.default_rng(0), 2), senseradius=16), num_envs=2)
    with pytest.raises((ValueError, _lib.McovError)):
        VecDecGridRL([np.full((10, 10), np.inf)], base, num_envs=2)   # cell values must be finite
    # 32-bit plane offsets: refused with a message before any arena is allocated
    with pytest.raises(_lib.McovError, match="shard the batch"):
        VecDecGridRL([np.ones((1073, 1073))], dict(base, numrobot=16), num_envs=8000)
