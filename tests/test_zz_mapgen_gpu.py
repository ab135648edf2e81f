"""gridgen on the device (mcov_gen_maps): Bernoulli maps written straight into the packed
planes.  The draws are a Philox stream, not numpy's, so the checks are the properties the reference's
gridgen (Utils/gridmaker.py:107-139) guarantees plus exact agreement with the packer on the same maps."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

GC = dict(numgrids=6, gridwidth=37, gridlen=70, prob_obst=0.25)


@pytest.mark.parametrize("border", [False, True])
def test_generated_planes_equal_packed_planes(border):
    """generate -> host grids -> the ordinary packer must give back the very same words and free counts."""
    import torch
    from marl_coverage_b200 import MapTable
    gen = MapTable.generate(GC, add_border=border, seed=11)
    grids = gen.to_grids()
    assert len(grids) == 6 and grids[0].shape == (37, 70)
    assert all(set(np.unique(g)) <= {-1.0, 1.0} for g in grids)              # gridgen's value set
    packed = MapTable(grids, add_border=border, device="cuda:0")
    assert (gen.W, gen.L, gen.plane_words) == (packed.W, packed.L, packed.plane_words)
    assert torch.equal(gen.wall, packed.wall) and torch.equal(gen.pos, packed.pos)
    assert torch.equal(gen.free_counts, packed.free_counts)
    want = np.array([np.count_nonzero(g > 0) for g in grids])
    assert np.array_equal(gen.free_counts.cpu().numpy(), want)


def test_distribution_seed_and_independence():
    import torch
    from marl_coverage_b200 import MapTable
    a = MapTable.generate(GC, add_border=False, seed=3)
    b = MapTable.generate(GC, add_border=False, seed=3)
    c = MapTable.generate(GC, add_border=False, seed=4)
    assert torch.equal(a.wall, b.wall) and torch.equal(a.pos, b.pos)          # same seed, same maps
    assert not torch.equal(a.wall, c.wall)                                    # another seed, other maps
    more = MapTable.generate(dict(GC, numgrids=9), add_border=False, seed=3)
    assert torch.equal(more.wall[:6], a.wall)                                 # a map does not depend on numgrids
    bordered = MapTable.generate(GC, add_border=True, seed=3).to_grids()
    assert all(np.array_equal(x, y) for x, y in zip(bordered, a.to_grids()))  # nor on the border ring
    # obstacle fraction: 6 * 37 * 70 = 15 540 cells, sigma = sqrt(p (1 - p) / n) = 0.0035
    frac = np.mean([np.mean(g < 0) for g in a.to_grids()])
    assert abs(frac - 0.25) < 5 * 0.0035
    # rows and columns are not copies of one another
    g0 = a.to_grids()[0]
    assert not np.array_equal(g0[0], g0[1]) and not np.array_equal(g0[:, 0], g0[:, 1])
    none = MapTable.generate(dict(GC, prob_obst=0.0), add_border=False, seed=1).to_grids()
    every = MapTable.generate(dict(GC, prob_obst=1.0), add_border=False, seed=1).to_grids()
    assert all((g > 0).all() for g in none) and all((g < 0).all() for g in every)


def test_env_on_a_generated_table_matches_the_oracle_on_its_grids():
    import torch
    from marl_coverage_b200 import MapTable, VecDecGridRL
    from oracle import coverage_oracle as oc
    cfg = dict(numrobot=3, maxsteps=1000, collision_penalty=5, done_thresh=1, done_incr=0, terminal_reward=30,
               dist_reward=0, train_maxsteps=1000, test_maxsteps=1000, egoradius=4, mini_map_rad=0, comm_radius=0,
               allow_comm=0, map_sharing=0, single_square_tool=0, dijkstra_input=0, sensor_type="square_sensor",
               sensor_config={"range": 2})
    table = MapTable.generate(dict(numgrids=4, gridwidth=23, gridlen=41, prob_obst=0.15), add_border=True, seed=8)
    grids = table.to_grids()
    E, T, N = 8, 25, 3
    rng = np.random.RandomState(0)
    np.random.seed(0)
    oracles = [oc.OracleDecGridRL(grids, cfg) for _ in range(E)]
    start = np.zeros((E, N, 2), np.int16)
    maps = np.zeros((E,), np.int32)
    for e, o in enumerate(oracles):
        maps[e] = e % len(grids)
        o.reset(False, None, map_index=int(maps[e]))
        start[e, :, 0], start[e, :, 1] = o._xinds, o._yinds
        o.reset(False, None, start_xy=start[e], map_index=int(maps[e]))
    vec = VecDecGridRL(table, cfg, num_envs=E)
    vec.reset(map_idx=maps, start_xy=start)
    for t in range(T):
        act = rng.randint(0, 4, size=(E, N)).astype(np.uint8)
        obs, rew, done = vec.step(torch.from_numpy(act))
        for e, o in enumerate(oracles):
            oo, orr, od = o.step(act[e].astype(np.int64))
            assert np.array_equal(obs[e].cpu().numpy(), oo.astype(np.float32)), (t, e)
            assert float(vec.reward64[e].item()) == float(orr) and bool(done[e].item()) == bool(od)
    with pytest.raises(ValueError):
        from marl_coverage_b200 import VecSuperGridRL
        VecSuperGridRL(table, dict(numrobot=1, train_maxsteps=10, test_maxsteps=10, collision_penalty=5,
                                   senseradius=2, free_penalty=0, done_thresh=1, done_incr=0, terminal_reward=30,
                                   dist_reward=0, use_scanning=0), num_envs=2)      # bordered table, Super env


@pytest.mark.parametrize("border", [False, True])
@pytest.mark.parametrize("seed", [0, 3, (7 << 32) | 1])
def test_generated_maps_equal_the_numpy_model(border, seed):
    """Bit-exact pin of the kernel's draws: oracle/philox.py restates Philox4x32-10 (checked on CPU against the
    published known-answer vectors) and the generator's counter / threshold rule."""
    from marl_coverage_b200 import MapTable
    from oracle.philox import gen_maps
    cfg = dict(numgrids=5, gridwidth=33, gridlen=45, prob_obst=0.3)
    got = MapTable.generate(cfg, add_border=border, seed=seed).to_grids()
    want = gen_maps(5, 33, 45, 0.3, seed)
    assert all(np.array_equal(a, b) for a, b in zip(got, want))


@pytest.mark.parametrize("kind", ["dec", "super"])
def test_reset_draws_equal_the_numpy_model(kind):
    """The episode-start draws of the fused kernels (map index, distinct walkable start cells) are a pure function
    of (seed, GLOBAL env id, episode): oracle/philox.py:draw_starts restates them bit for bit."""
    from marl_coverage_b200 import VecDecGridRL, VecSuperGridRL, load_grid_set
    from oracle.philox import draw_starts
    grids = load_grid_set("bg2_100x100", 5)
    E, N, base = 48, 4, 1000
    if kind == "dec":
        cfg = dict(numrobot=N, maxsteps=1000, collision_penalty=5, done_thresh=1, done_incr=0, terminal_reward=30,
                   dist_reward=0, train_maxsteps=1000, test_maxsteps=1000, egoradius=5, mini_map_rad=0,
                   comm_radius=0, allow_comm=0, map_sharing=0, single_square_tool=0, dijkstra_input=0,
                   sensor_type="square_sensor", sensor_config={"range": 2})
        env = VecDecGridRL(grids, cfg, num_envs=E, env_id_base=base)
        walls = np.stack([np.pad(g < 0, 1, constant_values=True) for g in grids])      # env coordinates: with the ring
    else:
        cfg = dict(numrobot=N, train_maxsteps=1000, test_maxsteps=1000, collision_penalty=5, senseradius=2,
                   free_penalty=0, done_thresh=1, done_incr=0, terminal_reward=30, dist_reward=0, use_scanning=0)
        env = VecSuperGridRL(grids, cfg, num_envs=E, env_id_base=base)
        walls = np.stack([g < 0 for g in grids])
    for episode, seed in enumerate([12345, (9 << 32) | 6]):
        env.reset(seed=seed)
        got_map = env.head()["map_idx"].cpu().numpy()
        got_pos = env.positions().cpu().numpy()
        for e in range(E):
            mi, cells = draw_starts(walls, N, seed=seed, gid=base + e, episode=episode)
            assert int(got_map[e]) == mi, (episode, e)
            assert np.array_equal(got_pos[e].astype(np.int64), cells), (episode, e)
