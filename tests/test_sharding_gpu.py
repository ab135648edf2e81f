"""Multi-GPU semantics checked on one device: G logical shards of a batch (env_id_base = shard
offset) must reproduce the unsharded batch env for env — device-RNG resets included — and the
per-shard statistics must add up to the single-batch statistics (what the NCCL all-reduce sums)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

CFG = dict(numrobot=4, maxsteps=40, collision_penalty=5, done_thresh=0.5, done_incr=0.0, terminal_reward=30,
           dist_reward=0, train_maxsteps=40, test_maxsteps=40, egoradius=5, mini_map_rad=0, comm_radius=0,
           allow_comm=0, map_sharing=0, single_square_tool=0, dijkstra_input=0, sensor_type="square_sensor",
           sensor_config={"range": 2})


@pytest.mark.parametrize("shards", [2, 4])
def test_logical_shards_equal_single_batch(shards):
    import torch
    from marl_coverage_b200 import VecDecGridRL, load_grid_set
    from marl_coverage_b200.sharding import shard_range
    grids = load_grid_set("bg2_100x100", 6)
    E, N, T = 203, 4, 90
    g = torch.Generator(device="cuda")
    g.manual_seed(7)
    acts = torch.randint(0, 4, (T, E, N), dtype=torch.uint8, device="cuda", generator=g)
    whole = VecDecGridRL(grids, CFG, num_envs=E, auto_reset=True, truncate_steps=40, env_id_base=0, seed=5)
    parts = []
    for r in range(shards):
        lo, hi = shard_range(E, r, shards)
        parts.append((lo, hi, VecDecGridRL(grids, CFG, num_envs=hi - lo, auto_reset=True, truncate_steps=40,
                                           env_id_base=lo, seed=5)))
    o = whole.reset().clone()
    for lo, hi, p in parts:
        assert torch.equal(p.reset(), o[lo:hi])          # device Philox keyed by the global env id
    tot_whole = np.zeros(8)
    tot_parts = np.zeros(8)
    for t in range(T):
        o, r, d = whole.step(acts[t], want_stats=True)
        tot_whole += whole.stats.cpu().numpy()
        for lo, hi, p in parts:
            po, pr, pd = p.step(acts[t, lo:hi].contiguous(), want_stats=True)
            assert torch.equal(po, o[lo:hi]) and torch.equal(pd, d[lo:hi])
            assert torch.equal(p.reward64, whole.reward64[lo:hi])
            tot_parts += p.stats.cpu().numpy()
    assert tot_whole[0] > 0                               # episodes did finish and auto-reset
    idx = [0, 1, 2, 4, 5, 6]                              # integer-valued sums: exact in any order
    assert np.array_equal(tot_whole[idx], tot_parts[idx])
    assert np.allclose(tot_whole[3], tot_parts[3], rtol=1e-12)   # coverage fractions: order of summation differs


@pytest.mark.parametrize("group,variant", [(8, "MCOV_DEC_STORE=1"), (32, "MCOV_DEC_STORE=1"), (8, "MCOV_DEC_TEAM=16"),
                                           (8, "MCOV_DEC_TEAM=32"), (16, "MCOV_DEC_TEAM=32"), (8, "MCOV_DEC_SHAPE=1"),
                                           (8, "MCOV_DEC_SMT=1"), (16, "MCOV_DEC_SMT=1"), (32, "MCOV_DEC_SMT=1")])
def test_kernel_variants_equal_plain_path_with_auto_reset(group, variant, monkeypatch):
    """The variants of the DecGridRL step kernel — store warp (bit-stream slots -> TMA bulk stores), store teams
    (16 / 32 lanes write a block together), CTA-per-env shape, truth-map table staged into shared memory by TMA (one
    1024-thread CTA per SM) — against the plain lane-group path on the same seeded
    batch with fused auto-resets: a finished env re-runs its pass in RESET mode before its block is written.  Three
    CTAs loop over the batch (slots and teams are reused many times; the batch size leaves some teams incomplete)."""
    import torch
    from marl_coverage_b200 import VecDecGridRL, load_grid_set
    grids = load_grid_set("bg2_100x100", 6)
    E, N, T = 301, 4, 90
    g = torch.Generator(device="cuda")
    g.manual_seed(8)
    acts = torch.randint(0, 4, (T, E, N), dtype=torch.uint8, device="cuda", generator=g)
    monkeypatch.setenv("MCOV_DEC_GROUP", str(group))
    monkeypatch.setenv("MCOV_DEC_GRID", "3")
    monkeypatch.setenv("MCOV_DEC_SHAPE", "0")
    envs = []
    for on in (False, True):                                        # pins are read once, at mcov_create
        if on:
            k, v = variant.split("=")
            monkeypatch.setenv(k, v)
        envs.append(VecDecGridRL(grids, CFG, num_envs=E, auto_reset=True, truncate_steps=40, seed=5))
    a, b = envs
    assert torch.equal(a.reset(), b.reset())
    episodes = 0.0
    for t in range(T):
        oa, ra, da = a.step(acts[t], want_stats=True)
        ob, rb, db = b.step(acts[t], want_stats=True)
        assert torch.equal(oa, ob) and torch.equal(da, db) and torch.equal(a.reward64, b.reward64), "step %d" % t
        sa, sb = a.stats.cpu().numpy(), b.stats.cpu().numpy()
        assert np.array_equal(sa[[0, 2, 4, 5, 6]], sb[[0, 2, 4, 5, 6]]) and np.allclose(sa, sb, rtol=1e-12)   # float sums: group order differs
        episodes += float(a.stats[0].item())
    assert episodes > 0
    fa, fb = a.planes(), b.planes()
    assert all(np.array_equal(x, y) for x, y in zip(fa, fb))
