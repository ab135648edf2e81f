"""The C ABI promises that every call only enqueues work on the caller's stream (include/mcov.h): a rollout
of several steps must therefore be capturable in a CUDA graph and replay to the same bits as eager stepping."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

DEC = dict(numrobot=4, maxsteps=30, collision_penalty=5, done_thresh=0.6, done_incr=0, terminal_reward=30,
           dist_reward=0, train_maxsteps=30, test_maxsteps=30, egoradius=5, mini_map_rad=0, comm_radius=0,
           allow_comm=0, map_sharing=0, single_square_tool=0, dijkstra_input=0, sensor_type="square_sensor",
           sensor_config={"range": 2})
SUPER = dict(numrobot=3, train_maxsteps=30, test_maxsteps=30, collision_penalty=5, senseradius=2, free_penalty=0.2,
             done_thresh=0.5, done_incr=0, terminal_reward=30, dist_reward=0, use_scanning=0)


@pytest.mark.parametrize("kind", ["dec", "dec_lidar", "super"])
def test_rollout_in_a_cuda_graph_equals_eager(kind):
    import torch
    from marl_coverage_b200 import VecDecGridRL, VecSuperGridRL, load_grid_set
    grids = load_grid_set("bg2_100x100", 5)
    cfg = dict(SUPER) if kind == "super" else dict(DEC)
    if kind == "dec_lidar":
        cfg.update(sensor_type="lidar", sensor_config={"range": 6, "num_lasers": 15})
    cls = VecSuperGridRL if kind == "super" else VecDecGridRL
    E, N, T = 301, cfg["numrobot"], 12
    gen = torch.Generator(device="cuda")
    gen.manual_seed(3)
    acts = torch.randint(0, 4, (3 * T, E, N), dtype=torch.uint8, device="cuda", generator=gen)

    def make():
        env = cls(grids, cfg, num_envs=E, auto_reset=True, truncate_steps=30, seed=9)
        env.reset(seed=1234)
        return env

    eager, graphed = make(), make()
    obs_log = torch.empty((T,) + tuple(graphed.obs_shape), device="cuda")
    rew_log = torch.empty((T, E), dtype=torch.float64, device="cuda")
    done_log = torch.empty((T, E), dtype=torch.uint8, device="cuda")
    stats_log = torch.empty((T, 8), dtype=torch.float64, device="cuda")
    cursor = torch.zeros((), dtype=torch.int64, device="cuda")      # which block of T action rows to play
    act_buf = torch.empty((T, E, N), dtype=torch.uint8, device="cuda")

    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        graphed.step(acts[0], want_stats=True)                       # warm-up launch outside the capture
        graphed.reset(seed=1234)                                      # kernel parameters (the auto-reset seed
        g = torch.cuda.CUDAGraph()                                    # among them) are frozen into the graph
        with torch.cuda.graph(g, stream=side):
            for t in range(T):
                o, r, d = graphed.step(act_buf[t], want_stats=True)
                obs_log[t].copy_(o)
                rew_log[t].copy_(graphed.reward64)
                done_log[t].copy_(d)
                stats_log[t].copy_(graphed.stats)
    torch.cuda.current_stream().wait_stream(side)
    graphed.reset(seed=1234)
    # the device RNG is keyed by (seed of the reset call, env id, episode counter): give `eager` the same history
    eager.step(acts[0], want_stats=True)
    eager.reset(seed=1234)
    eager.reset(seed=1234)
    assert torch.equal(graphed.obs, eager.obs)
    episodes = 0.0
    for block in range(3):
        act_buf.copy_(acts[block * T:(block + 1) * T])
        g.replay()
        torch.cuda.synchronize()
        for t in range(T):
            o, r, d = eager.step(acts[block * T + t], want_stats=True)
            assert torch.equal(obs_log[t], o), (kind, block, t)
            assert torch.equal(rew_log[t], eager.reward64)
            assert torch.equal(done_log[t], d)
            assert torch.equal(stats_log[t], eager.stats)
            episodes += float(eager.stats[0].item())
    assert episodes > 0, "the rollout must cross auto-resets inside the graph"


def test_step_into_caller_result_buffers():
    """step(results=...) writes reward / reward64 / done / stats into the caller's second buffer set (what the
    double-buffered host loop of bench.py uses) and leaves the env's own tensors alone."""
    import torch
    from marl_coverage_b200 import VecDecGridRL, load_grid_set
    grids = load_grid_set("bg2_100x100", 3)
    a = VecDecGridRL(grids, DEC, num_envs=50, auto_reset=True, truncate_steps=30, seed=2)
    b = VecDecGridRL(grids, DEC, num_envs=50, auto_reset=True, truncate_steps=30, seed=2)
    a.reset(seed=5)
    b.reset(seed=5)
    res = [b.make_result_buffers() for _ in range(2)]
    gen = torch.Generator(device="cuda")
    gen.manual_seed(1)
    for t in range(40):
        act = torch.randint(0, 4, (50, 4), dtype=torch.uint8, device="cuda", generator=gen)
        oa, ra, da = a.step(act, want_stats=True)
        keep = (b.reward.clone(), b.done.clone(), b.stats.clone())
        ob, rb, db = b.step(act, want_stats=True, results=res[t & 1])
        assert rb is res[t & 1]["reward"] and db is res[t & 1]["done"]
        assert torch.equal(oa, ob) and torch.equal(ra, rb) and torch.equal(da, db)
        assert torch.equal(a.reward64, res[t & 1]["reward64"]) and torch.equal(a.stats, res[t & 1]["stats"])
        assert torch.equal(b.reward, keep[0]) and torch.equal(b.done, keep[1]) and torch.equal(b.stats, keep[2])


@pytest.mark.parametrize("kind", ["dec", "super"])
def test_get_set_state_resumes_a_rollout(kind):
    """Checkpoint / resume: a rollout continued from get_state() in ANOTHER env object repeats bit for bit,
    fused auto-resets (device RNG) included."""
    import torch
    from marl_coverage_b200 import VecDecGridRL, VecSuperGridRL, load_grid_set
    grids = load_grid_set("bg2_100x100", 4)
    cfg = dict(SUPER) if kind == "super" else dict(DEC)
    cls = VecSuperGridRL if kind == "super" else VecDecGridRL
    E, N = 97, cfg["numrobot"]
    gen = torch.Generator(device="cuda")
    gen.manual_seed(11)
    acts = torch.randint(0, 4, (70, E, N), dtype=torch.uint8, device="cuda", generator=gen)
    a = cls(grids, cfg, num_envs=E, auto_reset=True, truncate_steps=30, seed=3)
    a.reset()
    for t in range(20):
        a.step(acts[t], want_stats=True)
    snap = a.get_state()
    want = []
    for t in range(20, 70):
        o, r, d = a.step(acts[t], want_stats=True)
        want.append((o.clone(), a.reward64.clone(), d.clone(), a.stats.clone()))
    assert sum(float(w[3][0]) for w in want) > 0                      # episodes ended and were re-seeded
    b = cls(grids, cfg, num_envs=E, auto_reset=True, truncate_steps=30, seed=999)
    b.reset()
    b.step(acts[5])
    b.set_state(snap)
    assert torch.equal(b.obs, snap["obs"]) and torch.equal(b.positions(), a.__class__.positions(b))
    for k, t in enumerate(range(20, 70)):
        o, r, d = b.step(acts[t], want_stats=True)
        assert torch.equal(o, want[k][0]) and torch.equal(b.reward64, want[k][1]), (kind, t)
        assert torch.equal(d, want[k][2]) and torch.equal(b.stats, want[k][3]), (kind, t)


def test_caller_buffers_are_validated():
    """Buffers that reach the kernels as raw pointers are checked on the host: a wrong size, dtype, layout or
    device raises instead of overrunning device memory."""
    import torch
    from marl_coverage_b200 import VecDecGridRL, load_grid_set
    grids = load_grid_set("bg2_100x100", 2)
    env = VecDecGridRL(grids, DEC, num_envs=6, seed=1)
    env.reset(seed=3)
    act = torch.zeros((6, 4), dtype=torch.uint8, device="cuda")
    good = torch.empty_like(env.obs)
    assert env.step(act, out=good)[0] is good                                       # accepted
    bad_outs = [torch.empty(env.obs.numel() - 1, device="cuda"),                    # too small
                torch.empty_like(env.obs, dtype=torch.float64),                     # wrong dtype
                torch.empty(tuple(env.obs_shape) + (2,), device="cuda")[..., 0],    # not contiguous
                torch.empty(tuple(env.obs_shape))]                                  # host memory
    for bad in bad_outs:
        with pytest.raises(ValueError):
            env.step(act, out=bad)
        with pytest.raises(ValueError):
            env.observe(out=bad)
        with pytest.raises(ValueError):
            env.reset(out=bad)
    with pytest.raises(ValueError):
        env.step(torch.zeros((5,), dtype=torch.int64, device="cuda"))               # base-4 actions: wrong E
    res = env.make_result_buffers()
    res["done"] = torch.zeros((5,), dtype=torch.uint8, device="cuda")
    with pytest.raises(ValueError):
        env.step(act, results=res)
    with pytest.raises(ValueError):
        env.render(out=torch.empty((3,), dtype=torch.uint8, device="cuda"))
    twin = VecDecGridRL(grids, DEC, num_envs=6, seed=1)                             # the refusals touched nothing
    twin.reset(seed=3)
    twin.step(act)
    assert torch.equal(twin.step(act)[0], env.step(act)[0])


@pytest.mark.parametrize("kind,keep_obs", [("dec", False), ("dec", True), ("super", False)])
def test_rollout_graph_helper_equals_eager(kind, keep_obs):
    """marl_coverage_b200.RolloutGraph: K steps captured once, replayed with new actions each time, against the same
    steps taken one by one — rewards, done flags, statistics, observations, state; fused auto-resets included."""
    import torch
    from marl_coverage_b200 import RolloutGraph, VecDecGridRL, VecSuperGridRL, load_grid_set
    grids = load_grid_set("bg2_100x100", 5)
    cfg = dict(SUPER) if kind == "super" else dict(DEC)
    cls = VecSuperGridRL if kind == "super" else VecDecGridRL
    E, N, K, rounds = 97, cfg["numrobot"], 8, 5
    gen = torch.Generator(device="cuda")
    gen.manual_seed(11)
    acts = torch.randint(0, 4, (rounds, K, E, N), dtype=torch.uint8, device="cuda", generator=gen)
    envs = []
    for _ in range(2):
        env = cls(grids, cfg, num_envs=E, auto_reset=True, truncate_steps=30, seed=9)
        env.reset(seed=77)
        envs.append(env)
    eager, graphed = envs
    roll = RolloutGraph(graphed, K, keep_obs=keep_obs)
    ns = graphed.layout.off_scratch                                 # state proper; behind it: library scratch
    assert torch.equal(graphed.obs, eager.obs) and torch.equal(graphed._arena[:ns], eager._arena[:ns])   # capture left no trace
    episodes = 0.0
    for r in range(rounds):
        rew, done = roll.replay(acts[r])
        for k in range(K):
            o, _, d = eager.step(acts[r, k], want_stats=True)
            assert torch.equal(roll.reward64[k], eager.reward64) and torch.equal(done[k], d), "round %d step %d" % (r, k)
            assert torch.equal(roll.stats[k], eager.stats)
            if keep_obs:
                assert torch.equal(roll.obs_log[k], o)
            episodes += float(eager.stats[0].item())
        assert torch.equal(graphed.obs, eager.obs)
        assert torch.equal(graphed._arena[:graphed.layout.off_scratch], eager._arena[:eager.layout.off_scratch])
    assert episodes > 0
