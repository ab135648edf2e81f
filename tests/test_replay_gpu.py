"""GPU replay-buffer rows (mcov_replay_push / mcov_replay_gather through the C ABI) against the
reference-recorded golden ring and the oracle; the batched collector against a per-env restatement of
Utils/utils.py:generate_episode driven by the oracle env."""
import os

import numpy as np
import pytest

from conftest import GOLDEN_DIR
from oracle import coverage_oracle as oc

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("chunk", [1, 7, 23])
def test_ring_matches_reference_golden(chunk):
    import torch
    from marl_coverage_b200 import ReplayBuffer
    g = np.load(os.path.join(GOLDEN_DIR, "replay_ring.npz"))
    n, size = g["st"].shape[0], int(g["size"])
    buf = ReplayBuffer((3, 5, 5), 2, size)
    dev = lambda a, dt=torch.float32: torch.from_numpy(np.ascontiguousarray(a)).to("cuda", dt)
    k = 0
    while k < n:
        hi = min(k + chunk, n)
        if chunk == 1:
            buf.addtransition(g["st"][k], g["ac"][k], g["rw"][k], g["nx"][k], g["dn"][k])
        else:
            buf.addbatch(dev(g["st"][k:hi]), dev(g["ac"][k:hi], torch.uint8), dev(g["rw"][k:hi]),
                         dev(g["nx"][k:hi]), dev(g["dn"][k:hi], torch.uint8))
        k = hi
    assert buf._ptr == int(g["ptr"]) and buf._size == int(g["n_size"])
    for t, name in ((buf._state, "ring_state"), (buf._nextstate, "ring_next"), (buf._action, "ring_action"),
                    (buf._reward, "ring_reward"), (buf._done, "ring_done")):
        assert np.array_equal(t.cpu().numpy(), g[name]), name
    st, ac, rw, nx, dn = buf.gather(g["idx"])
    assert np.array_equal(st.cpu().numpy(), g["ring_state"][g["idx"]])
    assert np.array_equal(nx.cpu().numpy(), g["ring_next"][g["idx"]])
    assert np.array_equal(ac.cpu().numpy(), g["ring_action"][g["idx"]])
    assert np.array_equal(rw.cpu().numpy(), g["ring_reward"][g["idx"]])
    assert np.array_equal(dn.cpu().numpy(), g["ring_done"][g["idx"]])
    np.random.seed(5)
    out = buf.samplebatch(9)
    assert np.array_equal(out[0].cpu().numpy(), g["seed5_state"])
    assert np.array_equal(out[2].cpu().numpy(), g["seed5_reward"])


def test_vectorised_rows_and_errors():
    import torch
    from marl_coverage_b200 import ReplayBuffer, _lib
    rng = np.random.default_rng(3)
    buf = ReplayBuffer((4, 3, 11, 11), 4, 50)                  # 1452 floats per row: float4 path
    ref = oc.OracleReplayBuffer((4, 3, 11, 11), 4, 50)
    for _ in range(9):
        E = 13
        st = rng.random((E, 4, 3, 11, 11)).astype(np.float32)
        nx = rng.random((E, 4, 3, 11, 11)).astype(np.float32)
        ac = rng.integers(0, 4, (E, 4)).astype(np.uint8)
        rw = rng.integers(-20, 30, E).astype(np.float32)
        dn = (rng.random(E) < 0.2).astype(np.uint8)
        buf.addbatch(*(torch.from_numpy(a).cuda() for a in (st, ac, rw, nx, dn)))
        for e in range(E):
            ref.addtransition(st[e], ac[e], rw[e], nx[e], dn[e])
    assert np.array_equal(buf._state.cpu().numpy(), ref._state)
    assert np.array_equal(buf._nextstate.cpu().numpy(), ref._nextstate)
    assert np.array_equal(buf._action.cpu().numpy(), ref._action)
    assert np.array_equal(buf._done.cpu().numpy(), ref._done) and buf._ptr == ref._ptr
    with pytest.raises(ValueError):
        z = torch.zeros(60, device="cuda")
        buf.addbatch(torch.zeros(60, 4, 3, 11, 11, device="cuda"), torch.zeros(60, 4, device="cuda"), z,
                     torch.zeros(60, 4, 3, 11, 11, device="cuda"), z.to(torch.uint8))
    lib = _lib.load()
    assert lib.mcov_replay_push(None, None, 4, 4, None, 1, 1, None, None, 1, None, None, None, None, None, 0, 4,
                                None) == -2


@pytest.mark.parametrize("shape", [(3, 11, 11), (17,), (3, 7, 7), (5, 5, 5)])
def test_unaligned_rows_every_phase(shape):
    """Rows whose length is not a multiple of 4 floats (the reference's 363-float state[0] rows) go through
    the shifted 128-bit copy: every source / destination phase, strided sources, ring wrap, gather."""
    import torch
    from marl_coverage_b200 import ReplayBuffer
    row = int(np.prod(shape))
    g = torch.Generator(device="cuda")
    g.manual_seed(row)
    cap = 61
    buf = ReplayBuffer(shape, 1, cap)
    want_s = torch.zeros((cap, row), device="cuda")
    want_n = torch.zeros((cap, row), device="cuda")
    ptr = 0
    for it, E in enumerate((7, 13, 29, 31, 5)):
        stride = row + it                                        # 0..4 floats of padding: every source phase
        st = torch.rand((E, stride), device="cuda", generator=g)
        nx = torch.rand((E, stride), device="cuda", generator=g)
        ac = torch.randint(0, 4, (E, 1), dtype=torch.uint8, device="cuda", generator=g)
        rw = torch.rand((E,), device="cuda", generator=g)
        dn = torch.zeros((E,), dtype=torch.uint8, device="cuda")
        buf.addbatch(st, ac, rw, nx, dn, src_stride=stride)
        rows = (ptr + torch.arange(E, device="cuda")) % cap
        want_s[rows] = st[:, :row]
        want_n[rows] = nx[:, :row]
        ptr = (ptr + E) % cap
    assert buf._ptr == ptr
    assert torch.equal(buf._state.reshape(cap, row), want_s)
    assert torch.equal(buf._nextstate.reshape(cap, row), want_n)
    idx = np.array([0, 60, 17, 17, 33, 1, 59, 2, 44], np.int64)
    out = buf.gather(idx)
    assert torch.equal(out[0].reshape(len(idx), row), want_s[torch.from_numpy(idx).cuda()])
    assert torch.equal(out[3].reshape(len(idx), row), want_n[torch.from_numpy(idx).cuda()])


CFG = dict(numrobot=3, maxsteps=1000, collision_penalty=5, done_thresh=0.6, done_incr=0.0, terminal_reward=30,
           dist_reward=0, train_maxsteps=25, test_maxsteps=25, egoradius=3, mini_map_rad=0, comm_radius=0,
           allow_comm=0, map_sharing=0, single_square_tool=0, dijkstra_input=0, sensor_type="square_sensor",
           sensor_config={"range": 2})


def test_collector_matches_generate_episode_semantics():
    """collect() == E copies of the reference rollout loop (utils.py:17-42: step, max-step cut-off,
    append (state[0], action, reward, next_state[0], done), restart) run on the oracle env."""
    import torch
    from marl_coverage_b200 import ReplayBuffer, VecDecGridRL, collect, gridload
    train, _ = gridload(None)
    E, N, T = 6, CFG["numrobot"], 70
    env = VecDecGridRL(train, CFG, num_envs=E, auto_reset=False, seed=11)
    env.reset()
    np.random.seed(0)
    oracles = [oc.OracleDecGridRL(train, CFG) for _ in range(E)]

    def sync_oracle(e):
        pos = env.positions()[e].cpu().numpy()
        mi = int(env.head()["map_idx"][e].item())
        return oracles[e].reset(False, None, start_xy=pos, map_index=mi)[0]

    states = [sync_oracle(e) for e in range(E)]
    g = torch.Generator(device="cuda")
    g.manual_seed(3)
    acts = torch.randint(0, 4, (T, E, N), dtype=torch.uint8, device="cuda", generator=g)
    buf = ReplayBuffer((3, 7, 7), N, 1000)
    ref = oc.OracleReplayBuffer((3, 7, 7), N, 1000)
    step = {"t": 0}

    def policy(obs):
        a = acts[step["t"]]
        step["t"] += 1
        return a

    total = 0.0
    episodes = 0
    for t in range(T):
        r, ep = collect(env, policy, buf, 1, state0=True)
        total += r
        episodes += ep
        a = acts[t].cpu().numpy()
        for e, o in enumerate(oracles):
            nxt, rew, done = o.step(a[e].astype(np.int64))
            if o._currstep == o._train_maxsteps:                # utils.py:27-28
                done = True
            ref.addtransition(states[e][0], a[e], rew, nxt[0], done)      # state[0]: robot 0 only (utils.py:31)
            states[e] = sync_oracle(e) if done else nxt
    assert episodes > 0
    n = E * T
    assert buf._ptr == ref._ptr == n
    assert np.array_equal(buf._state[:n].cpu().numpy(), ref._state[:n])
    assert np.array_equal(buf._nextstate[:n].cpu().numpy(), ref._nextstate[:n])
    assert np.array_equal(buf._action[:n].cpu().numpy(), ref._action[:n])
    assert np.array_equal(buf._reward[:n].cpu().numpy(), ref._reward[:n])
    assert np.array_equal(buf._done[:n].cpu().numpy(), ref._done[:n])
