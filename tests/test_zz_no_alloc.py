"""mcov_step_static_scratch (include/mcov.h): the DecGridRL step whose launch scratch lives in a file-scope
``static __device__`` array instead of the state arena — no counterpart in the reference, added
for callers that cannot let a library near the allocator on the step path.

Called through the C ABI (ctypes -> libmcov_b200.so) on the handle of a batched env and compared with the existing entry
point, ``mcov_step``, on a twin env: observations, float64 rewards, done flags and positions bit for bit, statistics
exactly where they are integer sums and to 1e-12 where they are float sums (the tolerance the other step-variant tests
use), on several calls in a row — the static array is shared by every call and keeps the last call's contents.

This file sorts last on purpose: a failure here must not hide another test."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

CFG = dict(numrobot=4, maxsteps=40, collision_penalty=5, done_thresh=0.5, done_incr=0.0, terminal_reward=30,
           dist_reward=0, train_maxsteps=40, test_maxsteps=40, egoradius=5, mini_map_rad=0, comm_radius=0,
           allow_comm=0, map_sharing=0, single_square_tool=0, dijkstra_input=0, sensor_type="square_sensor",
           sensor_config={"range": 2})
INT_STATS = [0, 2, 4, 5, 6]                       # episodes, lengths, collisions, new cells, agent-steps: integer sums


def _static_step(env, act, out):
    """One call of the new entry point through the C ABI; every buffer is the caller's."""
    import torch
    from marl_coverage_b200 import _lib
    assert act.dtype == torch.uint8 and act.is_contiguous() and act.is_cuda
    with torch.cuda.device(env.device):
        _lib.check(env._lib.mcov_step_static_scratch(
            env._handle, act.data_ptr(), _lib.ACT_U8_PER_ROBOT, out["obs"].data_ptr(), out["reward"].data_ptr(),
            out["reward64"].data_ptr(), out["done"].data_ptr(), out["stats"].data_ptr(),
            torch.cuda.current_stream(env.device).cuda_stream))


def _buffers(env):
    import torch
    E, dev = env.num_envs, env.device
    return {"obs": torch.empty(tuple(env.obs_shape), dtype=torch.float32, device=dev),
            "reward": torch.zeros((E,), dtype=torch.float32, device=dev),
            "reward64": torch.zeros((E,), dtype=torch.float64, device=dev),
            "done": torch.zeros((E,), dtype=torch.uint8, device=dev),
            "stats": torch.zeros((8,), dtype=torch.float64, device=dev)}


@pytest.mark.parametrize("shape,E,lidar", [(0, 301, False), (1, 37, False), (0, 150, True)])
def test_static_scratch_step_equals_mcov_step(shape, E, lidar, monkeypatch):
    import torch
    from marl_coverage_b200 import VecDecGridRL, load_grid_set
    monkeypatch.setenv("MCOV_DEC_SHAPE", str(shape))          # lane-group kernel | CTA-per-env kernel (read at mcov_create)
    grids = load_grid_set("bg2_100x100", 6)
    cfg = dict(CFG)
    if lidar:
        cfg.update(sensor_type="lidar", sensor_config={"range": 6, "num_lasers": 15})
    N, T = cfg["numrobot"], 60
    gen = torch.Generator(device="cuda")
    gen.manual_seed(21)
    acts = torch.randint(0, 4, (T, E, N), dtype=torch.uint8, device="cuda", generator=gen)
    ref = VecDecGridRL(grids, cfg, num_envs=E, auto_reset=True, truncate_steps=40, seed=5)
    new = VecDecGridRL(grids, cfg, num_envs=E, auto_reset=True, truncate_steps=40, seed=5)
    assert torch.equal(ref.reset(seed=99), new.reset(seed=99))
    out = _buffers(new)
    _static_step(new, acts[0], out)                            # first call on this device resolves the symbol address
    ref.step(acts[0], want_stats=True)
    torch.cuda.synchronize()
    alloc0, free0 = torch.cuda.memory_allocated(), torch.cuda.mem_get_info()[0]
    episodes = 0.0
    for t in range(1, T):                                      # many calls in a row on the shared array
        o, _, d = ref.step(acts[t], want_stats=True)
        _static_step(new, acts[t], out)
        assert torch.equal(out["obs"], o), "obs step %d" % t
        assert torch.equal(out["reward64"], ref.reward64) and torch.equal(out["reward"], ref.reward), "reward step %d" % t
        assert torch.equal(out["done"], d), "done step %d" % t
        sa, sb = ref.stats.cpu().numpy(), out["stats"].cpu().numpy()
        assert np.array_equal(sa[INT_STATS], sb[INT_STATS]) and np.allclose(sa, sb, rtol=1e-12, atol=0), "stats step %d" % t
        episodes += float(sa[0])
    assert episodes > 0                                        # fused auto-resets happened on both paths
    assert torch.equal(ref.positions(), new.positions())
    ns = ref.layout.off_scratch                                # state proper; behind it: the arena scratch mcov_step uses
    assert torch.equal(ref._arena[:ns], new._arena[:ns])
    torch.cuda.synchronize()
    # nothing was allocated by the calls: neither through torch's allocator nor on the device behind its back
    assert torch.cuda.memory_allocated() == alloc0
    assert torch.cuda.mem_get_info()[0] >= free0


def test_static_scratch_refusals(monkeypatch):
    """Sizes beyond the stated maximum and unsupported handles are refused with a message, before any launch."""
    import torch
    from marl_coverage_b200 import VecDecGridRL, VecSuperGridRL, _lib, load_grid_set
    grids = load_grid_set("bg2_100x100", 2)
    # CTA-per-env shape with 2400 envs: one CTA per env up to 16 per SM -> more than MCOV_STATIC_SCRATCH_MAX_ROWS CTAs
    monkeypatch.setenv("MCOV_DEC_SHAPE", "1")
    big = VecDecGridRL(grids, CFG, num_envs=2400, auto_reset=True, truncate_steps=40, seed=1)
    big.reset()
    act = torch.zeros((2400, 4), dtype=torch.uint8, device="cuda")
    with pytest.raises(_lib.McovError, match="MCOV_STATIC_SCRATCH_MAX_ROWS"):
        _static_step(big, act, _buffers(big))
    obs, _, _ = big.step(act)                                  # the handle is untouched: mcov_step still works
    assert torch.isfinite(obs).all()
    monkeypatch.delenv("MCOV_DEC_SHAPE")
    gated = VecDecGridRL(grids, dict(CFG, dijkstra_input=1), num_envs=4, seed=1)
    gated.reset()
    with pytest.raises(_lib.McovError, match="use mcov_step"):
        _static_step(gated, torch.zeros((4, 4), dtype=torch.uint8, device="cuda"), _buffers(gated))
    sup = VecSuperGridRL(grids, dict(numrobot=2, train_maxsteps=40, test_maxsteps=40, collision_penalty=5, senseradius=2,
                                     free_penalty=0.2, done_thresh=1, done_incr=0, terminal_reward=30, dist_reward=0,
                                     use_scanning=0), num_envs=3, seed=1)
    sup.reset()
    with pytest.raises(_lib.McovError, match="DecGridRL handles only"):
        _lib.check(sup._lib.mcov_step_static_scratch(
            sup._handle, torch.zeros((3, 2), dtype=torch.uint8, device="cuda").data_ptr(), _lib.ACT_U8_PER_ROBOT,
            sup.obs.data_ptr(), sup.reward.data_ptr(), sup.reward64.data_ptr(), sup.done.data_ptr(), None,
            torch.cuda.current_stream().cuda_stream))
    assert _lib.STATIC_SCRATCH_MAX_ROWS == 2048                 # the binding states the same limit as include/mcov.h
