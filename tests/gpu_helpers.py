"""Shared helpers for the -m gpu parity tests: replay golden traces / oracle runs on the CUDA env."""
import numpy as np

from oracle.trace import OP_RESET, crc_f32, compare_snap


def vec_snapshot(vec, kind, e, reward, done, obs):
    """Builds the snapshot dict compare_snap() expects from env ``e`` of a batched CUDA env."""
    head = vec.head()
    s = dict(pos=vec.positions()[e].cpu().numpy(),
             reward=float(reward), done=int(done),
             pct=float(vec.percent_covered()[e].item()),
             done_thresh=float(head["done_thresh"][e].item()),
             currstep=int(head["currstep"][e].item()))
    if kind == "dec":
        free, obst, vis = vec.planes(envs=[e])
        s["obs"] = obs[e].cpu().numpy()
        s["adj"] = vec.adjacency()[e].astype(np.uint8)
        s["ckpt_a"] = free[0].astype(np.uint8)
        s["ckpt_b"] = obst[0].astype(np.uint8)
        s["ckpt_visited"] = vis[0].astype(np.uint8)
    else:
        cov, seen = vec.planes(envs=[e])
        st = obs[e].cpu().numpy()
        s["state_crc"] = crc_f32(st)
        s["ckpt_a"] = cov[0].astype(np.uint8)
        s["ckpt_b"] = seen[0].astype(np.uint8)
        s["ckpt_state"] = st
    return s


def replay_trace_on_gpu(trace, vec_cls, num_envs=1, check_envs=(0,), obs_tol=0.0, reward_rtol=0.0,
                        device="cuda:0"):
    """Every env of the batch replays the same golden session; envs in ``check_envs`` are compared
    op by op (positions, reward f64, done, coverage, thresholds, obs bits, planes at checkpoints)."""
    import torch
    vec = vec_cls(trace.float_maps(), trace.config, num_envs=num_envs, device=device)
    E, N = num_envs, trace.config["numrobot"]
    for k in range(trace.num_ops):
        if trace.op_kind[k] == OP_RESET:
            mi = np.full((E,), int(trace.op_map[k]), dtype=np.int32)
            sx = np.broadcast_to(trace.op_start[k][None], (E, N, 2)).astype(np.int16)
            obs = vec.reset(map_idx=mi, start_xy=sx)
            rew = np.zeros(E)
            done = np.zeros(E, dtype=np.uint8)
        else:
            act = torch.full((E,), int(trace.op_action[k]), dtype=torch.int64, device=device)
            obs, _, d = vec.step(act)
            rew = vec.reward64.cpu().numpy()
            done = d.cpu().numpy()
        for e in check_envs:
            snap = vec_snapshot(vec, trace.kind, e, rew[e], done[e], obs)
            compare_snap(trace, k, snap, obs_tol=obs_tol, reward_rtol=reward_rtol, where="gpu env %d" % e)
    return vec
