"""K env steps as one CUDA graph: the launch cost of a rollout is paid once per K steps.

The C ABI only enqueues work on the caller's stream (include/mcov.h), so a sequence of ``step`` calls captures as
is; this helper owns the buffers such a capture needs and keeps the env's own state consistent around it.  It is the
batched counterpart of the inner loop of the reference's ``generate_episode`` (Utils/utils.py:17-42) for callers
whose actions for the next K steps are already on the device (random exploration, a replayed action log, a policy
that is itself part of the captured stream).

    roll = RolloutGraph(env, steps=16)
    roll.actions.copy_(next_16_actions)        # uint8 [K, E, N], device
    roll.replay()                              # K fused step launches, no host work in between
    roll.reward64, roll.done, roll.stats       # [K, E], [K, E], [K, 8]; env.obs holds the last observation
"""
import numpy as np

from . import _lib


class RolloutGraph:
    """Captures ``steps`` consecutive ``env.step`` calls of a VecDecGridRL / VecSuperGridRL batch.

    ``keep_obs=True`` also logs every step's observation block ([K, *obs_shape] float32 — large: 5.8 kB per env
    and step at configs[1]); otherwise only the last one stays in ``env.obs``.  Everything a kernel launch freezes
    into the graph (buffer addresses, the seed the fused auto-resets derive their draws from) is the env's at
    capture time: call ``env.reset(seed=...)`` before building the graph, not between replays."""

    def __init__(self, env, steps, keep_obs=False, want_stats=True):
        import torch
        if steps < 1:
            raise ValueError("steps must be >= 1")
        self.env, self.steps = env, int(steps)
        K, E, N, dev = self.steps, env.num_envs, env.num_robots, env.device
        self.actions = torch.zeros((K, E, N), dtype=torch.uint8, device=dev)
        self.reward = torch.zeros((K, E), dtype=torch.float32, device=dev)
        self.reward64 = torch.zeros((K, E), dtype=torch.float64, device=dev)
        self.done = torch.zeros((K, E), dtype=torch.uint8, device=dev)
        self.stats = torch.zeros((K, _lib.STATS_LEN), dtype=torch.float64, device=dev)
        self.obs_log = torch.empty((K,) + tuple(env.obs_shape), dtype=torch.float32, device=dev) if keep_obs else None
        self._results = [{"reward": self.reward[k], "reward64": self.reward64[k], "done": self.done[k],
                          "stats": self.stats[k]} for k in range(K)]
        self._want_stats = bool(want_stats)
        self._stream = torch.cuda.Stream(dev)
        self._graph = torch.cuda.CUDAGraph()
        # a warm-up step outside the capture (first launches configure kernel attributes), then the state is put back
        saved = env.get_state()
        self._stream.wait_stream(torch.cuda.current_stream(dev))
        with torch.cuda.stream(self._stream):
            self._one(0)
            env.set_state(saved)
            with torch.cuda.graph(self._graph, stream=self._stream):
                for k in range(K):
                    self._one(k)
        torch.cuda.current_stream(dev).wait_stream(self._stream)

    def _one(self, k):
        env = self.env
        env.step(self.actions[k], out=None if self.obs_log is None else self.obs_log[k],
                 want_stats=self._want_stats, results=self._results[k])
        if self.obs_log is not None and k == self.steps - 1:
            env.obs.copy_(self.obs_log[k])       # env.obs keeps meaning "the latest observation"

    def replay(self, actions=None):
        """Runs the K captured steps on the current stream.  ``actions``: uint8 [K, E, N] (device or host), copied
        into ``self.actions`` first; None = whatever ``self.actions`` holds.  Returns (reward [K, E] float32,
        done [K, E] uint8) — views of this object's buffers, overwritten by the next replay."""
        import torch
        if actions is not None:
            a = actions if torch.is_tensor(actions) else torch.as_tensor(np.asarray(actions))
            if tuple(a.shape) != tuple(self.actions.shape):
                raise ValueError("actions must be [K, E, N] = %s" % (tuple(self.actions.shape),))
            self.actions.copy_(a.to(torch.uint8), non_blocking=True)
        self._graph.replay()
        return self.reward, self.done
