"""GPU replay buffer + batched rollout collector.

``ReplayBuffer`` keeps the reference's surface (Policies/replaybuffer.py:3-157: float32 ring arrays
``_state/_action/_reward/_nextstate/_done``, ``_ptr/_size/_maxsize``, ``addtransition``,
``addepisode``, ``samplebatch``, ``samplesequential``, ``samplebatchsequential``) but the arrays are
CUDA tensors and the row traffic goes through the C ABI (``mcov_replay_push`` / ``mcov_replay_gather``):
a batched env step lands in the ring with ONE launch, straight from the observation buffers the step
kernel wrote — no device->host->device bounce as in Policies/dqn.py:121 / replaybuffer.py:33-48.

``collect`` is the batched form of Utils/utils.py:6-44 ``generate_episode``: policy -> step ->
max-step cut-off -> append (state, action, reward, next_state, done) -> restart finished envs; with
``state0=True`` it stores ``state[0]`` (robot 0's observation) exactly as utils.py:31 does.
"""
import ctypes as C

import numpy as np

from . import _lib


def _shape(length, shape=None):
    if shape is None:
        return (length,)
    return (length, shape) if np.isscalar(shape) else (length, *shape)


class ReplayBuffer(object):
    def __init__(self, obs_dim, act_dim, size, graph_dim=None, device="cuda:0"):
        import torch
        if not torch.cuda.is_available():
            raise RuntimeError("marl_coverage_b200.ReplayBuffer needs a CUDA device; no CPU fallback")
        if graph_dim is not None:
            raise NotImplementedError("graph rows (GNN policies) are outside the env-step hot path")
        self._lib = _lib.load()
        self.device = torch.device(device)
        self._maxsize = int(size)
        self._size = 0
        f32 = dict(dtype=torch.float32, device=self.device)
        self._state = torch.zeros(_shape(size, obs_dim), **f32)
        self._action = torch.zeros(_shape(size, act_dim), **f32)
        self._reward = torch.zeros(size, **f32)
        self._nextstate = torch.zeros(_shape(size, obs_dim), **f32)
        self._done = torch.zeros(size, **f32)
        self._store_graph = False
        self._start_lis = []
        self._len_lis = []
        self._ptr = 0
        self._row = int(np.prod(self._state.shape[1:])) if self._state.dim() > 1 else 1
        self._adim = int(np.prod(self._action.shape[1:])) if self._action.dim() > 1 else 1

    def _stream(self):
        import torch
        return torch.cuda.current_stream(self.device).cuda_stream

    # ---- batched insert: E transitions, one launch -------------------------------------------------
    def addbatch(self, state, action, reward, next_state, done, src_stride=None):
        """state / next_state: float32 CUDA tensors holding E rows of ``obs_dim`` (row e starts
        ``src_stride`` floats after row e-1; default = contiguous).  action: uint8 or float32 [E, act_dim]
        (or [E]); reward float32 [E]; done uint8 / bool [E]."""
        import torch
        E = int(reward.shape[0])
        if E > self._maxsize:
            raise ValueError("batch of %d rows exceeds the buffer capacity %d" % (E, self._maxsize))
        stride = self._row if src_stride is None else int(src_stride)
        need = ((E - 1) * stride + self._row) * 4                   # bytes the kernel reads from each source
        for name, t in (("state", state), ("next_state", next_state)):
            if not (torch.is_tensor(t) and t.is_cuda and t.dtype == torch.float32):
                raise ValueError("%s must be a float32 CUDA tensor" % name)
            if stride < self._row or t.untyped_storage().nbytes() - t.storage_offset() * 4 < need:
                raise ValueError("%s holds fewer than %d rows of %d floats at stride %d" % (name, E, self._row, stride))
        if not (reward.is_cuda and reward.dtype == torch.float32 and reward.is_contiguous() and done.is_cuda
                and done.numel() == E and action.is_cuda and action.numel() == E * self._adim):
            raise ValueError("reward float32 [E], done [E] and action [E, %d] must be CUDA tensors" % self._adim)
        act = action if action.dtype in (torch.uint8, torch.float32) else action.to(torch.float32)
        act = act.contiguous()
        dn = done if done.dtype == torch.uint8 else done.to(torch.uint8)
        with torch.cuda.device(self.device):
            _lib.check(self._lib.mcov_replay_push(
                state.data_ptr(), next_state.data_ptr(), self._row, stride, act.data_ptr(),
                1 if act.dtype == torch.uint8 else 0, self._adim, reward.data_ptr(), dn.data_ptr(), E,
                self._state.data_ptr(), self._nextstate.data_ptr(), self._action.data_ptr(),
                self._reward.data_ptr(), self._done.data_ptr(), self._ptr, self._maxsize, self._stream()))
        self._keep = (act, dn)
        self._ptr = (self._ptr + E) % self._maxsize
        self._size = min(self._size + E, self._maxsize)

    # ---- reference surface -------------------------------------------------------------------------
    def _dev(self, x, dtype):
        import torch
        t = x if torch.is_tensor(x) else torch.as_tensor(np.asarray(x))
        return t.to(device=self.device, dtype=dtype).reshape(1, -1).contiguous()

    def addtransition(self, state, action, reward, next_state, done, graph=None, next_graph=None):
        import torch
        self.addbatch(self._dev(state, torch.float32), self._dev(action, torch.float32),
                      self._dev(reward, torch.float32).reshape(1), self._dev(next_state, torch.float32),
                      self._dev(done, torch.uint8).reshape(1))

    def addepisode(self, episode):
        self._start_lis.append(self._ptr)
        self._len_lis.append(len(episode))
        for tr in episode:
            self.addtransition(tr[0], tr[1], tr[2], tr[3], tr[4])
        while sum(self._len_lis) > self._maxsize:
            self._len_lis.pop(0)
            self._start_lis.pop(0)

    def gather(self, indices):
        """Rows ``indices`` (int64, host or device) of all five arrays, one launch."""
        import torch
        idx = torch.as_tensor(np.asarray(indices) if not torch.is_tensor(indices) else indices)
        if not idx.is_cuda and idx.numel() and (int(idx.min()) < 0 or int(idx.max()) >= self._maxsize):
            raise IndexError("replay indices outside [0, %d)" % self._maxsize)   # host values: checked here for free
        idx = idx.to(device=self.device, dtype=torch.int64).reshape(-1).contiguous()   # device values: clamped by the kernel
        B = int(idx.numel())
        f32 = dict(dtype=torch.float32, device=self.device)
        st = torch.empty((B,) + tuple(self._state.shape[1:]), **f32)
        nx = torch.empty_like(st)
        ac = torch.empty((B,) + tuple(self._action.shape[1:]), **f32)
        rw = torch.empty((B,), **f32)
        dn = torch.empty((B,), **f32)
        with torch.cuda.device(self.device):
            _lib.check(self._lib.mcov_replay_gather(
                idx.data_ptr(), self._maxsize, B, self._row, self._adim, self._state.data_ptr(), self._nextstate.data_ptr(),
                self._action.data_ptr(), self._reward.data_ptr(), self._done.data_ptr(), st.data_ptr(),
                nx.data_ptr(), ac.data_ptr(), rw.data_ptr(), dn.data_ptr(), self._stream()))
        self._keep_idx = idx
        return st, ac, rw, nx, dn

    def samplebatch(self, N):
        return self.gather(np.random.randint(0, self._size, size=N))

    def samplesequential(self, N):
        assert N > 0, "you need to sample at least 1 step"
        ep_num = np.random.randint(0, len(self._start_lis))
        start_incr = np.random.randint(0, self._len_lis[ep_num] - N + 1)
        start_ind = self._start_lis[ep_num] + start_incr
        return self.gather(np.remainder(np.arange(start_ind, start_ind + N), self._maxsize))

    def samplebatchsequential(self, batchsize, N):
        assert N > 0, "you need to sample at least 1 step"
        rows = []
        for _ in range(batchsize):
            ep_num = np.random.randint(0, len(self._start_lis))
            start_incr = np.random.randint(0, self._len_lis[ep_num] - N + 1)
            start_ind = self._start_lis[ep_num] + start_incr
            rows.append(np.remainder(np.arange(start_ind, start_ind + N), self._maxsize))
        idx = np.stack(rows, axis=1).reshape(-1)               # time-major, as np.swapaxes(., 0, 1) yields
        out = self.gather(idx)
        return tuple(t.reshape((N, batchsize) + tuple(t.shape[1:])) for t in out)


def collect(env, policy, buffer, steps, state0=False, testing=False):
    """Batched generate_episode (Utils/utils.py:6-44) for ``steps`` env steps of every env.

    env     VecDecGridRL / VecSuperGridRL built with ``auto_reset=False`` (the terminal observation
            must reach the buffer as next_state before the env restarts, as in the reference loop)
    policy  callable: float32 obs CUDA tensor -> uint8 [E, N] (or int64 [E]) CUDA actions
    Returns (sum of rewards, number of finished episodes)."""
    import torch
    maxsteps = env.env_config["test_maxsteps" if testing else "train_maxsteps"]
    E = env.num_envs
    bufs = [torch.empty_like(env.obs), torch.empty_like(env.obs)]
    cur = 0
    env.observe(out=bufs[cur])
    row_stride = int(np.prod(env.obs_shape[1:]))
    total = torch.zeros((), dtype=torch.float64, device=env.device)
    episodes = torch.zeros((), dtype=torch.int64, device=env.device)
    for _ in range(steps):
        state = bufs[cur]
        action = policy(state)
        nxt, reward, done = env.step(action, out=bufs[cur ^ 1])
        cut = env.head()["currstep"] == maxsteps                # utils.py:25-28
        done_all = (done.bool() | cut).to(torch.uint8)
        act = action if action.dim() == 2 else action.reshape(E, 1).to(torch.float32)
        buffer.addbatch(state, act, reward, nxt, done_all, src_stride=row_stride if state0 else None)
        total += env.reward64.sum()
        episodes += done_all.sum()
        env.reset(mask=done_all, out=bufs[cur ^ 1])             # masked: no host sync on the loop
        cur ^= 1
    return float(total.item()), int(episodes.item())
