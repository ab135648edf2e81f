"""SuperGrid_Sim: the stateless one-robot, one-step lookahead DRQN / HA* use to filter or roll out
candidate actions (Environments/sim_env.py:16-259), batched on the GPU.

``VecSuperGridSim`` evaluates B (state, action) pairs per launch through ``mcov_sim_step``;
``SuperGrid_Sim`` keeps the reference's constructor and ``step(state, action, grid)`` signature on top of
it (an B = 1 view that marshals numpy <-> device).  No CPU implementation exists here: without the CUDA
library and a device the constructors raise.

The reference function cannot run at its HEAD (``np.asscalar`` was removed from numpy; ``percent_covered``
is called without its ``grid`` argument, sim_env.py:146-147,187,248); both are read the way the
surrounding code intends, and ``tests/golden/pin_sim_render.py`` pins that reading against the reference with
exactly those two shims.
"""
import ctypes as C

import numpy as np

from . import _lib
from .vec_env import VecSuperGridRL, _ptr


class VecSuperGridSim:
    """Batched lookahead over a fixed map table.

    grids       list of (W, L) arrays (<0 obstacle, >0 free) — the ``grid`` arguments the reference passes
    env_config  SuperGridRL keys (sim_env.py:34-45)"""

    def __init__(self, grids, env_config, device="cuda:0"):
        cfg = dict(env_config)
        cfg["numrobot"] = 1                      # the lookahead moves one robot (sim_env.py:78,146)
        self._env = VecSuperGridRL(grids, cfg, num_envs=1, device=device)
        self._lib = self._env._lib
        self.device = self._env.device
        self.W, self.L = self._env.W, self._env.L
        self.done_thresh = float(env_config["done_thresh"])
        self.done_incr = float(env_config["done_incr"])

    @property
    def launch_count(self):
        return self._env.launch_count

    def step(self, states, actions, map_idx=None, pos_xy=None, out=None, inplace=False):
        """states float32 CUDA [B, 4, W, L]; actions uint8 [B]; map_idx int32 [B] (None: map 0);
        pos_xy int16 [B, 2] (None: read from the position plane).
        Returns (next_states [B,4,W,L] float32, reward float64 [B], terminal uint8 [B]) on the device.
        ``inplace=True`` patches ``states`` instead of writing a copy (the reference's side effect)."""
        import torch
        env = self._env
        st = states
        if not torch.is_tensor(st):
            st = torch.as_tensor(np.asarray(st, dtype=np.float32))
        st = st.to(device=self.device, dtype=torch.float32)
        if st.dim() != 4 or tuple(st.shape[1:]) != (4, self.W, self.L):
            raise ValueError("states must be [B, 4, %d, %d], got %s" % (self.W, self.L, tuple(st.shape)))
        if not st.is_contiguous():
            if inplace:
                raise ValueError("inplace lookahead needs a contiguous state tensor")
            st = st.contiguous()
        B = st.shape[0]
        act = env._as_dev(actions, torch.uint8, (B,))
        mi = env._as_dev(map_idx, torch.int32, (B,))
        pxy = env._as_dev(pos_xy, torch.int16, (B, 2))
        if out is not None and not inplace and not (
                torch.is_tensor(out) and out.is_cuda and out.dtype == torch.float32 and out.is_contiguous()
                and out.numel() == st.numel()):
            raise ValueError("out must be a contiguous float32 CUDA tensor %s" % (tuple(st.shape),))
        nxt = st if inplace else (out if out is not None else torch.empty_like(st))
        r64 = torch.empty((B,), dtype=torch.float64, device=self.device)
        term = torch.empty((B,), dtype=torch.uint8, device=self.device)
        with torch.cuda.device(self.device):
            _lib.check(self._lib.mcov_sim_step(env._handle, st.data_ptr(), act.data_ptr(), _ptr(mi), _ptr(pxy),
                                               int(B), C.c_double(self.done_thresh), nxt.data_ptr(), None,
                                               r64.data_ptr(), term.data_ptr(), env._stream()))
        self._keep = (st, act, mi, pxy)
        return nxt, r64, term


class SuperGrid_Sim(object):
    """Drop-in for Environments/sim_env.py:SuperGrid_Sim (same ctor / step / isTerminal / percent_covered)."""

    def __init__(self, obs_dim, env_config):
        c = env_config
        self._numrobot = c["numrobot"]
        self._train_maxsteps = c["train_maxsteps"]
        self._test_maxsteps = c["test_maxsteps"]
        self._collision_penalty = c["collision_penalty"]
        self._senseradius = c["senseradius"]
        self._free_penalty = c["free_penalty"]
        self._done_thresh = c["done_thresh"]
        self._done_incr = c["done_incr"]
        self._terminal_reward = c["terminal_reward"]
        self._dist_r = c["dist_reward"]
        self._use_scanning = c["use_scanning"]
        self._obs_dim = obs_dim
        self._num_actions = 4 ** self._numrobot
        self.prev_a = None
        self._cfg = dict(c)
        self._sims = {}                      # one packed map table per distinct grid the caller passes
        from .vec_env import _torch
        _torch()                             # no CUDA device -> raise now, not at the first step

    def _sim_for(self, grid):
        g = np.ascontiguousarray(grid)
        key = (g.shape, g.tobytes())
        sim = self._sims.get(key)
        if sim is None:
            sim = self._sims[key] = VecSuperGridSim([np.array(g)], self._cfg)
        return sim

    def step(self, state, action, grid):
        planes = state[0]                                             # sim_env.py:75
        sim = self._sim_for(grid)
        sim.done_thresh = float(self._done_thresh)
        st = np.asarray(planes, dtype=np.float32)[None]
        nxt, r64, _ = sim.step(st, np.array([int(action)], dtype=np.uint8))
        out = nxt[0].double().cpu().numpy()
        # the reference updates the caller's observed-obstacle and free planes in place (:157,169)
        planes[1][...] = out[1]
        planes[2][...] = out[2]
        return out, float(r64[0].item())

    def isInBounds(self, x, y, width, length):
        return x >= 0 and x < width and y >= 0 and y < length

    def isOccupied(self, x, y, grid):
        return bool(grid[x, y] < 0)

    def percent_covered(self, state, grid):
        pos_img, observed_obstacles, free, dist = state
        return np.count_nonzero(free < 1) / np.count_nonzero(grid > 0)

    def isTerminal(self, state, grid=None):
        if grid is None:
            raise TypeError("isTerminal needs the grid (sim_env.py:243 calls percent_covered without it)")
        if min(self._done_thresh, 1) <= self.percent_covered(state, grid):
            self._done_thresh += self._done_incr
            return True
        return False
