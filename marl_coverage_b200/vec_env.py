"""Batched B200 environments: E independent DecGridRL / SuperGridRL instances stepped by one fused
CUDA kernel through the C ABI of include/mcov.h.

Host code is plumbing only (torch owns device memory and streams); there is no Python or torch
implementation of the dynamics here and no fallback: without libmcov_b200.so + a CUDA device the
constructors raise.

Semantics follow the reference classes (Environments/dec_grid_rl.py, super_grid_rl.py) env by env;
the batched API adds what a vectorised rollout needs: per-env ``done`` tensors, optional in-place
auto-reset with a device RNG, truncation at ``train_maxsteps`` (Utils/utils.py:25-28) and per-step
episode statistics.
"""
import ctypes as C

import numpy as np

from . import _lib
from .maps import MapTable, unpack_plane_host
from .sensors import lidar_beam_table


def _torch():
    import torch
    if not torch.cuda.is_available():
        raise RuntimeError("marl_coverage_b200 needs a CUDA device (B200, sm_100a); no CPU fallback")
    return torch


class _VecEnv:
    ENV_TYPE = None

    def __init__(self, grids, env_config, num_envs, device="cuda:0", auto_reset=False,
                 truncate_steps=0, env_id_base=0, seed=0):
        torch = _torch()
        self._lib = _lib.load()
        self.device = torch.device(device)
        self.num_envs = int(num_envs)
        self.env_config = dict(env_config)
        self.num_robots = int(env_config["numrobot"])
        self.seed = int(seed)
        self._reset_calls = 0
        self._handle = C.c_void_p()
        cfg = self._make_config(env_config, auto_reset, truncate_steps, env_id_base)
        self._cfg = cfg
        with torch.cuda.device(self.device):
            _lib.check(self._lib.mcov_create(C.byref(cfg), C.byref(self._handle)))
            self._post_create(env_config)
            if isinstance(grids, MapTable):                       # a table built on the device (MapTable.generate)
                if grids.add_border != (self.ENV_TYPE == _lib.ENV_DEC) or grids.device != self.device:
                    raise ValueError("MapTable was built with add_border=%s on %s; this env needs add_border=%s on %s"
                                     % (grids.add_border, grids.device, self.ENV_TYPE == _lib.ENV_DEC, self.device))
                self.maps = grids
            else:
                self.maps = MapTable(grids, add_border=(self.ENV_TYPE == _lib.ENV_DEC), device=self.device)
            _lib.check(self._lib.mcov_set_maps(self._handle, self.maps.wall.data_ptr(),
                                               self.maps.pos.data_ptr(), self.maps.free_counts.data_ptr(),
                                               self.maps.n_maps, self.maps.W, self.maps.L))
            if self.ENV_TYPE == _lib.ENV_SUPER and self.maps.values is not None:
                # free cells worth something other than 1 (super_grid_rl.py:189); DecGridRL's reward is a count
                _lib.check(self._lib.mcov_set_cell_values(self._handle, self.maps.values.data_ptr()))
            lay = _lib.McovLayout()
            _lib.check(self._lib.mcov_state_layout(self._handle, C.byref(lay)))
            self.layout = lay
            self._arena = torch.empty((int(lay.total_bytes),), dtype=torch.uint8, device=self.device)
            _lib.check(self._lib.mcov_bind_state(self._handle, self._arena.data_ptr(),
                                                 self._arena.numel(), self._stream()))
            dims = (C.c_int32 * 5)()
            nd = C.c_int32()
            _lib.check(self._lib.mcov_obs_shape(self._handle, C.byref(dims), C.byref(nd)))
            self.obs_shape = tuple(int(dims[k]) for k in range(nd.value))
            E = self.num_envs
            self.obs = torch.zeros(self.obs_shape, dtype=torch.float32, device=self.device)
            self.reward = torch.zeros((E,), dtype=torch.float32, device=self.device)
            self.reward64 = torch.zeros((E,), dtype=torch.float64, device=self.device)
            self.done = torch.zeros((E,), dtype=torch.uint8, device=self.device)
            self.stats = torch.zeros((_lib.STATS_LEN,), dtype=torch.float64, device=self.device)
        self.W, self.L = int(lay.W), int(lay.L)

    # -- subclasses fill the struct from the reference's env_config keys -----------------------
    def _make_config(self, c, auto_reset, truncate_steps, env_id_base):
        raise NotImplementedError

    def _post_create(self, env_config):
        pass

    def _base_config(self, c, auto_reset, truncate_steps, env_id_base):
        cfg = _lib.McovConfig()
        cfg.struct_size = C.sizeof(_lib.McovConfig)
        cfg.env_type = self.ENV_TYPE
        cfg.num_envs = self.num_envs
        cfg.num_robots = self.num_robots
        cfg.env_id_base = int(env_id_base)
        cfg.truncate_steps = int(truncate_steps)
        cfg.auto_reset = int(bool(auto_reset))
        cfg.collision_penalty = float(c["collision_penalty"])
        cfg.terminal_reward = float(c["terminal_reward"])
        cfg.done_thresh = float(c["done_thresh"])
        cfg.done_incr = float(c["done_incr"])
        cfg.dist_reward = int(bool(c["dist_reward"]))
        return cfg

    def _stream(self):
        import torch
        return torch.cuda.current_stream(self.device).cuda_stream

    def __del__(self):
        try:
            if getattr(self, "_handle", None) is not None and self._handle.value:
                self._lib.mcov_destroy(self._handle)
                self._handle = C.c_void_p()
        except Exception:
            pass

    def _check_obs_out(self, t):
        """A caller-supplied observation buffer goes to the kernel as a raw pointer: refuse anything the kernel
        would overrun or misread instead of corrupting device memory."""
        import torch
        n = int(np.prod(self.obs_shape))
        if not (torch.is_tensor(t) and t.is_cuda and t.dtype == torch.float32 and t.is_contiguous()
                and t.numel() == n and (self.device.index is None or t.device.index == self.device.index)):
            raise ValueError("out must be a contiguous float32 CUDA tensor of %d elements %s on %s"
                             % (n, self.obs_shape, self.device))
        return t

    def _check_results(self, res):
        import torch
        seen = self.__dict__.setdefault("_results_ok", [])       # validated once per buffer set, not per step
        if any(res is r for r in seen):
            return res
        E = self.num_envs
        want = (("reward", torch.float32, E), ("reward64", torch.float64, E), ("done", torch.uint8, E),
                ("stats", torch.float64, _lib.STATS_LEN))
        for key, dt, n in want:
            t = res[key]
            if not (t.is_cuda and t.dtype == dt and t.is_contiguous() and t.numel() == n):
                raise ValueError("results[%r] must be a contiguous %s CUDA tensor of %d elements" % (key, dt, n))
        del seen[3:]
        seen.insert(0, res)
        return res

    # -- the hot path -----------------------------------------------------------------------------
    def reset(self, map_idx=None, start_xy=None, mask=None, seed=None, out=None):
        """Starts new episodes (all envs, or those where ``mask`` is non-zero).

        map_idx  int32 [E] or None (device RNG: uniform over the map table)
        start_xy int16 [E, N, 2] env coordinates or None (device RNG: uniform distinct walkable cells)
        Returns the float32 CUDA observation tensor (written in place into ``out`` or ``self.obs``)."""
        import torch
        obs = self.obs if out is None else self._check_obs_out(out)
        mi = self._as_dev(map_idx, torch.int32, (self.num_envs,), lo=0, hi=self.maps.n_maps)
        sx = self._as_dev(start_xy, torch.int16, (self.num_envs, self.num_robots, 2), lo=0,
                          hi=(self.W, self.L))
        mk = self._as_dev(mask, torch.uint8, (self.num_envs,))
        if seed is None:
            seed = (self.seed + 0x9E3779B97F4A7C15 * self._reset_calls) & 0xFFFFFFFFFFFFFFFF
        self._reset_calls += 1
        self._last_seed = int(seed)
        with torch.cuda.device(self.device):
            _lib.check(self._lib.mcov_reset(self._handle, _ptr(mk), _ptr(mi), _ptr(sx),
                                            C.c_uint64(int(seed)), obs.data_ptr(), self._stream()))
        self._keep = (mi, sx, mk)           # keep inputs alive until the stream has consumed them
        return obs

    def get_state(self):
        """Checkpoint of the whole batch: the state arena (planes, headers, poses, thresholds, episode
        counters), the last observation and the seed the fused auto-resets derive their draws from.
        ``set_state`` of it resumes the rollout bit for bit (tests/test_graph_gpu.py)."""
        return {"arena": self._arena.clone(), "obs": self.obs.clone(), "seed": getattr(self, "_last_seed", None),
                "reset_calls": self._reset_calls}

    def set_state(self, state):
        import torch
        if state["arena"].numel() != self._arena.numel() or tuple(state["obs"].shape) != tuple(self.obs.shape):
            raise ValueError("state belongs to an env batch of a different shape")
        if state["seed"] is not None:
            # a reset with an all-zero mask touches no env but re-arms the library's auto-reset seed
            self.reset(mask=torch.zeros((self.num_envs,), dtype=torch.uint8, device=self.device),
                       seed=state["seed"])
        self._arena.copy_(state["arena"])
        self.obs.copy_(state["obs"])
        self._reset_calls = int(state["reset_calls"])

    def make_result_buffers(self):
        """A second set of per-step outputs for ``step(results=...)``: lets a caller double-buffer the
        device->host read of step t against the kernel of step t+1."""
        import torch
        E = self.num_envs
        return {"reward": torch.zeros((E,), dtype=torch.float32, device=self.device),
                "reward64": torch.zeros((E,), dtype=torch.float64, device=self.device),
                "done": torch.zeros((E,), dtype=torch.uint8, device=self.device),
                "stats": torch.zeros((_lib.STATS_LEN,), dtype=torch.float64, device=self.device)}

    def step(self, actions, out=None, want_stats=False, results=None):
        """One transition of every env.

        actions: uint8 [E, N] per-robot moves (0..3) or int64 [E] base-4 joint actions (the
        reference's integer form; -1 = sentinel).  CUDA tensors are used as they are; host
        tensors/arrays are copied to the device first.
        ``results`` (from make_result_buffers) receives reward / reward64 / done / stats instead of the
        env's own tensors.
        Returns (obs float32 CUDA, reward float32 [E] CUDA, done uint8 [E] CUDA)."""
        import torch
        obs = self.obs if out is None else self._check_obs_out(out)
        if results is not None:
            self._check_results(results)
        rew = self.reward if results is None else results["reward"]
        rew64 = self.reward64 if results is None else results["reward64"]
        dn = self.done if results is None else results["done"]
        stats = self.stats if results is None else results["stats"]
        act = actions
        if not torch.is_tensor(act):
            act = torch.as_tensor(np.asarray(act))
        if act.dtype == torch.int64 and act.dim() == 1:
            fmt = _lib.ACT_I64_BASE4
            if act.shape[0] != self.num_envs:
                raise ValueError("base-4 joint actions must be int64 [E] = [%d]" % self.num_envs)
        elif act.dim() == 2:
            fmt = _lib.ACT_U8_PER_ROBOT
            if act.dtype != torch.uint8:
                act = act.to(torch.uint8)
            if tuple(act.shape) != (self.num_envs, self.num_robots):
                raise ValueError("actions must be [E, N] = %s" % ((self.num_envs, self.num_robots),))
        else:
            raise ValueError("actions must be int64 [E] (base-4 joint) or uint8 [E, N]")
        if act.device != self.device:
            act = act.to(self.device, non_blocking=True)
        act = act.contiguous()
        with torch.cuda.device(self.device):
            _lib.check(self._lib.mcov_step(self._handle, act.data_ptr(), fmt, obs.data_ptr(),
                                           rew.data_ptr(), rew64.data_ptr(), dn.data_ptr(),
                                           stats.data_ptr() if want_stats else None,
                                           self._stream()))
        self._keep_act = act
        return obs, rew, dn

    def observe(self, out=None):
        import torch
        obs = self.obs if out is None else self._check_obs_out(out)
        with torch.cuda.device(self.device):
            _lib.check(self._lib.mcov_observe(self._handle, obs.data_ptr(), self._stream()))
        return obs

    def percent_covered(self):
        import torch
        out = torch.empty((self.num_envs,), dtype=torch.float64, device=self.device)
        with torch.cuda.device(self.device):
            _lib.check(self._lib.mcov_percent_covered(self._handle, out.data_ptr(), self._stream()))
        return out

    def render(self, env_ids=None, out=None):
        """Headless render() frames (super_grid_rl.py:423-463, dec_grid_rl.py:554-616): uint8 CUDA
        tensor [n, H, Wd, 3] for the listed envs (None: all), composed from the bit planes on the GPU."""
        import torch
        h, w = C.c_int32(), C.c_int32()
        _lib.check(self._lib.mcov_render_shape(self._handle, C.byref(h), C.byref(w)))
        ids = None if env_ids is None else self._as_dev(env_ids, torch.int32, (len(env_ids),))
        n = self.num_envs if ids is None else int(ids.numel())
        if out is None:
            out = torch.empty((n, h.value, w.value, 3), dtype=torch.uint8, device=self.device)
        elif not (torch.is_tensor(out) and out.is_cuda and out.dtype == torch.uint8 and out.is_contiguous()
                  and out.numel() == n * h.value * w.value * 3):
            raise ValueError("out must be a contiguous uint8 CUDA tensor [%d, %d, %d, 3]" % (n, h.value, w.value))
        with torch.cuda.device(self.device):
            _lib.check(self._lib.mcov_render(self._handle, _ptr(ids), n, out.data_ptr(), self._stream()))
        self._keep_ids = ids
        return out

    @property
    def launch_count(self):
        return int(self._lib.mcov_launch_count(self._handle))

    # -- state access (get_state / set_state through the arena layout) ----------------------------
    def _as_dev(self, x, dtype, shape, lo=None, hi=None):
        """Host values are range-checked here, where it is free ([lo, hi), hi per last-axis column when a tuple);
        values that already live on the device cannot be looked at without a sync and are clamped by the kernels."""
        import torch
        if x is None:
            return None
        t = x if torch.is_tensor(x) else torch.as_tensor(np.asarray(x))
        if lo is not None and not t.is_cuda and t.numel():
            his = hi if isinstance(hi, tuple) else (hi,)
            for k, h in enumerate(his):
                col = t[..., k] if isinstance(hi, tuple) else t
                if int(col.min()) < lo or int(col.max()) >= h:
                    raise ValueError("value outside [%d, %d) in %s input" % (lo, h, tuple(shape)))
        t = t.to(device=self.device, dtype=dtype).contiguous()
        if tuple(t.shape) != tuple(shape):
            raise ValueError("expected shape %s, got %s" % (tuple(shape), tuple(t.shape)))
        return t

    def _hdr(self):
        lay = self.layout
        n = self.num_envs * lay.hdr_stride
        return self._arena[lay.off_hdr:lay.off_hdr + n].view(self.num_envs, lay.hdr_stride)

    def _hdr_field(self, off, nbytes, dtype):
        return self._hdr()[:, off:off + nbytes].contiguous().view(dtype)

    def _set_hdr_field(self, off, value):
        import torch
        v = value.contiguous().view(torch.uint8).reshape(self.num_envs, -1)
        self._hdr()[:, off:off + v.shape[1]] = v

    def positions(self):
        """int16 [E, N, 2] (x, y) in env coordinates."""
        import torch
        lay = self.layout
        return self._hdr_field(lay.hdr_off_pos, 4 * self.num_robots, torch.int16).view(
            self.num_envs, self.num_robots, 2)

    def head(self):
        """dict of per-env scalars: done_thresh f64, ep_return f64, currstep, map_idx, flags, episode."""
        import torch
        return {
            "done_thresh": self._hdr_field(0, 8, torch.float64).view(-1),
            "ep_return": self._hdr_field(8, 8, torch.float64).view(-1),
            "currstep": self._hdr_field(16, 4, torch.int32).view(-1),
            "map_idx": self._hdr_field(20, 4, torch.int32).view(-1),
            "flags": self._hdr_field(24, 4, torch.int32).view(-1),
            "episode": self._hdr_field(28, 4, torch.int32).view(-1),
        }

    def set_done_thresh(self, value):
        import torch
        t = torch.as_tensor(value, dtype=torch.float64, device=self.device).expand(self.num_envs)
        self._set_hdr_field(0, t.contiguous().view(-1, 1))

    def needs_reset(self):
        lay = self.layout
        return self._arena[lay.off_needs_reset:lay.off_needs_reset + self.num_envs]

    def _plane_words(self, off, per_env):
        import torch
        n = self.num_envs * per_env * 4
        return self._arena[off:off + n].view(torch.int32).view(self.num_envs, per_env)

    def _unpack(self, words_i32):
        w = words_i32.cpu().numpy().view(np.uint32)
        return unpack_plane_host(w, self.W, self.L)


def _ptr(t):
    return None if t is None else t.data_ptr()


class VecDecGridRL(_VecEnv):
    """E x DecGridRL (Environments/dec_grid_rl.py:21-552).

    obs: float32 [E, N, C, 2e+1, 2e+1] — layer 0 robots, 1 own known-free, 2 own known-obstacle
    (+ gated layers).  ``env_config`` takes the reference's keys verbatim."""
    ENV_TYPE = _lib.ENV_DEC

    def _make_config(self, c, auto_reset, truncate_steps, env_id_base):
        cfg = self._base_config(c, auto_reset, truncate_steps, env_id_base)
        cfg.maxsteps = int(c["maxsteps"])
        st = c["sensor_type"]
        if st == "lidar":
            cfg.sensor_type = _lib.SENSOR_LIDAR
            cfg.num_lasers = int(c["sensor_config"]["num_lasers"])
        elif st == "square_sensor":
            cfg.sensor_type = _lib.SENSOR_SQUARE
        else:
            raise ValueError("sensor_type must be 'lidar' or 'square_sensor' (dec_grid_rl.py:71-75)")
        cfg.sensor_range = int(c["sensor_config"]["range"])
        cfg.egoradius = int(c["egoradius"])
        cfg.mini_map_rad = int(c["mini_map_rad"])
        cfg.comm_radius = int(c["comm_radius"])
        cfg.allow_comm = int(bool(c["allow_comm"]))
        cfg.map_sharing = int(bool(c["map_sharing"]))
        cfg.single_square_tool = int(bool(c["single_square_tool"]))
        cfg.dijkstra_input = int(bool(c["dijkstra_input"]))
        return cfg

    def _post_create(self, env_config):
        if env_config["sensor_type"] == "lidar":
            beams = np.ascontiguousarray(lidar_beam_table(int(env_config["sensor_config"]["num_lasers"])))
            _lib.check(self._lib.mcov_set_lidar_beams(
                self._handle, beams.ctypes.data_as(C.POINTER(C.c_double)), beams.shape[0]))

    def adjacency(self):
        """float64 [E, N, N] 0/1 adjacency (dec_grid_rl.py:374-391) as numpy."""
        import torch
        lay, N = self.layout, self.num_robots
        rows = self._hdr_field(lay.hdr_off_adj, 4 * N, torch.int32).view(self.num_envs, N).cpu().numpy()
        rows = rows.view(np.uint32)
        return ((rows[:, :, None] >> np.arange(N, dtype=np.uint32)) & 1).astype(np.float64)

    def coverage_counts(self):
        import torch
        lay, N = self.layout, self.num_robots
        return self._hdr_field(lay.hdr_off_cov, 4 * N, torch.int32).view(self.num_envs, N)

    def planes(self, envs=None):
        """(free [E,N,W,L], obst [E,N,W,L], visited [E,W,L]) boolean numpy arrays; ``envs`` (a list of env
        indices) restricts the unpacking to those envs, in that order."""
        lay, N, PW = self.layout, self.num_robots, self.layout.plane_words
        fo = self._plane_words(lay.off_fo, N * PW * 2).view(self.num_envs, N, PW, 2)
        vis = self._plane_words(lay.off_visited, PW)
        if envs is not None:
            fo, vis = fo[list(envs)], vis[list(envs)]
        free = self._unpack(fo[..., 0].contiguous())
        obst = self._unpack(fo[..., 1].contiguous())
        return free, obst, self._unpack(vis)


class VecSuperGridRL(_VecEnv):
    """E x SuperGridRL (Environments/super_grid_rl.py:18-421).

    obs: float32 [E, N+3, W, L] (or [E, 4, W, L] with use_scanning): one-hot robot planes,
    observed obstacles, not-yet-covered, inverted L1 distance channel."""
    ENV_TYPE = _lib.ENV_SUPER

    def _make_config(self, c, auto_reset, truncate_steps, env_id_base):
        cfg = self._base_config(c, auto_reset, truncate_steps, env_id_base)
        cfg.free_penalty = float(c["free_penalty"])
        cfg.senseradius = int(c["senseradius"])
        cfg.use_scanning = int(bool(c["use_scanning"]))
        return cfg

    def planes(self, envs=None):
        """(covered [E,W,L], observed_obstacles [E,W,L]) boolean numpy arrays; ``envs`` restricts the unpacking."""
        lay, PW = self.layout, self.layout.plane_words
        cov, seen = self._plane_words(lay.off_covered, PW), self._plane_words(lay.off_seen_obst, PW)
        if envs is not None:
            cov, seen = cov[list(envs)], seen[list(envs)]
        return self._unpack(cov), self._unpack(seen)
