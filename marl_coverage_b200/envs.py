"""Reference-compatible single-environment classes: an E = 1 view of the batched CUDA env.

``DecGridRL`` / ``SuperGridRL`` keep the constructor signatures, methods, attributes and return
types of Environments/dec_grid_rl.py:21-616 and Environments/super_grid_rl.py:18-463 so that
grid_rl_main.py, Utils/utils.py:generate_episode and the Policies/* consumers run unchanged
(SURVEY.md §8b).  Every transition is computed by the CUDA kernels; this file only marshals:
it draws the map / start poses from the global ``np.random`` stream in the reference's order
(so a seeded run walks the same episodes as the reference), converts the action form, and copies
the float32 device tensors back as float64 numpy arrays.
"""
import numpy as np

from .sensors import LidarSensor, SquareSensor
from .vec_env import VecDecGridRL, VecSuperGridRL


def _scalar_action(action):
    """Normalises the reference's action forms.  Returns ('sentinel' | 'joint' | 'per_robot', value)."""
    if action is None:
        return "sentinel", -1
    try:
        import torch
        if torch.is_tensor(action):
            action = action.item() if action.dim() == 0 or action.numel() == 1 else action.cpu().numpy()
    except ImportError:
        pass
    if isinstance(action, (list, tuple)):
        action = np.asarray(action)
    if isinstance(action, np.ndarray):
        if action.ndim == 0:
            action = action.item()
        else:
            return "per_robot", action.astype(np.int64).ravel()
    a = int(action)
    if a == -1:
        return "sentinel", -1
    return "joint", a


class _CompatBase(object):
    def _build(self, vec_cls, train_set, test_set, env_config):
        grids = list(train_set)
        self._n_train = len(grids)
        if test_set is not None and test_set is not train_set:
            grids = grids + list(test_set)
            self._test_base = self._n_train
        else:
            self._test_base = 0
        self._vec = vec_cls(grids, env_config, num_envs=1)

    def _map_choice(self, testing, ind):
        if testing and self._test_gridlis is not None:
            return self._test_base + int(ind), self._test_gridlis[ind]
        k = np.random.randint(len(self._train_gridlis))
        return k, self._train_gridlis[k]

    def _sync_head(self):
        h = self._vec.head()
        self._currstep = int(h["currstep"][0])
        self._done_thresh_dev = float(h["done_thresh"][0])
        pos = self._vec.positions()[0].cpu().numpy()
        self._xinds = pos[:, 0].astype(int)
        self._yinds = pos[:, 1].astype(int)

    @property
    def _done_thresh(self):
        return self._done_thresh_host

    @_done_thresh.setter
    def _done_thresh(self, v):
        self._done_thresh_host = v
        if getattr(self, "_vec", None) is not None:
            self._vec.set_done_thresh(float(v))

    def _absorb_thresh(self, done, kind):
        """The kernel bumps done_thresh on coverage exactly where the reference's done() does; mirror
        it on the host and report whether this step ended by coverage (the reference prints then)."""
        before = self._done_thresh_host
        changed = self._done_thresh_dev != float(before)
        if changed:
            self._done_thresh_host = self._done_thresh_dev
        if not done or kind == "sentinel":
            return False
        return changed or min(before, 1) <= self.percent_covered()

    def isInBounds(self, x, y):
        return x >= 0 and x < self._gridwidth and y >= 0 and y < self._gridlen

    def isOccupied(self, x, y):
        if self._grid[x][y] < 0:
            return True
        return bool(np.any((self._xinds == x) & (self._yinds == y)))

    def percent_covered(self):
        return float(self._vec.percent_covered()[0].item())

    def done(self):
        """dec_grid_rl.py:533-546 / super_grid_rl.py:401-414 — note the side effect on done_thresh."""
        if min(self._done_thresh, 1) <= self.percent_covered():
            print("Full Environment Covered")
            self._done_thresh = self._done_thresh + self._done_incr
            return True
        if getattr(self, "_maxsteps", None) is not None and self._currstep == self._maxsteps:
            return True
        return False

    def _frame(self):
        """render() without the window: the frame is composed by the CUDA render kernel."""
        return self._vec.render()[0].cpu().numpy()


class DecGridRL(_CompatBase):
    """Drop-in for Environments/dec_grid_rl.py:DecGridRL, stepped on the GPU."""

    def __init__(self, train_set, env_config, use_graph=False, test_set=None):
        super().__init__()
        c = env_config
        self._vec = None
        self._train_gridlis = train_set
        self._test_gridlis = test_set
        self._numrobot = c["numrobot"]
        self._maxsteps = c["maxsteps"]
        self._collision_penalty = c["collision_penalty"]
        self._done_thresh_host = c["done_thresh"]
        self._done_incr = c["done_incr"]
        self._terminal_reward = c["terminal_reward"]
        self._dist_r = c["dist_reward"]
        self._train_maxsteps = c["train_maxsteps"]
        self._test_maxsteps = c["test_maxsteps"]
        self._egoradius = c["egoradius"]
        self._mini_map_rad = c["mini_map_rad"]
        self._comm_radius = c["comm_radius"]
        self._allow_comm = c["allow_comm"]
        self._map_sharing = c["map_sharing"]
        self._use_graph = use_graph
        self._single_square_tool = c["single_square_tool"]
        self._dijkstra_input = c["dijkstra_input"]
        if c["sensor_type"] == "lidar":
            self._sensor = LidarSensor(c["sensor_config"])
        elif c["sensor_type"] == "square_sensor":
            self._sensor = SquareSensor(c["sensor_config"])
        self._pad = max(self._egoradius, self._mini_map_rad)
        self._build(VecDecGridRL, train_set, test_set, c)
        self.reset(False, None)
        self._obs_dim = self.get_egocentric_observations()[0].shape
        self._num_actions = 4

    def reset(self, testing, ind):
        # dec_grid_rl.py:449-531 — host draws in the reference's order, device does the rest
        k, base = self._map_choice(testing, ind)
        self._grid = np.pad(base, (1,), "constant", constant_values=(-1,))
        self._gridwidth, self._gridlen = self._grid.shape
        n = self._numrobot
        start = np.zeros((1, n, 2), dtype=np.int16)
        taken = set()
        count = 0
        while count != n:
            x = np.random.randint(self._gridwidth)
            y = np.random.randint(self._gridlen)
            if self._grid[x][y] >= 0 and (x, y) not in taken:
                taken.add((x, y))
                start[0, count] = (x, y)
                count += 1
        obs = self._vec.reset(map_idx=np.array([k], dtype=np.int32), start_xy=start)
        self._sync_head()
        self._numfree = int(np.count_nonzero(self._grid > 0))
        self._adjacency_matrix = self._vec.adjacency()[0]
        observations = obs[0].double().cpu().numpy()
        if self._allow_comm and self._use_graph:
            return observations, self._grid, self._adjacency_matrix
        return observations, self._grid

    def step(self, action):
        import torch
        kind, val = _scalar_action(action)
        if kind == "per_robot":
            act = torch.as_tensor(val.reshape(1, -1).astype(np.uint8))
        else:
            act = torch.tensor([val], dtype=torch.int64)
        obs, _, done = self._vec.step(act)
        reward = float(self._vec.reward64[0].item())
        done = bool(done[0].item())
        self._sync_head()
        if self._absorb_thresh(done, kind):
            print("Full Environment Covered")
        self._adjacency_matrix = self._vec.adjacency()[0]
        observations = obs[0].double().cpu().numpy()
        if self._allow_comm and self._use_graph:
            return [observations, self._adjacency_matrix], reward, done
        return observations, reward, done

    def get_egocentric_observations(self):
        return self._vec.observe()[0].double().cpu().numpy()

    # views of device state under the reference's attribute names (read-only copies)
    @property
    def _visited(self):
        return self._vec.planes()[2][0].astype(np.float64)

    @property
    def _observed_obstacles(self):
        return self._vec.planes()[1][0].any(axis=0).astype(np.float64)

    @property
    def _robot_pos_map(self):
        m = np.zeros((self._gridwidth, self._gridlen))
        m[self._xinds, self._yinds] = 1
        return m

    def _padded(self, planes):
        p = self._pad
        return np.pad(planes.astype(np.float64), [(0, 0)] * (planes.ndim - 2) + [(p, p), (p, p)])

    @property
    def _free_pad(self):
        return self._padded(self._vec.planes()[0][0])

    @property
    def _obst_pad(self):
        return self._padded(self._vec.planes()[1][0])

    def render(self):
        # dec_grid_rl.py:554-616 returns pygame.surfarray.array3d(surf): uint8
        return self._frame()


class SuperGridRL(_CompatBase):
    """Drop-in for Environments/super_grid_rl.py:SuperGridRL, stepped on the GPU."""

    def __init__(self, train_set, env_config, test_set=None):
        super().__init__()
        c = env_config
        self._vec = None
        self._train_gridlis = train_set
        self._test_gridlis = test_set
        self._numrobot = c["numrobot"]
        self._train_maxsteps = c["train_maxsteps"]
        self._test_maxsteps = c["test_maxsteps"]
        self._collision_penalty = c["collision_penalty"]
        self._senseradius = c["senseradius"]
        self._free_penalty = c["free_penalty"]
        self._done_thresh_host = c["done_thresh"]
        self._done_incr = c["done_incr"]
        self._terminal_reward = c["terminal_reward"]
        self._dist_r = c["dist_reward"]
        self._use_scanning = c["use_scanning"]
        self._prev_states = []
        self.a_prev = None
        self._build(VecSuperGridRL, train_set, test_set, c)
        self.reset(False, False)
        self.a_inv = {2: 3, 3: 2, 0: 1, 1: 0}
        state = self.get_state()
        self._obs_dim = state[0].shape
        self._num_actions = 4 ** self._numrobot

    def reset(self, testing, ind):
        # super_grid_rl.py:343-399
        k, self._grid = self._map_choice(testing, ind)
        self._gridwidth, self._gridlen = self._grid.shape
        n = self._numrobot
        start = np.zeros((1, n, 2), dtype=np.int16)
        taken = set()
        count = 0
        while count != n:
            x = np.random.randint(self._gridwidth)
            y = np.random.randint(self._gridlen)
            if self._grid[x][y] >= 0 and (x, y) not in taken:
                taken.add((x, y))
                start[0, count] = (x, y)
                count += 1
        obs = self._vec.reset(map_idx=np.array([k], dtype=np.int32), start_xy=start)
        self._sync_head()
        return (obs[0].double().cpu().numpy(), self._currstep), self._grid

    def step(self, action):
        import torch
        kind, val = _scalar_action(action)
        if kind == "per_robot":
            act = torch.as_tensor(val.reshape(1, -1).astype(np.uint8))
        else:
            act = torch.tensor([val], dtype=torch.int64)
        obs, _, done = self._vec.step(act)
        reward = float(self._vec.reward64[0].item())
        done = bool(done[0].item())
        self._sync_head()
        if kind != "sentinel":
            self.a_prev = 0
        if self._absorb_thresh(done, kind):
            print("Full Environment Covered")
        return (obs[0].double().cpu().numpy(), self._currstep), reward, done

    def get_state(self):
        return (self._vec.observe()[0].double().cpu().numpy(), self._currstep)

    @property
    def _free(self):
        return (~self._vec.planes()[0][0]).astype(np.float64)

    @property
    def _observed_obstacles(self):
        return self._vec.planes()[1][0].astype(np.float64)

    def get_pos_image(self):
        if self._use_scanning:
            ret = np.zeros((self._gridwidth, self._gridlen))
            ret[self._xinds, self._yinds] = 1
            return [ret]
        out = []
        for x, y in zip(self._xinds, self._yinds):
            layer = np.zeros((self._gridwidth, self._gridlen))
            layer[x, y] = 1
            out.append(layer)
        return out

    def render(self):
        # super_grid_rl.py:423-463 returns the float64 image it blits
        return self._frame().astype(np.float64)
