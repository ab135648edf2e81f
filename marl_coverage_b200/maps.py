"""Map supply for the batched env: the reference's ``gridload`` / ``gridgen`` contract
(Utils/gridmaker.py:7-139) plus the device-side bit packer.

A "grid" is what the reference hands to the env ctor: a float array [W, L] with <0 obstacle,
>0 free (PNG maps: exactly -1 / +1), axis 0 = x.  ``MapTable`` packs a list of grids into the
strip-major bit planes of include/mcov.h with the CUDA packer ``mcov_pack_maps``.
"""
import os

import numpy as np

from . import _lib

DATA_DIR = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "data", "grids")


def load_grid_set(name, limit=None):
    """Committed copy of the reference's Grids/<name> PNG set, SORTED by filename, as the float64
    +-1 grids ``gridload`` produces (gridmaker.py:89-92)."""
    z = np.load(os.path.join(DATA_DIR, name + ".npz"))
    W, L = (int(v) for v in z["shape"])
    bits = z["bits"] if limit is None else z["bits"][:limit]
    out = []
    for row in bits:
        free = np.unpackbits(row)[:W * L].reshape(W, L)
        out.append(np.where(free > 0, 1.0, -1.0))
    return out


def gridload(grid_config=None):
    """Same contract as Utils/gridmaker.py:7-104: returns (train_set, test_set).

    ``grid_config=None`` -> the nine hand-made 15x15 maps (6 train / 3 test).  Otherwise
    ``grid_dir`` may be a directory of PNGs (read like the reference, but in SORTED order so runs
    are reproducible — the reference uses raw os.listdir order) or the name of a committed set."""
    if grid_config is None:
        z = np.load(os.path.join(DATA_DIR, "builtin15.npz"))
        grids = load_grid_set("builtin15")
        n_train = int(z["n_train"])
        return grids[:n_train], grids[n_train:]
    grid_dir = grid_config["grid_dir"]
    numgrids = grid_config["numgrids"]
    base = os.path.basename(os.path.normpath(grid_dir))
    if os.path.isdir(grid_dir):
        from PIL import Image
        gridlis = []
        for fname in sorted(os.listdir(grid_dir))[:numgrids]:
            img = np.array(Image.open(os.path.join(grid_dir, fname))).astype(float)
            gridlis.append(np.clip(img - 1, -1, 1))
    elif os.path.exists(os.path.join(DATA_DIR, base + ".npz")):
        gridlis = load_grid_set(base, numgrids)
    else:
        raise FileNotFoundError(grid_dir)
    half = len(gridlis) // 2
    if half == 1:
        return gridlis, gridlis
    return gridlis[:half], gridlis[half:]


def gridgen(grid_config):
    """Bernoulli obstacle maps — Utils/gridmaker.py:107-139 (same np.random call, same split)."""
    p = grid_config["prob_obst"]
    gridlis = [np.random.choice(a=[1.0, -1.0], size=(grid_config["gridwidth"], grid_config["gridlen"]),
                                p=[1 - p, p]) for _ in range(grid_config["numgrids"])]
    n = len(gridlis)
    if n == 1:
        return gridlis, gridlis
    return gridlis[:n], gridlis[n:]


def grids_to_cells(grids):
    """float grids -> (int8 sign cells [M, W, L], float64 cell values [M, W, L] or None).

    The bit planes carry what every rule of the envs needs — obstacle (< 0), walkable (>= 0), counted (> 0).
    The one place the reference reads a cell's VALUE is SuperGridRL's coverage reward (``reward += grid[j][k]``,
    super_grid_rl.py:189, sim_env.py:159): maps with free cells worth something other than 1 keep their values
    in a float64 table next to the planes (PNG maps give exactly -1 / +1 and need none)."""
    arr = np.stack([np.asarray(g, dtype=np.float64) for g in grids])
    if not np.all(np.isfinite(arr)):
        raise ValueError("maps must hold finite cell values")
    cells = np.sign(arr).astype(np.int8)
    unit = np.array_equal(np.where(arr > 0, 1.0, np.where(arr < 0, -1.0, 0.0)), arr)
    return cells, (None if unit else np.where(arr > 0, arr, 0.0))


# ---- host-side (numpy) view of the strip-major layout: used by tests and get_state --------------

def pack_plane_host(mask):
    """bool [W, L] -> uint32 [strips * Wp] strip-major words (include/mcov.h)."""
    W, L = mask.shape
    strips, Wp, PW = _geometry(W, L)
    out = np.zeros((strips, Wp), dtype=np.uint32)
    pad = np.zeros((W, strips * 32), dtype=np.uint64)
    pad[:, :L] = mask
    weights = (np.uint64(1) << np.arange(32, dtype=np.uint64))
    for s in range(strips):
        out[s, :W] = (pad[:, 32 * s:32 * s + 32] * weights).sum(axis=1).astype(np.uint32)
    return out.reshape(-1)


def unpack_plane_host(words, W, L):
    """uint32 [..., strips * Wp] -> bool [..., W, L]."""
    strips, Wp, PW = _geometry(W, L)
    w = np.asarray(words, dtype=np.uint32).reshape(words.shape[:-1] + (strips, Wp))
    bits = (w[..., None] >> np.arange(32, dtype=np.uint32)) & np.uint32(1)       # [..., s, x, 32]
    bits = np.moveaxis(bits, -3, -2)                                              # [..., x, s, 32]
    bits = bits.reshape(bits.shape[:-2] + (strips * 32,))
    return bits[..., :W, :L].astype(bool)


def _geometry(W, L):
    strips = (L + 31) // 32
    Wp = (W + 3) // 4 * 4
    return strips, Wp, strips * Wp


class MapTable:
    """Device table of packed maps: ``wall`` / ``pos`` uint32 [M, plane_words] and ``free_counts``
    int32 [M] (= np.count_nonzero(grid > 0), dec_grid_rl.py:517), built by the CUDA packer."""

    def __init__(self, grids, add_border, device):
        import torch
        lib = _lib.load()
        cells, values = grids_to_cells(grids)
        if add_border:
            # DecGridRL never reads a cell's value for its reward (a count, dec_grid_rl.py:243-256), but its square
            # sensor stores np.clip(cell, 0, 1) / np.clip(-cell, 0, 1) in the robots' float maps
            # (squaresensor.py:34-35): magnitudes >= 1 clip to exactly 1 and are fine, magnitudes inside (0, 1)
            # would make those maps fractional, which bit planes cannot hold
            mag = np.abs(np.stack([np.asarray(g, dtype=np.float64) for g in grids]))
            if np.any((mag > 0) & (mag < 1)):
                raise ValueError("DecGridRL maps must not hold cell values of magnitude inside (0, 1): the reference's "
                                 "square sensor would store them as fractional known-free / known-obstacle values")
            values = None
        M, W, L = cells.shape
        self._alloc(M, W, L, add_border, device)
        with torch.cuda.device(self.device):
            # float64 [M, W, L] reward of covering each free cell (map coordinates, no border), or None = all 1.0
            self.values = None if values is None else torch.from_numpy(values).to(self.device).contiguous()
            cells_dev = torch.from_numpy(cells).to(self.device)
            b = 1 if add_border else 0
            st = torch.cuda.current_stream(self.device).cuda_stream
            _lib.check(lib.mcov_pack_maps(cells_dev.data_ptr(), M, W, L, b, self.wall.data_ptr(),
                                          self.pos.data_ptr(), self.free_counts.data_ptr(), st))
            torch.cuda.current_stream(self.device).synchronize()   # cells_dev may be freed now

    def _alloc(self, M, W, L, add_border, device):
        import torch
        b = 1 if add_border else 0
        self.add_border = bool(add_border)
        self.n_maps, self.W, self.L = M, W + 2 * b, L + 2 * b
        self.strips, self.Wp, self.plane_words = _lib.plane_geometry(self.W, self.L)
        self.device = torch.device(device)
        self.values = None
        with torch.cuda.device(self.device):
            self.wall = torch.empty((M, self.plane_words), dtype=torch.int32, device=self.device)
            self.pos = torch.empty((M, self.plane_words), dtype=torch.int32, device=self.device)
            self.free_counts = torch.empty((M,), dtype=torch.int32, device=self.device)

    @classmethod
    def generate(cls, grid_config, add_border, device="cuda:0", seed=0):
        """``gridgen`` on the device (Utils/gridmaker.py:107-139): ``numgrids`` Bernoulli(prob_obst) maps of
        gridwidth x gridlen written straight into the packed planes by ``mcov_gen_maps`` (Philox stream keyed
        by ``seed``; same distribution as the reference, not numpy's draws).  The env constructors take the
        table in place of a list of grids; ``add_border`` must be True for DecGridRL, False for SuperGridRL."""
        import torch
        lib = _lib.load()
        self = cls.__new__(cls)
        M, W, L = int(grid_config["numgrids"]), int(grid_config["gridwidth"]), int(grid_config["gridlen"])
        self._alloc(M, W, L, add_border, device)
        with torch.cuda.device(self.device):
            st = torch.cuda.current_stream(self.device).cuda_stream
            _lib.check(lib.mcov_gen_maps(M, W, L, 1 if add_border else 0, float(grid_config["prob_obst"]),
                                         int(seed) & 0xFFFFFFFFFFFFFFFF, self.wall.data_ptr(), self.pos.data_ptr(),
                                         self.free_counts.data_ptr(), st))
        return self

    def to_grids(self):
        """The table back as the reference's host grids (float64 +1 free / -1 obstacle / 0 walkable-uncounted,
        border ring removed): what ``gridgen`` would have returned for these maps."""
        wall = unpack_plane_host(self.wall.cpu().numpy().view(np.uint32), self.W, self.L)
        pos = unpack_plane_host(self.pos.cpu().numpy().view(np.uint32), self.W, self.L)
        b = 1 if self.add_border else 0
        g = np.where(wall, -1.0, np.where(pos, 1.0, 0.0))
        if b:
            g = g[:, 1:-1, 1:-1]
        if self.values is not None:
            g = np.where(g > 0, self.values.cpu().numpy(), g)
        return [np.array(m) for m in g]
