"""Test infrastructure (never imported by the product): numpy model of the device RNG and of the on-device
map generator, so the kernel's draws can be pinned bit for bit.

``philox4x32_10`` is Random123's Philox4x32-10 (Salmon et al., SC'11) — the generator
marl_coverage_b200/csrc/mcov_common.cuh implements for reset draws and ``mcov_gen_maps``; it is pinned here
against the published known-answer vectors (tests/test_philox_cpu.py).  ``gen_maps`` restates
``gen_maps_kernel`` (mcov_api.cu): cell (m, x, y) takes lane ``y & 3`` of Philox(counter = (m, x, y >> 2,
0x47454E4D), key = (seed low, seed high)) and is an obstacle iff that 32-bit draw < round(prob_obst * 2^32) —
the distribution of the reference's ``np.random.choice(a=[1.0, -1.0], p=[1 - p, p])`` (Utils/gridmaker.py:128-129).
"""
import math

import numpy as np

GEN_TAG = 0x47454E4D


def philox4x32_10(ctr, key):
    """ctr uint32 [..., 4], key uint32 [..., 2] -> uint32 [..., 4]."""
    m0, m1 = np.uint64(0xD2511F53), np.uint64(0xCD9E8D57)
    w0, w1 = np.uint32(0x9E3779B9), np.uint32(0xBB67AE85)
    ctr = np.asarray(ctr, dtype=np.uint32)
    key = np.asarray(key, dtype=np.uint32)
    c = [ctr[..., i].copy() for i in range(4)]
    k0, k1 = key[..., 0].copy(), key[..., 1].copy()
    with np.errstate(over="ignore"):
        for _ in range(10):
            p0 = m0 * c[0].astype(np.uint64)
            p1 = m1 * c[2].astype(np.uint64)
            hi0, lo0 = (p0 >> np.uint64(32)).astype(np.uint32), p0.astype(np.uint32)
            hi1, lo1 = (p1 >> np.uint64(32)).astype(np.uint32), p1.astype(np.uint32)
            c = [hi1 ^ c[1] ^ k0, lo1, hi0 ^ c[3] ^ k1, lo0]
            k0 = k0 + w0
            k1 = k1 + w1
    return np.stack(c, axis=-1)


def gen_maps(n_maps, W, L, prob_obst, seed):
    """The grids ``mcov_gen_maps`` draws: list of float64 [W, L] arrays, +1 free / -1 obstacle."""
    thresh = int(math.floor(prob_obst * 4294967296.0 + 0.5))
    m, x, q = np.meshgrid(np.arange(n_maps, dtype=np.uint32), np.arange(W, dtype=np.uint32),
                          np.arange((L + 3) // 4, dtype=np.uint32), indexing="ij")
    ctr = np.stack([m, x, q, np.full_like(m, GEN_TAG)], axis=-1)
    key = np.broadcast_to(np.array([seed & 0xFFFFFFFF, (seed >> 32) & 0xFFFFFFFF], dtype=np.uint32),
                          ctr.shape[:-1] + (2,))
    draws = philox4x32_10(ctr, key).reshape(n_maps, W, -1)[:, :, :L]          # [m, x, y]: lane y & 3 of group y >> 2
    grids = np.where(draws.astype(np.uint64) < np.uint64(thresh), -1.0, 1.0)
    return [np.array(g) for g in grids]


RESET_TAG = 0x4D434F56


def _bounded(u, n):
    return (int(u) * int(n)) >> 32


def draw_starts(walls, n_robots, seed, gid, episode):
    """Restates ``draw_starts`` (mcov_common.cuh): the episode-start draw of the fused kernels for the env with
    GLOBAL id ``gid`` at its ``episode``-th reset.  walls: bool [M, W, L] obstacle planes in env coordinates
    (DecGridRL: with the border ring).  Returns (map index, int [N, 2] distinct walkable cells).
    Same distribution as the reference's draws (dec_grid_rl.py:462-502, super_grid_rl.py:354-388): a uniform
    map, then uniform cells by rejection; Philox counter = (gid, episode, draw number)."""
    key = np.array([seed & 0xFFFFFFFF, (seed >> 32) & 0xFFFFFFFF], dtype=np.uint32)

    def draw(k):
        return philox4x32_10(np.array([gid, episode, k, RESET_TAG], dtype=np.uint32), key)

    M, W, L = walls.shape
    mi = _bounded(draw(0)[0], M)
    wall = walls[mi]
    cells = []
    k = 1
    while len(cells) < n_robots and k < 8192:
        r = draw(k)
        for h in range(2):
            if len(cells) == n_robots:
                break
            x, y = _bounded(r[2 * h], W), _bounded(r[2 * h + 1], L)
            if not wall[x, y] and (x, y) not in cells:
                cells.append((x, y))
        k += 1
    for x in range(W):                                  # pathological maps: deterministic scan
        for y in range(L):
            if len(cells) < n_robots and not wall[x, y] and (x, y) not in cells:
                cells.append((x, y))
    while len(cells) < n_robots:
        cells.append((0, 0))
    return mi, np.array(cells, dtype=np.int64)
