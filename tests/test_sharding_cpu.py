"""world_size-2 gloo test of the multi-GPU host logic (env sharding + statistics all-reduce)."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from marl_coverage_b200.sharding import StatsReducer, shard_range

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_shard_ranges_partition():
    for total in (1, 7, 8, 4096, 1048576, 1000003):
        for world in (1, 2, 3, 4, 8):
            spans = [shard_range(total, r, world) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == total
            for a, b in zip(spans, spans[1:]):
                assert a[1] == b[0]
            sizes = [hi - lo for lo, hi in spans]
            assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        shard_range(10, 2, 2)


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _worker(rank, world, port, total_envs, steps, out_dir, block):
    sys.path.insert(0, ROOT)
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    lo, hi = shard_range(total_envs, rank, world)
    rows = []
    red = StatsReducer("cpu", world=world, block=block, on_reduced=lambda view, first: rows.append((first, view.clone())))
    rng = np.random.default_rng(0)                     # same stream on every rank: global per-env stats
    for _ in range(steps):
        per_env = rng.integers(0, 5, size=(total_envs, 8)).astype(np.float64)
        buf = red.buffer()
        buf.copy_(torch.from_numpy(per_env[lo:hi].sum(axis=0)))   # what the step kernel folds per shard
        red.submit()
    tot = red.finish()
    np.save(os.path.join(out_dir, "tot_%d.npy" % rank), np.array(list(tot.values())))
    # the hook saw every step's GLOBAL row exactly once, in order (warm-up reductions do not reach it)
    np.save(os.path.join(out_dir, "rows_%d.npy" % rank), torch.cat([v for _, v in rows]).numpy())
    dist.destroy_process_group()


@pytest.mark.parametrize("block", [1, 3, 8])
def test_stats_allreduce_gloo_world2(tmp_path, block):
    world, total_envs, steps = 2, 37, 7                # 7 steps: blocks of 3 and 8 end on a partial block
    mp.spawn(_worker, args=(world, _free_port(), total_envs, steps, str(tmp_path), block), nprocs=world,
             join=True)
    rng = np.random.default_rng(0)
    want = np.zeros(8)
    per_step = []
    for _ in range(steps):
        per_step.append(rng.integers(0, 5, size=(total_envs, 8)).astype(np.float64).sum(axis=0))
        want += per_step[-1]
    for r in range(world):
        got = np.load(os.path.join(str(tmp_path), "tot_%d.npy" % r))
        assert np.array_equal(got, want)               # every rank holds the global sums
        assert np.array_equal(np.load(os.path.join(str(tmp_path), "rows_%d.npy" % r)), np.stack(per_step))
