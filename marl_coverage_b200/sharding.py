"""Multi-GPU plumbing for the batched env: one process per GPU, envs sharded by index.

The env step has no cross-env data dependence, so the only collective is the
all-reduce(SUM) of the 8-double episode-statistics vector the step kernel folds per launch.  The
device RNG is keyed by the GLOBAL env id (``env_id_base + local index``), which makes per-env
results independent of how many GPUs the batch is spread over.

Works with any ``torch.distributed`` backend: NCCL over NVLink on the B200 box, gloo in the CPU
tests (tests/test_sharding_cpu.py, world_size 2).
"""
import os

STATS_FIELDS = ("episodes", "return_sum", "length_sum", "coverage_sum", "collisions", "new_cells",
                "agent_steps", "reserved")


def shard_range(total_envs, rank, world):
    """[lo, hi) of the global env ids owned by ``rank``: contiguous, sizes differ by at most one."""
    if not (0 <= rank < world):
        raise ValueError("rank %d outside world of %d" % (rank, world))
    base, rem = divmod(int(total_envs), int(world))
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def env_from_torchrun():
    """(rank, local_rank, world) from the torchrun environment (1-process defaults)."""
    return (int(os.environ.get("RANK", "0")), int(os.environ.get("LOCAL_RANK", "0")),
            int(os.environ.get("WORLD_SIZE", "1")))


class StatsReducer:
    """Double-buffered all-reduce of the per-step statistics vector on a side stream, so the tiny,
    latency-bound collective overlaps the next step kernel (CUDA) — or a plain blocking all-reduce
    (CPU / gloo)."""

    def __init__(self, device, world=None, group=None):
        import torch
        import torch.distributed as dist
        self.torch, self.dist, self.group = torch, dist, group
        self.device = torch.device(device)
        self.world = world if world is not None else (dist.get_world_size(group) if dist.is_initialized() else 1)
        self.cuda = self.device.type == "cuda"
        self.bufs = [torch.zeros(8, dtype=torch.float64, device=self.device) for _ in range(2)]
        self.total = torch.zeros(8, dtype=torch.float64, device=self.device)
        self.k = 0
        if self.cuda and self.world > 1:
            self.side = torch.cuda.Stream(device=self.device)
            self.ev_ready = [torch.cuda.Event() for _ in range(2)]
            self.ev_done = [torch.cuda.Event() for _ in range(2)]

    def buffer(self):
        """The statistics buffer the next step must write (waits until its previous all-reduce is done)."""
        b = self.k & 1
        if self.cuda and self.world > 1 and self.k >= 2:
            self.torch.cuda.current_stream(self.device).wait_event(self.ev_done[b])
        return self.bufs[b]

    def submit(self):
        """Call after the step that wrote ``buffer()`` has been enqueued."""
        b = self.k & 1
        if self.world > 1 and self.cuda:
            main = self.torch.cuda.current_stream(self.device)
            self.ev_ready[b].record(main)
            with self.torch.cuda.stream(self.side):
                self.side.wait_event(self.ev_ready[b])
                self.dist.all_reduce(self.bufs[b], group=self.group)
                self.total.add_(self.bufs[b])
                self.ev_done[b].record(self.side)
        else:
            if self.world > 1:
                self.dist.all_reduce(self.bufs[b], group=self.group)
            self.total.add_(self.bufs[b])
        self.k += 1

    def finish(self):
        """Joins the side stream; returns the running global totals (dict of python floats)."""
        if self.cuda and self.world > 1:
            self.torch.cuda.current_stream(self.device).wait_stream(self.side)
        if self.cuda:
            self.torch.cuda.synchronize(self.device)
        vals = self.total.cpu().tolist()
        return dict(zip(STATS_FIELDS, vals))
