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
    """All-reduce of every step's statistics vector.

    CUDA, world > 1: the vectors of ``block`` consecutive steps are collected in a [block, 8] device block
    and reduced with ONE collective on a side stream while the next block's steps run (two blocks, event
    hand-over), so the latency-bound collective neither sits between two step kernels nor costs a launch per
    step; ``block=1`` reduces after every step.  Every step's vector is reduced either way — ``last_block()``
    exposes the per-step global rows of the block that completed last, ``total`` their running sum.
    CPU / gloo: a plain blocking all-reduce with the same block semantics."""

    def __init__(self, device, world=None, group=None, block=1, on_reduced=None, warm=True):
        import torch
        import torch.distributed as dist
        self.torch, self.dist, self.group = torch, dist, group
        self.device = torch.device(device)
        self.world = world if world is not None else (dist.get_world_size(group) if dist.is_initialized() else 1)
        self.cuda = self.device.type == "cuda"
        self.block = max(1, int(block))
        self.blocks = [torch.zeros((self.block, 8), dtype=torch.float64, device=self.device) for _ in range(2)]
        self.total = torch.zeros(8, dtype=torch.float64, device=self.device)
        self.k = 0
        self._last = None
        self.overlap = self.cuda and self.world > 1
        # on_reduced(view, first_step): called right after a block's reduction has been enqueued, inside the
        # stream that ran it (side stream when overlapping) — e.g. to copy the global rows to pinned host memory
        self.on_reduced = on_reduced
        if self.overlap:
            self.side = torch.cuda.Stream(device=self.device)
            self.ev_ready = [torch.cuda.Event() for _ in range(2)]
            self.ev_done = [torch.cuda.Event() for _ in range(2)]
        if warm:
            self.warm()

    def warm(self):
        """Runs one reduction of an all-zero block through every code path a real one takes (the collective on
        the side stream, the row sum, the accumulation), so that the first real reduction — which lands in the
        middle of a rollout — carries no one-off cost: lazy module loads of the two torch kernels, the first
        NCCL launch on a brand-new stream.  Collective: every rank constructs its reducer at the same point."""
        hook, self.on_reduced = self.on_reduced, None
        for blk in (0, 1):
            self._reduce(blk, self.block)
            if self.block > 1:
                self._reduce(blk, 1)
        if self.overlap:
            self.torch.cuda.current_stream(self.device).wait_stream(self.side)
        if self.cuda:
            self.torch.cuda.synchronize(self.device)
        self.total.zero_()
        self._last = None
        self.on_reduced = hook

    def buffer(self):
        """The 8-double row the next step must write (waits until the block's previous reduction is done)."""
        blk, row = (self.k // self.block) & 1, self.k % self.block
        if self.overlap and row == 0 and self.k >= 2 * self.block:
            self.torch.cuda.current_stream(self.device).wait_event(self.ev_done[blk])
        return self.blocks[blk][row]

    def _reduce(self, blk, rows):
        view = self.blocks[blk][:rows]
        if self.overlap:
            main = self.torch.cuda.current_stream(self.device)
            self.ev_ready[blk].record(main)
            with self.torch.cuda.stream(self.side):
                self.side.wait_event(self.ev_ready[blk])
                self.dist.all_reduce(view, group=self.group)
                self.total.add_(view[0] if rows == 1 else view.sum(dim=0))
                if self.on_reduced is not None:
                    self.on_reduced(view, self.k - rows)
                self.ev_done[blk].record(self.side)
        else:
            if self.world > 1:
                self.dist.all_reduce(view, group=self.group)
            self.total.add_(view[0] if rows == 1 else view.sum(dim=0))
            if self.on_reduced is not None:
                self.on_reduced(view, self.k - rows)
        self._last = view

    def submit(self):
        """Call after the step that wrote ``buffer()`` has been enqueued."""
        blk, row = (self.k // self.block) & 1, self.k % self.block
        self.k += 1
        if row == self.block - 1:
            self._reduce(blk, self.block)

    def last_block(self):
        """[rows, 8] global per-step statistics of the most recently reduced block (None before the first);
        on CUDA valid once the side stream has passed ``ev_done`` — ``finish()`` or a stream wait."""
        return self._last

    def flush(self):
        """Reduces a partially filled block now (end of a rollout); the next step starts a fresh block."""
        rows = self.k % self.block
        if rows:
            self._reduce((self.k // self.block) & 1, rows)
            self.k += self.block - rows

    def finish(self):
        """Flushes, joins the side stream; returns the running global totals."""
        self.flush()
        if self.overlap:
            self.torch.cuda.current_stream(self.device).wait_stream(self.side)
        if self.cuda:
            self.torch.cuda.synchronize(self.device)
        vals = self.total.cpu().tolist()
        return dict(zip(STATS_FIELDS, vals))
