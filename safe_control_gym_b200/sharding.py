"""Multi-GPU layout of the batched env step.

Envs are independent, so the path shards with no data-path exchange: rank r of G owns the contiguous
global env ids [r * n_total / G, (r + 1) * n_total / G); its env is created with ``seed_offset`` = the first id,
and because Philox is keyed by (seed, GLOBAL env id) a given env's trajectory does not depend on G.  The only
collective is the sum of the 16-double episode-statistics vector (SCG_STAT_*), once per reporting interval:
NCCL over NVLink on the GPUs, gloo in the CPU tests; ``StatsMerger`` overlaps it with the next launch.
"""
import torch
import torch.distributed as dist

from . import _lib as L


def shard_range(n_total, rank, world):
    """Global env ids [lo, hi) of ``rank``; sizes differ by at most one and cover [0, n_total) exactly."""
    if not (0 <= rank < world):
        raise ValueError("rank out of range")
    base, rem = divmod(int(n_total), int(world))
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def merge_stats(stats, group=None):
    """All-reduce (sum) of an episode-statistics vector (float64[SCG_STATS_LEN], any device) in place and
    return it as the ``env.stats()`` dict.  A no-op without an initialised process group."""
    assert stats.dtype == torch.float64 and stats.numel() == L.STATS_LEN
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(stats, op=dist.ReduceOp.SUM, group=group)
    vals = stats.detach().cpu().tolist()
    return {name: float(vals[i]) for i, name in enumerate(L.STAT_NAMES)}


class StatsMerger:
    """Overlapped merge for stepping loops: ``submit(stats)`` snapshots the vector into one of two buffers and starts an
    asynchronous all-reduce on the backend's own stream / thread, so the caller's next launch does not wait for it;
    ``drain()`` waits for what is in flight and returns the most recent merged vector (``None`` before the first submit).
    Without a process group the snapshot is the result."""

    def __init__(self, device, group=None):
        self.group = group
        self.bufs = [torch.zeros(L.STATS_LEN, dtype=torch.float64, device=device) for _ in range(2)]
        self.pending = [None, None]
        self.turn = 0
        self.last = None
        self.active = dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1

    def submit(self, stats):
        assert stats.dtype == torch.float64 and stats.numel() == L.STATS_LEN
        j = self.turn & 1
        self.turn += 1
        if self.pending[j] is not None:   # the merge issued two submits ago: long finished
            self.pending[j].wait()
        self.bufs[j].copy_(stats)
        self.pending[j] = dist.all_reduce(self.bufs[j], op=dist.ReduceOp.SUM, group=self.group, async_op=True) if self.active else None
        self.last = j

    def drain(self):
        for j in range(2):
            if self.pending[j] is not None:
                self.pending[j].wait()
                self.pending[j] = None
        return None if self.last is None else self.bufs[self.last]


def make_sharded(env_cls, n_total, rank=None, world=None, device=None, **kwargs):
    """This rank's slice of an ``n_total``-env batch: ``env_cls(num_envs=hi - lo, seed_offset=lo, ...)``."""
    if rank is None:
        rank = dist.get_rank() if dist.is_initialized() else 0
    if world is None:
        world = dist.get_world_size() if dist.is_initialized() else 1
    lo, hi = shard_range(n_total, rank, world)
    if device is None:
        device = torch.device("cuda", torch.cuda.current_device())
    return env_cls(num_envs=hi - lo, seed_offset=lo, device=device, **kwargs)
