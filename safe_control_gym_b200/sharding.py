"""Multi-GPU layout of the batched env step.

Envs are independent, so the path shards with no data-path exchange: rank r of G owns the contiguous
global env ids [r * n_total / G, (r + 1) * n_total / G); its env is created with ``seed_offset`` = the first id,
and because Philox is keyed by (seed, GLOBAL env id) a given env's trajectory does not depend on G.  The only
collective is the sum of the 16-double episode-statistics vector (SCG_STAT_*), once per reporting interval:
NCCL over NVLink on the GPUs, gloo in the CPU tests.
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
