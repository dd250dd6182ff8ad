"""Host-side multi-GPU logic on CPU: world_size-2 gloo run of the shard layout and of the one collective of
the path (sum of the episode-statistics vector), plus the shard arithmetic."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from safe_control_gym_b200 import _lib as L
from safe_control_gym_b200.sharding import merge_stats, shard_range


def test_shard_ranges_partition_the_batch():
    for n in (1, 7, 65536, 16777216, 1000003):
        for world in (1, 2, 3, 4, 8):
            spans = [shard_range(n, r, world) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == n
            assert all(spans[i][1] == spans[i + 1][0] for i in range(world - 1))
            sizes = [hi - lo for lo, hi in spans]
            assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        shard_range(10, 2, 2)


def _worker(rank, world, port, n_total, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    lo, hi = shard_range(n_total, rank, world)
    # each rank's statistics as its kernels would leave them: here every env contributes its global id
    ids = torch.arange(lo, hi, dtype=torch.float64)
    stats = torch.zeros(L.STATS_LEN, dtype=torch.float64)
    stats[L.STAT_NAMES.index("env_steps")] = float(hi - lo) * 10
    stats[L.STAT_NAMES.index("episodes")] = float(hi - lo)
    stats[L.STAT_NAMES.index("return")] = float(ids.sum())
    merged = merge_stats(stats)
    if rank == 0:
        torch.save(merged, out)
    dist.barrier()
    dist.destroy_process_group()


def test_stats_allreduce_world2(tmp_path):
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    n_total, out = 1001, str(tmp_path / "merged.pt")
    mp.spawn(_worker, args=(2, port, n_total, out), nprocs=2, join=True)
    merged = torch.load(out)
    assert merged["env_steps"] == n_total * 10 and merged["episodes"] == n_total
    assert merged["return"] == n_total * (n_total - 1) / 2      # every global id exactly once


def test_merge_stats_without_process_group_is_identity():
    s = torch.arange(L.STATS_LEN, dtype=torch.float64)
    m = merge_stats(s.clone())
    assert [m[k] for k in L.STAT_NAMES] == list(range(len(L.STAT_NAMES)))
