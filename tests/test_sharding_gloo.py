"""Host-side multi-GPU logic on CPU: world_size-2 gloo run of the shard layout and of the one collective of
the path (sum of the episode-statistics vector), plus the shard arithmetic."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from safe_control_gym_b200 import _lib as L
from safe_control_gym_b200.sharding import StatsMerger, merge_stats, shard_range


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


def _merger_worker(rank, world, port, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    m = StatsMerger(torch.device("cpu"))
    assert m.drain() is None
    stats = torch.zeros(L.STATS_LEN, dtype=torch.float64)
    seen = []
    for step in range(1, 6):   # a stepping loop: the running statistics grow, every step submits a snapshot
        stats[L.STAT_NAMES.index("env_steps")] = float(step * (rank + 1))
        m.submit(stats)
        stats[L.STAT_NAMES.index("episodes")] += 1.0   # the caller keeps writing: the snapshot must not change
        if step == 3:
            seen.append(m.drain().clone())
    seen.append(m.drain().clone())
    if rank == 0:
        torch.save(seen, out)
    dist.barrier()
    dist.destroy_process_group()


def test_stats_merger_overlapped_world2(tmp_path):
    """StatsMerger (what bench.py uses at N > 1): snapshots are merged asynchronously, two buffers in turn; drain() returns
    the latest merged vector and later writes of the caller do not leak into a submitted snapshot."""
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    out = str(tmp_path / "seen.pt")
    mp.spawn(_merger_worker, args=(2, port, out), nprocs=2, join=True)
    seen = torch.load(out)
    i_steps, i_ep = L.STAT_NAMES.index("env_steps"), L.STAT_NAMES.index("episodes")
    assert seen[0][i_steps] == 3 * (1 + 2) and seen[0][i_ep] == 2 * 2     # step 3: two increments had happened on each rank
    assert seen[1][i_steps] == 5 * (1 + 2) and seen[1][i_ep] == 4 * 2


def test_stats_merger_without_process_group():
    m = StatsMerger(torch.device("cpu"))
    s = torch.arange(L.STATS_LEN, dtype=torch.float64)
    m.submit(s)
    s += 1
    assert torch.equal(m.drain(), torch.arange(L.STATS_LEN, dtype=torch.float64))
