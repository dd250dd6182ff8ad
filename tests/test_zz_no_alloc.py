"""scg_rollout_noalloc (include/scg_b200.h; no counterpart in the reference):
the fused rollout for callers that cannot call cudaMalloc.  Its ticket-queue scratch is a static device array inside the
library; results must equal scg_rollout's.  Collected last on purpose."""
import ctypes as C

import pytest
import torch

import bench

pytestmark = pytest.mark.gpu


def _pair(n):
    from safe_control_gym_b200.envs.quadrotor import BatchedQuadrotor
    kw = bench.env_kwargs()
    a = BatchedQuadrotor(num_envs=n, dtype=torch.float32, attitude="fast_intrinsic", **kw)
    b = BatchedQuadrotor(num_envs=n, dtype=torch.float32, attitude="fast_intrinsic", **kw)
    return a, b


@pytest.mark.parametrize("n", [65536, 5037])
def test_noalloc_rollout_equals_rollout_on_two_calls_in_a_row(n):
    k = 192
    a, b = _pair(n)
    acts = bench.make_actions(torch, n, k, torch.device("cuda:0"), 7, 0.005)
    ra, rb = torch.empty(k, n, device="cuda"), torch.empty(k, n, device="cuda")
    da, db = torch.empty(k, n, dtype=torch.uint8, device="cuda"), torch.empty(k, n, dtype=torch.uint8, device="cuda")
    a.reset(); b.reset()
    b.set_rollout_queue(0)   # the existing entry point on the static grid: no scratch involved at all
    for call in range(2):
        torch.cuda.synchronize()
        free0 = torch.cuda.mem_get_info()[0]
        a.rollout_noalloc(acts, reward_out=ra, done_out=da)
        torch.cuda.synchronize()
        if call == 1:   # (the first call may load the kernel's module)
            assert torch.cuda.mem_get_info()[0] >= free0, "device memory was allocated during scg_rollout_noalloc"
        assert a.last_rollout_queued
        b.rollout(acts, reward_out=rb, done_out=db)
        torch.cuda.synchronize()
        # integer outputs bit-exact; floats bit-identical as well (a K-step rollout equals any split of it into ranges)
        assert torch.equal(da, db) and torch.equal(ra, rb), call
        for name in ("state", "ctrl_step", "episode", "flags", "obs", "env_state", "ep_return", "ep_length", "reward", "done"):
            assert torch.equal(getattr(a, "buf_" + name), getattr(b, "buf_" + name)), (call, name)
    sa, sb = a.stats(), b.stats()
    assert sa["env_steps"] == 2 * n * k and sa["episodes"] > 0
    for key in sa:
        if key in ("mse", "return"):   # float sums: the order of the additions differs between the schedules
            assert abs(sa[key] - sb[key]) <= 1e-5 * abs(sb[key]) + 1e-9, key
        else:
            assert sa[key] == sb[key], key
    # the handle's own (cudaMalloc'ed) queue still works after the static one was lent to it
    a.set_rollout_queue(1); b.set_rollout_queue(0)
    a.rollout(acts); b.rollout(acts)
    torch.cuda.synchronize()
    assert a.last_rollout_queued and torch.equal(a.buf_state, b.buf_state)
    a.close(); b.close()


def test_noalloc_rejects_what_does_not_fit():
    from safe_control_gym_b200 import _lib as L
    n, k = 65536, 128
    a, _b = _pair(n)
    _b.close()
    acts = bench.make_actions(torch, n, k, torch.device("cuda:0"), 7, 0.005)
    a.reset()
    a.set_rollout_queue(-1, 1)   # 1-step chunks: 128 x 1,024 ranges + 2 words > SCG_NOALLOC_QUEUE_WORDS
    assert 2 + k * (n // 64) > L.NOALLOC_QUEUE_WORDS
    state = a.buf_state.clone()
    with pytest.raises(L.ScgError, match="queue scratch"):
        a.rollout_noalloc(acts)
    with pytest.raises(L.ScgError, match="k_steps must be >= 64"):
        a.rollout_noalloc(acts[:32].contiguous())
    torch.cuda.synchronize()
    assert torch.equal(a.buf_state, state)   # nothing ran
    args = L.RolloutArgs()
    assert a.lib.scg_rollout_noalloc(a._h, C.byref(args), None) == -2   # SCG_ERR_NULL: no actions
    a.set_rollout_queue(-1, 0)
    a.rollout_noalloc(acts)      # default chunk: fits
    torch.cuda.synchronize()
    assert a.last_rollout_queued
    a.close()
