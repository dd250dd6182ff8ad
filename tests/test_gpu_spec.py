"""The SPEC instantiations (csrc/scg_params.cuh: SpecCfg -- one compile-time family of scenario switches per env kind) against
the non-SPEC kernels of the same config (``set_spec_enabled(False)``): bit-identical buffers and per-step outputs for every
BASELINE config that takes one, static grid and ticket queue, with auto-resets."""
import pytest
import torch

import bench

pytestmark = pytest.mark.gpu


def _make(cid, n):
    from safe_control_gym_b200.envs.cartpole import BatchedCartPole
    from safe_control_gym_b200.envs.quadrotor import BatchedQuadrotor
    g = torch.Generator(device="cuda"); g.manual_seed(11)
    if cid == "3":
        kw = bench.config3_kwargs(); kw.pop("info_in_reset"); kw["episode_len_sec"] = 1
        env = BatchedCartPole(num_envs=n, dtype=torch.float32, attitude="fast_intrinsic", auto_reset=True, **kw)
        acts = lambda k: torch.empty(k, n, 1, device="cuda").normal_(0.0, 3.0, generator=g).clamp_(-10, 10)
    elif cid == "4":
        kw = bench.config4_kwargs(); kw["episode_len_sec"] = 1
        env = BatchedQuadrotor(num_envs=n, dtype=torch.float32, attitude="fast_intrinsic", auto_reset=True, **kw)
        hover = env.MASS * 9.8 / 2
        acts = lambda k: torch.empty(k, n, 2, device="cuda").normal_(hover, 0.01, generator=g).clamp_(0, 4 * hover)
    else:
        kw = {"2": bench.env_kwargs, "2l": bench.level2_like_kwargs, "5": bench.config5_kwargs}[cid]()
        env = BatchedQuadrotor(num_envs=n, dtype=torch.float32, attitude="fast_intrinsic", auto_reset=True, **kw)
        acts = lambda k: bench.make_actions(torch, n, k, torch.device("cuda:0"), 5, 0.01)
    return env, acts


@pytest.mark.parametrize("cid", ["2", "2l", "3", "4", "5"])
def test_spec_kernels_equal_the_generic_ones(cid):
    n, k = 3000, 130
    a, acts = _make(cid, n)
    b, _ = _make(cid, n)
    b.set_spec_enabled(False)
    A = acts(k)
    a.reset(); b.reset()
    ra, rb = torch.empty(k, n, device="cuda"), torch.empty(k, n, device="cuda")
    da, db = torch.empty(k, n, dtype=torch.uint8, device="cuda"), torch.empty(k, n, dtype=torch.uint8, device="cuda")
    for queue in (0, 1):
        a.set_rollout_queue(queue, 32); b.set_rollout_queue(queue, 32)
        a.rollout(A, reward_out=ra, done_out=da); b.rollout(A, reward_out=rb, done_out=db)
        torch.cuda.synchronize()
        assert a.last_rollout_spec and not b.last_rollout_spec      # the bench config takes its kind's SPEC instantiation
        assert torch.equal(ra, rb) and torch.equal(da, db)
        for name in ("state", "ctrl_step", "episode", "flags", "obs", "ep_return", "ep_length", "inertia", "dist_offset"):
            assert torch.equal(getattr(a, "buf_" + name), getattr(b, "buf_" + name)), (cid, queue, name)
    if cid == "2":   # the host-driven step has its own instantiation of the family (SPEC = 2: per-step outputs compiled in)
        for t in range(40):
            oa, rwa, dna, ia = a.step(A[t]); ob, rwb, dnb, ib = b.step(A[t])
            assert int(a.lib.scg_last_rollout_spec(a._h)) == 2 and int(b.lib.scg_last_rollout_spec(b._h)) == 0
            assert torch.equal(oa, ob) and torch.equal(rwa, rwb) and torch.equal(dna, dnb)
            for name in ("info", "c_values", "mse", "terminal_obs", "env_state", "state", "flags"):
                assert torch.equal(getattr(a, "buf_" + name), getattr(b, "buf_" + name)), (t, name)
    sa, sb = a.stats(), b.stats()
    assert sa["episodes"] > 0
    for key in sa:   # counts are exact; the two fp64 sums are formed by atomics in launch order (last bits may differ)
        if key in ("return", "mse"):
            assert abs(sa[key] - sb[key]) <= 1e-9 * max(1.0, abs(sb[key])), key
        else:
            assert sa[key] == sb[key], key
    if cid == "5":   # the logged form (TMA store) too
        la, lb = torch.empty(16, 16, n, device="cuda"), torch.empty(16, 16, n, device="cuda")
        a.set_rollout_queue(-1); b.set_rollout_queue(-1)
        a.rollout(A[:16], log_out=la); b.rollout(A[:16], log_out=lb)
        torch.cuda.synchronize()
        assert a.last_rollout_spec and not b.last_rollout_spec and a.last_rollout_tlog and b.last_rollout_tlog
        assert torch.equal(la, lb)
    a.close(); b.close()
