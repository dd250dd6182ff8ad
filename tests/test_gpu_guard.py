"""Out-of-bounds and hand-over checks of our own (compute-sanitizer is closed on this GPU pool: "runs under it have
left GPUs needing a reset", gpurun_out/sanitizer_*_r2a.log).

* guard bands: every batch buffer the kernels write is re-bound as the middle of a larger allocation whose margins
  hold a canary; after resets, single steps, fused rollouts (static grid and ticket queue), closed-loop rollouts and
  logged rollouts on RAGGED batches (N % 64 != 0, N % 4 != 0) every canary word must be intact.
* hand-over: the ticket queue passes env state between thread blocks through global memory with st.release.gpu /
  ld.acquire.gpu; a missing fence shows up as a stale read, i.e. a result that differs from the static grid or between
  repetitions.  The queued launch is repeated many times with short chunks (maximal hand-over rate) and compared bit for
  bit with the static grid every time.
"""
import pytest
import torch

import bench
from test_gpu_config4 import CFG as CFG4
from test_gpu_pid import CFG3 as PID3

pytestmark = pytest.mark.gpu
PAD = 4096   # elements on either side of every buffer
BUFS = ["state", "ctrl_step", "episode", "flags", "dist_offset", "inertia", "obs", "env_state", "reward", "done", "info", "mse",
        "c_values", "terminal_obs", "ep_return", "ep_length", "ep_violations", "stats"]


def _canary(dtype):
    return {torch.float32: 1.2345678e30, torch.float64: 1.2345678e300, torch.int32: 0x5A5A5A5A, torch.uint8: 0xA5}[dtype]


def _guard(t):
    big = torch.full((t.numel() + 2 * PAD,), _canary(t.dtype), dtype=t.dtype, device=t.device)
    mid = big[PAD:PAD + t.numel()].view(t.shape)
    mid.copy_(t)
    return big, mid


def _guard_env(env):
    """Re-bind every batch buffer of `env` inside a canary-filled allocation; returns the list of (name, big)."""
    bigs = []
    for name in BUFS:
        big, mid = _guard(getattr(env, "buf_" + name))
        setattr(env, "buf_" + name, mid)
        bigs.append((name, big))
    env._obs_view = env.buf_obs.t()
    env._bind()
    return bigs


def _check(bigs, what):
    for name, big in bigs:
        c = _canary(big.dtype)
        lo, hi = big[:PAD], big[-PAD:]
        assert bool((lo == c).all()) and bool((hi == c).all()), f"{what}: write outside buf_{name}"


def _extra(shape, dtype=torch.float32):
    big, mid = _guard(torch.zeros(*shape, dtype=dtype, device="cuda"))
    return big, mid


@pytest.mark.parametrize("n", [5037, 5036, 64 * 3 + 1, 2 * 1184 * 64 + 4])
def test_guard_bands_config2(n):
    from safe_control_gym_b200.envs.quadrotor import BatchedQuadrotor
    env = BatchedQuadrotor(num_envs=n, dtype=torch.float32, attitude="fast_intrinsic", **bench.env_kwargs())
    bigs = _guard_env(env)
    k = 70
    acts = bench.make_actions(torch, n, k, torch.device("cuda:0"), 3, 0.01)
    env.reset()
    _check(bigs, "reset")
    env.step(acts[0])
    _check(bigs, "step")
    for mode in (0, 1):
        env.set_rollout_queue(mode, 32)
        rbig, rew = _extra((k, n)); dbig, dn = _extra((k, n), torch.uint8)
        lbig, log = _extra(((k + 2) // 3, 16, n))
        env.rollout(acts, reward_out=rew, done_out=dn, log_out=log, log_stride=3)
        assert env.last_rollout_tlog == (mode == 0 and n % 4 == 0)   # TMA-staged log store on the static grid
        env.rollout(acts)
        torch.cuda.synchronize()
        _check(bigs + [("reward_out", rbig), ("done_out", dbig), ("log_out", lbig)], f"rollout queue={mode}")
        assert bool(torch.isfinite(log).all()) and bool((log[:, 12:] > 0).all())   # every record was written (rpm > 0)
    mask = (torch.arange(n, device="cuda") % 3 == 0).to(torch.uint8)
    env.reset(mask=mask)
    _check(bigs, "masked reset")
    env.close()


@pytest.mark.parametrize("n", [64 * 2 + 5, 3001])
def test_guard_bands_closed_loop_and_disturbed(n):
    from safe_control_gym_b200.envs.quadrotor import BatchedQuadrotor
    env = BatchedQuadrotor(num_envs=n, dtype=torch.float32, auto_reset=True, **PID3)
    bigs = _guard_env(env)
    pbig, env.pid_mem = _extra((9, n))
    k = 2 * env.CTRL_STEPS + 9
    abig, aout = _extra((k, n, 4))
    env.reset()
    for mode in (0, 1):
        env.set_rollout_queue(mode, 32)
        env.rollout_pid(k, actions_out=aout)
        torch.cuda.synchronize()
        _check(bigs + [("pid_mem", pbig), ("actions_out", abig)], f"rollout_pid queue={mode}")
    env.close()
    env = BatchedQuadrotor(num_envs=n, dtype=torch.float32, auto_reset=True, **CFG4)
    bigs = _guard_env(env)
    acts = (0.16 + 0.01 * torch.randn(150, n, 2, generator=torch.Generator().manual_seed(5))).cuda()
    env.reset()
    for mode in (0, 1):
        env.set_rollout_queue(mode, 32)
        env.rollout(acts)
        env.step(acts[0])
        torch.cuda.synchronize()
        _check(bigs, f"config 4 queue={mode}")
    env.close()


def test_queue_handover_is_repeatable():
    """30 queued launches with 16-step chunks (1,024 env blocks x 16 chunks = 16,384 hand-overs per launch) against the
    static grid: state, counters and flags bit-identical after every launch."""
    from safe_control_gym_b200.envs.quadrotor import BatchedQuadrotor
    n, k = 65536, 256
    kw = bench.env_kwargs()
    a = BatchedQuadrotor(num_envs=n, dtype=torch.float32, attitude="fast_intrinsic", **kw)
    b = BatchedQuadrotor(num_envs=n, dtype=torch.float32, attitude="fast_intrinsic", **kw)
    a.set_rollout_queue(1, 16); b.set_rollout_queue(0)
    acts = bench.make_actions(torch, n, k, torch.device("cuda:0"), 11, 0.005)
    a.reset(); b.reset()
    for rep in range(30):
        a.rollout(acts); b.rollout(acts)
        torch.cuda.synchronize()
        assert a.last_rollout_queued and not b.last_rollout_queued
        for name in ("state", "ctrl_step", "episode", "flags", "ep_length", "ep_return"):
            assert torch.equal(getattr(a, "buf_" + name), getattr(b, "buf_" + name)), (rep, name)
    sa, sb = a.stats(), b.stats()
    assert sa["env_steps"] == sb["env_steps"] == 30 * n * k and sa["episodes"] == sb["episodes"]
    a.close(); b.close()
