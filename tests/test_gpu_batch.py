"""Batch-level properties of the CUDA path that do not need a per-env reference: Philox-mode parity against
the oracle on the same counter-based streams, fused rollout == host-driven stepping (bit-identical), shard
equivalence (N envs on one GPU == two halves with global env ids), episode statistics, the Logger record, and
determinism at BASELINE config-2 size."""
import numpy as np
import pytest
import torch

import bench
from oracle.quad_oracle import PhiloxEnvDraws, QuadOracle

pytestmark = pytest.mark.gpu


def _quad(n, dtype=torch.float32, attitude="fast", **over):
    from safe_control_gym_b200.envs.quadrotor import BatchedQuadrotor
    kw = bench.env_kwargs()
    kw.update(over)
    return BatchedQuadrotor(num_envs=n, dtype=dtype, attitude=attitude, **kw)


def _actions(n, k, seed=3, std=0.005, dtype=torch.float32):
    return bench.make_actions(torch, n, k, torch.device("cuda:0"), seed, std).to(dtype)


@pytest.mark.parametrize("dtype,tol,attitude", [(torch.float64, 1e-10, "fast"), (torch.float32, 1e-3, "fast"),
                                                (torch.float32, 1e-3, "fast_intrinsic")])
def test_philox_mode_matches_oracle(dtype, tol, attitude):
    """BASELINE config 2 in small (level0 constraints, Philox init randomisation, auto-reset on violation):
    the kernel and the numpy oracle draw from the same (seed, env id, episode) counters.  Free-running (the fp32
    bar is a drift bar; the per-step 1e-5 check of this config is test_gpu_teacher_forced.py)."""
    n, steps, base = 64, 90, 7000
    env = _quad(n, dtype, attitude, seed_offset=base, info_in_reset=True)
    acts = _actions(n, steps, dtype=dtype).cpu().numpy().astype(np.float64)
    lanes = [0, 1, 33, 63]
    kw = bench.env_kwargs()
    seed = kw.pop("seed")
    refs = {i: QuadOracle(draws=PhiloxEnvDraws(seed, base + i), info_in_reset=True, **kw) for i in lanes}
    obs, _ = env.reset()
    for i in lanes:
        o_obs, _ = refs[i].reset()
        assert np.max(np.abs(obs[i].double().cpu().numpy() - o_obs)) <= tol
    resets, worst = 0, 0.0
    for t in range(steps):
        obs, rew, done, info = env.step(acts[t])
        term = info["terminal_observation"]
        for i in lanes:
            o_obs, o_rew, o_done, o_info = refs[i].step(acts[t, i])
            assert bool(done[i]) == bool(o_done) and float(rew[i]) == float(o_rew), (t, i)
            assert int(info["constraint_violation"][i]) == int(o_info["constraint_violation"])
            got = (term if o_done else obs)[i].double().cpu().numpy()
            err = float(np.max(np.abs(got - o_obs) / np.maximum(1.0, np.abs(o_obs))))
            worst = max(worst, err)
            assert err <= tol, (str(dtype), t, i, err)
            if o_done:
                resets += 1
                r_obs, _ = refs[i].reset()
                assert np.max(np.abs(obs[i].double().cpu().numpy() - r_obs)) <= tol
    assert resets >= 2
    print(f"philox parity {dtype}: worst {worst:.2e}, {resets} auto-resets")
    env.close()


@pytest.mark.parametrize("attitude", ["fast", "fast_intrinsic", "roundtrip"])
def test_fused_rollout_equals_host_driven_steps(attitude):
    """The K-step kernel (state in registers, spare initial states drawn ahead) and K single-step launches are
    the same computation: bit-identical state, counters and statistics."""
    n, k = 8192, 48
    acts = _actions(n, k)
    a, b = _quad(n, attitude=attitude), _quad(n, attitude=attitude)
    a.reset(); b.reset()
    rew = torch.empty(k, n, device="cuda"); dn = torch.empty(k, n, dtype=torch.uint8, device="cuda")
    a.rollout(acts, reward_out=rew, done_out=dn)
    for t in range(k):
        _, r, d, _ = b.step(acts[t])
        assert torch.equal(rew[t], r) and torch.equal(dn[t].bool(), d), t
    torch.cuda.synchronize()
    for name in ("state", "ctrl_step", "episode", "flags", "obs", "env_state", "ep_return", "ep_length"):
        assert torch.equal(getattr(a, "buf_" + name), getattr(b, "buf_" + name)), name
    sa, sb = a.stats(), b.stats()
    # integer-valued statistics agree exactly; the mse sum is accumulated in fp32 per thread, so its rounding
    # depends on how many steps one launch covers
    assert all(sa[key] == sb[key] for key in sa if key != "mse") and abs(sa["mse"] - sb["mse"]) <= 1e-6 * sb["mse"]
    assert sa["env_steps"] == n * k and sa["episodes"] > 0
    # VecRecordEpisodeStatistics semantics: return / length of finished episodes
    fin = dn.bool()
    assert sa["episodes"] == int(fin.sum())
    a.close(); b.close()


def test_shard_equivalence():
    """N envs on one GPU == N/2 + N/2 with seed_offset (global env ids): identical trajectories, summed stats."""
    n, k = 4096, 64
    acts = _actions(n, k)
    whole = _quad(n)
    lo, hi = _quad(n // 2, seed_offset=0), _quad(n // 2, seed_offset=n // 2)
    for e in (whole, lo, hi):
        e.reset()
    whole.rollout(acts)
    lo.rollout(acts[:, :n // 2].contiguous())
    hi.rollout(acts[:, n // 2:].contiguous())
    torch.cuda.synchronize()
    assert torch.equal(whole.buf_state[:, :n // 2], lo.buf_state) and torch.equal(whole.buf_state[:, n // 2:], hi.buf_state)
    assert torch.equal(whole.buf_episode[n // 2:], hi.buf_episode)
    sw, sl, sh = whole.stats(), lo.stats(), hi.stats()
    for key in ("episodes", "env_steps", "violation_steps", "length", "return"):
        assert sw[key] == sl[key] + sh[key], key
    for e in (whole, lo, hi):
        e.close()


def test_logged_rollout_record():
    """log_out carries the Logger record [pos3 vel3 rpy3 ang_vel3 rpm4] (Logger.py:50-66) per env and step."""
    n, k = 512, 20
    acts = _actions(n, k)
    a, b = _quad(n, auto_reset=False, done_on_violation=False), _quad(n, auto_reset=False, done_on_violation=False)
    a.reset(); b.reset()
    log = torch.zeros(k // 2, 16, n, device="cuda")
    a.rollout(acts, log_out=log, log_stride=2)
    for t in range(k):
        b.step(acts[t])
        if t % 2 == 0:
            s = b.buf_env_state
            rec = log[t // 2]
            assert torch.equal(rec[0], s[0]) and torch.equal(rec[1], s[2]) and torch.equal(rec[2], s[4])   # pos
            assert torch.equal(rec[3], s[1]) and torch.equal(rec[5], s[5])                                   # vel
            assert torch.equal(rec[6:9], s[6:9])                                                             # rpy
            assert torch.equal(rec[9:12], b.buf_state[9:12])                                                 # rpy_rates (ang_vel in DYN)
            kf = 3.16e-10
            thrust = kf * (rec[12:16] ** 2)                                                                  # rpm -> per-motor thrust
            assert torch.allclose(thrust, acts[t].t().clamp(0.0281616878, 0.1483414471), rtol=2e-6)
    a.close(); b.close()


def test_config2_size_is_deterministic():
    """BASELINE config 2 at full batch size (65,536 envs), 100 steps: two runs are bit-identical and every
    env-step is accounted for."""
    n, k = 65536, 100
    acts = _actions(n, k)
    out = []
    for _ in range(2):
        e = _quad(n, attitude="fast_intrinsic")
        e.reset()
        e.rollout(acts)
        torch.cuda.synchronize()
        out.append((e.buf_state.clone(), e.buf_episode.clone(), e.stats()))
        e.close()
    assert torch.equal(out[0][0], out[1][0]) and torch.equal(out[0][1], out[1][1]) and out[0][2] == out[1][2]
    st = out[0][2]
    assert st["env_steps"] == n * k and st["length"] <= st["env_steps"] and st["episodes"] == st["violated_episodes"] + st["truncated"] + st["completed"]
    assert not torch.isnan(out[0][0]).any()


def test_kernel_variant_selection():
    """scg_kernel_variant: 3 = PLAIN + SPEC for the headline workload (competition switches compiled in), 2 = PLAIN for
    other extra-free fp32 configs, 1 = LEAN as soon as a per-step extra or a
    reset-time draw beyond uniform init offsets is configured (the spare reset carries those too), 0 = GENERAL for the
    literal round-trip attitude (validation), 4 = PLAIN + DS for fp32 configs whose extras are inline action disturbances and
    spare-carried reset draws (3D: with the nominal inertia in the integrator)."""
    from safe_control_gym_b200.envs.cartpole import BatchedCartPole
    lib_variant = lambda env: int(env.lib.scg_kernel_variant(env._h))
    imp = lambda off: {"action": [dict(disturbance_func="impulse", magnitude=0.01, step_offset=off, duration=2, decay_rate=0.9)]}
    cases = [(_quad(64), 3), (_quad(64, done_on_completion=False), 2), (_quad(64, cost="rl_reward"), 2), (_quad(64, gate_geometry=True), 4),
             (_quad(64, disturbances={"action": [dict(disturbance_func="white_noise", std=0.01)]}), 4),
             (_quad(64, disturbances=imp(None)), 4), (_quad(64, disturbances=imp(3)), 4),
             (_quad(64, disturbances={"observation": [dict(disturbance_func="white_noise", std=0.01)]}), 3),   # applied where obs is written
             (_quad(64, randomized_inertial_prop=True), 4), (_quad(64, randomized_inertial_prop=True, dyn_uses_episode_inertia=True), 1),
             (_quad(64, attitude="roundtrip"), 0),
             (_quad(64, dtype=torch.float64), 1),
             (BatchedCartPole(num_envs=64, disturbances={"action": [dict(disturbance_func="white_noise", std=0.2)]}), 4),
             (BatchedCartPole(num_envs=64, dtype=torch.float64, disturbances={"action": [dict(disturbance_func="white_noise", std=0.2)]}), 1),
             (BatchedCartPole(num_envs=64, disturbances={"action": [dict(disturbance_func="uniform", low=-0.1, high=0.1)]}), 1),
             (_quad(64, quad_type=2, task_info=dict(stabilization_goal=[0, 1], stabilization_goal_tolerance=0.05), cost="rl_reward",
                    init_state=None, init_state_randomization_info=None, constraints=None, disturbances=imp(None)), 4)]
    got = [lib_variant(env) for env, _ in cases]
    for env, _ in cases:
        env.close()
    assert got == [want for _, want in cases], got


def test_spare_reset_with_inertia_and_offsets_equals_stepping():
    """The general spare (inertial properties + disturbance offsets drawn ahead) in a fused rollout against K
    single-step launches of the same env: bit-identical state, inertia and offset buffers after many auto-resets."""
    U = lambda lo, hi: dict(distrib="uniform", low=lo, high=hi)
    over = dict(randomized_inertial_prop=True, dyn_uses_episode_inertia=True,
                inertial_prop_randomization_info={"M": U(-0.005, 0.005), "Ixx": U(-1e-6, 1e-6), "Iyy": dict(distrib="normal", loc=0.0, scale=3e-7)},
                disturbances={"action": [dict(disturbance_func="impulse", magnitude=0.01, step_offset=None, duration=2, decay_rate=0.9),
                                         dict(disturbance_func="step", magnitude=0.002, step_offset=None)],
                              "observation": [dict(disturbance_func="step", magnitude=0.001, step_offset=None)]},
                episode_len_sec=1)
    n, k = 4096, 150
    acts = _actions(n, k)
    a, b = _quad(n, **over), _quad(n, **over)
    a.reset(); b.reset()
    a.rollout(acts)
    for t in range(k):
        b.step(acts[t])
    torch.cuda.synchronize()
    for name in ("state", "ctrl_step", "episode", "flags", "obs", "inertia", "dist_offset", "ep_length"):
        assert torch.equal(getattr(a, "buf_" + name), getattr(b, "buf_" + name)), name
    assert int(a.buf_episode.max()) >= 3   # 60-step episodes: every env restarted at least twice
    a.close(); b.close()


def test_ragged_batch_and_masked_reset():
    """Edge cases of the batch layout: a batch that does not fill its last block (100 envs in blocks of 64: the idle
    lanes shadow the last env and stay out of the statistics) equals the first 100 envs of a 128-env batch, and
    reset(mask) restarts exactly the masked envs (DummyVecEnv-style partial reset) leaving the others untouched."""
    k = 40
    small, big = _quad(100, attitude="fast_intrinsic"), _quad(128, attitude="fast_intrinsic")
    small.reset(); big.reset()
    acts = _actions(128, k, std=0.02)
    small.rollout(acts[:, :100].contiguous()); big.rollout(acts)
    assert torch.equal(small.state, big.state[:100])
    assert torch.equal(small.buf_ctrl_step, big.buf_ctrl_step[:100]) and torch.equal(small.buf_episode, big.buf_episode[:100])
    st = small.stats()
    assert st["env_steps"] == 100 * k and st["episodes"] > 0
    # one env, one step: the smallest launch
    one = _quad(1, attitude="fast_intrinsic")
    one.reset()
    one.step(acts[0, :1])
    assert torch.equal(one.state[0], _first_step_state(acts))
    # masked reset
    before_state, before_t, before_ep = big.state.clone(), big.buf_ctrl_step.clone(), big.buf_episode.clone()
    mask = (torch.arange(128, device="cuda:0") % 3 == 0)
    big.reset(mask=mask.to(torch.uint8))
    keep = ~mask
    assert torch.equal(big.state[keep], before_state[keep]) and torch.equal(big.buf_ctrl_step[keep], before_t[keep])
    assert int(big.buf_ctrl_step[mask].abs().sum()) == 0 and torch.equal(big.buf_episode[mask], before_ep[mask] + 1)
    assert torch.equal(big.buf_episode[keep], before_ep[keep])
    for e in (small, big, one):
        e.close()


@pytest.mark.parametrize("n,k,chunk,over", [(5000, 200, 64, {}), (70000, 150, 32, {}), (1000, 130, 48, {"randomized_inertial_prop": True})])
def test_queued_rollout_equals_static_rollout(n, k, chunk, over):
    """The time-chunk ticket queue (resident blocks take (chunk, env block) tickets, state handed over through the
    buffers) and the static grid (one block per 64 envs for all K steps) are the same computation: bit-identical
    state, counters, per-step rewards / dones, exact integer statistics -- on a ragged batch (n % 64 != 0), with more
    env blocks than resident slots (70,000 envs) and on the LEAN kernel with per-episode inertia."""
    if over:
        over = dict(over, inertial_prop_randomization_info=dict(M=dict(distrib="uniform", low=-0.005, high=0.005),
                                                                Ixx=dict(distrib="uniform", low=-1e-6, high=1e-6)))
    acts = _actions(n, k, std=0.01)
    a, b = _quad(n, attitude="fast_intrinsic", **over), _quad(n, attitude="fast_intrinsic", **over)
    a.set_rollout_queue(1, chunk); b.set_rollout_queue(0)
    a.reset(); b.reset()
    ra, rb = torch.empty(k, n, device="cuda"), torch.empty(k, n, device="cuda")
    da, db = torch.empty(k, n, dtype=torch.uint8, device="cuda"), torch.empty(k, n, dtype=torch.uint8, device="cuda")
    for _ in range(2):   # the second launch continues from the first one's buffers
        a.rollout(acts, reward_out=ra, done_out=da); b.rollout(acts, reward_out=rb, done_out=db)
        torch.cuda.synchronize()
        assert a.last_rollout_queued and not b.last_rollout_queued
        assert torch.equal(ra, rb) and torch.equal(da, db)
        for name in ("state", "ctrl_step", "episode", "flags", "obs", "env_state", "ep_return", "ep_length", "reward", "done"):
            assert torch.equal(getattr(a, "buf_" + name), getattr(b, "buf_" + name)), name
    sa, sb = a.stats(), b.stats()
    assert sa["env_steps"] == 2 * n * k and sa["episodes"] > 0
    for key in sa:
        if key in ("mse", "return"):
            assert abs(sa[key] - sb[key]) <= 1e-5 * abs(sb[key]) + 1e-9, key
        else:
            assert sa[key] == sb[key], key
    a.close(); b.close()


def _first_step_state(acts):
    ref = _quad(64, attitude="fast_intrinsic")
    ref.reset()
    ref.step(acts[0, :64].contiguous())
    s = ref.state[0].clone()
    ref.close()
    return s


@pytest.mark.parametrize("k", [128, 200])
def test_bench_path_matches_oracle_on_sampled_lanes(k):
    """The EXACT kernel bench.py times -- BASELINE config 2 at full width (65,536 envs), attitude fast_intrinsic, SPEC
    specialisation, AUTOMATIC time-chunk queue, K >= 128, no per-step outputs -- against QuadOracle(PhiloxEnvDraws) on
    sampled lanes: after the launch the lanes' env.state, step / episode counters, lagged violation flag and running
    episode length must be what the fp64 oracle reaches through the same actions and the same counter-based draws
    (free-running fp32 over K steps with auto-resets: the drift bar 1e-3 of this config, integers exact)."""
    n, base = 65536, 0
    kw = bench.env_kwargs()
    seed = kw.pop("seed")
    env = _quad(n, torch.float32, "fast_intrinsic")
    acts = _actions(n, k, seed=1337)
    env.reset()
    env.rollout(acts)
    torch.cuda.synchronize()
    assert env.last_rollout_queued, "the automatic schedule must pick the ticket queue for this shape"
    assert int(env.lib.scg_kernel_variant(env._h)) == 3   # LEAN + PLAIN + SPEC
    lanes = [0, 1, 31, 32, 63, 64, 4097, 20000, 40001, 65535]
    a_host = acts[:, lanes].double().cpu().numpy()
    state = env.buf_env_state.t()[lanes].double().cpu().numpy()
    steps, eps = env.buf_ctrl_step[lanes].cpu().numpy(), env.buf_episode[lanes].cpu().numpy()
    flags, ep_len = env.buf_flags[lanes].cpu().numpy(), env.buf_ep_length[lanes].cpu().numpy()
    worst, resets = 0.0, 0
    for j, i in enumerate(lanes):
        o = QuadOracle(draws=PhiloxEnvDraws(seed, base + i), **kw)
        o.reset()
        length = 0
        for t in range(k):
            _, _, d, _ = o.step(a_host[t, j])
            length += 1
            if d:
                o.reset()
                resets += 1
                length = 0
        assert int(steps[j]) == o.ctrl_step_counter and int(eps[j]) == o.episode + 1, (i, steps[j], o.ctrl_step_counter)
        assert bool(flags[j] & 1) == bool(o.cnstr_violation) and int(ep_len[j]) == length, i
        err = float(np.max(np.abs(state[j] - o.state) / np.maximum(1.0, np.abs(o.state))))
        worst = max(worst, err)
        assert err <= 1e-3, (i, err)
    assert resets >= 5
    print(f"bench path (fast_intrinsic, SPEC, automatic queue, K = {k}): worst state err {worst:.2e} on {len(lanes)} lanes, {resets} resets")
    env.close()


@pytest.mark.parametrize("variant", ["spec", "plain", "ds_gates", "lean"])
@pytest.mark.parametrize("n,k,stride", [(4100, 37, 1), (4100, 40, 3), (64 * 1184 + 8, 9, 1)])
def test_tma_staged_log_equals_plain_log(variant, n, k, stride):
    """Row X1: the trajectory log written through the TMA unit (records staged in shared memory, one bulk tensor store per
    warp and LOG_T logged steps, scg_step.cuh: rollout_range<..., TLOG>) carries the same bytes as the plain strided
    stores -- ragged last warp (columns clipped by the hardware), an odd number of logged steps (the tail goes out one
    step at a time), log strides, more than one wave of blocks, and the PLAIN, SPEC, PLAIN + DS (BASELINE config 5: gate geometry) and LEAN loops."""
    over = {"spec": {}, "plain": dict(done_on_violation=False), "ds_gates": dict(gate_geometry=True),
            "lean": dict(randomized_inertial_prop=True, dyn_uses_episode_inertia=True)}[variant]
    a, b = _quad(n, attitude="fast_intrinsic", **over), _quad(n, attitude="fast_intrinsic", **over)
    assert int(a.lib.scg_kernel_variant(a._h)) == {"spec": 3, "plain": 2, "ds_gates": 4, "lean": 1}[variant]
    b.set_rollout_queue(1)   # a forced queue keeps the plain store path (k < 64: static grid, plain stores)
    acts = _actions(n, k, std=0.01)
    rows = (k + stride - 1) // stride
    la = torch.full((rows, 16, n), float("nan"), device="cuda"); lb = torch.full((rows, 16, n), float("nan"), device="cuda")
    a.reset(); b.reset()
    for _ in range(2):
        a.rollout(acts, log_out=la, log_stride=stride); b.rollout(acts, log_out=lb, log_stride=stride)
        torch.cuda.synchronize()
        assert a.last_rollout_tlog == (variant != "lean") and not b.last_rollout_tlog   # the LEAN kernels keep plain stores
        assert not torch.isnan(la).any() and torch.equal(la, lb)
        for name in ("state", "ctrl_step", "episode", "flags", "obs", "ep_length"):
            assert torch.equal(getattr(a, "buf_" + name), getattr(b, "buf_" + name)), name
    assert a.stats()["episodes"] == b.stats()["episodes"]
    # a log tensor the TMA unit cannot address (row pitch not a multiple of 16 bytes) takes the plain path
    c = _quad(4099, attitude="fast_intrinsic", **over)
    c.reset()
    c.rollout(_actions(4099, 8), log_out=torch.empty(8, 16, 4099, device="cuda"))
    assert not c.last_rollout_tlog
    a.close(); b.close(); c.close()


def test_stats_read_folds_the_replicas():
    """The device accumulators are SCG_STAT_REPLICAS copies of the statistics vector (a warp adds to copy `warp index mod 64`);
    scg_stats_read returns their totals -- equal to the facade's own fold, with every warp-group copy in use, and zeroed by
    scg_stats_reset."""
    import ctypes as C
    from safe_control_gym_b200 import _lib as L
    n, k = 64 * 200 + 40, 70
    env = _quad(n, attitude="fast_intrinsic")
    env.reset()
    env.rollout(_actions(n, k, std=0.01))
    for t in range(3):
        env.step(_actions(n, 1, std=0.01)[0])
    out = (C.c_double * L.STATS_LEN)()
    L.check(env.lib.scg_stats_read(env._h, out, env._stream()))
    torch.cuda.synchronize()
    want = env.stats_vector().cpu().numpy()
    assert np.array_equal(np.array(out[:]), want)
    assert want[L.STAT_NAMES.index("env_steps")] == n * (k + 3)
    assert int((env.buf_stats[:, L.STAT_NAMES.index("env_steps")] > 0).sum()) == L.STAT_REPLICAS   # every replica took its share
    L.check(env.lib.scg_stats_reset(env._h, env._stream()))
    torch.cuda.synchronize()
    assert float(env.buf_stats.abs().sum()) == 0.0
    env.close()
