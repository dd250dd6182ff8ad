"""BASELINE config 4 in small: Quadrotor 2D DYN with impulse + step action disturbances (offsets drawn per episode)
and per-episode mass / Iyy that reach the integrator (dyn_uses_episode_inertia extension) -- the
GENERAL kernel against the numpy oracle on the same Philox counters, free-running with auto-reset."""
import numpy as np
import pytest
import torch

from oracle.quad_oracle import PhiloxEnvDraws, QuadOracle

pytestmark = pytest.mark.gpu

U = lambda lo, hi: dict(distrib="uniform", low=lo, high=hi)
CFG = dict(quad_type=2, ctrl_freq=60, pyb_freq=240, episode_len_sec=1, task="stabilization", cost="rl_reward",
           task_info=dict(stabilization_goal=[0, 1], stabilization_goal_tolerance=0.05),
           init_state=dict(init_x=0, init_x_dot=0, init_z=1.0, init_z_dot=0, init_theta=0, init_theta_dot=0),
           randomized_init=True,
           init_state_randomization_info={"init_x": U(-0.3, 0.3), "init_z": U(-0.3, 0.3), "init_theta": U(-0.2, 0.2),
                                          "init_theta_dot": U(-0.5, 0.5)},
           randomized_inertial_prop=True, dyn_uses_episode_inertia=True,
           inertial_prop_randomization_info={"M": U(-0.005, 0.005), "Iyy": U(-1e-6, 1e-6)},
           disturbances={"action": [dict(disturbance_func="impulse", magnitude=0.01, step_offset=None, duration=2, decay_rate=0.9),
                                    dict(disturbance_func="step", magnitude=0.005, step_offset=None)]},
           constraints=[dict(constraint_form="default_constraint", constrained_variable="input"),
                        dict(constraint_form="default_constraint", constrained_variable="state")],
           done_on_violation=False, done_on_out_of_bound=True, rew_exponential=True, seed=44)


@pytest.mark.parametrize("dtype,tol", [(torch.float64, 1e-10), (torch.float32, 2e-4)])
def test_config4_matches_oracle(dtype, tol):
    from safe_control_gym_b200.envs.quadrotor import BatchedQuadrotor
    n, steps, base = 64, 150, 500
    env = BatchedQuadrotor(num_envs=n, dtype=dtype, auto_reset=True, seed_offset=base, info_in_reset=True, **CFG)
    assert int(env.lib.scg_kernel_variant(env._h)) == (4 if dtype == torch.float32 else 1)   # fp32: PLAIN + DS (inline impulse / step, spare restart carrying inertia + offsets)
    kw = {k: v for k, v in CFG.items() if k != "seed"}
    lanes = [0, 7, 31, 32, 63]
    refs = {i: QuadOracle(draws=PhiloxEnvDraws(CFG["seed"], base + i), info_in_reset=True, **kw) for i in lanes}
    hover = env.MASS * 9.8 / 2
    rng = np.random.default_rng(9)
    acts = np.clip(hover + rng.normal(0, 0.01, size=(steps, n, 2)), 0, 4 * hover)
    obs, _ = env.reset()
    for i in lanes:
        o_obs, _ = refs[i].reset()
        assert np.max(np.abs(obs[i].double().cpu().numpy() - o_obs)) <= tol
    worst, resets = 0.0, 0
    for t in range(steps):
        obs, rew, done, info = env.step(acts[t])
        term = info["terminal_observation"]
        for i in lanes:
            o_obs, o_rew, o_done, o_info = refs[i].step(acts[t, i])
            assert bool(done[i]) == bool(o_done), (t, i)
            assert int(info["constraint_violation"][i]) == int(o_info["constraint_violation"]), (t, i)
            got = (term if o_done else obs)[i].double().cpu().numpy()
            err = float(np.max(np.abs(got - o_obs) / np.maximum(1.0, np.abs(o_obs))))
            err = max(err, abs(float(rew[i]) - o_rew) / max(1.0, abs(o_rew)))
            worst = max(worst, err)
            assert err <= tol, (str(dtype), t, i, err)
            if o_done:
                resets += 1
                r_obs, _ = refs[i].reset()
                assert np.max(np.abs(obs[i].double().cpu().numpy() - r_obs)) <= tol
    assert resets >= 5   # episodes are 60 steps long: every lane restarts at least twice, with new mass / offsets
    print(f"config 4 {dtype}: worst rel err {worst:.2e}, {resets} auto-resets")
    env.close()


def test_config4_queued_rollout_equals_static_rollout():
    """The ready-list queue on the 2D LEAN kernel with impulse + step disturbances and per-episode mass / inertia (offsets
    and inertial properties travel through the buffers between ranges): forced queue == static grid bit for bit."""
    from safe_control_gym_b200.envs.quadrotor import BatchedQuadrotor
    n, k = 2500, 140
    acts = (0.16 + 0.01 * torch.randn(k, n, 2, generator=torch.Generator().manual_seed(3))).cuda()
    a = BatchedQuadrotor(num_envs=n, dtype=torch.float32, auto_reset=True, **CFG)
    b = BatchedQuadrotor(num_envs=n, dtype=torch.float32, auto_reset=True, **CFG)
    a.set_rollout_queue(1, 32); b.set_rollout_queue(0)
    a.reset(); b.reset()
    for _ in range(2):
        a.rollout(acts); b.rollout(acts)
        torch.cuda.synchronize()
        assert a.last_rollout_queued and not b.last_rollout_queued
        for name in ("state", "ctrl_step", "episode", "flags", "obs", "ep_return", "ep_length", "inertia", "dist_offset"):
            assert torch.equal(getattr(a, "buf_" + name), getattr(b, "buf_" + name)), name
    sa, sb = a.stats(), b.stats()
    assert sa["env_steps"] == 2 * n * k and sa["episodes"] == sb["episodes"] > 0
    a.close(); b.close()


@pytest.mark.parametrize("kind", ["quad2d", "cartpole"])
def test_ds_kernel_equals_lean_kernel(kind):
    """The PLAIN + DS specialisation (variant 4: inline impulse / step / white-noise disturbances, cheap restart) against the LEAN
    kernel of the same config -- reached by handing the env a constraint tolerance it never uses (CFG_COLD_CONSTRAINTS is
    a LEAN-only feature; the tolerance only feeds is_almost_active, which this cost does not read): bit-identical."""
    from test_gpu_cartpole import CFG as CFG3
    n, k = 3000, 150
    if kind == "quad2d":
        from safe_control_gym_b200.envs.quadrotor import BatchedQuadrotor as Env
        cfg, nu = dict(CFG), 2
        acts = (0.16 + 0.01 * torch.randn(k, n, nu, generator=torch.Generator().manual_seed(3))).cuda()
        ntol = 4
    else:
        from safe_control_gym_b200.envs.cartpole import BatchedCartPole as Env
        cfg, nu = dict(CFG3), 1
        acts = torch.randn(k, n, nu, generator=torch.Generator().manual_seed(3)).mul_(3.0).cuda()
        ntol = 2
    lean = dict(cfg)
    lean["constraints"] = [dict(c) for c in cfg["constraints"]]
    lean["constraints"][0]["tolerance"] = [1e-3] * ntol
    a, b = Env(num_envs=n, dtype=torch.float32, auto_reset=True, **cfg), Env(num_envs=n, dtype=torch.float32, auto_reset=True, **lean)
    assert int(a.lib.scg_kernel_variant(a._h)) == 4 and int(b.lib.scg_kernel_variant(b._h)) == 1
    a.reset(); b.reset()
    ra, rb = torch.empty(k, n, device="cuda"), torch.empty(k, n, device="cuda")
    for queue in (0, 1):
        a.set_rollout_queue(queue, 32); b.set_rollout_queue(queue, 32)
        a.rollout(acts, reward_out=ra); b.rollout(acts, reward_out=rb)
        torch.cuda.synchronize()
        assert torch.equal(ra, rb)
        for name in ("state", "ctrl_step", "episode", "flags", "obs", "ep_return", "ep_length", "inertia", "dist_offset"):
            assert torch.equal(getattr(a, "buf_" + name), getattr(b, "buf_" + name)), name
    # single steps take the same kernels (K = 1)
    oa, *_ = a.step(acts[0]); ob, *_ = b.step(acts[0])
    assert torch.equal(oa, ob)
    sa, sb = a.stats(), b.stats()
    assert sa["episodes"] == sb["episodes"] > 0 and sa["violation_steps"] == sb["violation_steps"]
    a.close(); b.close()


def test_ds_kernel_3d_level2_like_matches_lean_kernel():
    """quad_type 3 takes the PLAIN + DS kernel too (variant 4: the packed substep on nominal inertia -- the reference's DYN integrator
    never reads the per-episode draws, SURVEY.md F9 -- plus inline white noise / impulse on the action channel and the cheap restart
    that carries the per-episode inertia and offset draws): competition/level2.yaml's randomisation set.  Checked against the LEAN
    kernel of the same config (forced by an unused constraint tolerance).  The packed substep folds -g dt into one fma where the
    LEAN kernel multiplies and adds, so states agree to fp32 rounding accumulated over the rollout, not bit for bit; the integer
    state (episode, step counter, per-episode draws) must be identical on every lane that did not sit on a termination edge."""
    import bench
    from safe_control_gym_b200.envs.quadrotor import BatchedQuadrotor as Env
    n, k = 4096, 120
    cfg = bench.env_kwargs()
    cfg.update(episode_len_sec=1, randomized_inertial_prop=True,
               inertial_prop_randomization_info={"M": U(-0.01, 0.01), "Ixx": U(-1e-6, 1e-6), "Iyy": U(-1e-6, 1e-6), "Izz": U(-1e-6, 1e-6)},
               randomized_gates_and_obstacles=True, gates_and_obstacles_randomization_info=dict(gates=U(-0.1, 0.1), obstacles=U(-0.2, 0.2)),
               disturbances={"action": [dict(disturbance_func="white_noise", std=0.001),
                                        dict(disturbance_func="impulse", magnitude=0.01, step_offset=None, duration=2, decay_rate=0.9)]})
    lean = dict(cfg)
    lean["constraints"] = [dict(c) for c in cfg["constraints"]]
    lean["constraints"][0]["tolerance"] = [1e-3] * 8
    a = Env(num_envs=n, dtype=torch.float32, attitude="fast_intrinsic", auto_reset=True, **cfg)
    b = Env(num_envs=n, dtype=torch.float32, attitude="fast_intrinsic", auto_reset=True, **lean)
    assert int(a.lib.scg_kernel_variant(a._h)) == 4 and int(b.lib.scg_kernel_variant(b._h)) == 1
    a.reset(); b.reset()
    acts = bench.make_actions(torch, n, k, torch.device("cuda:0"), 5, 0.005)
    ra, rb = torch.empty(k, n, device="cuda"), torch.empty(k, n, device="cuda")
    for queue in (0, 1):
        a.set_rollout_queue(queue, 32); b.set_rollout_queue(queue, 32)
        a.rollout(acts, reward_out=ra); b.rollout(acts, reward_out=rb)
        torch.cuda.synchronize()
        same = (a.buf_episode == b.buf_episode) & (a.buf_ctrl_step == b.buf_ctrl_step)
        assert float(same.float().mean()) > 0.999, float(same.float().mean())
        assert int(a.buf_episode.max()) >= 2          # restarts happened inside the launch
        for name in ("inertia", "dist_offset"):
            x, y = getattr(a, "buf_" + name), getattr(b, "buf_" + name)
            assert torch.equal(x[..., same], y[..., same]), name
        err = (a.buf_state[:, same] - b.buf_state[:, same]).abs().max()
        assert float(err) < 2e-3, float(err)
        assert float((ra[:, same] - rb[:, same]).abs().max()) < 2e-3
    a.close(); b.close()
