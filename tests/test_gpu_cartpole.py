"""Batched CartPole (kernel k_rollout<SCG_CARTPOLE>) against oracle/cartpole_oracle.py on the same Philox
streams (key = (seed, env id)): BASELINE config 3 in small -- white-noise action disturbance, default input +
state constraints, randomised init and inertial properties, done_on_violation with auto-reset.  The env is
absent from the reference fork (parity unpinned); this pins kernel == oracle."""
import numpy as np
import pytest
import torch

from oracle.cartpole_oracle import CartPoleOracle
from oracle.quad_oracle import PhiloxEnvDraws

pytestmark = pytest.mark.gpu

CFG = dict(ctrl_freq=50, pyb_freq=50, episode_len_sec=2, cost="quadratic", task="traj_tracking",
           task_info=dict(stabilization_goal=[0], stabilization_goal_tolerance=0.01, trajectory_type="circle", num_cycles=1,
                          trajectory_plane="zx", trajectory_position_offset=[0, 0], trajectory_scale=0.2),
           init_state=dict(init_x=0, init_x_dot=0, init_theta=0, init_theta_dot=0), randomized_init=True,
           init_state_randomization_info={k: dict(distrib="uniform", low=-0.05, high=0.05)
                                          for k in ["init_x", "init_x_dot", "init_theta", "init_theta_dot"]},
           inertial_prop=dict(pole_length=1.0, pole_mass=0.1, cart_mass=1.0), randomized_inertial_prop=True,
           inertial_prop_randomization_info=dict(pole_length=dict(distrib="uniform", low=-0.25, high=0.0),
                                                 cart_mass=dict(distrib="uniform", low=0.5, high=1.5),
                                                 pole_mass=dict(distrib="uniform", low=0.05, high=0.15)),
           constraints=[dict(constraint_form="default_constraint", constrained_variable="input", strict=False),
                        dict(constraint_form="default_constraint", constrained_variable="state")],
           disturbances=dict(action=[dict(disturbance_func="white_noise", std=0.1)]),
           done_on_violation=True, done_on_out_of_bound=True, info_in_reset=True, seed=1337)


@pytest.mark.parametrize("dtype,tol", [(torch.float64, 1e-11), (torch.float32, 2e-4)])
def test_cartpole_matches_oracle_free_running(dtype, tol):
    from safe_control_gym_b200.envs.cartpole import BatchedCartPole
    n, steps, base = 96, 130, 1000
    env = BatchedCartPole(num_envs=n, dtype=dtype, auto_reset=True, seed_offset=base, **CFG)
    rng = np.random.default_rng(5)
    acts = rng.normal(0, 4.0, size=(steps, n, 1))
    acts[17, 3, 0] = 25.0    # raw action beyond the input bound: violation -> done -> auto-reset
    lanes = [0, 3, 31, 32, 95]
    refs = {}
    for i in lanes:
        kw = {k: v for k, v in CFG.items() if k != "seed"}
        refs[i] = CartPoleOracle(draws=PhiloxEnvDraws(CFG["seed"], base + i), **kw)
    obs, info = env.reset()
    for i in lanes:
        o_obs, o_info = refs[i].reset()
        assert np.max(np.abs(obs[i].double().cpu().numpy() - o_obs)) <= tol
    worst, n_done = 0.0, 0
    for t in range(steps):
        obs, rew, done, info = env.step(acts[t])
        term = info["terminal_observation"]
        for i in lanes:
            o_obs, o_rew, o_done, o_info = refs[i].step(acts[t, i])
            assert bool(done[i]) == bool(o_done), (t, i)
            assert int(info["constraint_violation"][i]) == int(o_info["constraint_violation"]), (t, i)
            got_obs = (term if o_done else obs)[i].double().cpu().numpy()
            err = np.max(np.abs(got_obs - o_obs) / np.maximum(1.0, np.abs(o_obs)))
            err = max(err, abs(float(rew[i]) - o_rew) / max(1.0, abs(o_rew)))
            c = info["constraint_values"][i].double().cpu().numpy()
            big = np.abs(o_info["constraint_values"]) > 1e30
            err = max(err, float(np.max(np.abs(c[~big] - o_info["constraint_values"][~big]))))
            worst = max(worst, err)
            assert err <= tol, (str(dtype), t, i, err)
            if o_done:   # DummyVecEnv.step_wait: the returned obs is the first obs of the new episode
                n_done += 1
                r_obs, _ = refs[i].reset()
                assert np.max(np.abs(obs[i].double().cpu().numpy() - r_obs)) <= tol
    assert n_done >= 5
    print(f"cartpole {dtype}: worst rel err {worst:.2e} over {steps} steps x {len(lanes)} lanes, {n_done} auto-resets")
    env.close()


def test_cartpole_rollout_equals_stepping():
    """K fused steps == K host-driven steps (bit-identical state), and the statistics count every env-step."""
    from safe_control_gym_b200.envs.cartpole import BatchedCartPole
    n, k = 4096, 64
    acts = torch.randn(k, n, 1, device="cuda", generator=torch.Generator(device="cuda").manual_seed(1)) * 3.0
    a = BatchedCartPole(num_envs=n, dtype=torch.float32, **CFG)
    b = BatchedCartPole(num_envs=n, dtype=torch.float32, **CFG)
    a.reset(); b.reset()
    a.rollout(acts)
    for t in range(k):
        b.step(acts[t])
    torch.cuda.synchronize()
    assert torch.equal(a.buf_state, b.buf_state) and torch.equal(a.buf_episode, b.buf_episode)
    sa, sb = a.stats(), b.stats()
    assert sa["env_steps"] == sb["env_steps"] == n * k and sa["episodes"] == sb["episodes"] > 0
    a.close(); b.close()


def test_cartpole_queued_rollout_equals_static_rollout():
    """The ready-list queue on the CartPole kernel (LEAN with the general spare: randomised inertia, white noise per
    step): forced queue == static grid bit for bit, on a ragged batch, across two launches."""
    from safe_control_gym_b200.envs.cartpole import BatchedCartPole
    n, k = 3000, 150
    acts = torch.randn(k, n, 1, generator=torch.Generator().manual_seed(11)).mul_(3.0).cuda()
    a = BatchedCartPole(num_envs=n, dtype=torch.float32, auto_reset=True, **CFG)
    b = BatchedCartPole(num_envs=n, dtype=torch.float32, auto_reset=True, **CFG)
    a.set_rollout_queue(1, 48); b.set_rollout_queue(0)
    a.reset(); b.reset()
    for _ in range(2):
        a.rollout(acts); b.rollout(acts)
        torch.cuda.synchronize()
        assert a.last_rollout_queued and not b.last_rollout_queued
        for name in ("state", "ctrl_step", "episode", "flags", "obs", "ep_return", "ep_length", "inertia"):
            assert torch.equal(getattr(a, "buf_" + name), getattr(b, "buf_" + name)), name
    sa, sb = a.stats(), b.stats()
    assert sa["env_steps"] == 2 * n * k and sa["episodes"] == sb["episodes"] > 0 and sa["violation_steps"] == sb["violation_steps"]
    assert abs(sa["return"] - sb["return"]) <= 1e-9 * abs(sb["return"])
    a.close(); b.close()
