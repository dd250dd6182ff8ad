"""Row N4's optional part, the ``dyn_aero`` extension: the drag and ground-effect terms of the reference's PYB_* modes
(base_aviary.py:437-497) folded into the DYN integrator as one more external force / body torque per substep -- kernel against
oracle/quad_oracle.py (``_aero``: a restatement of those two functions; parity unpinned, PyBullet's link states are what the
reference evaluates them on) with drones that start low (ground effect active, clipped heights), move sideways (drag) and
restart (the lagged rotor-speed sum is zero after a reset)."""
import numpy as np
import pytest
import torch

from oracle.quad_oracle import PhiloxEnvDraws, QuadOracle

pytestmark = pytest.mark.gpu
U = lambda lo, hi: dict(distrib="uniform", low=lo, high=hi)
CFG = dict(quad_type=3, ctrl_freq=60, pyb_freq=240, episode_len_sec=0.5, cost="rl_reward", task="stabilization",
           task_info=dict(stabilization_goal=[0, 0, 0.3], stabilization_goal_tolerance=0.0),
           init_state=dict(init_x=0, init_y=0, init_z=0.06, init_x_dot=0.6, init_y_dot=-0.3), randomized_init=True,
           init_state_randomization_info={"init_z": U(-0.03, 0.2), "init_x_dot": U(-0.5, 0.5), "init_phi": U(-0.3, 0.3),
                                          "init_theta": U(-0.3, 0.3), "init_psi": U(-1.0, 1.0)},
           constraints=[dict(constraint_form="default_constraint", constrained_variable="input")],
           done_on_out_of_bound=True, seed=9)


@pytest.mark.parametrize("terms", [("drag",), ("ground_effect",), ("drag", "ground_effect")])
@pytest.mark.parametrize("dtype,attitude,tol", [(torch.float64, "fast", 1e-11), (torch.float64, "roundtrip", 1e-11), (torch.float32, "fast", 2e-4)])
def test_aero_terms_match_oracle(terms, dtype, attitude, tol):
    from safe_control_gym_b200.envs.quadrotor import BatchedQuadrotor
    n, steps, base = 48, 75, 40
    env = BatchedQuadrotor(num_envs=n, dtype=dtype, attitude=attitude, auto_reset=True, seed_offset=base, dyn_aero=terms, **CFG)
    plain = BatchedQuadrotor(num_envs=n, dtype=dtype, attitude=attitude, auto_reset=True, seed_offset=base, **CFG)
    assert int(env.lib.scg_kernel_variant(env._h)) in (0, 1)          # LEAN / GENERAL: the extension is not in the PLAIN loops
    kw = {k: v for k, v in CFG.items() if k != "seed"}
    lanes = [0, 5, 17, 47]
    refs = {i: QuadOracle(draws=PhiloxEnvDraws(CFG["seed"], base + i), dyn_aero=terms, **kw) for i in lanes}
    hover = env.MASS * 9.8 / 4
    rng = np.random.default_rng(3)
    acts = np.clip(hover * (1.0 + rng.normal(0, 0.05, size=(steps, n, 4))), 0, 2 * hover)
    obs = env.reset(); plain.reset()
    for i in lanes:
        refs[i].reset()
    worst, resets = 0.0, 0
    for t in range(steps):
        obs, rew, done, info = env.step(acts[t])
        pobs, *_ = plain.step(acts[t])
        term = info["terminal_observation"]
        for i in lanes:
            o_obs, o_rew, o_done, _ = refs[i].step(acts[t, i])
            assert bool(done[i]) == bool(o_done), (t, i)
            got = (term if o_done else obs)[i].double().cpu().numpy()
            err = float(np.max(np.abs(got - o_obs) / np.maximum(1.0, np.abs(o_obs))))
            worst = max(worst, err)
            assert err <= tol, (terms, str(dtype), attitude, t, i, err)
            if o_done:
                resets += 1
                refs[i].reset()
    assert resets >= 4                                                   # 30-step episodes: the lagged speed sum restarts at zero
    # the terms do something: the trajectories differ from the plain DYN ones
    assert float((env.buf_state - plain.buf_state).abs().max()) > 1e-3
    print(f"dyn_aero {terms} {dtype} {attitude}: worst rel err {worst:.2e}, {resets} resets")
    env.close(); plain.close()


def test_aero_fused_rollout_equals_stepping():
    """The rotor-speed sum of the previous substep travels through the aero_mem buffer between launches and ranges."""
    from safe_control_gym_b200.envs.quadrotor import BatchedQuadrotor
    n, k = 700, 90
    a = BatchedQuadrotor(num_envs=n, dtype=torch.float32, auto_reset=True, dyn_aero=("drag", "ground_effect"), **CFG)
    b = BatchedQuadrotor(num_envs=n, dtype=torch.float32, auto_reset=True, dyn_aero=("drag", "ground_effect"), **CFG)
    hover = a.MASS * 9.8 / 4
    acts = (hover * (1 + 0.05 * torch.randn(k, n, 4, generator=torch.Generator().manual_seed(1)))).cuda()
    a.reset(); b.reset()
    a.set_rollout_queue(1, 32)
    a.rollout(acts)
    for t in range(k):
        b.step(acts[t])
    torch.cuda.synchronize()
    for name in ("state", "ctrl_step", "episode", "aero"):
        assert torch.equal(getattr(a, "buf_" + name), getattr(b, "buf_" + name)), name
    a.close(); b.close()
