"""closed-loop fused rollouts with the on-device DSL PID against the host loop the reference
runs (env oracle + oracle/pid_oracle.py, itself pinned to the reference's PIDController by
tests/golden/pid_controller.npz): same actions, same trajectories, auto-resets included."""
import numpy as np
import pytest
import torch

from oracle.pid_oracle import PIDOracle
from oracle.quad_oracle import PhiloxEnvDraws, QuadOracle

pytestmark = pytest.mark.gpu

U = lambda lo, hi: dict(distrib="uniform", low=lo, high=hi)
CFG3 = dict(quad_type=3, ctrl_freq=60, pyb_freq=240, episode_len_sec=1, cost="rl_reward", task="traj_tracking",
            task_info=dict(trajectory_type="figure8", num_cycles=1, trajectory_plane="xz", trajectory_position_offset=[0, 1.0],
                           trajectory_scale=0.4, proj_point=[0, 0, 1.0], proj_normal=[0, 1, 1]),
            init_state=dict(init_x=0, init_y=0, init_z=1.0), randomized_init=True,
            init_state_randomization_info={"init_x": U(-0.2, 0.2), "init_y": U(-0.2, 0.2), "init_z": U(-0.2, 0.2),
                                           "init_phi": U(-0.2, 0.2), "init_theta": U(-0.2, 0.2), "init_psi": U(-0.3, 0.3)},
            constraints=[dict(constraint_form="default_constraint", constrained_variable="input")],
            gates=[[4.0, 4.0, 0, 0, 0, 0, 0]],   # (the reference's goal-hold test cannot index a trajectory goal: keep a gate pending)
            done_on_violation=False, done_on_out_of_bound=True, seed=21)
CFG2 = dict(quad_type=2, ctrl_freq=50, pyb_freq=200, episode_len_sec=1, cost="rl_reward", task="stabilization",
            task_info=dict(stabilization_goal=[0.3, 1.2], stabilization_goal_tolerance=0.0),
            init_state=dict(init_x=0, init_z=1.0), randomized_init=True,
            init_state_randomization_info={"init_x": U(-0.3, 0.3), "init_z": U(-0.3, 0.3), "init_theta": U(-0.3, 0.3)},
            done_on_out_of_bound=True, seed=22)


# The oracle env is driven by the kernel's own actions, so every PID evaluation is compared on (almost) the same
# state: a free-running comparison only measures how fast the 70,000-gain attitude loop amplifies rounding
# (1e-16 -> 4e-6 within 100 steps on the 2D case).
@pytest.mark.parametrize("cfg", [CFG3, CFG2], ids=["3d_tracking", "2d_stabilization"])
@pytest.mark.parametrize("dtype,tol", [(torch.float64, 1e-9), (torch.float32, 2e-4)])
def test_closed_loop_rollout_matches_host_loop(cfg, dtype, tol):
    from safe_control_gym_b200.envs.quadrotor import BatchedQuadrotor
    n, base = 64, 300
    env = BatchedQuadrotor(num_envs=n, dtype=dtype, auto_reset=True, seed_offset=base, **cfg)
    k = 2 * env.CTRL_STEPS + 7            # crosses two truncations: the controller restarts with the env
    nu = env.action_dim
    acts = torch.empty(k, n, nu, dtype=dtype, device="cuda")
    rew = torch.empty(k, n, dtype=dtype, device="cuda")
    dn = torch.empty(k, n, dtype=torch.uint8, device="cuda")
    env.reset()
    env.rollout_pid(k, actions_out=acts, reward_out=rew, done_out=dn)
    final = env.state.double().cpu().numpy()
    acts, rew, dn = acts.double().cpu().numpy(), rew.double().cpu().numpy(), dn.cpu().numpy()
    kw = {a: b for a, b in cfg.items() if a != "seed"}
    dt = 1.0 / cfg["ctrl_freq"]
    worst = 0.0
    for i in (0, 13, 63):
        o = QuadOracle(draws=PhiloxEnvDraws(cfg["seed"], base + i), **kw)
        pid = PIDOracle()
        o.reset()
        for t in range(k):
            st, step = o.state, o.ctrl_step_counter
            if cfg["task"] == "traj_tracking":
                row = o.X_GOAL[min(step, o.X_GOAL.shape[0] - 1)]
                tp, tv = ([row[0], row[2], row[4]], [row[1], row[3], row[5]]) if cfg["quad_type"] == 3 else ([row[0], 0, row[2]], [row[1], 0, row[3]])
            else:
                g = o.X_GOAL
                tp, tv = ([g[0], g[2], g[4]] if cfg["quad_type"] == 3 else [g[0], 0, g[2]]), [0, 0, 0]
            a = pid.thrusts_3d(dt, st, tp, tv) if cfg["quad_type"] == 3 else pid.thrusts_2d(dt, st, tp, tv)
            err = float(np.max(np.abs(acts[t, i] - a) / np.maximum(1e-2, np.abs(a))))
            worst = max(worst, err)
            assert err <= tol, (str(dtype), t, i, acts[t, i], a)
            _, r, d, _ = o.step(acts[t, i])
            assert bool(dn[t, i]) == bool(d), (t, i)
            assert abs(rew[t, i] - r) <= tol * max(1.0, abs(r)), (t, i)
            if d:
                o.reset()
                pid.reset()
        assert np.max(np.abs(final[i] - o.state) / np.maximum(1.0, np.abs(o.state))) <= tol
    print(f"closed loop {cfg['quad_type']}D {dtype}: worst relative action error {worst:.2e} over {k} steps")
    env.close()
