"""Generate golden vectors by running the REFERENCE's own Quadrotor code (physics=dyn).

Run in the build container only:  ``python tests/golden/make_golden.py``
(needs /root/reference; the GPU box only reads the committed ``*.npz``).

The reference classes are imported from /root/reference unchanged; third-party modules
that are not installed are replaced by tests/golden/ref_stubs.py.  Each scenario is a
kwargs dict (the reference's YAML keys) + a seeded action sequence.  Recorded per event
(reset or step): obs, reward, done, constraint values / violation flag, mse, truncation,
``env.state`` and the raw kinematics; plus every RNG draw in call order, so that the
oracle and the CUDA kernels can be fed the identical numbers.
"""
import json
import os
import sys

import numpy as np
import yaml

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
import ref_stubs  # noqa: E402

REF = "/root/reference"


def level0(**over):
    cfg = yaml.safe_load(open(os.path.join(REF, "competition/level0.yaml")))["quadrotor_config"]
    cfg.pop("gui", None)
    cfg.pop("camera_view", None)
    cfg["physics"] = "dyn"
    cfg.update(over)
    return cfg


def hover_actions(rng, T, nu, mass=0.03454, std=0.005, lo=None, hi=None):
    a = mass * 9.8 / nu + rng.normal(0, std, size=(T, nu))
    if lo is not None:
        a = np.clip(a, lo, hi)
    return a


def scenarios():
    rng = np.random.default_rng(1337)
    out = {}
    # 1/2: BASELINE config 1 -- level0 + physics=dyn at 60/240 (S=4) and 500/500 (S=1).
    a = hover_actions(rng, 120, 4, lo=0.03, hi=0.145)
    a[40] = 0.0          # raw action below the input bound: violation -> done -> reset
    a[41:] += 0.004      # climb a little afterwards
    out["level0_dyn_s4"] = dict(cfg=level0(ctrl_freq=60, pyb_freq=240), actions=a)
    out["level0_dyn_s1"] = dict(cfg=level0(ctrl_freq=500, pyb_freq=500),
                                actions=hover_actions(rng, 80, 4, lo=0.03, hi=0.145))
    # 3: 3D tumbling -- short differential-thrust pulses spin the body up to ~15 rad/s in pitch (plus
    #    slower roll / yaw), then equal thrust: the attitude keeps running through the asin fold, the
    #    +-pi wraps and the gimbal band of getEulerFromQuaternion for 2.5 s.
    a = np.tile(np.array([0.085, 0.085, 0.085, 0.085]), (150, 1)) + rng.normal(0, 0.0005, size=(150, 4))
    a[0:6] = np.array([0.12, 0.05, 0.05, 0.12])     # pitch torque
    a[6:9] = np.array([0.11, 0.11, 0.06, 0.06])     # roll torque
    a[9:14] = np.array([0.13, 0.04, 0.13, 0.04])    # yaw torque
    out["quad3d_tumble"] = dict(cfg=dict(physics="dyn", quad_type=3, ctrl_freq=60, pyb_freq=240, seed=7,
                                         episode_len_sec=5, cost="rl_reward", task="stabilization",
                                         task_info=dict(stabilization_goal=[0, 0, 1], stabilization_goal_tolerance=0.05),
                                         randomized_init=False, done_on_out_of_bound=False,
                                         init_state=dict(init_z=1.0, init_phi=0.2, init_psi=3.0, init_q=0.5)),
                                actions=a)
    # 4: 2D, normalised RL action space, disturbances on action + observation, default
    #    constraints with tolerance + penalty, randomised init + inertia, out-of-bound done.
    cfg = dict(physics="dyn", quad_type=2, ctrl_freq=60, pyb_freq=240, seed=11, episode_len_sec=2,
               normalized_rl_action_space=True, cost="rl_reward", task="stabilization",
               task_info=dict(stabilization_goal=[0.3, 1.0], stabilization_goal_tolerance=0.05),
               init_state=dict(init_x=0.1, init_z=0.9, init_theta=0.05, init_theta_dot=0.2),
               randomized_init=True,
               init_state_randomization_info={k: dict(distrib="uniform", low=-0.05, high=0.05)
                                              for k in ["init_x", "init_z", "init_theta", "init_x_dot"]},
               randomized_inertial_prop=True,
               inertial_prop_randomization_info=dict(M=dict(distrib="uniform", low=-0.01, high=0.01),
                                                     Iyy=dict(distrib="uniform", low=-1e-6, high=1e-6)),
               constraints=[dict(constraint_form="default_constraint", constrained_variable="input",
                                 tolerance=[0.1] * 4),
                            dict(constraint_form="bounded_constraint", constrained_variable="state",
                                 active_dims=[2, 0], lower_bounds=[0.5, -0.25], upper_bounds=[1.3, 0.45],
                                 strict=True, tolerance=[0.05] * 4)],
               done_on_violation=True, use_constraint_penalty=True, constraint_penalty=-2.5,
               disturbances=dict(
                   observation=[dict(disturbance_func="white_noise", std=0.01)],
                   action=[dict(disturbance_func="impulse", magnitude=0.02, step_offset=5, duration=6, decay_rate=0.8),
                           dict(disturbance_func="step", magnitude=-0.01),
                           dict(disturbance_func="uniform", low=-0.002, high=0.004, mask=[1, 0])]),
               rew_state_weight=[1, 0.1, 1, 0.1, 0.5, 0.01], rew_act_weight=0.1, done_on_out_of_bound=True)
    out["quad2d_disturbed"] = dict(cfg=cfg, actions=rng.uniform(-1.3, 1.3, size=(260, 2)))
    # 5: 1D trajectory tracking with goal-horizon observations and periodic action noise;
    #    runs into the time limit (truncation) and wraps into a second episode.
    cfg = dict(physics="dyn", quad_type=1, ctrl_freq=50, pyb_freq=100, seed=3, episode_len_sec=1,
               cost="rl_reward", task="traj_tracking", obs_goal_horizon=2, rew_exponential=False,
               task_info=dict(trajectory_type="figure8", num_cycles=1, trajectory_plane="xz",
                              trajectory_position_offset=[0, 1.0], trajectory_scale=0.4),
               init_state=dict(init_x=0.2, init_x_dot=0.1), randomized_init=False,
               disturbances=dict(action=[dict(disturbance_func="periodic", scale=0.01, frequency=2.0)]),
               done_on_out_of_bound=False)
    out["quad1d_tracking"] = dict(cfg=cfg, actions=hover_actions(rng, 70, 1, std=0.05))
    # 6: 3D trajectory tracking (projected circle), dynamics-channel disturbance (consumes
    #    draws but never reaches the DYN integrator, F5) and an action adversary.
    cfg = dict(physics="dyn", quad_type=3, ctrl_freq=60, pyb_freq=240, seed=5, episode_len_sec=1,
               cost="rl_reward", task="traj_tracking", rew_state_weight=0.5, rew_act_weight=2.0,
               task_info=dict(trajectory_type="circle", num_cycles=1, trajectory_plane="xy",
                              trajectory_position_offset=[0.0, 0.0], trajectory_scale=0.5,
                              proj_point=[0, 0, 1.0], proj_normal=[0, 1, 2]),
               init_state=dict(init_x=0.5, init_z=1.0), randomized_init=True,
               # explicit table: the class-level default is pruned in place per quad type
               # (quadrotor.py:226-229), which would make this scenario order-dependent.
               init_state_randomization_info={k: dict(distrib="uniform", low=-w, high=w) for k, w in
                                              [("init_x", 0.05), ("init_y", 0.05), ("init_z", 0.05),
                                               ("init_phi", 0.1), ("init_r", 0.1)]},
               constraints=[dict(constraint_form="default_constraint", constrained_variable="state"),
                            dict(constraint_form="linear_constraint", constrained_variable="state",
                                 A=[[1, 0, 1, 0, 0, 0, 0, 0, 0, 0, 0, 0], [0, 0, 0, 0, -1, 0, 2, 0, 0, 0, 0, 0]],
                                 b=[0.9, -0.6]),
                            dict(constraint_form="quadratic_constraint", constrained_variable="input",
                                 P=np.diag([1.0, 2.0, 1.0, 0.5]).tolist(), b=0.035)],
               disturbances=dict(dynamics=[dict(disturbance_func="white_noise", std=0.1)]),
               adversary_disturbance="action", adversary_disturbance_scale=0.01,
               adversary_disturbance_offset=0.001, done_on_out_of_bound=True,
               # one far-away gate: with zero gates the reference's goal check indexes the
               # (T, 12) X_GOAL table and raises (quadrotor.py:1117-1121).
               gates=[[5.0, 5.0, 0, 0, 0, 0, 0]])
    out["quad3d_tracking"] = dict(cfg=cfg, actions=hover_actions(rng, 90, 4, std=0.01),
                                  adv=rng.uniform(-1.5, 1.5, size=(90, 4)))
    # ---- round 2.  Closed-loop scenarios: the actions come from the DSL PID (oracle/pid_oracle.py, pinned to the
    # reference's PIDController by pid_controller.npz) fed with the reference env's own observations, plus seeded
    # noise; what is recorded is the action sequence actually applied, so replays are open loop like the others.
    rng2 = np.random.default_rng(2024)
    # 7: cost "quadratic", stabilisation: reward = -SymbolicModel.loss (quadrotor.py:915-922) with a non-identity
    #    diagonal Q / R (set_cost_function_param), and the goal_reached termination (quadrotor.py:963-967): the PID
    #    brings the full 12-state within the tolerance ball, the episode ends, a second one follows (F10: the flag of
    #    info["goal_reached"] lags by one step and survives the reset).
    cfg = dict(physics="dyn", quad_type=3, ctrl_freq=60, pyb_freq=240, seed=17, episode_len_sec=6, cost="quadratic",
               task="stabilization", task_info=dict(stabilization_goal=[0.2, -0.1, 1.0], stabilization_goal_tolerance=0.12),
               init_state=dict(init_x=0.2, init_y=-0.1, init_z=1.3), randomized_init=True,
               init_state_randomization_info={k: dict(distrib="uniform", low=-w, high=w) for k, w in
                                              [("init_x", 0.05), ("init_y", 0.05), ("init_z", 0.05), ("init_theta", 0.02)]},
               constraints=[dict(constraint_form="default_constraint", constrained_variable="input"),
                            dict(constraint_form="bounded_constraint", constrained_variable="state", active_dims=[4, 5],
                                 lower_bounds=[0.5, -1.0], upper_bounds=[1.6, 1.0])],
               done_on_out_of_bound=True, info_in_reset=True)
    out["quad3d_quadratic_stab"] = dict(cfg=cfg, steps=400, stop_after_dones=2, tail=12,
                                        QR=(np.diag(0.5 * np.arange(1.0, 13.0)), np.diag([0.1, 0.2, 0.3, 0.4])),
                                        policy=pid_policy(rng2, 0.0001, lambda env: ([0.2, -0.1, 1.0], [0, 0, 0])))
    # 8: cost "quadratic", trajectory tracking: the reward reads X_GOAL[ctrl_step_counter] un-clamped
    #    (quadrotor.py:924-930) with dense (off-diagonal, symmetric) Q and R; runs into the time limit and on into a
    #    second episode.
    Qd = np.diag(np.linspace(2.0, 0.2, 12))
    Qd[0, 1] = Qd[1, 0] = 0.3; Qd[4, 5] = Qd[5, 4] = -0.2; Qd[2, 7] = Qd[7, 2] = 0.15; Qd[6, 9] = Qd[9, 6] = 0.05
    Rd = np.diag([1.0, 0.5, 2.0, 1.5]); Rd[0, 3] = Rd[3, 0] = 0.25; Rd[1, 2] = Rd[2, 1] = -0.4
    cfg = dict(physics="dyn", quad_type=3, ctrl_freq=60, pyb_freq=240, seed=23, episode_len_sec=1, cost="quadratic",
               task="traj_tracking", task_info=dict(trajectory_type="figure8", num_cycles=1, trajectory_plane="xz",
                                                    trajectory_position_offset=[0.0, 1.0], trajectory_scale=0.3,
                                                    proj_point=[0, 0, 1.0], proj_normal=[0, 1, 1]),
               init_state=dict(init_x=0.0, init_y=0.0, init_z=1.0), randomized_init=True,
               init_state_randomization_info={k: dict(distrib="uniform", low=-w, high=w) for k, w in
                                              [("init_x", 0.05), ("init_z", 0.05), ("init_phi", 0.05), ("init_psi", 0.1)]},
               done_on_out_of_bound=True, gates=[[5.0, 5.0, 0, 0, 0, 0, 0]])
    out["quad3d_quadratic_track"] = dict(cfg=cfg, steps=85, QR=(Qd, Rd),
                                         policy=pid_policy(rng2, 0.001, lambda env: track_target(env)))
    # 9: level0 + physics=dyn for 140 control steps WITHOUT a reset (north_star: "trajectory agreement reported over 100
    #    steps"): the PID flies from the start pad towards the first gate, every constraint stays inactive.
    out["level0_dyn_long"] = dict(cfg=level0(ctrl_freq=60, pyb_freq=240), steps=140, expect_dones=0,
                                  policy=pid_policy(rng2, 0.0005, lambda env: ([-0.6, -2.7, 0.8], [0, 0, 0])))
    # 10: competition/level2.yaml unchanged (+ physics=dyn at 60/240): randomized_gates_and_obstacles +
    #     randomized_inertial_prop + randomized init pin the ORDER of the reset draws -- obstacles, gates, inertial
    #     properties, initial state (quadrotor.py:371-488) -- over three episodes (raw actions below the input bound
    #     end the first two).  The poses the reference hands to loadURDF and EFFECTIVE_GATES_POSITIONS are recorded.
    cfg = yaml.safe_load(open(os.path.join(REF, "competition/level2.yaml")))["quadrotor_config"]
    cfg.pop("gui", None); cfg.pop("camera_view", None)
    cfg.update(physics="dyn", ctrl_freq=60, pyb_freq=240)
    a = hover_actions(rng2, 45, 4, std=0.002, lo=0.03, hi=0.145)
    a[14] = 0.0; a[29, 2] = 0.01
    out["level2_dyn"] = dict(cfg=cfg, actions=a)
    return out


def pid_policy(rng, noise_std, target_fn):
    """action(env, obs) = DSL PID thrusts towards target_fn(env) + N(0, noise_std^2), clipped to the action space."""
    from oracle.pid_oracle import PIDOracle
    pid = PIDOracle(m=0.03454)   # the env's mass (the class default is the CF2.x 0.027 kg)

    def policy(env, obs, fresh):
        if fresh:
            pid.reset()
        tp, tv = target_fn(env)
        a = pid.thrusts_3d(env.CTRL_TIMESTEP, np.asarray(obs)[:12], tp, tv) + rng.normal(0, noise_std, size=4)
        return np.clip(a, env.action_space.low.astype(np.float64), env.action_space.high.astype(np.float64))
    return policy


def track_target(env):
    row = env.X_GOAL[min(env.ctrl_step_counter, env.X_GOAL.shape[0] - 1)]
    return [row[0], row[2], row[4]], [row[1], row[3], row[5]]


def run(name, spec, Quadrotor):
    cfg = spec["cfg"]
    kwargs = {k: v for k, v in cfg.items()}
    ref_stubs.DRAW_LOG.clear()
    env = Quadrotor(**kwargs)
    if "QR" in spec:   # benchmark_env.py:237-254, before the first reset
        env.set_cost_function_param(np.array(spec["QR"][0], dtype=float), np.array(spec["QR"][1], dtype=float))
    ev = {k: [] for k in ["type", "obs", "rew", "done", "viol", "trunc", "mse", "cvals", "state", "kin", "quat", "ndraw", "goal"]}
    rs = {k: [] for k in ["gates", "obstacles", "inertia"]}   # per reset: what the draws turned into
    nc = env.num_constraints

    def record(kind, obs, rew, done, info):
        ev["type"].append(kind)
        ev["obs"].append(np.asarray(obs, dtype=np.float64))
        ev["rew"].append(float(rew))
        ev["done"].append(int(bool(done)))
        ev["viol"].append(int(info.get("constraint_violation", -1)))
        ev["trunc"].append(int(info.get("TimeLimit.truncated", -1)))
        ev["mse"].append(float(info.get("mse", np.nan)))
        ev["goal"].append(int(info["goal_reached"]) if "goal_reached" in info else -1)
        c = np.full(max(nc, 1), np.nan)
        if "constraint_values" in info:
            v = np.asarray(info["constraint_values"], dtype=np.float64)
            c[:len(v)] = v
        ev["cvals"].append(c)
        ev["state"].append(np.asarray(env.state, dtype=np.float64))
        ev["kin"].append(np.hstack([env.pos[0], env.vel[0], env.rpy[0], env.rpy_rates[0]]))
        ev["quat"].append(np.array(env.quat[0], dtype=np.float64, copy=True))   # the attitude PyBullet holds (x, y, z, w)
        ev["ndraw"].append(len(ref_stubs.DRAW_LOG))

    def do_reset():
        ret = env.reset()
        obs, info = ret if env.INFO_IN_RESET else (ret, {})
        if not env.INFO_IN_RESET and env.constraints is not None:
            info = {"constraint_values": env.constraints.get_values(env, only_state=True)}
        record(0, obs, 0.0, False, info)
        import pybullet as fake
        from oracle import bullet_math as bm
        rs["gates"].append(np.array(env.EFFECTIVE_GATES_POSITIONS, dtype=np.float64).reshape(len(env.GATES), 6))
        ob = []
        for i in env.OBSTACLES_IDS:   # pose handed to loadURDF (quadrotor.py:385-388): position + yaw of the quaternion
            b = fake.bodies[i]
            ob.append(list(b["pos"]) + [bm.euler_from_quat(b["quat"])[2]])
        rs["obstacles"].append(np.array(ob, dtype=np.float64).reshape(len(env.OBSTACLES), 4))
        rs["inertia"].append(np.array([env.OVERRIDDEN_QUAD_MASS] + list(env.OVERRIDDEN_QUAD_INERTIA), dtype=np.float64))
        return obs

    obs = do_reset()
    applied, dones, fresh, stop_at = [], 0, True, None
    steps = len(spec["actions"]) if "actions" in spec else spec["steps"]
    for t in range(steps):
        if stop_at is not None and t >= stop_at:
            break
        a = np.array(spec["actions"][t]) if "actions" in spec else spec["policy"](env, obs, fresh)
        fresh = False
        applied.append(np.array(a, dtype=np.float64))
        if "adv" in spec:
            env.set_adversary_control(spec["adv"][t])
        obs, rew, done, info = env.step(np.array(a))
        record(1, obs, rew, done, info)
        if done:  # DummyVecEnv.step_wait auto-reset (dummy_vec_env.py:38-49)
            obs = do_reset()
            dones += 1
            fresh = True
            if dones == spec.get("stop_after_dones", -1):
                stop_at = t + 1 + spec.get("tail", 0)
    if "expect_dones" in spec:
        assert dones == spec["expect_dones"], (name, dones)
    if "stop_after_dones" in spec:
        assert dones >= spec["stop_after_dones"], (name, dones)
    spec = dict(spec, actions=np.array(applied))
    draws = list(ref_stubs.DRAW_LOG)
    flat = np.concatenate([d[1].reshape(-1) for d in draws]) if draws else np.zeros(0)
    out = dict(config_json=np.array(json.dumps(cfg)), actions=np.asarray(spec["actions"], dtype=np.float64),
               draw_flat=flat, draw_len=np.array([d[1].size for d in draws], dtype=np.int64),
               draw_kind=np.array([d[0] for d in draws]), X_GOAL=np.asarray(env.X_GOAL, dtype=np.float64),
               U_GOAL=np.asarray(env.U_GOAL, dtype=np.float64),
               action_low=env.action_space.low, action_high=env.action_space.high,
               state_low=env.state_space.low, state_high=env.state_space.high)
    if "adv" in spec:
        out["adv"] = np.asarray(spec["adv"], dtype=np.float64)
    if "QR" in spec:
        out["Q"], out["R"] = np.array(spec["QR"][0], dtype=np.float64), np.array(spec["QR"][1], dtype=np.float64)
    for k, v in rs.items():
        out["reset_" + k] = np.array(v)
    for k, v in ev.items():
        out["ev_" + k] = np.array(v)
    np.savez_compressed(os.path.join(HERE, name + ".npz"), **out)
    print(f"{name}: {len(ev['type'])} events, {int(np.sum(ev['done']))} dones, {len(draws)} draws, "
          f"final state {np.round(ev['state'][-1][:6], 4)}")


if __name__ == "__main__":
    Quadrotor = ref_stubs.install(REF)
    only = sys.argv[1:]
    for name, spec in scenarios().items():
        if not only or name in only:
            run(name, spec, Quadrotor)
