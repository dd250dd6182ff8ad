"""Known-answer cases of SURVEY.md section 8(c) ("golden vectors the build must create itself"), written once against
the reference's scalar env signatures and run on both implementations: the numpy oracle (test_oracle_kat.py, CPU) and
the CUDA path with ``num_envs=1, squeeze=True`` through the C ABI (test_gpu_kat.py).  Every expected value is a closed
form of the cited reference lines, not an output of either implementation."""
import numpy as np

KF, M, L_ARM, G = 3.16e-10, 0.03454, 0.0397, 9.8               # assets/cf2x.urdf:5,12-13; base_aviary.py:90
IXX = IYY = 1.4e-5
IZZ = 2.17e-5
KM = 7.94e-12
SCALE, CONST, PWM_MIN, PWM_MAX = 0.2685, 4070.3, 20000.0, 65535.0   # base_aviary.py:96-99

REST = dict(init_x=0, init_x_dot=0, init_y=0, init_y_dot=0, init_z=1.0, init_z_dot=0, init_phi=0, init_theta=0, init_psi=0,
            init_p=0, init_q=0, init_r=0)


def motor_force(thrust):
    """quadrotor_utils.py:21-67 then KF rpm^2 (base_aviary.py:534)."""
    pwm = np.clip((np.sqrt(thrust / KF) - CONST) / SCALE, PWM_MIN, PWM_MAX)
    return KF * (SCALE * pwm + CONST) ** 2


def base(quad_type=3, freq=50, **over):
    goal = {1: [0, 1], 2: [0, 1], 3: [0, 0, 1]}[quad_type]
    kw = dict(quad_type=quad_type, ctrl_freq=freq, pyb_freq=freq, episode_len_sec=2, task="stabilization", cost="rl_reward",
              task_info=dict(stabilization_goal=goal, stabilization_goal_tolerance=0.0),
              init_state=dict(REST), randomized_init=False, done_on_out_of_bound=False, seed=0)
    kw.update(over)
    return kw


def kat_hover(make, tol):
    """(1) at rest, action = U_GOAL: acceleration = KF sum rpm^2 / m - g = 0 up to the PWM <-> thrust round trip, zero torque."""
    env = make(**base())
    env.reset()
    env.step(np.full(4, M * G / 4))
    s = env.state
    assert abs(s[5]) <= tol and abs(s[4] - 1.0) <= tol, s
    assert np.max(np.abs(np.delete(s, 4))) <= tol, s


def kat_free_fall(make, tol):
    """(2) action = action_space.low: every motor at (float32-rounded) MIN_PWM; z_dot after one step = dt (4 f_min / m - g) (base_aviary.py:552-563)."""
    env = make(**base())
    env.reset()
    low = env.action_space.low if hasattr(env, "action_space") else env.action_low     # quadrotor.py:699-733
    env.step(np.asarray(low, dtype=np.float64))
    dt = 1.0 / 50
    f_min = KF * (SCALE * PWM_MIN + CONST) ** 2
    f = motor_force(np.asarray(low, dtype=np.float64))       # the float32-rounded bound sits a hair above MIN_PWM's force
    assert np.all(np.abs(f - f_min) <= 1e-7 * f_min)
    acc = f.sum() / M - G
    s = env.state
    assert abs(s[5] - dt * acc) <= tol * abs(dt * acc) + tol, (s[5], dt * acc)
    assert abs(s[4] - (1.0 + dt * dt * acc)) <= tol, s[4]          # semi-implicit: pos += dt * new vel (base_aviary.py:562-563)


def kat_imbalance(make, tol):
    """(3) one stronger motor: torque signs and sizes of base_aviary.py:546-551 after one step from rest
    (rates = dt J^-1 tau; cross term zero at rest)."""
    env = make(**base())
    env.reset()
    hover = M * G / 4
    a = np.array([1.2 * hover, hover, hover, hover])
    env.step(a)
    f = motor_force(a)
    arm = L_ARM / np.sqrt(2.0)
    tx = (f[0] + f[1] - f[2] - f[3]) * arm
    ty = (-f[0] + f[1] + f[2] - f[3]) * arm
    tz = KM / KF * (f[0] - f[1] + f[2] - f[3])
    dt = 1.0 / 50
    want = dt * np.array([tx / IXX, ty / IYY, tz / IZZ])
    s = env.state
    # at small angles the body rates p, q, r of the observation equal the Euler-rate slots the integrator carries
    got_angles = np.array([s[6], s[7], s[8]])
    assert tx > 0 and ty < 0 and tz > 0
    assert np.max(np.abs(got_angles - dt * want) / np.abs(dt * want)) <= max(tol, 1e-12) * 10, (got_angles, dt * want)
    return got_angles


LEVEL0_CONSTRAINTS = [dict(constraint_form="default_constraint", constrained_variable="input"),
                      dict(constraint_form="bounded_constraint", constrained_variable="state", active_dims=[0, 2, 4],
                           upper_bounds=[3, 3, 2], lower_bounds=[-3, -3, -0.1])]


def kat_level0_constraints(make, tol):
    """(4) level0's initial state [-0.9, -2.9, 0.03]: state-constraint values [lb - x; x - ub]
    = [-2.1, -0.1, -0.13, -3.9, -5.9, -1.97], not violated (level0.yaml:39-43,136-155, constraints.py:273-304);
    a raw action of 0 is below the input lower bound => violated (constraints.py:307-368)."""
    init = dict(REST, init_x=-0.9, init_y=-2.9, init_z=0.03)
    env = make(**base(init_state=init, constraints=LEVEL0_CONSTRAINTS, info_in_reset=True, freq=60, pyb_freq=240))
    _, info = env.reset()
    want = np.array([-2.1, -0.1, -0.13, -3.9, -5.9, -1.97])
    assert np.max(np.abs(np.asarray(info["constraint_values"], dtype=float) - want)) <= max(tol, 1e-15) * 10
    _, _, _, info = env.step(np.full(4, M * G / 4))
    assert int(info["constraint_violation"]) == 0
    c = np.asarray(info["constraint_values"], dtype=float)
    assert c.shape == (14,) and np.all(c < 0)
    _, _, _, info = env.step(np.zeros(4))
    assert int(info["constraint_violation"]) == 1
    assert np.all(np.asarray(info["constraint_values"], dtype=float)[:4] > 0)      # lb - 0 > 0 on all four inputs


def kat_impulse_window(make, tol):
    """(5) ImpulseDisturbance(step_offset=2, duration=1, magnitude=10) on the observation: peak = int(2 + 1/2) = 2, non-zero
    only at ctrl step 2 (disturbances.py:89-143).  The env is at rest under hover thrust, so obs - state is the disturbance."""
    env = make(**base(quad_type=1, disturbances={"observation": [dict(disturbance_func="impulse", magnitude=10, step_offset=2,
                                                                      duration=1, decay_rate=1)]}))
    env.reset()
    seen = []
    for _ in range(5):
        obs, _, _, _ = env.step(np.array([M * G]))
        seen.append(float(np.asarray(obs)[0] - env.state[0]))
    hit = [abs(v) > 1.0 for v in seen]
    assert hit.count(True) == 1, seen
    k = hit.index(True)
    assert abs(seen[k] - 10.0) <= max(tol, 1e-12) * 100, seen
    return k


def kat_2d_mapping(make, tol):
    """(6) 2D: rpm = [p0, p1, p1, p0] (quadrotor.py:838-840) => no roll torque, pitch torque (f1 + f2 - f0 - f3) L / sqrt 2."""
    env = make(**base(quad_type=2))
    env.reset()
    hover = M * G / 2
    a = np.array([hover, 1.1 * hover])
    env.step(a)
    f = motor_force(a / 2)                                                      # per motor of each pair
    ty = 2 * (f[1] - f[0]) * L_ARM / np.sqrt(2.0)
    dt = 1.0 / 50
    s = env.state                                                               # [x, x_dot, z, z_dot, theta, theta_dot]
    assert abs(s[5] - dt * ty / IYY) <= max(tol, 1e-12) * abs(dt * ty / IYY) * 10, (s[5], dt * ty / IYY)
    assert abs(s[4] - dt * dt * ty / IYY) <= max(tol, 1e-12) * abs(dt * dt * ty / IYY) * 10


def kat_time_limit(make, tol):
    """(7) done and TimeLimit.truncated exactly at ctrl step CTRL_STEPS (quadrotor.py:997-1001, benchmark_env.py:457-461)."""
    env = make(**base(quad_type=1, freq=10, episode_len_sec=1))
    env.reset()
    flags = []
    for _ in range(10):
        _, _, done, info = env.step(np.array([M * G]))
        flags.append((bool(done), bool(info.get("TimeLimit.truncated", False))))
    assert flags[:9] == [(False, False)] * 9 and flags[9] == (True, True), flags


def kat_gimbal_snap(make, tol):
    """(8) theta -> pi/2 exactly: bullet's eulerFromQuat snaps roll to 0 and folds it into yaw (yaw' = psi - phi); the
    state read back after one step from (phi, pi/2, psi) at rest with hover thrust has roll = 0, pitch = pi/2."""
    init = dict(REST, init_phi=0.3, init_theta=np.pi / 2, init_psi=0.5)
    env = make(**base(init_state=init))
    env.reset()
    env.step(np.full(4, M * G / 4))
    s = env.state
    assert abs(s[6]) <= max(tol, 1e-15) * 10 and abs(s[7] - np.pi / 2) <= max(tol, 1e-7), s[6:9]
    assert abs(s[8] - (0.5 - 0.3)) <= max(tol * 100, 1e-7), s[8]


ALL = [kat_hover, kat_free_fall, kat_imbalance, kat_level0_constraints, kat_impulse_window, kat_2d_mapping,
       kat_time_limit, kat_gimbal_snap]
