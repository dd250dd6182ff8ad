"""ctypes wrapper of oracle/scg_oracle.c (TEST INFRASTRUCTURE, see oracle/__init__.py).

``COracle3D`` takes a ``QuadOracle`` (so every constant comes from the pinned numpy
restatement) and runs the same 3D control step over [12][n] state arrays with OpenMP."""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB = os.path.join(HERE, "libscg_oracle.so")


class OracleCfg(C.Structure):
    _fields_ = [("substeps", C.c_int32), ("ctrl_steps", C.c_int32), ("ctrl_freq", C.c_int32),
                ("dt", C.c_double), ("mass", C.c_double), ("gravity_force", C.c_double),
                ("J", C.c_double * 3), ("J_inv", C.c_double * 3), ("kf", C.c_double), ("km", C.c_double), ("arm", C.c_double),
                ("pwm_scale", C.c_double), ("pwm_const", C.c_double), ("pwm_min", C.c_double), ("pwm_max", C.c_double),
                ("act_low", C.c_double * 4), ("act_high", C.c_double * 4),
                ("n_bounded", C.c_int32), ("var", C.c_int32 * 4), ("strict", C.c_int32 * 4),
                ("lo", (C.c_double * 12) * 4), ("hi", (C.c_double * 12) * 4),
                ("goal", C.c_double * 12), ("goal_tol", C.c_double),
                ("n_gates", C.c_int32), ("done_on_violation", C.c_int32), ("done_on_completion", C.c_int32),
                ("auto_reset", C.c_int32),
                ("init_state", C.c_double * 12), ("init_lo", C.c_double * 12), ("init_hi", C.c_double * 12),
                ("seed", C.c_uint32), ("env_id_base", C.c_int64)]


def load(build=True):
    if build and (not os.path.exists(LIB) or os.path.getmtime(LIB) < os.path.getmtime(os.path.join(HERE, "scg_oracle.c"))):
        subprocess.run(["make", "-s", "-C", HERE], check=True)
    lib = C.CDLL(LIB)
    p = C.c_void_p
    lib.oracle_quad3d_rollout.argtypes = [C.POINTER(OracleCfg), C.c_int64, C.c_int32, p, p, p, p, p, p, p, p, p, C.c_int32]
    lib.oracle_quad3d_reset.argtypes = [C.POINTER(OracleCfg), C.c_int64, p, p, p, p, p]
    return lib


class COracle3D:
    def __init__(self, oracle, n, seed=0, env_id_base=0, auto_reset=True, init_rand=None, threads=0):
        assert oracle.QUAD_TYPE == 3 and oracle.COST == "competition" and not oracle.NORMALIZED_RL_ACTION_SPACE
        assert not oracle.disturbances
        self.lib = load()
        c = OracleCfg()
        c.substeps, c.ctrl_steps, c.ctrl_freq = oracle.PYB_STEPS_PER_CTRL, int(oracle.CTRL_STEPS), int(oracle.CTRL_FREQ)
        c.dt, c.mass, c.gravity_force = oracle.PYB_TIMESTEP, oracle.MASS, oracle.GRAVITY
        for j in range(3):
            c.J[j], c.J_inv[j] = oracle.J[j, j], oracle.J_INV[j, j]
        c.kf, c.km, c.arm = oracle.KF, oracle.KM, oracle.L
        c.pwm_scale, c.pwm_const, c.pwm_min, c.pwm_max = oracle.PWM2RPM_SCALE, oracle.PWM2RPM_CONST, oracle.MIN_PWM, oracle.MAX_PWM
        for j in range(4):
            c.act_low[j], c.act_high[j] = float(oracle.action_low[j]), float(oracle.action_high[j])
        cons = oracle.constraints or []
        assert len(cons) <= 4
        c.n_bounded = len(cons)
        for k, con in enumerate(cons):
            assert con.form in ("bounded_constraint", "default_constraint")
            c.var[k], c.strict[k] = (0 if con.var == "state" else 1), int(con.strict)
            for d in range(12):
                c.lo[k][d], c.hi[k][d] = -np.inf, np.inf
            active = [int(np.argmax(row)) for row in con.F]
            for j, d in enumerate(active):
                c.lo[k][d], c.hi[k][d] = float(con.lower_bounds[j]), float(con.upper_bounds[j])
        for d in range(12):
            c.goal[d] = float(oracle.X_GOAL[d])
        c.goal_tol = float(oracle.TASK_INFO["stabilization_goal_tolerance"])
        c.n_gates, c.done_on_violation = len(oracle.GATES), int(oracle.DONE_ON_VIOLATION)
        c.done_on_completion, c.auto_reset = int(oracle.DONE_ON_COMPLETION), int(auto_reset)
        from .quad_oracle import STATE_LABELS
        for d, name in enumerate(STATE_LABELS[3]):
            c.init_state[d] = float(oracle.INIT[name])
            if init_rand and name in init_rand:
                c.init_lo[d], c.init_hi[d] = float(init_rand[name]["low"]), float(init_rand[name]["high"])
        c.seed, c.env_id_base = int(seed) & 0xFFFFFFFF, int(env_id_base)
        self.cfg, self.n, self.threads = c, int(n), int(threads)
        self.state = np.zeros((12, n))
        self.ctrl_step = np.zeros(n, dtype=np.int32)
        self.episode = np.zeros(n, dtype=np.int32)
        self.flags = np.zeros(n, dtype=np.uint32)
        self.obs = np.zeros((12, n))
        self.stats = np.zeros(16)

    @staticmethod
    def _p(a):
        return None if a is None else C.c_void_p(a.ctypes.data)

    def reset(self):
        self.lib.oracle_quad3d_reset(C.byref(self.cfg), self.n, self._p(self.state), self._p(self.ctrl_step),
                                     self._p(self.episode), self._p(self.flags), self._p(self.obs))
        return self.obs

    def rollout(self, actions, reward_out=None, done_out=None):
        a = np.ascontiguousarray(actions, dtype=np.float64)
        assert a.shape[1:] == (self.n, 4)
        self.lib.oracle_quad3d_rollout(C.byref(self.cfg), self.n, a.shape[0], self._p(a), self._p(self.state),
                                       self._p(self.ctrl_step), self._p(self.episode), self._p(self.flags), self._p(self.obs),
                                       self._p(reward_out), self._p(done_out), self._p(self.stats), self.threads)
        return self.obs

    def num_threads(self):
        return self.lib.oracle_num_threads() if self.threads <= 0 else self.threads


def _fill_base(c, oracle, nx, nu, labels, seed, env_id_base, auto_reset, init_rand):
    """The oracle_cfg part shared by the 3-D port and the extended port."""
    c.substeps, c.ctrl_steps, c.ctrl_freq = oracle.PYB_STEPS_PER_CTRL, int(oracle.CTRL_STEPS), int(oracle.CTRL_FREQ)
    c.dt, c.mass, c.gravity_force = oracle.PYB_TIMESTEP, oracle.MASS, oracle.GRAVITY
    for j in range(3):
        c.J[j], c.J_inv[j] = oracle.J[j, j], oracle.J_INV[j, j]
    c.kf, c.km, c.arm = oracle.KF, oracle.KM, oracle.L
    c.pwm_scale, c.pwm_const, c.pwm_min, c.pwm_max = oracle.PWM2RPM_SCALE, oracle.PWM2RPM_CONST, oracle.MIN_PWM, oracle.MAX_PWM
    for j in range(nu):
        c.act_low[j], c.act_high[j] = float(oracle.action_low[j]), float(oracle.action_high[j])
    cons = oracle.constraints or []
    assert len(cons) <= 4
    c.n_bounded = len(cons)
    for k, con in enumerate(cons):
        assert con.form in ("bounded_constraint", "default_constraint") and con.tolerance is None
        c.var[k], c.strict[k] = (0 if con.var == "state" else 1), int(con.strict)
        for d in range(12):
            c.lo[k][d], c.hi[k][d] = -np.inf, np.inf
        active = [int(np.argmax(row)) for row in con.F]
        for j, d in enumerate(active):
            c.lo[k][d], c.hi[k][d] = float(con.lower_bounds[j]), float(con.upper_bounds[j])
    for d in range(nx):
        c.goal[d] = float(oracle.X_GOAL[d])
    c.goal_tol = float(oracle.TASK_INFO["stabilization_goal_tolerance"])
    c.n_gates, c.done_on_violation = len(oracle.GATES), int(oracle.DONE_ON_VIOLATION)
    c.done_on_completion, c.auto_reset = int(oracle.DONE_ON_COMPLETION), int(auto_reset)
    for d, name in enumerate(labels):
        c.init_state[d] = float(oracle.INIT[name])
        if init_rand and name in init_rand:
            assert init_rand[name]["distrib"] == "uniform"
            c.init_lo[d], c.init_hi[d] = float(init_rand[name]["low"]), float(init_rand[name]["high"])
    c.seed, c.env_id_base = int(seed) & 0xFFFFFFFF, int(env_id_base)


class ExtCfg(C.Structure):
    _fields_ = [("base", OracleCfg), ("quad_type", C.c_int32),
                ("cost_rl", C.c_int32), ("rew_exponential", C.c_int32), ("done_on_out_of_bound", C.c_int32), ("done_on_collision", C.c_int32),
                ("rew_state_weight", C.c_double * 12), ("rew_act_weight", C.c_double * 4), ("u_goal", C.c_double * 4),
                ("mse_weight", C.c_double * 12), ("state_low", C.c_double * 12), ("state_high", C.c_double * 12),
                ("n_dist", C.c_int32), ("dist_kind", C.c_int32 * 4), ("dist_offset", C.c_int32 * 4), ("dist_slot", C.c_int32 * 4),
                ("max_step", C.c_int32),
                ("dist_magnitude", C.c_double * 4), ("dist_duration", C.c_double * 4), ("dist_decay", C.c_double * 4),
                ("dyn_uses_episode_inertia", C.c_int32),
                ("prop", C.c_double * 4), ("prop_lo", C.c_double * 4), ("prop_hi", C.c_double * 4), ("gravity_acc", C.c_double),
                ("gate_geometry", C.c_int32), ("n_obstacles", C.c_int32),
                ("gates", (C.c_double * 7) * 8), ("obstacles", (C.c_double * 6) * 8), ("pose_lo", C.c_double * 2), ("pose_hi", C.c_double * 2),
                ("pose_rand", C.c_int32 * 2)]


class COracleQuadExt:
    """BASELINE configs 4 / 5 (and anything in between) on the host cores: quad_type 2 or 3, stabilisation, cost rl_reward
    or competition, impulse / step action disturbances, per-episode inertia, gate geometry, trajectory log.  Constants come
    from a ``QuadOracle`` built with the same kwargs."""

    def __init__(self, oracle, n, seed=0, env_id_base=0, auto_reset=True, threads=0):
        from .quad_oracle import STATE_LABELS
        assert oracle.QUAD_TYPE in (2, 3) and oracle.TASK == "stabilization" and oracle.COST in ("rl_reward", "competition")
        assert not oracle.NORMALIZED_RL_ACTION_SPACE and oracle.adversary_disturbance is None
        self.lib = load()
        p = C.c_void_p
        self.lib.oracle_quad_ext_rollout.argtypes = [C.POINTER(ExtCfg), C.c_int64, C.c_int32, p, p, p, p, p, p, p, p, p, C.c_int32, p, C.c_int32]
        self.lib.oracle_quad_ext_reset.argtypes = [C.POINTER(ExtCfg), C.c_int64, p, p, p, p, p]
        qt = oracle.QUAD_TYPE
        self.nx, self.nu = (6, 2) if qt == 2 else (12, 4)
        c = ExtCfg()
        _fill_base(c.base, oracle, self.nx, self.nu, STATE_LABELS[qt], seed, env_id_base, auto_reset,
                   oracle.INIT_STATE_RAND_INFO if oracle.RANDOMIZED_INIT else None)
        c.quad_type, c.cost_rl, c.rew_exponential = qt, int(oracle.COST == "rl_reward"), int(bool(oracle.rew_exponential))
        c.done_on_out_of_bound, c.done_on_collision = int(bool(oracle.done_on_out_of_bound)), int(bool(oracle.DONE_ON_COLLISION))
        rsw, raw_ = np.broadcast_to(oracle.rew_state_weight, (self.nx,)), np.broadcast_to(oracle.rew_act_weight, (self.nu,))
        for d in range(self.nx):
            c.rew_state_weight[d], c.mse_weight[d] = float(rsw[d]), float(oracle.info_mse_metric_state_weight[d])
            c.state_low[d], c.state_high[d] = float(oracle.state_low[d]), float(oracle.state_high[d])
        for j in range(self.nu):
            c.rew_act_weight[j], c.u_goal[j] = float(raw_[j]), float(oracle.U_GOAL[j])
        slot = 0
        for mode, lst in oracle.disturbances.items():
            for d in lst:
                if mode == "action":
                    assert d.kind in ("impulse", "step") and d.mask is None and c.n_dist < 4
                    k = c.n_dist
                    c.dist_kind[k], c.dist_slot[k] = (0 if d.kind == "impulse" else 1), slot
                    c.dist_offset[k] = -1 if d.step_offset is None else int(d.step_offset)
                    c.dist_magnitude[k] = float(d.magnitude)
                    c.dist_duration[k], c.dist_decay[k] = float(getattr(d, "duration", 1)), float(getattr(d, "decay_rate", 1))
                    c.max_step = int(d.max_step)
                    c.n_dist += 1
                else:
                    raise AssertionError("the C port carries action-channel impulse / step disturbances only")
                slot += 1
        c.dyn_uses_episode_inertia, c.gravity_acc = int(bool(oracle.dyn_uses_episode_inertia)), oracle.GRAVITY_ACC
        nominal = [oracle.MASS, oracle.J[0, 0], oracle.J[1, 1], oracle.J[2, 2]]
        for j, key in enumerate(["M", "Ixx", "Iyy", "Izz"]):
            c.prop[j] = float(nominal[j])
            info = oracle.INERTIAL_PROP_RAND_INFO.get(key) if oracle.RANDOMIZED_INERTIAL_PROP else None
            if info:
                assert info["distrib"] == "uniform"
                c.prop_lo[j], c.prop_hi[j] = float(info["low"]), float(info["high"])
        c.gate_geometry, c.n_obstacles = int(bool(oracle.gate_geometry)), len(oracle.OBSTACLES)
        for g, gate in enumerate(oracle.GATES):
            for j in range(7):
                c.gates[g][j] = float(gate[j])
        for g, ob in enumerate(oracle.OBSTACLES):
            for j in range(6):
                c.obstacles[g][j] = float(ob[j])
        if oracle.RANDOMIZED_GATES_AND_OBS:
            for q, which in enumerate(("gates", "obstacles")):
                info = oracle.GATES_AND_OBS_RAND_INFO[which]
                assert info["distrib"] == "uniform"
                c.pose_rand[q], c.pose_lo[q], c.pose_hi[q] = 1, float(info["low"]), float(info["high"])
        self.cfg, self.n, self.threads = c, int(n), int(threads)
        self.state = np.zeros((self.nx, n))
        self.ctrl_step = np.zeros(n, dtype=np.int32)
        self.episode = np.zeros(n, dtype=np.int32)
        self.flags = np.zeros(n, dtype=np.uint32)
        self.obs = np.zeros((self.nx, n))
        self.stats = np.zeros(16)

    _p = staticmethod(COracle3D._p)

    def reset(self):
        self.lib.oracle_quad_ext_reset(C.byref(self.cfg), self.n, self._p(self.state), self._p(self.ctrl_step), self._p(self.episode),
                                       self._p(self.flags), self._p(self.obs))
        return self.obs

    def rollout(self, actions, reward_out=None, done_out=None, log_out=None, log_stride=1):
        a = np.ascontiguousarray(actions, dtype=np.float64)
        assert a.shape[1:] == (self.n, self.nu)
        self.lib.oracle_quad_ext_rollout(C.byref(self.cfg), self.n, a.shape[0], self._p(a), self._p(self.state), self._p(self.ctrl_step),
                                         self._p(self.episode), self._p(self.flags), self._p(self.obs), self._p(reward_out),
                                         self._p(done_out), self._p(log_out), int(log_stride), self._p(self.stats), self.threads)
        return self.obs

    def num_threads(self):
        return self.lib.oracle_num_threads() if self.threads <= 0 else self.threads


class CartPoleCfg(C.Structure):
    _fields_ = [("substeps", C.c_int32), ("ctrl_steps", C.c_int32), ("task_tracking", C.c_int32), ("n_goal_rows", C.c_int32),
                ("dt", C.c_double), ("gravity", C.c_double), ("act_low", C.c_double), ("act_high", C.c_double),
                ("state_low", C.c_double * 4), ("state_high", C.c_double * 4),
                ("x_threshold", C.c_double), ("theta_threshold", C.c_double), ("noise_std", C.c_double), ("goal_tol", C.c_double),
                ("done_on_violation", C.c_int32), ("done_on_out_of_bound", C.c_int32), ("auto_reset", C.c_int32),
                ("randomized_inertia", C.c_int32),
                ("init_state", C.c_double * 4), ("init_lo", C.c_double * 4), ("init_hi", C.c_double * 4),
                ("prop", C.c_double * 3), ("prop_lo", C.c_double * 3), ("prop_hi", C.c_double * 3),
                ("x_goal", C.c_void_p), ("seed", C.c_uint32), ("env_id_base", C.c_int64)]


class COracleCartPole:
    """BASELINE config 3 on the host cores (oracle/cartpole_oracle.py restated in C): cost quadratic with Q = R = I, one
    white-noise action disturbance, default input + state constraints."""

    def __init__(self, oracle, n, seed=0, env_id_base=0, auto_reset=True, threads=0):
        from .cartpole_oracle import LABELS
        assert oracle.COST == "quadratic" and not oracle.NORMALIZED_RL_ACTION_SPACE and oracle.PYB_STEPS_PER_CTRL >= 1
        assert np.array_equal(oracle.Q, np.eye(4)) and np.array_equal(oracle.R, np.eye(1))
        self.lib = load()
        p = C.c_void_p
        self.lib.oracle_cartpole_rollout.argtypes = [C.POINTER(CartPoleCfg), C.c_int64, C.c_int32, p, p, p, p, p, p, p, p, p, C.c_int32]
        self.lib.oracle_cartpole_reset.argtypes = [C.POINTER(CartPoleCfg), C.c_int64, p, p, p, p]
        c = CartPoleCfg()
        c.substeps, c.ctrl_steps, c.task_tracking = oracle.PYB_STEPS_PER_CTRL, int(oracle.CTRL_STEPS), int(oracle.TASK == "traj_tracking")
        self._goal = np.ascontiguousarray(np.atleast_2d(oracle.X_GOAL), dtype=np.float64)
        c.n_goal_rows, c.x_goal = self._goal.shape[0], self._goal.ctypes.data
        c.dt, c.gravity = oracle.PYB_TIMESTEP, oracle.GRAVITY_ACC
        c.act_low, c.act_high = float(oracle.action_low[0]), float(oracle.action_high[0])
        cons = oracle.constraints or []
        forms = sorted((k.form, k.var, bool(k.strict)) for k in cons)
        assert forms == [("default_constraint", "input", False), ("default_constraint", "state", False)], forms
        for d in range(4):
            c.state_low[d], c.state_high[d] = float(oracle.state_low[d]), float(oracle.state_high[d])
        c.x_threshold, c.theta_threshold = oracle.x_threshold, oracle.theta_threshold_radians
        dist = oracle.disturbances
        assert set(dist) <= {"action"} and len(dist.get("action", [])) <= 1
        if dist.get("action"):
            d = dist["action"][0]
            assert d.kind == "white_noise" and d.mask is None
            c.noise_std = float(d.std[0])
        c.goal_tol = float(oracle.TASK_INFO["stabilization_goal_tolerance"])
        c.done_on_violation, c.done_on_out_of_bound = int(bool(oracle.DONE_ON_VIOLATION)), int(bool(oracle.done_on_out_of_bound))
        c.auto_reset, c.randomized_inertia = int(auto_reset), int(bool(oracle.RANDOMIZED_INERTIAL_PROP))
        for d, name in enumerate(LABELS):
            c.init_state[d] = float(oracle.INIT[name])
            info = oracle.INIT_STATE_RAND_INFO.get(name) if oracle.RANDOMIZED_INIT else None
            if info:
                assert info["distrib"] == "uniform"
                c.init_lo[d], c.init_hi[d] = float(info["low"]), float(info["high"])
        nominal = [oracle.EFFECTIVE_POLE_LENGTH, oracle.POLE_MASS, oracle.CART_MASS]
        for j, key in enumerate(["pole_length", "pole_mass", "cart_mass"]):
            c.prop[j] = float(nominal[j])
            info = oracle.INERTIAL_PROP_RAND_INFO.get(key) if oracle.RANDOMIZED_INERTIAL_PROP else None
            if info:
                assert info["distrib"] == "uniform"
                c.prop_lo[j], c.prop_hi[j] = float(info["low"]), float(info["high"])
        c.seed, c.env_id_base = int(seed) & 0xFFFFFFFF, int(env_id_base)
        self.cfg, self.n, self.threads = c, int(n), int(threads)
        self.state, self.inertia = np.zeros((4, n)), np.zeros((3, n))
        self.ctrl_step, self.episode = np.zeros(n, dtype=np.int32), np.zeros(n, dtype=np.int32)
        self.flags = np.zeros(n, dtype=np.uint32)
        self.stats = np.zeros(16)

    _p = staticmethod(COracle3D._p)

    def reset(self):
        self.lib.oracle_cartpole_reset(C.byref(self.cfg), self.n, self._p(self.state), self._p(self.inertia), self._p(self.ctrl_step),
                                       self._p(self.episode))
        return self.state

    def rollout(self, actions, reward_out=None, done_out=None):
        a = np.ascontiguousarray(np.asarray(actions, dtype=np.float64).reshape(-1, self.n))
        self.lib.oracle_cartpole_rollout(C.byref(self.cfg), self.n, a.shape[0], self._p(a), self._p(self.state), self._p(self.inertia),
                                         self._p(self.ctrl_step), self._p(self.episode), self._p(self.flags), self._p(reward_out),
                                         self._p(done_out), self._p(self.stats), self.threads)
        return self.state

    def num_threads(self):
        return self.lib.oracle_num_threads() if self.threads <= 0 else self.threads
