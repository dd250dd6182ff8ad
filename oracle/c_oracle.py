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
