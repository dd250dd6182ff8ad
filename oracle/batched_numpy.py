"""TEST INFRASTRUCTURE (oracle) -- CPU baseline C2 of SURVEY.md section 8(d): the Quadrotor 3D DYN step of
oracle/quad_oracle.py vectorised over N envs with numpy ([N] arrays per field), i.e. the best a "reference + numpy"
implementation could do on a CPU.  Restricted to BASELINE config 2 (level0: stabilisation, cost "competition", bounded
position + default input constraints, done on violation / completion, uniform init randomisation, auto-reset); the
arithmetic follows the same reference lines as QuadOracle and is checked against it in tests/test_oracle_batched.py.
Randomness is numpy's (a timing baseline: the draws are not the Philox streams of the product path).
"""
import numpy as np

from .quad_oracle import CF2X, GRAVITY_ACC

PI = np.pi
GIMBAL_SARG = 0.99999


def quat_from_euler(r, p, y):
    """bullet_math.quat_from_euler over arrays."""
    sph, cph, sth, cth, sps, cps = np.sin(r / 2), np.cos(r / 2), np.sin(p / 2), np.cos(p / 2), np.sin(y / 2), np.cos(y / 2)
    q = np.stack([sph * cth * cps - cph * sth * sps, cph * sth * cps + sph * cth * sps,
                  cph * cth * sps - sph * sth * cps, cph * cth * cps + sph * sth * sps])
    return q / np.sqrt((q * q).sum(0))


def euler_from_quat(q):
    """bullet_math.euler_from_quat over arrays (gimbal branch by mask)."""
    x, y, z, w = q
    sarg = -2.0 * (x * z - w * y)
    lo, hi = sarg <= -GIMBAL_SARG, sarg >= GIMBAL_SARG
    roll = np.arctan2(2.0 * (y * z + w * x), w * w - x * x - y * y + z * z)
    pitch = np.arcsin(np.clip(sarg, -1.0, 1.0))
    yaw = np.arctan2(2.0 * (x * y + w * z), w * w + x * x - y * y - z * z)
    roll = np.where(lo | hi, 0.0, roll)
    pitch = np.where(lo, -0.5 * PI, np.where(hi, 0.5 * PI, pitch))
    yaw = np.where(lo, 2.0 * np.arctan2(x, -y), np.where(hi, 2.0 * np.arctan2(-x, y), yaw))
    return roll, pitch, yaw


def third_column(q):
    """Third column of bullet_math.matrix_from_quat over arrays (the thrust direction)."""
    x, y, z, w = q
    s = 2.0 / (x * x + y * y + z * z + w * w)
    return (x * z + w * y) * s, (y * z - w * x) * s, 1.0 - (x * x + y * y) * s


class BatchedQuadOracle:
    """N lock-step envs of BASELINE config 2; `kw` = bench.env_kwargs()."""

    def __init__(self, n, kw, seed=0):
        self.n = n
        self.S = kw["pyb_freq"] // kw["ctrl_freq"]
        self.dt = 1.0 / kw["pyb_freq"]
        self.ctrl_freq = kw["ctrl_freq"]
        self.ctrl_steps = kw["episode_len_sec"] * kw["ctrl_freq"]
        self.m, self.kf, self.km, self.L = CF2X["M"], CF2X["KF"], CF2X["KM"], CF2X["L"]
        self.J = np.array([CF2X["IXX"], CF2X["IYY"], CF2X["IZZ"]])
        rmin = CF2X["PWM2RPM_SCALE"] * CF2X["MIN_PWM"] + CF2X["PWM2RPM_CONST"]
        rmax = CF2X["PWM2RPM_SCALE"] * CF2X["MAX_PWM"] + CF2X["PWM2RPM_CONST"]
        self.a_lo = np.float64(np.float32(self.kf * rmin ** 2))
        self.a_hi = np.float64(np.float32(self.kf * rmax ** 2))
        init = kw["init_state"]
        self.init = np.array([init[k] for k in ("init_x", "init_x_dot", "init_y", "init_y_dot", "init_z", "init_z_dot",
                                                "init_phi", "init_theta", "init_psi")], dtype=float)
        info = kw["init_state_randomization_info"]
        names = ("init_x", "init_x_dot", "init_y", "init_y_dot", "init_z", "init_z_dot", "init_phi", "init_theta", "init_psi")
        self.rand_lo = np.array([info.get(k, {}).get("low", 0.0) for k in names])
        self.rand_hi = np.array([info.get(k, {}).get("high", 0.0) for k in names])
        c = kw["constraints"][1]
        self.box_dims = np.array(c["active_dims"])
        self.box_lo, self.box_hi = np.array(c["lower_bounds"], dtype=float), np.array(c["upper_bounds"], dtype=float)
        self.goal = np.array(kw["task_info"]["stabilization_goal"], dtype=float)
        self.rng = np.random.default_rng(seed)
        self.pos = np.zeros((3, n)); self.vel = np.zeros((3, n)); self.rpy = np.zeros((3, n)); self.rates = np.zeros((3, n))
        self.quat = np.zeros((4, n)); self.t = np.zeros(n, dtype=np.int64); self.viol = np.zeros(n, dtype=bool)
        self.episodes = 0

    def reset(self, mask=None, init=None):
        """quadrotor.py:356-504 for the masked envs; `init` ([9][k]) injects the randomised initial state."""
        idx = np.arange(self.n) if mask is None else np.nonzero(mask)[0]
        if idx.size == 0:
            return
        if init is None:
            init = self.init[:, None] + self.rng.uniform(self.rand_lo[:, None], self.rand_hi[:, None], size=(9, idx.size))
        self.pos[:, idx] = init[[0, 2, 4]]; self.vel[:, idx] = init[[1, 3, 5]]
        q = quat_from_euler(init[6], init[7], init[8])
        self.quat[:, idx] = q
        r, p, y = euler_from_quat(q)
        self.rpy[:, idx] = np.stack([r, p, y]); self.rates[:, idx] = 0.0
        self.t[idx] = 0
        self.episodes += idx.size

    def state(self):
        """env.state with p, q, r left out: [x, x_dot, y, y_dot, z, z_dot, phi, theta, psi] x N."""
        return np.stack([self.pos[0], self.vel[0], self.pos[1], self.vel[1], self.pos[2], self.vel[2], *self.rpy])

    def step(self, action, auto_reset=True):
        """action [4][N] -> (reward [N], done [N], violation [N]); quadrotor.py:507-558 + benchmark_env.py:422-463."""
        thrust = np.clip(action, self.a_lo, self.a_hi)
        pwm = (np.sqrt(np.maximum(thrust, 0.0) / self.kf) - CF2X["PWM2RPM_CONST"]) / CF2X["PWM2RPM_SCALE"]
        rpm = CF2X["PWM2RPM_SCALE"] * np.clip(pwm, CF2X["MIN_PWM"], CF2X["MAX_PWM"]) + CF2X["PWM2RPM_CONST"]
        f = rpm ** 2 * self.kf
        zt = rpm ** 2 * self.km
        T = f.sum(0)
        arm = self.L / np.sqrt(2.0)
        tau = np.stack([(f[0] + f[1] - f[2] - f[3]) * arm, (-f[0] + f[1] + f[2] - f[3]) * arm, zt[0] - zt[1] + zt[2] - zt[3]])
        g = GRAVITY_ACC * self.m
        J = self.J[:, None]
        for _ in range(self.S):     # base_aviary.py:522-576
            cx, cy, cz = third_column(self.quat)
            force = np.stack([cx * T, cy * T, cz * T - g])
            w = self.rates
            jw = J * w
            cross = np.stack([w[1] * jw[2] - w[2] * jw[1], w[2] * jw[0] - w[0] * jw[2], w[0] * jw[1] - w[1] * jw[0]])
            self.vel = self.vel + self.dt * (force / self.m)
            self.rates = w + self.dt * ((tau - cross) / J)
            self.pos = self.pos + self.dt * self.vel
            pre = self.rpy + self.dt * self.rates
            self.quat = quat_from_euler(pre[0], pre[1], pre[2])
            self.rpy = np.stack(euler_from_quat(self.quat))
        self.t += 1
        # constraints (constraints.py:273-368): input bounds on the raw action, position box on the state
        c_in = np.maximum(self.a_lo - action, action - self.a_hi).max(0)
        p3 = self.pos[[0, 1, 2]] if list(self.box_dims) == [0, 2, 4] else None
        c_st = np.maximum(self.box_lo[:, None] - p3, p3 - self.box_hi[:, None]).max(0)
        violated = np.round(np.maximum(c_in, c_st), 8) > 0.0
        reward = np.where(self.viol, -100.0, 0.0)      # previous step's flag (F8); no gates / goal hold in DYN (F5)
        self.viol = violated
        done = violated | (self.t >= self.ctrl_steps)
        if auto_reset:
            self.reset(mask=done)
        return reward, done, violated
