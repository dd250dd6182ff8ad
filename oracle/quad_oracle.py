"""numpy fp64 restatement of the reference Quadrotor ``physics=dyn`` env step.

TEST INFRASTRUCTURE (see oracle/__init__.py) -- never imported by the product.

One env, reference execution style (scalar numpy, one call per ctrl step).
Every method cites the reference code it follows (paths under
/root/reference/safe_control_gym/).  Documented deviations:
``TIMESTEP := PYB_TIMESTEP`` in the integrator and ``np.round_ -> np.round``.

Randomness goes through a small provider object (``SequentialDraws`` wraps a
numpy ``RandomState`` and reproduces the reference draw ORDER; ``PhiloxEnvDraws``
implements the counter-based contract of oracle/philox.py; ``ReplayDraws``
feeds back recorded numbers).
"""
import copy
import math

import numpy as np

from . import bullet_math as bm
from . import philox as px

# envs/gym_pybullet_drones/assets/cf2x.urdf:5,12-13 (parsed at base_aviary.py:612-651)
CF2X = dict(M=0.03454, L=0.0397, THRUST2WEIGHT=2.25, IXX=1.4e-5, IYY=1.4e-5, IZZ=2.17e-5,
            KF=3.16e-10, KM=7.94e-12, PWM2RPM_SCALE=0.2685, PWM2RPM_CONST=4070.3,
            MIN_PWM=20000.0, MAX_PWM=65535.0, COLLISION_R=0.06, COLLISION_H=0.025)
GRAVITY_ACC = 9.8  # base_aviary.py:90
F32MAX = float(np.finfo(np.float32).max)
# cf2x.urdf <properties> (aerodynamic coefficients) and the propeller links' positions in the body frame
AERO = dict(GND_EFF_COEFF=11.36859, PROP_RADIUS=2.31348e-2, DRAG_COEFF=(9.1785e-7, 9.1785e-7, 10.311e-7),
            PROP_XY=((0.028, 0.028), (-0.028, 0.028), (-0.028, -0.028), (0.028, -0.028)))

STATE_LABELS = {
    1: ["init_x", "init_x_dot"],
    2: ["init_x", "init_x_dot", "init_z", "init_z_dot", "init_theta", "init_theta_dot"],
    3: ["init_x", "init_x_dot", "init_y", "init_y_dot", "init_z", "init_z_dot",
        "init_phi", "init_theta", "init_psi", "init_p", "init_q", "init_r"],
}  # quadrotor.py:207-212

# quadrotor.py:45-134 class-level default randomisation tables.
DEFAULT_INERTIAL_RAND = {
    "M": dict(distrib="uniform", low=0.022, high=0.032),
    "Ixx": dict(distrib="uniform", low=1.3e-5, high=1.5e-5),
    "Iyy": dict(distrib="uniform", low=1.3e-5, high=1.5e-5),
    "Izz": dict(distrib="uniform", low=2.07e-5, high=2.27e-5),
}
DEFAULT_INIT_RAND = {k: dict(distrib="uniform", low=lo, high=hi) for k, lo, hi in [
    ("init_x", -0.5, 0.5), ("init_x_dot", -0.01, 0.01), ("init_y", -0.5, 0.5),
    ("init_y_dot", -0.01, 0.01), ("init_z", 0.1, 1.5), ("init_z_dot", -0.01, 0.01),
    ("init_phi", -0.3, 0.3), ("init_theta", -0.3, 0.3), ("init_psi", -0.3, 0.3),
    ("init_p", -0.01, 0.01), ("init_theta_dot", -0.01, 0.01), ("init_q", -0.01, 0.01),
    ("init_r", -0.01, 0.01)]}
DEFAULT_TASK_INFO = dict(stabilization_goal=[0, 1], stabilization_goal_tolerance=0.05,
                         trajectory_type="circle", num_cycles=1, trajectory_plane="zx",
                         trajectory_position_offset=[0.5, 0], trajectory_scale=-0.5,
                         proj_point=[0, 0, 0.5], proj_normal=[0, 1, 1])  # quadrotor.py:136-146


# --------------------------------------------------------------------------- draws
class SequentialDraws:
    """Reference semantics: one numpy RandomState consumed in call order."""

    def __init__(self, rng):
        self.rng = rng

    def distrib(self, name, args, kwargs, **_ctx):
        return getattr(self.rng, name)(*args, **kwargs)

    def randint(self, n, **_ctx):
        return int(self.rng.randint(n))

    def uniform_vec(self, low, high, dim, **_ctx):
        return self.rng.uniform(low, high, size=dim)

    def normal_vec(self, std, dim, **_ctx):
        return self.rng.normal(0, std, size=dim)


class ReplayDraws:
    """Feeds back a recorded list of draws, in order (golden-vector replays)."""

    def __init__(self, values):
        self.values = list(values)
        self.pos = 0

    def _next(self):
        v = self.values[self.pos]
        self.pos += 1
        return v

    def distrib(self, name, args, kwargs, **_ctx):
        return float(np.asarray(self._next()).reshape(-1)[0])

    def randint(self, n, **_ctx):
        return int(np.asarray(self._next()).reshape(-1)[0])

    def uniform_vec(self, low, high, dim, **_ctx):
        return np.asarray(self._next(), dtype=np.float64).reshape(dim)

    def normal_vec(self, std, dim, **_ctx):
        return np.asarray(self._next(), dtype=np.float64).reshape(dim)


class PhiloxEnvDraws:
    """Counter-based contract (oracle/philox.py) for env ``env_id``."""

    def __init__(self, seed, env_id):
        self.p = px.PhiloxDraws(seed, env_id)

    def distrib(self, name, args, kwargs, stream, slot, episode):
        u = self.p.uniforms(slot + 1, 0, episode, stream)[slot]
        if name == "uniform":
            low = kwargs.get("low", args[0] if len(args) > 0 else 0.0)
            high = kwargs.get("high", args[1] if len(args) > 1 else 1.0)
            return low + (high - low) * u
        if name == "normal":
            loc = kwargs.get("loc", args[0] if len(args) > 0 else 0.0)
            scale = kwargs.get("scale", args[1] if len(args) > 1 else 1.0)
            return loc + scale * self.p.normals(slot + 1, 0, episode, stream)[slot]
        raise NotImplementedError(name)

    def randint(self, n, slot, episode):
        return self.p.randint(n, slot, episode, px.STREAM_RESET_OFFSETS)

    def uniform_vec(self, low, high, dim, channel, k, index, episode):
        u = self.p.uniforms(dim, index, episode, px.STREAM_DISTURB_BASE + 8 * channel + k)
        return np.asarray(low) + (np.asarray(high) - np.asarray(low)) * u

    def normal_vec(self, std, dim, channel, k, index, episode):
        return np.asarray(std) * self.p.normals(dim, index, episode,
                                                px.STREAM_DISTURB_BASE + 8 * channel + k)


# --------------------------------------------------------------------------- disturbances
CHANNELS = {"observation": px.CHANNEL_OBSERVATION, "action": px.CHANNEL_ACTION,
            "dynamics": px.CHANNEL_DYNAMICS}


class _Disturbance:
    """envs/disturbances.py restated as data + two functions (reset / apply)."""

    def __init__(self, env, channel, k, dim, spec):
        self.kind = spec["disturbance_func"]
        assert self.kind in ("impulse", "step", "uniform", "white_noise", "periodic"), \
            "[ERROR] in BenchmarkEnv._setup_disturbances(), disturbance type not available."
        self.channel, self.k, self.dim = channel, k, dim
        mask = spec.get("mask")
        self.mask = None if mask is None else np.asarray(mask)
        if self.mask is not None:
            assert dim == len(self.mask)
        self.max_step = int(env.EPISODE_LEN_SEC / env.CTRL_TIMESTEP)  # disturbances.py:111,162
        if self.kind == "impulse":  # disturbances.py:99-119
            self.magnitude = spec.get("magnitude", 1)
            self.step_offset = spec.get("step_offset")
            self.duration = spec.get("duration", 1)
            self.decay_rate = spec.get("decay_rate", 1)
            assert self.duration >= 1 and 0 < self.decay_rate <= 1
        elif self.kind == "step":  # :153-162
            self.magnitude = spec.get("magnitude", 1)
            self.step_offset = spec.get("step_offset")
        elif self.kind == "uniform":  # :189-204
            low, high = spec.get("low", 0.0), spec.get("high", 1.0)
            self.low = np.asarray([low] * dim) if isinstance(low, float) else np.asarray(low)
            self.high = np.asarray([high] * dim) if isinstance(high, float) else np.asarray(high)
        elif self.kind == "white_noise":  # :221-235
            std = spec.get("std", 1.0)
            self.std = np.asarray([std] * dim) if isinstance(std, float) else np.asarray(std)
            assert dim == len(self.std), "std shape should be the same as dim."
        elif self.kind == "periodic":  # :265-276 (mask is dropped by the reference ctor)
            self.scale = spec.get("scale", 1.0)
            self.frequency = spec.get("frequency", 1.0)
            self.mask = None
        self.offset = 0
        self.peak = 0

    def reset(self, env, slot):
        if self.kind in ("impulse", "step"):
            if self.step_offset is None:  # disturbances.py:122-126,169-172
                self.offset = env.draws.randint(self.max_step, slot=slot, episode=env.episode)
            else:
                self.offset = self.step_offset
            if self.kind == "impulse":
                self.peak = int(self.offset + self.duration / 2)

    def apply(self, target, env, index):
        t = env.ctrl_step_counter
        ctx = dict(channel=self.channel, k=self.k, index=index, episode=env.episode)
        if self.kind == "impulse":  # disturbances.py:128-143
            noise = 0
            if t >= self.offset:
                k = abs(t - self.peak)
                decay = self.decay_rate ** k if k < self.duration / 2 else 0
                noise = self.magnitude * decay
        elif self.kind == "step":  # :174-184
            noise = self.magnitude if t >= self.offset else 0
        elif self.kind == "uniform":  # :206-212
            noise = env.draws.uniform_vec(self.low, self.high, self.dim, **ctx)
        elif self.kind == "white_noise":  # :237-248
            noise = env.draws.normal_vec(self.std, self.dim, **ctx)
        else:  # periodic :278-288
            phase = env.draws.uniform_vec(-np.pi, np.pi, self.dim, **ctx)
            tt = env.pyb_step_counter * env.PYB_TIMESTEP
            noise = self.scale * np.sin(2 * np.pi * self.frequency * tt + phase)
        if self.mask is not None:
            noise = noise * self.mask
        return target + noise


# --------------------------------------------------------------------------- constraints
class _Constraint:
    """envs/constraints.py restated: one object per YAML entry, value = rounded g(x)."""

    def __init__(self, env, spec):
        form = spec["constraint_form"]
        assert form in ("linear_constraint", "quadratic_constraint", "bounded_constraint",
                        "default_constraint"), "[ERROR]. constraint not in list of available constraints"
        self.form = form
        self.var = spec["constrained_variable"]
        assert self.var in ("state", "input")
        full_dim = env.state_dim if self.var == "state" else env.action_dim
        self.strict = spec.get("strict", False)
        self.rounding = 8  # constraints.py:41
        active = spec.get("active_dims") if form != "default_constraint" else None
        if active is not None:
            if isinstance(active, int):
                active = [active]
            self.F = np.eye(full_dim)[active]  # constraints.py:77
        else:
            self.F = np.eye(full_dim)
        dim = self.F.shape[0]
        tol = spec.get("tolerance")
        self.tolerance = None if tol is None else np.array(tol, ndmin=1)
        if form == "quadratic_constraint":  # constraints.py:204-213
            self.P = np.array(spec["P"], ndmin=1)
            assert self.P.shape == (dim, dim)
            self.b = spec["b"]
            self.num_constraints = 1
        else:
            if form == "linear_constraint":  # :252-260
                self.A = np.array(spec["A"], ndmin=1)
                self.b = np.array(spec["b"], ndmin=1)
            else:
                if form == "default_constraint":  # :335-368 (float32 Box bounds promoted)
                    space_low, space_high = ((env.state_low, env.state_high) if self.var == "state"
                                             else (env.action_low, env.action_high))
                    ub = spec.get("upper_bounds")
                    lb = spec.get("lower_bounds")
                    ub = space_high if ub is None else np.array(ub, ndmin=1)
                    lb = space_low if lb is None else np.array(lb, ndmin=1)
                    lb, ub = lb.astype(np.float64), ub.astype(np.float64)
                else:  # bounded :295-303
                    lb = np.array(spec["lower_bounds"], ndmin=1)
                    ub = np.array(spec["upper_bounds"], ndmin=1)
                self.lower_bounds, self.upper_bounds = lb, ub
                n = lb.shape[0]
                self.A = np.vstack((-np.eye(n), np.eye(n)))
                self.b = np.hstack((-lb, ub))
            assert self.A.shape[1] == dim, "[ERROR] A has the wrong dimension!"
            self.num_constraints = self.A.shape[0]
        if self.tolerance is not None and len(self.tolerance) != self.num_constraints:
            raise ValueError("[ERROR] the tolerance dimension does not match the number of constraints.")

    def value(self, env):
        x = np.array(env.state if self.var == "state" else env.current_raw_input_action, ndmin=1)
        if self.form == "quadratic_constraint":
            c = x.T @ self.F.T @ self.P @ self.F @ x - self.b
        else:
            c = self.A @ self.F @ x - self.b
        return np.round(np.atleast_1d(np.squeeze(c)), decimals=self.rounding)  # constraints.py:119

    def violated(self, c):  # constraints.py:137-141
        return bool(np.any(c >= 0.0) if self.strict else np.any(c > 0.0))

    def almost_active(self, c):  # :151-158
        if self.tolerance is None:
            return False
        return bool(np.any(c + self.tolerance > 0.0))


# --------------------------------------------------------------------------- trajectories
def generate_trajectory(traj_type, traj_length, num_cycles, traj_plane, position_offset,
                        scaling, sample_time):
    """benchmark_env.py:465-519 (+ _figure8/_circle/_square :568-674)."""
    if traj_type not in ("circle", "square", "figure8"):
        raise ValueError("Trajectory type should be one of [circle, square, figure8].")
    period = traj_length / num_cycles
    axes = ["x", "y", "z"]
    if not (traj_plane[0] in axes and traj_plane[1] in axes and traj_plane[0] != traj_plane[1]):
        raise ValueError("Trajectory plane should be in form of ab, where a and b can be {x, y, z}.")
    ia, ib = axes.index(traj_plane[0]), axes.index(traj_plane[1])
    times = np.arange(0, traj_length, sample_time)
    pos = np.zeros((len(times), 3))
    vel = np.zeros((len(times), 3))
    w = 2.0 * np.pi / period
    for i, t in enumerate(times):
        if traj_type == "figure8":
            a, b = scaling * np.sin(w * t), scaling * np.sin(w * t) * np.cos(w * t)
            ad, bd = scaling * w * np.cos(w * t), scaling * w * (np.cos(w * t) ** 2 - np.sin(w * t) ** 2)
        elif traj_type == "circle":
            a, b = scaling * np.cos(w * t), scaling * np.sin(w * t)
            ad, bd = -scaling * w * np.sin(w * t), scaling * w * np.cos(w * t)
        else:
            seg_period = period / 4.0
            speed = scaling / seg_period
            cyc = t % period
            seg_t = cyc % seg_period
            seg = int(np.floor(cyc / seg_period))
            s = speed * seg_t
            a, b, ad, bd = [(0.0, s, 0.0, speed), (-s, scaling, -speed, 0.0),
                            (-scaling, scaling - s, 0.0, -speed), (-scaling + s, 0.0, speed, 0.0)][seg]
        pos[i, ia], vel[i, ia] = a + position_offset[0], ad
        pos[i, ib], vel[i, ib] = b + position_offset[1], bd
    return pos, vel


def transform_trajectory(pos, vel, point, normal):
    """transformations.py:52-123: orthogonal projection onto the plane (point, normal)."""
    n = np.array(normal[:3], dtype=np.float64)
    n = n / math.sqrt(float(np.sum(n * n)))
    M = np.identity(4)
    M[:3, :3] -= np.outer(n, n)
    M[:3, 3] = np.dot(np.array(point[:3], dtype=np.float64), n) * n
    ap = np.concatenate([pos, np.ones((pos.shape[0], 1))], -1)
    av = np.concatenate([vel, np.ones((vel.shape[0], 1))], -1)
    return np.matmul(ap, M.T)[:, :3], np.matmul(av, M.T)[:, :3]


# --------------------------------------------------------------------------- the env
class QuadOracle:
    """Single-env restatement of Quadrotor(BaseAviary(BenchmarkEnv)) with physics=dyn.

    Constructor kwargs are the reference's YAML keys (quadrotor.yaml:1-40,
    competition/level0.yaml:16-182).  ``draws`` is the randomness provider.
    ``dyn_uses_episode_inertia=True`` switches on the extension where the per-episode
    randomised mass / inertia reach the integrator; the reference
    behaviour (and the default) is False.
    """

    NAME = "quadrotor"

    def __init__(self, draws=None, seed=None, info_in_reset=False, normalized_rl_action_space=False,
                 task="stabilization", task_info=None, cost="rl_reward", pyb_freq=50, ctrl_freq=50,
                 episode_len_sec=5, init_state=None, randomized_init=True,
                 init_state_randomization_info=None, inertial_prop=None,
                 randomized_inertial_prop=False, inertial_prop_randomization_info=None,
                 constraints=None, done_on_violation=False, use_constraint_penalty=False,
                 constraint_penalty=-1, disturbances=None, adversary_disturbance=None,
                 adversary_disturbance_offset=0.0, adversary_disturbance_scale=0.01,
                 quad_type=2, norm_act_scale=0.1, obs_goal_horizon=0, rew_state_weight=1.0,
                 rew_act_weight=0.0001, rew_exponential=True, done_on_out_of_bound=True,
                 info_mse_metric_state_weight=None, gates=None, obstacles=None,
                 randomized_gates_and_obstacles=False, gates_and_obstacles_randomization_info=None,
                 done_on_collision=False, done_on_completion=False, reseed_on_reset=False,
                 dyn_uses_episode_inertia=False, gate_geometry=False, dyn_aero=(), **_ignored):
        # extension: the drag / ground-effect terms of the PYB_* modes folded into DYN (see _aero)
        self.dyn_aero = tuple(dyn_aero)
        # ---- Quadrotor.__init__ pre-super part (quadrotor.py:176-201)
        self.QUAD_TYPE = int(quad_type)
        self.norm_act_scale = norm_act_scale
        self.obs_goal_horizon = obs_goal_horizon
        self.rew_state_weight = np.array(rew_state_weight, ndmin=1, dtype=float)
        self.rew_act_weight = np.array(rew_act_weight, ndmin=1, dtype=float)
        self.rew_exponential = rew_exponential
        self.done_on_out_of_bound = done_on_out_of_bound
        nx = {1: 2, 2: 6, 3: 12}[self.QUAD_TYPE]
        if info_mse_metric_state_weight is None:
            self.info_mse_metric_state_weight = np.array(
                {1: [1, 0], 2: [1, 0, 1, 0, 0, 0], 3: [1, 0, 1, 0, 1, 0, 0, 0, 0, 0, 0, 0]}[self.QUAD_TYPE],
                dtype=float)
        else:
            if len(info_mse_metric_state_weight) != nx:
                raise ValueError("[ERROR] in Quadrotor.__init__(), wrong info_mse_metric_state_weight argument size.")
            self.info_mse_metric_state_weight = np.array(info_mse_metric_state_weight, ndmin=1, dtype=float)
        # ---- BaseAviary.__init__ constants (base_aviary.py:90-147)
        self.GRAVITY_ACC = GRAVITY_ACC
        self.MASS, self.L = CF2X["M"], CF2X["L"]
        self.J = np.diag([CF2X["IXX"], CF2X["IYY"], CF2X["IZZ"]])
        self.J_INV = np.linalg.inv(self.J)
        self.KF, self.KM = CF2X["KF"], CF2X["KM"]
        self.PWM2RPM_SCALE, self.PWM2RPM_CONST = CF2X["PWM2RPM_SCALE"], CF2X["PWM2RPM_CONST"]
        self.MIN_PWM, self.MAX_PWM = CF2X["MIN_PWM"], CF2X["MAX_PWM"]
        self.GRAVITY = self.GRAVITY_ACC * self.MASS  # frozen before any override (F9)
        # ---- BenchmarkEnv.__init__ (benchmark_env.py:139-218)
        self.TASK = task
        self.TASK_INFO = copy.deepcopy(DEFAULT_TASK_INFO if task_info is None else task_info)
        self.CTRL_FREQ, self.PYB_FREQ = ctrl_freq, pyb_freq
        if self.PYB_FREQ % self.CTRL_FREQ != 0:
            raise ValueError("[ERROR] in BenchmarkEnv.__init__(), pyb_freq is not divisible by env_freq.")
        self.PYB_STEPS_PER_CTRL = int(self.PYB_FREQ / self.CTRL_FREQ)
        self.CTRL_TIMESTEP = 1.0 / self.CTRL_FREQ
        self.PYB_TIMESTEP = 1.0 / self.PYB_FREQ
        self.EPISODE_LEN_SEC = episode_len_sec
        self.CTRL_STEPS = self.EPISODE_LEN_SEC * self.CTRL_FREQ
        self.RANDOMIZED_INIT = randomized_init
        self.INIT_STATE_RAND_INFO = copy.deepcopy(
            DEFAULT_INIT_RAND if init_state_randomization_info is None else init_state_randomization_info)
        self.RANDOMIZED_INERTIAL_PROP = randomized_inertial_prop
        self.INERTIAL_PROP_RAND_INFO = copy.deepcopy(
            DEFAULT_INERTIAL_RAND if inertial_prop_randomization_info is None else inertial_prop_randomization_info)
        self.NORMALIZED_RL_ACTION_SPACE = normalized_rl_action_space
        self.COST = cost
        assert cost in ("rl_reward", "quadratic", "competition")
        self._set_action_space()
        self._set_observation_space()
        self.action_dim, self.state_dim, self.obs_dim = len(self.action_low), nx, len(self.obs_low)
        if self.COST in ("quadratic", "competition"):
            self.Q, self.R = np.eye(self.obs_dim), np.eye(self.action_dim)
        self.DONE_ON_VIOLATION = done_on_violation
        self.use_constraint_penalty = use_constraint_penalty
        self.constraint_penalty = constraint_penalty
        self.constraints = None if constraints is None else [_Constraint(self, c) for c in constraints]
        self.num_constraints = 0 if constraints is None else sum(c.num_constraints for c in self.constraints)
        self.adversary_disturbance = adversary_disturbance
        self.adversary_disturbance_offset = adversary_disturbance_offset
        self.adversary_disturbance_scale = adversary_disturbance_scale
        self.adv_action = None
        dims = {"observation": self.obs_dim, "action": self.action_dim, "dynamics": self.QUAD_TYPE}
        self.disturbances = {}
        if disturbances is not None:  # benchmark_env.py:314-330, quadrotor.py:808-813
            for mode, specs in disturbances.items():
                assert mode in dims, "[ERROR] in BenchmarkEnv._setup_disturbances(), disturbance mode not available."
                self.disturbances[mode] = [_Disturbance(self, CHANNELS[mode], k, dims[mode], s)
                                           for k, s in enumerate(specs)]
        self.RND_SEED = seed
        self.RESEED_ON_RESET = reseed_on_reset
        self._user_draws = draws
        self.seed(seed)
        self.initial_reset = False
        self.INFO_IN_RESET = info_in_reset
        self.cnstr_violation = False  # benchmark_env.py:218 -- NOT cleared by reset()
        # F10: step() calls _get_info() before _get_done() (quadrotor.py:553-555) and _get_info reads self.goal_reached
        # (:1013), which only _get_done assigns (:965) -- an AttributeError on the first step as shipped.  Minimal
        # reading: False until the first _get_done; info["goal_reached"] is then the PREVIOUS step's flag and
        # survives reset() (nothing clears it).
        self.goal_reached = False
        self.episode = -1
        # ---- Quadrotor.__init__ post-super part (quadrotor.py:206-354)
        labels = STATE_LABELS[self.QUAD_TYPE]
        self.INIT = {k: 0.0 for k in DEFAULT_INIT_RAND}
        if init_state is not None:
            if isinstance(init_state, np.ndarray):
                for i, name in enumerate(labels):
                    self.INIT[name] = init_state[i]
            elif isinstance(init_state, dict):
                for name in labels:
                    self.INIT[name] = init_state.get(name, 0.0)
            else:
                raise ValueError("[ERROR] in Quadrotor.__init__(), init_state incorrect format.")
        for name in list(self.INIT_STATE_RAND_INFO.keys()):
            if name not in labels:
                self.INIT_STATE_RAND_INFO.pop(name, None)
        if self.QUAD_TYPE == 1:
            for k in ("Ixx", "Iyy", "Izz"):
                self.INERTIAL_PROP_RAND_INFO.pop(k, None)
        elif self.QUAD_TYPE == 2:
            for k in ("Ixx", "Izz"):
                self.INERTIAL_PROP_RAND_INFO.pop(k, None)
        if inertial_prop is None:
            pass
        elif self.QUAD_TYPE == 1 and np.array(inertial_prop).shape == (1,):
            self.MASS = inertial_prop[0]
        elif self.QUAD_TYPE == 2 and np.array(inertial_prop).shape == (2,):
            self.MASS, self.J[1, 1] = inertial_prop
        elif self.QUAD_TYPE == 3 and np.array(inertial_prop).shape == (4,):
            self.MASS, self.J[0, 0], self.J[1, 1], self.J[2, 2] = inertial_prop
        elif isinstance(inertial_prop, dict):
            self.MASS = inertial_prop.get("M", self.MASS)
            self.J[0, 0] = inertial_prop.get("Ixx", self.J[0, 0])
            self.J[1, 1] = inertial_prop.get("Iyy", self.J[1, 1])
            self.J[2, 2] = inertial_prop.get("Izz", self.J[2, 2])
        else:
            raise ValueError("[ERROR] in Quadrotor.__init__(), inertial_prop incorrect format.")
        self.U_GOAL = np.ones(self.action_dim) * self.MASS * self.GRAVITY_ACC / self.action_dim
        self._make_goal()
        self.OBSTACLES = [] if obstacles is None else obstacles
        self.GATES = [] if gates is None else gates
        self.RANDOMIZED_GATES_AND_OBS = bool(randomized_gates_and_obstacles)
        if self.RANDOMIZED_GATES_AND_OBS:
            if gates_and_obstacles_randomization_info is None:
                raise ValueError("[ERROR] Missing 'gates_and_obstacles_randomization_info' in YAML configuration.")
            self.GATES_AND_OBS_RAND_INFO = gates_and_obstacles_randomization_info
        self.DONE_ON_COLLISION = done_on_collision
        self.DONE_ON_COMPLETION = done_on_completion
        self.dyn_uses_episode_inertia = dyn_uses_episode_inertia
        self.gate_geometry = gate_geometry

    # ------------------------------------------------------------------ setup helpers
    def seed(self, seed=None):
        """benchmark_env.py:220-235.  gym's hashed MT seeding is replaced by RandomState(seed)
        unless the caller supplied a provider (F7: streams are injected in parity tests)."""
        if self._user_draws is not None:
            self.draws = self._user_draws
        else:
            self.draws = SequentialDraws(np.random.RandomState(seed))
        return [seed]

    def _set_action_space(self):
        """quadrotor.py:699-733: float32 Box bounds, promoted to float64 when used."""
        nu = {1: 1, 2: 2, 3: 4}[self.QUAD_TYPE]
        if self.NORMALIZED_RL_ACTION_SPACE:
            self.hover_thrust = self.GRAVITY_ACC * self.MASS / nu
            low, high = -np.ones(nu), np.ones(nu)
        else:
            n_motors = 4 / nu
            a_low = self.KF * n_motors * (self.PWM2RPM_SCALE * self.MIN_PWM + self.PWM2RPM_CONST) ** 2
            a_high = self.KF * n_motors * (self.PWM2RPM_SCALE * self.MAX_PWM + self.PWM2RPM_CONST) ** 2
            low, high = np.full(nu, a_low, np.float32), np.full(nu, a_high, np.float32)
        self.action_low, self.action_high = low.astype(np.float32), high.astype(np.float32)

    def _set_observation_space(self):
        """quadrotor.py:735-806."""
        xt, yt, zt = 5, 5, 2.5
        ph = th = 85 * math.pi / 180
        ps = 180 * math.pi / 180
        if self.QUAD_TYPE == 1:
            low, high = [0, -F32MAX], [zt, F32MAX]
        elif self.QUAD_TYPE == 2:
            low = [-xt, -F32MAX, 0, -F32MAX, -th, -F32MAX]
            high = [xt, F32MAX, zt, F32MAX, th, F32MAX]
        else:
            low = [-xt, -F32MAX, -yt, -F32MAX, 0, -F32MAX, -ph, -th, -ps, -F32MAX, -F32MAX, -F32MAX]
            high = [xt, F32MAX, yt, F32MAX, zt, F32MAX, ph, th, ps, F32MAX, F32MAX, F32MAX]
        low, high = np.array(low, dtype=np.float64), np.array(high, dtype=np.float64)
        self.state_low, self.state_high = low.astype(np.float32), high.astype(np.float32)
        if self.COST == "rl_reward" and self.obs_goal_horizon > 0:
            mul = 1 + self.obs_goal_horizon if self.TASK == "traj_tracking" else 2
            low, high = np.concatenate([low] * mul), np.concatenate([high] * mul)
        self.obs_low, self.obs_high = low.astype(np.float32), high.astype(np.float32)

    def _make_goal(self):
        """quadrotor.py:261-325."""
        ti = self.TASK_INFO
        if self.TASK == "stabilization":
            g = ti["stabilization_goal"]
            if self.QUAD_TYPE == 1:
                self.X_GOAL = np.hstack([g[1], 0.0])
            elif self.QUAD_TYPE == 2:
                self.X_GOAL = np.hstack([g[0], 0.0, g[1], 0.0, 0.0, 0.0])
            else:
                self.X_GOAL = np.hstack([g[0], 0.0, g[1], 0.0, g[2], 0.0, 0.0, 0.0, 0.0, 0.0, 0.0, 0.0])
        else:
            pos, vel = generate_trajectory(ti["trajectory_type"], self.EPISODE_LEN_SEC, ti["num_cycles"],
                                           ti["trajectory_plane"], ti["trajectory_position_offset"],
                                           ti["trajectory_scale"], self.CTRL_TIMESTEP)
            z = np.zeros(pos.shape[0])
            if self.QUAD_TYPE == 1:
                self.X_GOAL = np.vstack([pos[:, 2], vel[:, 2]]).T
            elif self.QUAD_TYPE == 2:
                self.X_GOAL = np.vstack([pos[:, 0], vel[:, 0], pos[:, 2], vel[:, 2], z, z]).T
            else:
                pt, vt = transform_trajectory(pos, vel, ti["proj_point"], ti["proj_normal"])
                self.X_GOAL = np.vstack([pt[:, 0], vt[:, 0], pt[:, 1], vt[:, 1], pt[:, 2], vt[:, 2],
                                         z, z, z, z, z, z]).T

    def set_cost_function_param(self, Q, R):
        """benchmark_env.py:237-254."""
        if self.initial_reset:
            raise RuntimeError("[ERROR] env.set_cost_function_param() cannot be called after the first reset of the environment.")
        self.Q, self.R = np.asarray(Q, dtype=float), np.asarray(R, dtype=float)

    def set_adversary_control(self, action):
        """benchmark_env.py:256-266 (adversary_action_space = Box(-1, 1))."""
        if self.adversary_disturbance is None:
            raise RuntimeError("[ERROR] adversary_disturbance does not exist, env.set_adversary_control() cannot be called.")
        self.adv_action = np.clip(action, -1.0, 1.0) * self.adversary_disturbance_scale + self.adversary_disturbance_offset

    # ------------------------------------------------------------------ reset
    def _randomize(self, values, info, stream):
        """benchmark_env.py:277-308: value += distrib(*args, **kwargs), in dict order of ``values``."""
        out = copy.deepcopy(values)
        info = copy.deepcopy(info)
        for slot, key in enumerate(values):
            if key in info:
                name = info[key].pop("distrib")
                args = info[key].pop("args", [])
                out[key] += self.draws.distrib(name, args, info[key], stream=stream, slot=slot,
                                               episode=self.episode)
        return out

    def reset(self):
        """quadrotor.py:356-504 with before_reset/after_reset (benchmark_env.py:356-383) and
        _reset_simulation (base_aviary.py:208-260)."""
        if self.RESEED_ON_RESET:
            self.seed(self.RND_SEED)
            self.episode = 0  # identical draws every episode
        else:
            self.episode += 1
        self.initial_reset = True
        self.pyb_step_counter = 0
        self.ctrl_step_counter = 0
        self.current_raw_input_action = None
        self.current_preprocessed_action = None
        slot = 0
        for mode in self.disturbances:
            for d in self.disturbances[mode]:
                d.reset(self, slot)
                slot += 1
        if self.adversary_disturbance is not None:
            self.adv_action = None
        self.obs_call_index = 0
        # _reset_simulation: zero kinematics, rpy_rates = 0.
        self.rpy_rates = np.zeros(3)
        self.last_clipped_action = np.zeros(4)
        self.last_substep_rpm = np.zeros(4)
        # Gates / obstacles (quadrotor.py:371-434): x, y, yaw offsets, 3 draws each.
        self.EFFECTIVE_OBSTACLES = []
        for i, ob in enumerate(self.OBSTACLES):
            off = self._pose_offsets("obstacles", px.STREAM_RESET_OBSTACLES, i)
            self.EFFECTIVE_OBSTACLES.append([ob[0] + off[0], ob[1] + off[1], ob[2] + 0.525,
                                             ob[3], ob[4], ob[5] + off[2]])
        self.EFFECTIVE_GATES_POSITIONS = []
        for i, gate in enumerate(self.GATES):
            if gate[6] == 0:
                height = 1.0
            elif gate[6] == 1:
                height = 0.525
            else:
                raise ValueError("[ERROR] Unknown gate type.")
            off = self._pose_offsets("gates", px.STREAM_RESET_GATES, i)
            self.EFFECTIVE_GATES_POSITIONS.append([gate[0] + off[0], gate[1] + off[1], gate[2] + height,
                                                   gate[3], gate[4], gate[5] + off[2]])
        self.NUM_GATES = len(self.GATES)
        self.current_gate = 0
        self.stepped_through_gate = False
        self.currently_collided = False
        self.at_goal_pos = False
        self.steps_at_goal_pos = 0
        self.task_completed = False
        # Inertial properties (quadrotor.py:447-467).
        prop = {"M": self.MASS, "Ixx": self.J[0, 0], "Iyy": self.J[1, 1], "Izz": self.J[2, 2]}
        if self.RANDOMIZED_INERTIAL_PROP:
            prop = self._randomize(prop, self.INERTIAL_PROP_RAND_INFO, px.STREAM_RESET_INERTIA)
            if any(v < 0 for v in prop.values()):
                raise ValueError("[ERROR] in Quadrotor.reset(), negative randomized inertial properties.")
        self.OVERRIDDEN_QUAD_MASS = prop["M"]
        self.OVERRIDDEN_QUAD_INERTIA = [prop["Ixx"], prop["Iyy"], prop["Izz"]]
        # Initial state (quadrotor.py:477-493).
        labels = STATE_LABELS[self.QUAD_TYPE]
        init = {name: self.INIT[name] for name in labels}
        if self.RANDOMIZED_INIT:
            init = self._randomize(init, self.INIT_STATE_RAND_INFO, px.STREAM_RESET_INIT)
        self.pos = np.array([init.get("init_" + k, 0.0) for k in "xyz"], dtype=float)
        self.vel = np.array([init.get("init_" + k + "_dot", 0.0) for k in "xyz"], dtype=float)
        rpy0 = [init.get("init_" + k, 0.0) for k in ("phi", "theta", "psi")]
        if self.QUAD_TYPE == 2:
            self.ang_v = np.array([0.0, init.get("init_theta_dot", 0.0), 0.0])
        else:
            self.ang_v = np.array([init.get("init_" + k, 0.0) for k in "pqr"], dtype=float)
        self.quat = bm.quat_from_euler(*rpy0)          # p.resetBasePositionAndOrientation
        self.rpy = np.array(bm.euler_from_quat(self.quat))  # _update_and_store_kinematic_information
        obs = self._get_observation()
        info = {"x_reference": self.X_GOAL, "u_reference": self.U_GOAL}
        if self.constraints is not None:  # after_reset: state constraints only
            vals = [c.value(self) for c in self.constraints if c.var == "state"]
            info["constraint_values"] = np.concatenate(vals) if vals else np.zeros(0)
        return (obs, info) if self.INFO_IN_RESET else obs

    def _pose_offsets(self, which, stream, i):
        if not self.RANDOMIZED_GATES_AND_OBS:
            return (0.0, 0.0, 0.0)
        info = copy.deepcopy(self.GATES_AND_OBS_RAND_INFO[which])
        name = info.pop("distrib")
        args = info.pop("args", [])
        return tuple(self.draws.distrib(name, args, info, stream=stream + 0, slot=4 * i + j,
                                        episode=self.episode) for j in range(3))

    # ------------------------------------------------------------------ step
    def _apply_disturbances(self, mode, target, index):
        for d in self.disturbances[mode]:
            target = d.apply(target, self, index)
        return target

    def _preprocess_control(self, action):
        """quadrotor.py:815-842 + cmd2pwm / pwm2rpm (quadrotor_utils.py:21-67)."""
        low, high = self.action_low.astype(np.float64), self.action_high.astype(np.float64)
        if self.NORMALIZED_RL_ACTION_SPACE:
            action = np.clip(action, low, high)
            thrust = (1 + self.norm_act_scale * action) * self.hover_thrust
        else:
            thrust = np.clip(action, low, high)
        self.current_preprocessed_action = thrust
        if "action" in self.disturbances:
            thrust = self._apply_disturbances("action", thrust, self.ctrl_step_counter)
        if self.adversary_disturbance == "action":
            thrust = thrust + self.adv_action
        thrust = np.asarray(thrust, dtype=float)
        n_motor = 4 // int(thrust.size)
        thrust = np.clip(thrust, np.zeros_like(thrust), None)
        pwm = (np.sqrt(thrust / n_motor / self.KF) - self.PWM2RPM_CONST) / self.PWM2RPM_SCALE
        if thrust.size == 1:
            pwm = np.repeat(pwm, 4)
        elif thrust.size == 2:
            pwm = np.concatenate([pwm, pwm[::-1]], 0)
        pwm = np.clip(pwm, self.MIN_PWM, self.MAX_PWM)
        return self.PWM2RPM_SCALE * pwm + self.PWM2RPM_CONST

    def _aero(self, rpm):
        """Extension (opt-in, off in parity mode): the aerodynamic terms the reference's PYB_* modes add with
        applyExternalForce (base_aviary.py:437-497), as world force and body torque for the explicit integrator.

        ``_ground_effect`` (:437-470): per propeller, height of its link (URDF offset rotated into the world) clipped at
        GND_EFF_H_CLIP (:145-147), force rpm^2 KF GND_EFF_COEFF (PROP_RADIUS / (4 h))^2 along the link's z axis, applied at
        the link, only while |roll| and |pitch| are below pi / 2.  ``_drag`` (:472-497): R (k * v) with k = -DRAG_COEFF *
        sum(2 pi rpm / 60) of ``last_clipped_action`` (the PREVIOUS substep's RPMs: zeros after a reset, :224, :297, :320),
        handed to applyExternalForce with LINK_FRAME, i.e. rotated into the world once more.  ``_downwash`` (:499-520) is
        zero for one drone per env."""
        R = np.array(bm.matrix_from_quat(self.quat)).reshape(3, 3)
        force, torque = np.zeros(3), np.zeros(3)
        if "ground_effect" in self.dyn_aero:
            max_rpm = np.sqrt((CF2X["THRUST2WEIGHT"] * self.GRAVITY) / (4 * self.KF))      # base_aviary.py:140-141
            max_thrust = 4 * self.KF * max_rpm ** 2
            h_clip = 0.25 * AERO["PROP_RADIUS"] * np.sqrt((15 * max_rpm ** 2 * self.KF * AERO["GND_EFF_COEFF"]) / max_thrust)
            heights = np.array([self.pos[2] + (R @ np.array([x, y, 0.0]))[2] for x, y in AERO["PROP_XY"]])
            heights = np.clip(heights, h_clip, np.inf)
            gnd = np.array(rpm ** 2) * self.KF * AERO["GND_EFF_COEFF"] * (AERO["PROP_RADIUS"] / (4 * heights)) ** 2
            if abs(self.rpy[0]) < np.pi / 2 and abs(self.rpy[1]) < np.pi / 2:
                for (x, y), g in zip(AERO["PROP_XY"], gnd):
                    force = force + R @ np.array([0.0, 0.0, g])
                    torque = torque + np.cross(np.array([x, y, 0.0]), np.array([0.0, 0.0, g]))
        if "drag" in self.dyn_aero:
            k = -1 * np.array(AERO["DRAG_COEFF"]) * np.sum(np.array(2 * np.pi * self.last_substep_rpm / 60))
            force = force + R @ (R @ (k * self.vel))
        return force, torque

    def _dynamics(self, rpm, ext_force=None):
        """base_aviary.py:522-576, TIMESTEP := PYB_TIMESTEP (F3).  One explicit substep followed
        by the PyBullet store/readback that re-derives rpy from the quaternion (:565-574, :362-374)."""
        dt = self.PYB_TIMESTEP
        mass, gravity, J, J_inv = self.MASS, self.GRAVITY, self.J, self.J_INV
        if self.dyn_uses_episode_inertia:  # extension (F9): per-episode values reach the integrator
            mass = self.OVERRIDDEN_QUAD_MASS
            gravity = self.GRAVITY_ACC * mass
            J = np.diag(self.OVERRIDDEN_QUAD_INERTIA)
            J_inv = np.diag(1.0 / np.array(self.OVERRIDDEN_QUAD_INERTIA))
        rotation = np.array(bm.matrix_from_quat(self.quat)).reshape(3, 3)
        forces = np.array(rpm ** 2) * self.KF
        thrust = np.array([0, 0, np.sum(forces)])
        force_world = np.dot(rotation, thrust) - np.array([0, 0, gravity])
        if ext_force is not None:  # extension (F5): the reference's applyExternalForce is a no-op in DYN
            force_world = force_world + ext_force
        zt = np.array(rpm ** 2) * self.KM
        z_torque = zt[0] - zt[1] + zt[2] - zt[3]
        x_torque = (forces[0] + forces[1] - forces[2] - forces[3]) * (self.L / np.sqrt(2))
        y_torque = (-forces[0] + forces[1] + forces[2] - forces[3]) * (self.L / np.sqrt(2))
        torques = np.array([x_torque, y_torque, z_torque])
        if self.dyn_aero:
            f_aero, t_aero = self._aero(rpm)
            force_world = force_world + f_aero
            torques = torques + t_aero
            self.last_substep_rpm = np.array(rpm, dtype=float)     # base_aviary.py:320
        torques = torques - np.cross(self.rpy_rates, np.dot(J, self.rpy_rates))
        rates_deriv = np.dot(J_inv, torques)
        acc = force_world / mass
        self.vel = self.vel + dt * acc
        self.rpy_rates = self.rpy_rates + dt * rates_deriv
        self.pos = self.pos + dt * self.vel
        rpy = self.rpy + dt * self.rpy_rates
        self.quat = bm.quat_from_euler(*rpy)
        self.ang_v = self.rpy_rates.copy()  # "ang_vel not computed by DYN" (:570-574)
        self.rpy = np.array(bm.euler_from_quat(self.quat))

    def _get_observation(self):
        """quadrotor.py:844-884."""
        if self.QUAD_TYPE == 1:
            self.state = np.array([self.pos[2], self.vel[2]])
        elif self.QUAD_TYPE == 2:
            self.state = np.array([self.pos[0], self.vel[0], self.pos[2], self.vel[2], self.rpy[1], self.ang_v[1]])
        else:
            Rob = np.array(bm.matrix_from_quat(self.quat)).reshape(3, 3)
            body = Rob.T @ self.ang_v
            self.state = np.hstack([self.pos[0], self.vel[0], self.pos[1], self.vel[1], self.pos[2],
                                    self.vel[2], self.rpy, body]).reshape(12,)
        obs = copy.deepcopy(self.state)
        if "observation" in self.disturbances:
            obs = self._apply_disturbances("observation", obs, self.obs_call_index)
        self.obs_call_index += 1
        return self.extend_obs(obs, self.ctrl_step_counter + 1)

    def extend_obs(self, obs, next_step):
        """benchmark_env.py:406-420."""
        if self.COST == "rl_reward" and self.TASK == "traj_tracking" and self.obs_goal_horizon > 0:
            idx = [min(next_step + i, self.X_GOAL.shape[0] - 1) for i in range(self.obs_goal_horizon)]
            return np.concatenate([obs, self.X_GOAL[idx].flatten()])
        if self.COST == "rl_reward" and self.TASK == "stabilization" and self.obs_goal_horizon > 0:
            return np.concatenate([obs, self.X_GOAL.flatten()])
        return obs

    def step(self, action):
        """quadrotor.py:507-558 with before_step / after_step (benchmark_env.py:394-463)."""
        if not self.initial_reset:
            raise RuntimeError("[ERROR] You must call env.reset() at least once before using env.step().")
        action = np.asarray(action, dtype=float)
        self.current_raw_input_action = action
        rpm = self._preprocess_control(action)
        force = None
        passive = "dynamics" in self.disturbances
        adv = self.adversary_disturbance == "dynamics"
        if passive or adv:
            force = np.zeros(self.QUAD_TYPE)
        if passive:
            force = self._apply_disturbances("dynamics", force, self.ctrl_step_counter)
        if adv and self.adv_action is not None:
            force = force + self.adv_action
            self.adv_action = None
        ext = None
        if force is not None and self.dyn_uses_episode_inertia:
            f = np.asarray(force, dtype=float).flatten()
            ext = {1: lambda: np.array([0, 0, f[0]]), 2: lambda: np.array([f[0], 0, f[1]]),
                   3: lambda: f}[self.QUAD_TYPE]()
        # _advance_simulation (base_aviary.py:266-322): S substeps; the rpy readback before each
        # substep (S > 1) and after the loop is folded into _dynamics.
        for _ in range(self.PYB_STEPS_PER_CTRL):
            self._dynamics(rpm, ext)
        self.last_clipped_action = rpm
        obs = self._get_observation()
        info = self._get_info()
        done = self._get_done()
        rew = self._get_reward()
        # after_step
        self.pyb_step_counter += self.PYB_STEPS_PER_CTRL
        self.ctrl_step_counter += 1
        if self.constraints is not None:
            cs_ = [c.value(self) for c in self.constraints]
            info["constraint_values"] = np.concatenate(cs_)
            if any(c.violated(v) for c, v in zip(self.constraints, cs_)):
                self.cnstr_violation = True
                info["constraint_violation"] = 1
                if self.DONE_ON_VIOLATION:
                    done = True
            else:
                self.cnstr_violation = False
                info["constraint_violation"] = 0
            if self.COST == "rl_reward" and self.use_constraint_penalty and \
                    any(c.almost_active(v) for c, v in zip(self.constraints, cs_)):
                rew += self.constraint_penalty
        if self.ctrl_step_counter >= self.CTRL_STEPS:
            info["TimeLimit.truncated"] = not done
            done = True
        return obs, rew, done, info

    def _goal_at(self, idx):
        return self.X_GOAL if self.TASK == "stabilization" else self.X_GOAL[idx]

    def _get_reward(self):
        """quadrotor.py:886-954 (quadratic: -SymbolicModel.loss, cost quadrotor.py:682)."""
        t = self.ctrl_step_counter
        if self.COST == "rl_reward":
            act_err = np.asarray(self.current_preprocessed_action) - self.U_GOAL
            err = self.state - self._goal_at(min(t, self.X_GOAL.shape[0] - 1))
            dist = np.sum(self.rew_state_weight * err * err) + np.sum(self.rew_act_weight * act_err * act_err)
            rew = -dist
            return np.exp(rew) if self.rew_exponential else rew
        if self.COST == "quadratic":
            ex = self.state - self._goal_at(t)
            eu = np.asarray(self.current_preprocessed_action) - self.U_GOAL
            return float(-1 * (0.5 * ex @ self.Q @ ex + 0.5 * eu @ self.R @ eu))
        reward = 0
        if self.stepped_through_gate:
            reward += 100
        if self.at_goal_pos:
            reward += 100
        if self.currently_collided:
            reward -= 1000
        if self.cnstr_violation:  # previous step's flag (F8)
            reward -= 100
        return reward

    def _get_done(self):
        """quadrotor.py:956-1002."""
        if self.TASK == "stabilization" and self.COST == "quadratic":
            self.goal_reached = bool(np.linalg.norm(self.state - self.X_GOAL) <
                                     self.TASK_INFO["stabilization_goal_tolerance"])
            if self.goal_reached:
                return True
        if self.done_on_out_of_bound:
            mask = np.array({1: [1, 0], 2: [1, 0, 1, 0, 1, 0], 3: [1, 0, 1, 0, 1, 0, 1, 1, 1, 0, 0, 0]}[self.QUAD_TYPE])
            oob = np.logical_or(self.state < self.state_low, self.state > self.state_high)
            if np.any(oob * mask):
                return True
        if self.DONE_ON_COLLISION and self.currently_collided:
            return True
        if self.DONE_ON_COMPLETION and self.task_completed:
            return True
        return False

    def _get_info(self):
        """quadrotor.py:1004-1134.  Collision scan is dead in DYN (F5): never collided.  Gate
        progress uses the analytic geometry of oracle/gate_geometry.py when ``gate_geometry`` is on
        (extension; with it off the gate index never advances)."""
        info = {}
        if self.TASK == "stabilization" and self.COST == "quadratic":
            info["goal_reached"] = self.goal_reached
        t = self.ctrl_step_counter
        err = self.state - self._goal_at(min(t, self.X_GOAL.shape[0] - 1))
        err = err * self.info_mse_metric_state_weight
        info["mse"] = np.sum(err ** 2)
        info["collision"] = (None, False)
        self.currently_collided = False
        if self.gate_geometry and self.QUAD_TYPE == 3:   # extension: the contact scan on the integrated pose
            from .gate_geometry import drone_collides
            self.currently_collided = drone_collides(self.pos, np.array(bm.matrix_from_quat(self.quat)).reshape(3, 3),
                                                     self.EFFECTIVE_GATES_POSITIONS, [g[6] for g in self.GATES],
                                                     self.EFFECTIVE_OBSTACLES)
            info["collision"] = (None, self.currently_collided)
        if self.pyb_step_counter > 0.5 * self.PYB_FREQ and self.NUM_GATES > 0 and self.current_gate < self.NUM_GATES:
            self.stepped_through_gate = False
            if self.gate_geometry:
                from .gate_geometry import gate_rays_hit_drone
                g = self.EFFECTIVE_GATES_POSITIONS[self.current_gate]
                height = 1.0 if self.GATES[self.current_gate][6] == 0 else 0.525
                if gate_rays_hit_drone(g[0], g[1], g[5], height, self.pos,
                                       np.array(bm.matrix_from_quat(self.quat)).reshape(3, 3)):
                    self.current_gate += 1
                    self.stepped_through_gate = True
        if self.current_gate < self.NUM_GATES:
            info["current_target_gate_id"] = self.current_gate
            info["current_target_gate_type"] = self.GATES[self.current_gate][6]
            if self.gate_geometry:   # quadrotor.py:1093-1108 on the integrated pose
                from .gate_geometry import gate_in_range
                g = self.EFFECTIVE_GATES_POSITIONS[self.current_gate]
                rng = gate_in_range(self.pos, g, self.GATES[self.current_gate][6])
                info["current_target_gate_in_range"] = rng
                info["current_target_gate_pos"] = list(g) if rng else list(self.GATES[self.current_gate][0:6])
        else:
            info["current_target_gate_id"] = -1
            info["current_target_gate_type"] = -1
            if self.gate_geometry:
                info["current_target_gate_in_range"] = False
                info["current_target_gate_pos"] = []
        info["at_goal_position"] = False
        info["task_completed"] = False
        if self.current_gate == self.NUM_GATES and self.QUAD_TYPE == 3:
            quad_xyz = np.array([self.state[0], self.state[2], self.state[4]])
            goal_xyz = np.array([self.X_GOAL[0], self.X_GOAL[2], self.X_GOAL[4]])
            if np.linalg.norm(quad_xyz - goal_xyz) < self.TASK_INFO["stabilization_goal_tolerance"]:
                self.at_goal_pos = True
                self.steps_at_goal_pos += 1
            else:
                self.at_goal_pos = False
                self.steps_at_goal_pos = 0
            if self.steps_at_goal_pos > self.CTRL_FREQ * 2:
                self.task_completed = True
            info["at_goal_position"] = self.at_goal_pos
            info["task_completed"] = self.task_completed
        return info
