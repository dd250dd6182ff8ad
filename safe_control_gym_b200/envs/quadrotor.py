"""Batched Quadrotor (quad_type 1D / 2D / 3D, physics = DYN) behind the reference's env API.

Accepts the reference's constructor kwargs / YAML keys unchanged
(safe_control_gym/envs/gym_pybullet_drones/quadrotor.yaml, competition/level*.yaml) plus the
batch kwargs ``num_envs, device, dtype, attitude, auto_reset, squeeze, seed_offset`` and
the opt-in extensions ``dyn_uses_episode_inertia``, ``dyn_applies_force``
(F5) and ``gate_geometry``.  All constants are derived exactly as Quadrotor.__init__ /
BaseAviary.__init__ / BenchmarkEnv.__init__ derive them (file:line cited inline) and
handed to libscg_b200.so through ``scg_config``; the per-step work is one CUDA launch.
"""
import copy
import ctypes as C
import math

import numpy as np
import torch

from .. import _lib as L
from ..constraints import GENERAL_CONSTRAINTS, create_constraint_list
from ..disturbances import create_disturbance_list
from ..spaces import Box
from ..trajectories import planar_reference, project_onto_plane
from .batched_env import BatchedEnvBase

# assets/cf2x.urdf:5,12-13 as parsed by BaseAviary._parse_urdf_parameters (base_aviary.py:612-651)
CF2X = dict(M=0.03454, L=0.0397, THRUST2WEIGHT=2.25, IXX=1.4e-5, IYY=1.4e-5, IZZ=2.17e-5, KF=3.16e-10,
            KM=7.94e-12, COLLISION_H=0.025, COLLISION_R=0.06, COLLISION_Z_OFFSET=0.0, MAX_SPEED_KMH=30.0,
            GND_EFF_COEFF=11.36859, PROP_RADIUS=2.31348e-2, DRAG_COEFF_XY=9.1785e-7, DRAG_COEFF_Z=10.311e-7,
            DW_COEFF_1=2267.18, DW_COEFF_2=0.16, DW_COEFF_3=-0.11, PWM2RPM_SCALE=0.2685, PWM2RPM_CONST=4070.3,
            MIN_PWM=20000.0, MAX_PWM=65535.0)

_LABELS = {1: ["init_x", "init_x_dot"],
           2: ["init_x", "init_x_dot", "init_z", "init_z_dot", "init_theta", "init_theta_dot"],
           3: ["init_x", "init_x_dot", "init_y", "init_y_dot", "init_z", "init_z_dot",
               "init_phi", "init_theta", "init_psi", "init_p", "init_q", "init_r"]}
_F32MAX = float(np.finfo(np.float32).max)


def _uniform(lo, hi):
    return {"distrib": "uniform", "low": lo, "high": hi}


def _rand_record(info):
    """One randomisation entry (benchmark_env.py:296-307) -> scg_rand."""
    r = L.Rand()
    if info is None:
        r.distrib = L.RAND_NONE
        return r
    info = dict(info)
    name = info.pop("distrib")
    args = list(info.pop("args", []))
    if name == "uniform":
        r.distrib = L.RAND_UNIFORM
        r.a = float(info.get("low", args[0] if len(args) > 0 else 0.0))
        r.b = float(info.get("high", args[1] if len(args) > 1 else 1.0))
    elif name == "normal":
        r.distrib = L.RAND_NORMAL
        r.a = float(info.get("loc", args[0] if len(args) > 0 else 0.0))
        r.b = float(info.get("scale", args[1] if len(args) > 1 else 1.0))
    else:
        raise NotImplementedError(f"randomisation distribution '{name}' is not available on the device "
                                  "(supported: uniform, normal)")
    return r


class BatchedQuadrotor(BatchedEnvBase):
    NAME = "quadrotor"
    AVAILABLE_CONSTRAINTS = copy.deepcopy(GENERAL_CONSTRAINTS)

    INERTIAL_PROP_RAND_INFO = {"M": _uniform(0.022, 0.032), "Ixx": _uniform(1.3e-5, 1.5e-5),
                               "Iyy": _uniform(1.3e-5, 1.5e-5), "Izz": _uniform(2.07e-5, 2.27e-5)}
    INIT_STATE_RAND_INFO = {
        "init_x": _uniform(-0.5, 0.5), "init_x_dot": _uniform(-0.01, 0.01), "init_y": _uniform(-0.5, 0.5),
        "init_y_dot": _uniform(-0.01, 0.01), "init_z": _uniform(0.1, 1.5), "init_z_dot": _uniform(-0.01, 0.01),
        "init_phi": _uniform(-0.3, 0.3), "init_theta": _uniform(-0.3, 0.3), "init_psi": _uniform(-0.3, 0.3),
        "init_p": _uniform(-0.01, 0.01), "init_theta_dot": _uniform(-0.01, 0.01), "init_q": _uniform(-0.01, 0.01),
        "init_r": _uniform(-0.01, 0.01)}
    TASK_INFO = {"stabilization_goal": [0, 1], "stabilization_goal_tolerance": 0.05, "trajectory_type": "circle",
                 "num_cycles": 1, "trajectory_plane": "zx", "trajectory_position_offset": [0.5, 0],
                 "trajectory_scale": -0.5, "proj_point": [0, 0, 0.5], "proj_normal": [0, 1, 1]}

    def __init__(self, num_envs=1, device="cuda", dtype=torch.float32, attitude="fast", auto_reset=True,
                 squeeze=False, seed_offset=0,
                 # BenchmarkEnv kwargs (benchmark_env.py:63-96)
                 output_dir=None, seed=None, info_in_reset=False, gui=False, verbose=False,
                 normalized_rl_action_space=False, task="stabilization", task_info=None, cost="rl_reward",
                 pyb_freq=50, ctrl_freq=50, episode_len_sec=5, init_state=None, randomized_init=True,
                 init_state_randomization_info=None, prior_prop=None, inertial_prop=None,
                 randomized_inertial_prop=False, inertial_prop_randomization_info=None, constraints=None,
                 done_on_violation=False, use_constraint_penalty=False, constraint_penalty=-1, disturbances=None,
                 adversary_disturbance=None, adversary_disturbance_offset=0.0, adversary_disturbance_scale=0.01,
                 # Quadrotor kwargs (quadrotor.py:148-163)
                 quad_type=2, norm_act_scale=0.1, obs_goal_horizon=0, rew_state_weight=1.0, rew_act_weight=0.0001,
                 rew_exponential=True, done_on_out_of_bound=True, info_mse_metric_state_weight=None,
                 # BaseAviary kwargs
                 physics="dyn",
                 # extensions (off = reference behaviour)
                 dyn_uses_episode_inertia=False, dyn_applies_force=False, gate_geometry=False, dyn_aero=(),
                 **kwargs):
        if str(getattr(physics, "value", physics)) != "dyn":
            raise NotImplementedError("only physics='dyn' is implemented (the PYB_* modes need Bullet's solver; "
                                      "SURVEY.md section 2, out of scope)")
        if gui:
            raise NotImplementedError("gui / rendering: unavailable on this box (no graphics stack)")
        # ---- Quadrotor.__init__ head (quadrotor.py:176-201)
        self.QUAD_TYPE = int(quad_type)
        if self.QUAD_TYPE not in (1, 2, 3):
            raise ValueError("[ERROR] in Quadrotor.__init__(), not implemented quad type.")
        nx, nu = {1: (2, 1), 2: (6, 2), 3: (12, 4)}[self.QUAD_TYPE]
        self.norm_act_scale = norm_act_scale
        self.obs_goal_horizon = int(obs_goal_horizon)
        self.rew_state_weight = np.array(rew_state_weight, ndmin=1, dtype=float)
        self.rew_act_weight = np.array(rew_act_weight, ndmin=1, dtype=float)
        self.rew_exponential = rew_exponential
        self.done_on_out_of_bound = done_on_out_of_bound
        if info_mse_metric_state_weight is None:
            w = {1: [1, 0], 2: [1, 0, 1, 0, 0, 0], 3: [1, 0, 1, 0, 1, 0, 0, 0, 0, 0, 0, 0]}[self.QUAD_TYPE]
        else:
            if len(info_mse_metric_state_weight) != nx:
                raise ValueError("[ERROR] in Quadrotor.__init__(), wrong info_mse_metric_state_weight argument size.")
            w = info_mse_metric_state_weight
        self.info_mse_metric_state_weight = np.array(w, ndmin=1, dtype=float)
        # ---- BaseAviary.__init__ constants (base_aviary.py:90-147)
        self.GRAVITY_ACC = 9.8
        self.RAD2DEG, self.DEG2RAD = 180 / np.pi, np.pi / 180
        self.MASS, self.L, self.THRUST2WEIGHT_RATIO = CF2X["M"], CF2X["L"], CF2X["THRUST2WEIGHT"]
        self.J = np.diag([CF2X["IXX"], CF2X["IYY"], CF2X["IZZ"]])
        self.J_INV = np.linalg.inv(self.J)
        self.KF, self.KM = CF2X["KF"], CF2X["KM"]
        self.COLLISION_H, self.COLLISION_R, self.COLLISION_Z_OFFSET = CF2X["COLLISION_H"], CF2X["COLLISION_R"], 0.0
        self.PWM2RPM_SCALE, self.PWM2RPM_CONST = CF2X["PWM2RPM_SCALE"], CF2X["PWM2RPM_CONST"]
        self.MIN_PWM, self.MAX_PWM = CF2X["MIN_PWM"], CF2X["MAX_PWM"]
        self.GROUND_PLANE_Z = 0
        self.GRAVITY = self.GRAVITY_ACC * self.MASS          # frozen here, before any override (F9)
        self.HOVER_RPM = np.sqrt(self.GRAVITY / (4 * self.KF))
        self.MAX_RPM = np.sqrt((self.THRUST2WEIGHT_RATIO * self.GRAVITY) / (4 * self.KF))
        self.MAX_THRUST = 4 * self.KF * self.MAX_RPM ** 2
        self.PHYSICS = "dyn"
        # ---- BenchmarkEnv.__init__ (benchmark_env.py:139-218)
        self.GUI, self.VERBOSE, self.output_dir = False, verbose, output_dir
        self.TASK = str(getattr(task, "value", task))
        if self.TASK not in ("stabilization", "traj_tracking"):
            raise ValueError(f"'{task}' is not a valid Task")
        self.TASK_INFO = copy.deepcopy(self.TASK_INFO if task_info is None else task_info)
        self.CTRL_FREQ, self.PYB_FREQ = ctrl_freq, pyb_freq
        if self.PYB_FREQ % self.CTRL_FREQ != 0:
            raise ValueError("[ERROR] in BenchmarkEnv.__init__(), pyb_freq is not divisible by env_freq.")
        self.PYB_STEPS_PER_CTRL = int(self.PYB_FREQ / self.CTRL_FREQ)
        self.CTRL_TIMESTEP, self.PYB_TIMESTEP = 1.0 / self.CTRL_FREQ, 1.0 / self.PYB_FREQ
        self.EPISODE_LEN_SEC = episode_len_sec
        self.CTRL_STEPS = self.EPISODE_LEN_SEC * self.CTRL_FREQ
        self.INIT_STATE, self.RANDOMIZED_INIT = init_state, randomized_init
        self.INIT_STATE_RAND_INFO = copy.deepcopy(self.INIT_STATE_RAND_INFO if init_state_randomization_info is None
                                                  else init_state_randomization_info)
        self.PRIOR_PROP, self.INERTIAL_PROP = prior_prop, inertial_prop
        self.RANDOMIZED_INERTIAL_PROP = randomized_inertial_prop
        self.INERTIAL_PROP_RAND_INFO = copy.deepcopy(self.INERTIAL_PROP_RAND_INFO if inertial_prop_randomization_info is None
                                                     else inertial_prop_randomization_info)
        self.NORMALIZED_RL_ACTION_SPACE = normalized_rl_action_space
        self.COST = str(getattr(cost, "value", cost))
        if self.COST not in ("rl_reward", "quadratic", "competition"):
            raise ValueError(f"'{cost}' is not a valid Cost")
        self._set_action_space(nu)
        self._set_observation_space()
        self.action_dim, self.obs_dim, self.state_dim = nu, self.observation_space.shape[0], nx
        if self.COST in ("quadratic", "competition"):
            self.Q, self.R = np.eye(self.obs_dim), np.eye(self.action_dim)
        self.CONSTRAINTS, self.DONE_ON_VIOLATION = constraints, done_on_violation
        self.use_constraint_penalty, self.constraint_penalty = use_constraint_penalty, constraint_penalty
        self.constraints, self.num_constraints = None, 0
        if constraints is not None:
            self.constraints = create_constraint_list(constraints, self.AVAILABLE_CONSTRAINTS, self)
            self.num_constraints = self.constraints.num_constraints
        self.DISTURBANCES = disturbances
        self.adversary_disturbance = adversary_disturbance
        self.adversary_disturbance_offset = adversary_disturbance_offset
        self.adversary_disturbance_scale = adversary_disturbance_scale
        self.DISTURBANCE_MODES = {"observation": {"dim": self.obs_dim}, "action": {"dim": nu},
                                  "dynamics": {"dim": self.QUAD_TYPE}}
        self.disturbances = {}
        slot = 0
        if disturbances is not None:
            for mode, specs in disturbances.items():
                assert mode in self.DISTURBANCE_MODES, "[ERROR] in BenchmarkEnv._setup_disturbances(), disturbance mode not available."
                shared = dict(self.DISTURBANCE_MODES[mode])
                if mode == "observation":
                    shared["dim"] = nx  # noise is added to the state part before extend_obs (quadrotor.py:878-884)
                self.disturbances[mode] = create_disturbance_list(specs, shared, self)
                for d in self.disturbances[mode]:
                    d.slot = slot
                    slot += 1
        if adversary_disturbance is not None:
            assert adversary_disturbance in self.DISTURBANCE_MODES, "[ERROR] in Cartpole._setup_disturbances()"
            dim = self.DISTURBANCE_MODES[adversary_disturbance]["dim"]
            self.adversary_action_space = Box(low=-1, high=1, shape=(dim,))
            self.adversary_observation_space = self.observation_space
        self.RND_SEED = seed
        self.RESEED_ON_RESET = bool(kwargs.get("reseed_on_reset", False))
        self.INFO_IN_RESET = info_in_reset
        self.seed(seed)
        # ---- Quadrotor.__init__ tail (quadrotor.py:206-354)
        self.INIT_STATE_LABELS = {1: _LABELS[1], 2: _LABELS[2], 3: _LABELS[3]}
        labels = _LABELS[self.QUAD_TYPE]
        init = {name: 0.0 for name in self.__class__.INIT_STATE_RAND_INFO}
        if init_state is not None:
            if isinstance(init_state, np.ndarray):
                for i, name in enumerate(labels):
                    init[name] = float(init_state[i])
            elif isinstance(init_state, dict):
                for name in labels:
                    init[name] = init_state.get(name, 0.0)
            else:
                raise ValueError("[ERROR] in Quadrotor.__init__(), init_state incorrect format.")
        for name, v in init.items():
            setattr(self, name.upper(), v)
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
        self._setup_symbolic()
        self.U_GOAL = np.ones(self.action_dim) * self.MASS * self.GRAVITY_ACC / self.action_dim
        self._make_goal()
        self.X_EQ, self.U_EQ = np.zeros(self.state_dim), self.U_GOAL
        self.OBSTACLES = kwargs.get("obstacles", []) or []
        self.GATES = kwargs.get("gates", []) or []
        self.NUM_GATES = len(self.GATES)
        self.RANDOMIZED_GATES_AND_OBS = bool(kwargs.get("randomized_gates_and_obstacles", False))
        if self.RANDOMIZED_GATES_AND_OBS:
            if "gates_and_obstacles_randomization_info" not in kwargs:
                raise ValueError("[ERROR] Missing 'gates_and_obstacles_randomization_info' in YAML configuration.")
            self.GATES_AND_OBS_RAND_INFO = kwargs["gates_and_obstacles_randomization_info"]
        for g in self.GATES:
            if g[6] not in (0, 1):
                raise ValueError("[ERROR] Unknown gate type.")
        self.DONE_ON_COLLISION = bool(kwargs.get("done_on_collision", False))
        self.DONE_ON_COMPLETION = bool(kwargs.get("done_on_completion", False))
        if self.RANDOMIZED_INERTIAL_PROP:   # reset() raises on negative values (quadrotor.py:464-465): check the support
            nominal = {"M": self.MASS, "Ixx": self.J[0, 0], "Iyy": self.J[1, 1], "Izz": self.J[2, 2]}
            for key, info in self.INERTIAL_PROP_RAND_INFO.items():
                if info.get("distrib") == "uniform" and nominal[key] + info.get("low", 0.0) < 0:
                    raise ValueError("[ERROR] in Quadrotor.reset(), negative randomized inertial properties.")
        # ---- batch
        self._dyn_uses_episode_inertia = bool(dyn_uses_episode_inertia)
        self._dyn_applies_force = bool(dyn_applies_force)
        self._gate_geometry = bool(gate_geometry)
        # quad_type 3: the PYB_* modes' aerodynamic terms folded into DYN (include/scg_b200.h: dyn_aero): "drag", "ground_effect"
        unknown = set(dyn_aero) - {"drag", "ground_effect"}
        if unknown:
            raise ValueError(f"dyn_aero: unknown term(s) {sorted(unknown)} (downwash is zero for one drone per env)")
        self._dyn_aero = (1 if "drag" in dyn_aero else 0) | (2 if "ground_effect" in dyn_aero else 0)
        self._auto_reset = bool(auto_reset)
        self._attitude = {"fast": L.ATT_FAST, "roundtrip": L.ATT_ROUNDTRIP, "fast_intrinsic": L.ATT_FAST_INTRINSIC}[attitude]
        if self._attitude == L.ATT_FAST_INTRINSIC and dtype != torch.float32:
            raise ValueError("attitude='fast_intrinsic' is an fp32-only mode")
        self._seed_offset = int(seed_offset)
        cfg = self._make_config(32 if dtype == torch.float32 else 64)
        rows = 0 if self.constraints is None else self.constraints.num_constraints
        self._n_rows = rows
        self._init_batch(cfg, num_envs, device, dtype, nx, nu, self.obs_dim, rows, slot, squeeze)
        idx, r = [], 0
        if self.constraints is not None:
            for c in self.constraints.constraints:
                if c.constrained_variable == "state":
                    idx.extend(range(r, r + c.num_constraints))
                r += c.num_constraints
        self._state_row_idx = torch.tensor(idx, dtype=torch.long, device=self.device)
        self.current_raw_input_action = None
        self.current_preprocessed_action = None
        self.adv_action = None

    # ------------------------------------------------------------------ spaces
    def _set_action_space(self, nu):
        """quadrotor.py:699-733."""
        labels = {1: ["T"], 2: ["T1", "T2"], 4: ["T1", "T2", "T3", "T4"]}[nu]
        self.ACTION_LABELS = labels
        self.ACTION_UNITS = ["N" if not self.NORMALIZED_RL_ACTION_SPACE else "-"] * nu
        if self.NORMALIZED_RL_ACTION_SPACE:
            self.hover_thrust = self.GRAVITY_ACC * self.MASS / nu
            self.action_space = Box(low=-np.ones(nu), high=np.ones(nu), dtype=np.float32)
        else:
            n_motors = 4 / nu
            a_low = self.KF * n_motors * (self.PWM2RPM_SCALE * self.MIN_PWM + self.PWM2RPM_CONST) ** 2
            a_high = self.KF * n_motors * (self.PWM2RPM_SCALE * self.MAX_PWM + self.PWM2RPM_CONST) ** 2
            self.action_space = Box(low=np.full(nu, a_low, np.float32), high=np.full(nu, a_high, np.float32), dtype=np.float32)

    def _set_observation_space(self):
        """quadrotor.py:735-806."""
        self.x_threshold, self.y_threshold, self.z_threshold = 5, 5, 2.5
        self.phi_threshold_radians = self.theta_threshold_radians = 85 * math.pi / 180
        self.psi_threshold_radians = 180 * math.pi / 180
        xt, yt, zt, ph, th, ps = 5, 5, 2.5, self.phi_threshold_radians, self.theta_threshold_radians, self.psi_threshold_radians
        M = _F32MAX
        if self.QUAD_TYPE == 1:
            low, high = [0, -M], [zt, M]
            self.STATE_LABELS, self.STATE_UNITS = ["z", "z_dot"], ["m", "m/s"]
        elif self.QUAD_TYPE == 2:
            low, high = [-xt, -M, 0, -M, -th, -M], [xt, M, zt, M, th, M]
            self.STATE_LABELS = ["x", "x_dot", "z", "z_dot", "theta", "theta_dot"]
            self.STATE_UNITS = ["m", "m/s", "m", "m/s", "rad", "rad/s"]
        else:
            low = [-xt, -M, -yt, -M, 0, -M, -ph, -th, -ps, -M, -M, -M]
            high = [xt, M, yt, M, zt, M, ph, th, ps, M, M, M]
            self.STATE_LABELS = ["x", "x_dot", "y", "y_dot", "z", "z_dot", "phi", "theta", "psi", "p", "q", "r"]
            self.STATE_UNITS = ["m", "m/s", "m", "m/s", "m", "m/s", "rad", "rad", "rad", "rad/s", "rad/s", "rad/s"]
        low, high = np.array(low, dtype=np.float64), np.array(high, dtype=np.float64)
        self.state_space = Box(low=low, high=high, dtype=np.float32)
        if self.COST == "rl_reward" and self.obs_goal_horizon > 0:
            mul = 1 + self.obs_goal_horizon if self.TASK == "traj_tracking" else 2
            low, high = np.concatenate([low] * mul), np.concatenate([high] * mul)
        self.observation_space = Box(low=low, high=high, dtype=np.float32)

    def _make_goal(self):
        """X_GOAL assembly (quadrotor.py:261-325)."""
        ti = self.TASK_INFO
        if self.TASK == "stabilization":
            g = ti["stabilization_goal"]
            if self.QUAD_TYPE == 1:
                self.X_GOAL = np.hstack([g[1], 0.0])
            elif self.QUAD_TYPE == 2:
                self.X_GOAL = np.hstack([g[0], 0.0, g[1], 0.0, 0.0, 0.0])
            else:
                self.X_GOAL = np.hstack([g[0], 0.0, g[1], 0.0, g[2], 0.0, 0.0, 0.0, 0.0, 0.0, 0.0, 0.0])
            return
        pos, vel, _ = planar_reference(ti["trajectory_type"], self.EPISODE_LEN_SEC, ti["num_cycles"], ti["trajectory_plane"],
                                       ti["trajectory_position_offset"], ti["trajectory_scale"], self.CTRL_TIMESTEP)
        zeros = np.zeros(pos.shape[0])
        if self.QUAD_TYPE == 1:
            cols = [pos[:, 2], vel[:, 2]]
        elif self.QUAD_TYPE == 2:
            cols = [pos[:, 0], vel[:, 0], pos[:, 2], vel[:, 2], zeros, zeros]
        else:
            pt, vt = project_onto_plane(pos, vel, ti["proj_point"], ti["proj_normal"])
            cols = [pt[:, 0], vt[:, 0], pt[:, 1], vt[:, 1], pt[:, 2], vt[:, 2]] + [zeros] * 6
        self.X_GOAL = np.vstack(cols).transpose()

    def _setup_symbolic(self):
        from ..symbolic import quadrotor_symbolic_model
        self.symbolic = quadrotor_symbolic_model(self)

    def seed(self, seed=None):
        """benchmark_env.py:220-235.  The device draws are counter-based Philox keyed by
        (seed, global env id); ``np_random`` is kept for host-side callers."""
        self.np_random = np.random.RandomState(None if seed is None else int(seed) % (2 ** 32))
        self._seed_value = 0 if seed is None else int(seed) % (2 ** 32)
        self.action_space.seed(seed)
        if getattr(self, "_h", None) is not None and self._h.value:   # after construction: re-key the device draws
            L.check(self.lib.scg_set_seed(self._h, self._seed_value))
            self._cfg.seed = self._seed_value
        return [seed]

    def set_cost_function_param(self, Q, R):
        """benchmark_env.py:237-254."""
        if self.initial_reset:
            raise RuntimeError("[ERROR] env.set_cost_function_param() cannot be called after the first reset of the environment.")
        self.Q, self.R = np.asarray(Q, dtype=float), np.asarray(R, dtype=float)
        self._rebuild()

    def set_adversary_control(self, action):
        """benchmark_env.py:256-266; consumed by the next step()."""
        if self.adversary_disturbance is None:
            raise RuntimeError("[ERROR] adversary_disturbance does not exist, env.set_adversary_control() cannot be called.")
        a = self._as_real(action, (self.num_envs, self.adversary_action_space.shape[0]))
        self.adv_action = torch.clamp(a, -1.0, 1.0) * self.adversary_disturbance_scale + self.adversary_disturbance_offset
        self._adv = self.adv_action.contiguous()
        L.check(self.lib.scg_set_injection(self._h, -1, 0, C.c_void_p(self._adv.data_ptr())))

    def _adv_consumed(self):
        if self.adversary_disturbance == "dynamics" and self._adv is not None:
            self._adv = None   # "Clear the adversary action, wait for the next one" (quadrotor.py:535-538)
            L.check(self.lib.scg_set_injection(self._h, -1, 0, None))

    def _rebuild(self):
        """Re-create the device handle after a host-side parameter change (before the first reset)."""
        cfg = self._make_config(32 if self.dtype == torch.float32 else 64)
        self.close()
        with torch.cuda.device(self.device):
            handle = C.c_void_p()
            L.check(self.lib.scg_create(C.byref(cfg), C.byref(handle)))
        self._h, self._cfg = handle, cfg
        self._bind()

    # ------------------------------------------------------------------ gate info (quadrotor.py:1093-1112)
    def _gate_info(self):
        """(pose [N, 6], in_range [N] bool) of the current target gate: one launch of scg_gate_info, cached per step."""
        tag = self.launch_count
        if getattr(self, "_gate_cache_tag", None) != tag:
            pose = torch.empty(6, self.num_envs, dtype=self.dtype, device=self.device)
            rng = torch.empty(self.num_envs, dtype=torch.uint8, device=self.device)
            with torch.cuda.device(self.device):
                L.check(self.lib.scg_gate_info(self._h, C.c_void_p(pose.data_ptr()), C.c_void_p(rng.data_ptr()), self._stream()))
            self._gate_cache = (pose.t(), rng.bool())
            self._gate_cache_tag = self.launch_count
        return self._gate_cache

    def _extra_info(self):
        if self.QUAD_TYPE != 3 or self.NUM_GATES == 0:
            return {}
        if getattr(self, "_gate_makers", None) is None:   # built once: step() must not touch the device for lazy keys
            def gate_type():
                if getattr(self, "_gate_types", None) is None:
                    self._gate_types = torch.tensor([int(g[6]) for g in self.GATES] + [-1], dtype=torch.int32, device=self.device)
                g = self._gate_id().long()
                return self._gate_types[torch.where(g < 0, torch.full_like(g, self.NUM_GATES), g)]
            self._gate_makers = {"current_target_gate_in_range": lambda: self._gate_info()[1],
                                 "current_target_gate_pos": lambda: self._gate_info()[0],
                                 "current_target_gate_type": gate_type}
        return self._gate_makers

    def _reset_info(self):
        """quadrotor.py:1136-1202 (host constants only; nothing here touches the device)."""
        info = {"symbolic_model": self.symbolic,
                "nominal_physical_parameters": {"quadrotor_mass": self.MASS, "quadrotor_ixx_inertia": self.J[0, 0],
                                                "quadrotor_iyy_inertia": self.J[1, 1], "quadrotor_izz_inertia": self.J[2, 2]},
                "x_reference": self.X_GOAL, "u_reference": self.U_GOAL,
                "symbolic_constraints": self.constraints.get_all_symbolic_models() if self.constraints is not None else {},
                "ctrl_timestep": self.CTRL_TIMESTEP, "ctrl_freq": self.CTRL_FREQ, "episode_len_sec": self.EPISODE_LEN_SEC,
                "quadrotor_kf": self.KF, "quadrotor_km": self.KM,
                "gate_dimensions": {"tall": {"shape": "square", "height": 1.0, "edge": 0.45},
                                    "low": {"shape": "square", "height": 0.525, "edge": 0.45}},
                "obstacle_dimensions": {"shape": "cylinder", "height": 1.05, "radius": 0.05},
                "nominal_gates_pos_and_type": self.GATES, "nominal_obstacles_pos": self.OBSTACLES,
                "initial_state_randomization": self.INIT_STATE_RAND_INFO if self.RANDOMIZED_INIT else {},
                "inertial_prop_randomization": self.INERTIAL_PROP_RAND_INFO if self.RANDOMIZED_INERTIAL_PROP else {},
                "gates_and_obs_randomization": self.GATES_AND_OBS_RAND_INFO if self.RANDOMIZED_GATES_AND_OBS else {},
                "disturbances": self.DISTURBANCES}
        return info

    # ------------------------------------------------------------------ scg_config
    def _make_config(self, precision):
        c = L.Config()
        nx, nu = self.state_dim, self.action_dim
        c.env_kind, c.precision, c.attitude_mode = self.QUAD_TYPE, precision, self._attitude
        c.substeps, c.ctrl_steps = self.PYB_STEPS_PER_CTRL, int(self.CTRL_STEPS)
        c.ctrl_freq, c.pyb_freq, c.pyb_dt = int(self.CTRL_FREQ), int(self.PYB_FREQ), self.PYB_TIMESTEP
        c.mass, c.gravity_acc, c.gravity_force = float(self.MASS), self.GRAVITY_ACC, float(self.GRAVITY)
        for j in range(3):
            c.J[j], c.J_inv[j] = float(self.J[j, j]), float(self.J_INV[j, j])   # J_INV stays the URDF inverse (F9)
        c.kf, c.km, c.arm = self.KF, self.KM, self.L
        c.pwm2rpm_scale, c.pwm2rpm_const, c.pwm_min, c.pwm_max = self.PWM2RPM_SCALE, self.PWM2RPM_CONST, self.MIN_PWM, self.MAX_PWM
        c.normalized_action = int(self.NORMALIZED_RL_ACTION_SPACE)
        c.norm_act_scale = float(self.norm_act_scale)
        c.hover_thrust = float(getattr(self, "hover_thrust", 0.0))
        for j in range(nu):
            c.act_low[j], c.act_high[j] = float(self.action_space.low[j]), float(self.action_space.high[j])
        for d in range(nx):
            c.state_low[d], c.state_high[d] = float(self.state_space.low[d]), float(self.state_space.high[d])
            c.mse_weight[d] = float(self.info_mse_metric_state_weight[d])
        c.done_on_out_of_bound = int(bool(self.done_on_out_of_bound))
        c.task = L.TASK_STABILIZATION if self.TASK == "stabilization" else L.TASK_TRAJ_TRACKING
        c.cost = {"rl_reward": L.COST_RL_REWARD, "quadratic": L.COST_QUADRATIC, "competition": L.COST_COMPETITION}[self.COST]
        c.rew_exponential, c.obs_goal_horizon = int(bool(self.rew_exponential)), self.obs_goal_horizon
        rsw = np.broadcast_to(self.rew_state_weight, (nx,)) if self.rew_state_weight.size in (1, nx) else None
        raw_ = np.broadcast_to(self.rew_act_weight, (nu,)) if self.rew_act_weight.size in (1, nu) else None
        if rsw is None or raw_ is None:
            raise ValueError("rew_state_weight / rew_act_weight must be scalars or match the state / action size")
        for d in range(nx):
            c.rew_state_weight[d] = float(rsw[d])
        for j in range(nu):
            c.rew_act_weight[j] = float(raw_[j])
            c.u_goal[j] = float(self.U_GOAL[j])
        if hasattr(self, "Q"):
            Q, R = np.asarray(self.Q, dtype=float), np.asarray(self.R, dtype=float)
            if Q.shape[0] < nx or R.shape[0] < nu:
                raise ValueError("Q / R have the wrong shape")
            for a in range(nx):
                for b in range(nx):
                    c.Q[a * L.MAX_NX + b] = float(Q[a, b])
            for a in range(nu):
                for b in range(nu):
                    c.R[a * L.MAX_NU + b] = float(R[a, b])
        c.goal_tolerance = float(self.TASK_INFO.get("stabilization_goal_tolerance", 0.05))
        goal = np.ascontiguousarray(np.atleast_2d(self.X_GOAL), dtype=np.float64)
        self._goal_host = goal   # keep alive across scg_create
        c.n_goal_rows = goal.shape[0]
        c.x_goal = goal.ctypes.data_as(C.POINTER(C.c_double))
        c.done_on_violation = int(bool(self.DONE_ON_VIOLATION))
        c.done_on_collision, c.done_on_completion = int(self.DONE_ON_COLLISION), int(self.DONE_ON_COMPLETION)
        c.use_constraint_penalty, c.constraint_penalty = int(bool(self.use_constraint_penalty)), float(self.constraint_penalty)
        if self.constraints is not None:
            recs = self.constraints.to_records()
            if len(recs) > L.MAX_CONSTRAINTS:
                raise ValueError(f"at most {L.MAX_CONSTRAINTS} constraint entries are supported")
            c.n_constraints = len(recs)
            for i, r in enumerate(recs):
                c.constraints[i] = r
        for mode, ch in (("observation", L.CH_OBSERVATION), ("action", L.CH_ACTION), ("dynamics", L.CH_DYNAMICS)):
            if mode in self.disturbances:
                lst = self.disturbances[mode].disturbances
                c.n_dist[ch] = len(lst)
                for k, d in enumerate(lst):
                    c.dist[ch][k] = d.to_record()
        c.adversary_channel = {None: -1, "action": L.CH_ACTION, "dynamics": L.CH_DYNAMICS}.get(self.adversary_disturbance, -2)
        if c.adversary_channel == -2:
            raise NotImplementedError("adversary disturbance on the observation channel is not used by the reference env")
        labels = _LABELS[self.QUAD_TYPE]
        for d, name in enumerate(labels):
            c.init_state[d] = float(getattr(self, name.upper()))
            info = self.INIT_STATE_RAND_INFO.get(name) if self.RANDOMIZED_INIT else None
            c.init_rand[d] = _rand_record(info)
        nominal = [self.MASS, self.J[0, 0], self.J[1, 1], self.J[2, 2]]
        for j, key in enumerate(["M", "Ixx", "Iyy", "Izz"]):
            c.inertia_nominal[j] = float(nominal[j])
            info = self.INERTIAL_PROP_RAND_INFO.get(key) if self.RANDOMIZED_INERTIAL_PROP else None
            c.inertia_rand[j] = _rand_record(info)
        c.reseed_on_reset = int(self.RESEED_ON_RESET)
        if len(self.GATES) > L.MAX_GATES or len(self.OBSTACLES) > L.MAX_OBSTACLES:
            raise ValueError("too many gates / obstacles for the device tables")
        c.n_gates, c.n_obstacles = len(self.GATES), len(self.OBSTACLES)
        for g, gate in enumerate(self.GATES):
            for j in range(7):
                c.gates[g][j] = float(gate[j])
        for g, ob in enumerate(self.OBSTACLES):
            for j in range(6):
                c.obstacles[g][j] = float(ob[j])
        if self.RANDOMIZED_GATES_AND_OBS:
            c.gate_rand = _rand_record(self.GATES_AND_OBS_RAND_INFO.get("gates"))
            c.obstacle_rand = _rand_record(self.GATES_AND_OBS_RAND_INFO.get("obstacles"))
        c.seed = self._seed_value
        c.env_id_base = self._seed_offset
        c.dyn_uses_episode_inertia = int(self._dyn_uses_episode_inertia)
        c.dyn_applies_force = int(self._dyn_applies_force)
        c.gate_geometry = int(self._gate_geometry)
        c.auto_reset = int(self._auto_reset)
        c.dyn_aero = int(self._dyn_aero) if self.QUAD_TYPE == 3 else 0
        for j, v in enumerate((CF2X["DRAG_COEFF_XY"], CF2X["DRAG_COEFF_XY"], CF2X["DRAG_COEFF_Z"])):   # base_aviary.py:639-641
            c.aero_drag_coeff[j] = v
        c.aero_gnd_eff_coeff, c.aero_prop_radius = CF2X["GND_EFF_COEFF"], CF2X["PROP_RADIUS"]
        c.aero_gnd_eff_h_clip = float(0.25 * CF2X["PROP_RADIUS"] * np.sqrt((15 * self.MAX_RPM ** 2 * self.KF * CF2X["GND_EFF_COEFF"])
                                                                          / self.MAX_THRUST))                       # base_aviary.py:145-147
        for j, (x, y) in enumerate(((0.028, 0.028), (-0.028, 0.028), (-0.028, -0.028), (0.028, -0.028))):   # cf2x.urdf prop links
            c.aero_prop_xy[j][0], c.aero_prop_xy[j][1] = x, y
        return c
