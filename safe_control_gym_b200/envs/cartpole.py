"""Batched CartPole behind the reference's env API.

PARITY UNPINNED: ``envs/gym_control/cartpole.py`` is absent from this fork of the reference.
The constructor follows the ``cartpole_config`` block that the fork still ships
(tests/scripts/verbose_api.yaml:1-91) and BenchmarkEnv (envs/benchmark_env.py:63-218): state
[x, x_dot, theta, theta_dot], one input (force on the cart; +-10 N, or [-1, 1] x 10 when
``normalized_rl_action_space``), ``inertial_prop`` / ``prior_prop`` with pole_length (effective length l of
the standard frictionless cart-pole ODE), pole_mass, cart_mass, out-of-bound at |x| > 2.4 or |theta| > 90 deg.
The true model (per-episode randomised properties) is integrated with RK4 at the simulation rate inside
the fused kernel (csrc/scg_env.cuh: cartpole_rk4); ``env.symbolic`` is the prior model built from
``prior_prop`` (csrc/scg_symbolic.cu).  Checked against oracle/cartpole_oracle.py.
"""
import copy
import ctypes as C
import math

import numpy as np
import torch

from .. import _lib as L
from ..constraints import GENERAL_CONSTRAINTS, SymmetricStateConstraint, create_constraint_list
from ..disturbances import create_disturbance_list
from ..spaces import Box
from ..symbolic import SymbolicModel
from ..trajectories import planar_reference
from .batched_env import BatchedEnvBase
from .quadrotor import _rand_record, _uniform

_LABELS = ["init_x", "init_x_dot", "init_theta", "init_theta_dot"]
_PROPS = ["pole_length", "pole_mass", "cart_mass"]
_F32MAX = float(np.finfo(np.float32).max)


class BatchedCartPole(BatchedEnvBase):
    NAME = "cartpole"
    AVAILABLE_CONSTRAINTS = dict(copy.deepcopy(GENERAL_CONSTRAINTS), abs_bound=SymmetricStateConstraint)
    INIT_STATE_RAND_INFO = {k: _uniform(-0.05, 0.05) for k in _LABELS}
    INERTIAL_PROP_RAND_INFO = {"pole_length": _uniform(-0.25, 0.0), "cart_mass": _uniform(0.5, 1.5),
                               "pole_mass": _uniform(0.05, 0.15)}
    TASK_INFO = {"stabilization_goal": [0], "stabilization_goal_tolerance": 0.05, "trajectory_type": "circle",
                 "num_cycles": 1, "trajectory_plane": "zx", "trajectory_position_offset": [0, 0], "trajectory_scale": 0.2}

    def __init__(self, num_envs=1, device="cuda", dtype=torch.float32, attitude="fast", auto_reset=True, squeeze=False, seed_offset=0,
                 output_dir=None, seed=None, info_in_reset=False, gui=False, verbose=False,
                 normalized_rl_action_space=False, task="stabilization", task_info=None, cost="rl_reward",
                 pyb_freq=50, ctrl_freq=50, episode_len_sec=5, init_state=None, randomized_init=True,
                 init_state_randomization_info=None, prior_prop=None, inertial_prop=None,
                 randomized_inertial_prop=False, inertial_prop_randomization_info=None, constraints=None,
                 done_on_violation=False, use_constraint_penalty=False, constraint_penalty=-1, disturbances=None,
                 adversary_disturbance=None, adversary_disturbance_offset=0.0, adversary_disturbance_scale=0.01,
                 obs_goal_horizon=0, rew_state_weight=1.0, rew_act_weight=0.0001, rew_exponential=True,
                 done_on_out_of_bound=True, info_mse_metric_state_weight=None, dyn_applies_force=False, **kwargs):
        if gui:
            raise NotImplementedError("gui / rendering: unavailable on this box (no graphics stack)")
        # how the fp32 kernels evaluate sin / cos of the pole angle: "fast" = ~1-ulp polynomial pair, "fast_intrinsic" = MUFU.SIN / MUFU.COS
        self._attitude = {"fast": L.ATT_FAST, "fast_intrinsic": L.ATT_FAST_INTRINSIC}[attitude]
        if self._attitude == L.ATT_FAST_INTRINSIC and dtype != torch.float32:
            raise ValueError("attitude='fast_intrinsic' is an fp32-only mode")
        nx, nu = 4, 1
        self.GRAVITY_ACC = 9.8
        self.x_threshold, self.theta_threshold_radians = 2.4, 90 * math.pi / 180
        self.action_scale, self.action_threshold = 10.0, 10.0
        self.obs_goal_horizon = int(obs_goal_horizon)
        self.rew_state_weight = np.array(rew_state_weight, ndmin=1, dtype=float)
        self.rew_act_weight = np.array(rew_act_weight, ndmin=1, dtype=float)
        self.rew_exponential, self.done_on_out_of_bound = rew_exponential, done_on_out_of_bound
        w = [1, 0, 1, 0] if info_mse_metric_state_weight is None else info_mse_metric_state_weight
        if len(w) != nx:
            raise ValueError("[ERROR] in CartPole.__init__(), wrong info_mse_metric_state_weight argument size.")
        self.info_mse_metric_state_weight = np.array(w, dtype=float)
        ip = dict(pole_length=0.5, pole_mass=0.1, cart_mass=1.0)
        if inertial_prop is not None:
            if not isinstance(inertial_prop, dict):
                raise ValueError("[ERROR] in CartPole.__init__(), inertial_prop incorrect format.")
            ip.update(inertial_prop)
        self.EFFECTIVE_POLE_LENGTH, self.POLE_MASS, self.CART_MASS = float(ip["pole_length"]), float(ip["pole_mass"]), float(ip["cart_mass"])
        self.PRIOR_PROP = dict(pole_length=self.EFFECTIVE_POLE_LENGTH, pole_mass=self.POLE_MASS, cart_mass=self.CART_MASS)
        self.PRIOR_PROP.update(prior_prop or {})
        self.GUI, self.VERBOSE, self.output_dir = False, verbose, output_dir
        self.TASK = str(getattr(task, "value", task))
        self.COST = str(getattr(cost, "value", cost))
        if self.TASK not in ("stabilization", "traj_tracking"):
            raise ValueError(f"'{task}' is not a valid Task")
        if self.COST not in ("rl_reward", "quadratic"):
            raise ValueError(f"'{cost}' is not a valid Cost")
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
        self.INERTIAL_PROP, self.RANDOMIZED_INERTIAL_PROP = inertial_prop, randomized_inertial_prop
        self.INERTIAL_PROP_RAND_INFO = copy.deepcopy(self.INERTIAL_PROP_RAND_INFO if inertial_prop_randomization_info is None
                                                     else inertial_prop_randomization_info)
        self.NORMALIZED_RL_ACTION_SPACE = normalized_rl_action_space
        lim = 1.0 if normalized_rl_action_space else self.action_threshold
        self.action_space = Box(low=np.full(nu, -lim, np.float32), high=np.full(nu, lim, np.float32), dtype=np.float32)
        self.ACTION_LABELS, self.ACTION_UNITS = ["U"], ["N" if not normalized_rl_action_space else "-"]
        self.STATE_LABELS, self.STATE_UNITS = ["x", "x_dot", "theta", "theta_dot"], ["m", "m/s", "rad", "rad/s"]
        low = np.array([-2 * self.x_threshold, -_F32MAX, -2 * self.theta_threshold_radians, -_F32MAX])
        self.state_space = Box(low=low, high=-low, dtype=np.float32)
        olow = low
        if self.COST == "rl_reward" and self.obs_goal_horizon > 0:
            olow = np.concatenate([low] * (1 + self.obs_goal_horizon if self.TASK == "traj_tracking" else 2))
        self.observation_space = Box(low=olow, high=-olow, dtype=np.float32)
        self.action_dim, self.obs_dim, self.state_dim = nu, self.observation_space.shape[0], nx
        if self.COST == "quadratic":
            self.Q, self.R = np.eye(nx), np.eye(nu)
        self.CONSTRAINTS, self.DONE_ON_VIOLATION = constraints, done_on_violation
        self.use_constraint_penalty, self.constraint_penalty = use_constraint_penalty, constraint_penalty
        self.constraints, self.num_constraints = None, 0
        if constraints is not None:
            self.constraints = create_constraint_list(constraints, self.AVAILABLE_CONSTRAINTS, self)
            self.num_constraints = self.constraints.num_constraints
        self.DISTURBANCES = disturbances
        self.DISTURBANCE_MODES = {"observation": {"dim": nx}, "action": {"dim": nu}, "dynamics": {"dim": 2}}
        self.disturbances, slot = {}, 0
        if disturbances is not None:
            for mode, specs in disturbances.items():
                assert mode in self.DISTURBANCE_MODES, "[ERROR] in BenchmarkEnv._setup_disturbances(), disturbance mode not available."
                self.disturbances[mode] = create_disturbance_list(specs, dict(self.DISTURBANCE_MODES[mode]), self)
                for d in self.disturbances[mode]:
                    d.slot = slot
                    slot += 1
        self.adversary_disturbance = adversary_disturbance
        self.adversary_disturbance_offset, self.adversary_disturbance_scale = adversary_disturbance_offset, adversary_disturbance_scale
        if adversary_disturbance is not None:
            assert adversary_disturbance in self.DISTURBANCE_MODES, "[ERROR] in Cartpole._setup_disturbances()"
            self.adversary_action_space = Box(low=-1, high=1, shape=(self.DISTURBANCE_MODES[adversary_disturbance]["dim"],))
            self.adversary_observation_space = self.observation_space
        self.RND_SEED, self.RESEED_ON_RESET, self.INFO_IN_RESET = seed, bool(kwargs.get("reseed_on_reset", False)), info_in_reset
        self.seed(seed)
        init = {k: 0.0 for k in _LABELS}
        if init_state is not None:
            if isinstance(init_state, np.ndarray):
                init = {k: float(init_state[i]) for i, k in enumerate(_LABELS)}
            elif isinstance(init_state, dict):
                init = {k: init_state.get(k, 0.0) for k in _LABELS}
            else:
                raise ValueError("[ERROR] in CartPole.__init__(), init_state incorrect format.")
        for k, v in init.items():
            setattr(self, k.upper(), v)
        self.U_GOAL = np.zeros(nu)
        ti = self.TASK_INFO
        if self.TASK == "stabilization":
            self.X_GOAL = np.hstack([ti["stabilization_goal"][0], 0.0, 0.0, 0.0])
        else:
            pos, vel, _ = planar_reference(ti["trajectory_type"], self.EPISODE_LEN_SEC, ti["num_cycles"], ti["trajectory_plane"],
                                           ti["trajectory_position_offset"], ti["trajectory_scale"], self.CTRL_TIMESTEP)
            z = np.zeros(pos.shape[0])
            self.X_GOAL = np.vstack([pos[:, 0], vel[:, 0], z, z]).transpose()
        self.X_EQ, self.U_EQ = np.zeros(nx), self.U_GOAL
        if self.RANDOMIZED_INERTIAL_PROP:
            nominal = dict(zip(_PROPS, [self.EFFECTIVE_POLE_LENGTH, self.POLE_MASS, self.CART_MASS]))
            for key, info in self.INERTIAL_PROP_RAND_INFO.items():
                if info.get("distrib") == "uniform" and nominal[key] + info.get("low", 0.0) < 0:
                    raise ValueError("[ERROR] in CartPole.reset(), negative randomized inertial properties.")
        self._dyn_applies_force, self._auto_reset, self._seed_offset = bool(dyn_applies_force), bool(auto_reset), int(seed_offset)
        precision = 32 if dtype == torch.float32 else 64
        cfg = self._make_config(precision, prior=False)
        rows = self.num_constraints
        self._n_rows = rows
        self._init_batch(cfg, num_envs, device, dtype, nx, nu, self.obs_dim, rows, slot, squeeze)
        idx, r = [], 0
        if self.constraints is not None:
            for c in self.constraints.constraints:
                if c.constrained_variable == "state":
                    idx.extend(range(r, r + c.num_constraints))
                r += c.num_constraints
        self._state_row_idx = torch.tensor(idx, dtype=torch.long, device=self.device)
        # the prior model lives in its own small handle (prior_prop instead of the true properties)
        prior_cfg = self._make_config(precision, prior=True)
        with torch.cuda.device(self.device):
            hp = C.c_void_p()
            L.check(self.lib.scg_create(C.byref(prior_cfg), C.byref(hp)))
        self._h_prior = hp
        self.symbolic = SymbolicModel(self, handle=hp)
        self.current_raw_input_action = self.current_preprocessed_action = self.adv_action = None

    def seed(self, seed=None):
        self.np_random = np.random.RandomState(None if seed is None else int(seed) % (2 ** 32))
        self._seed_value = 0 if seed is None else int(seed) % (2 ** 32)
        self.action_space.seed(seed)
        if getattr(self, "_h", None) is not None and self._h.value:   # after construction: re-key the device draws
            L.check(self.lib.scg_set_seed(self._h, self._seed_value))
            self._cfg.seed = self._seed_value
        return [seed]

    def set_adversary_control(self, action):
        if self.adversary_disturbance is None:
            raise RuntimeError("[ERROR] adversary_disturbance does not exist, env.set_adversary_control() cannot be called.")
        a = self._as_real(action, (self.num_envs, self.adversary_action_space.shape[0]))
        self._adv = (torch.clamp(a, -1.0, 1.0) * self.adversary_disturbance_scale + self.adversary_disturbance_offset).contiguous()
        L.check(self.lib.scg_set_injection(self._h, -1, 0, C.c_void_p(self._adv.data_ptr())))

    def close(self):
        if getattr(self, "_h_prior", None) is not None and self._h_prior.value:
            self.lib.scg_destroy(self._h_prior)
            self._h_prior = C.c_void_p()
        super().close()

    def _reset_info(self):
        return {"symbolic_model": self.symbolic, "x_reference": self.X_GOAL, "u_reference": self.U_GOAL,
                "physical_parameters": {"pole_effective_length": self.EFFECTIVE_POLE_LENGTH, "pole_mass": self.POLE_MASS,
                                        "cart_mass": self.CART_MASS},
                "symbolic_constraints": self.constraints.get_all_symbolic_models() if self.constraints is not None else {}}

    def _make_config(self, precision, prior):
        c = L.Config()
        nx, nu = 4, 1
        c.env_kind, c.precision, c.attitude_mode = L.CARTPOLE, precision, self._attitude
        c.substeps, c.ctrl_steps = self.PYB_STEPS_PER_CTRL, int(self.CTRL_STEPS)
        c.ctrl_freq, c.pyb_freq, c.pyb_dt = int(self.CTRL_FREQ), int(self.PYB_FREQ), self.PYB_TIMESTEP
        c.mass, c.gravity_acc, c.gravity_force = 1.0, self.GRAVITY_ACC, self.GRAVITY_ACC
        for j in range(3):
            c.J[j], c.J_inv[j] = 1.0, 1.0
        c.kf, c.km, c.arm = 1.0, 1.0, 1.0
        c.pwm2rpm_scale, c.pwm2rpm_const, c.pwm_min, c.pwm_max = 1.0, 0.0, 0.0, 1.0
        src = self.PRIOR_PROP if prior else dict(pole_length=self.EFFECTIVE_POLE_LENGTH, pole_mass=self.POLE_MASS, cart_mass=self.CART_MASS)
        c.cp_length, c.cp_pole_mass, c.cp_cart_mass = float(src["pole_length"]), float(src["pole_mass"]), float(src["cart_mass"])
        c.normalized_action, c.norm_act_scale, c.hover_thrust = int(self.NORMALIZED_RL_ACTION_SPACE), 0.0, self.action_scale
        c.act_low[0], c.act_high[0] = float(self.action_space.low[0]), float(self.action_space.high[0])
        thr = [self.x_threshold, _F32MAX, self.theta_threshold_radians, _F32MAX]   # out-of-bound thresholds
        for d in range(nx):
            c.state_low[d], c.state_high[d] = -thr[d], thr[d]
            c.mse_weight[d] = float(self.info_mse_metric_state_weight[d])
        c.done_on_out_of_bound = int(bool(self.done_on_out_of_bound))
        c.task = L.TASK_STABILIZATION if self.TASK == "stabilization" else L.TASK_TRAJ_TRACKING
        c.cost = L.COST_RL_REWARD if self.COST == "rl_reward" else L.COST_QUADRATIC
        c.rew_exponential, c.obs_goal_horizon = int(bool(self.rew_exponential)), self.obs_goal_horizon
        rsw = np.broadcast_to(self.rew_state_weight, (nx,))
        for d in range(nx):
            c.rew_state_weight[d] = float(rsw[d])
        c.rew_act_weight[0] = float(np.broadcast_to(self.rew_act_weight, (nu,))[0])
        c.u_goal[0] = 0.0
        if hasattr(self, "Q"):
            Q, R = np.asarray(self.Q, float), np.asarray(self.R, float)
            for a in range(nx):
                for b in range(nx):
                    c.Q[a * L.MAX_NX + b] = float(Q[a, b])
            c.R[0] = float(R[0, 0])
        c.goal_tolerance = float(self.TASK_INFO.get("stabilization_goal_tolerance", 0.05))
        goal = np.ascontiguousarray(np.atleast_2d(self.X_GOAL), dtype=np.float64)
        self._goal_host = goal
        c.n_goal_rows, c.x_goal = goal.shape[0], goal.ctypes.data_as(C.POINTER(C.c_double))
        c.done_on_violation = int(bool(self.DONE_ON_VIOLATION))
        c.use_constraint_penalty, c.constraint_penalty = int(bool(self.use_constraint_penalty)), float(self.constraint_penalty)
        if self.constraints is not None and not prior:
            recs = self.constraints.to_records()
            c.n_constraints = len(recs)
            for i, r in enumerate(recs):
                c.constraints[i] = r
        if not prior:
            for mode, ch in (("observation", L.CH_OBSERVATION), ("action", L.CH_ACTION), ("dynamics", L.CH_DYNAMICS)):
                if mode in self.disturbances:
                    lst = self.disturbances[mode].disturbances
                    c.n_dist[ch] = len(lst)
                    for k, d in enumerate(lst):
                        c.dist[ch][k] = d.to_record()
        c.adversary_channel = {None: -1, "action": L.CH_ACTION, "dynamics": L.CH_DYNAMICS}[self.adversary_disturbance] if not prior else -1
        for d, name in enumerate(_LABELS):
            c.init_state[d] = float(getattr(self, name.upper()))
            c.init_rand[d] = _rand_record(self.INIT_STATE_RAND_INFO.get(name) if self.RANDOMIZED_INIT else None)
        nominal = [self.EFFECTIVE_POLE_LENGTH, self.POLE_MASS, self.CART_MASS, 0.0]
        for j in range(4):
            c.inertia_nominal[j] = float(nominal[j])
            key = _PROPS[j] if j < 3 else None
            c.inertia_rand[j] = _rand_record(self.INERTIAL_PROP_RAND_INFO.get(key) if (self.RANDOMIZED_INERTIAL_PROP and key) else None)
        c.reseed_on_reset = int(self.RESEED_ON_RESET)
        c.seed, c.env_id_base = self._seed_value, self._seed_offset
        c.dyn_applies_force, c.auto_reset = int(self._dyn_applies_force), int(self._auto_reset)
        return c
