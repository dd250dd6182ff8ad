"""CPU restatement of the cart-pole env step -- TEST INFRASTRUCTURE ONLY.

PARITY UNPINNED: ``envs/gym_control/cartpole.py`` does not exist in this fork of the reference (SURVEY.md
F2).  The contract is inferred from what the fork still carries -- the ``cartpole_config`` block of
tests/scripts/verbose_api.yaml:1-91 (state [x, x_dot, theta, theta_dot], one input force, 50/50 Hz,
``prior_prop`` / ``inertial_prop`` with pole_length, pole_mass, cart_mass, disturbances on three channels,
default input + state constraints), SymmetricStateConstraint (envs/constraints.py:371-415) and BenchmarkEnv
(envs/benchmark_env.py) -- plus the standard frictionless cart-pole ODE, stepped with RK4 at the simulation
rate.  Everything BenchmarkEnv does around the dynamics (before/after_step, constraints, disturbances,
time limit, reward forms) follows the same reference lines as oracle/quad_oracle.py.
"""
import copy
import math

import numpy as np

from . import philox as px
from .quad_oracle import CHANNELS, SequentialDraws, _Constraint, _Disturbance, generate_trajectory

F32MAX = float(np.finfo(np.float32).max)
LABELS = ["init_x", "init_x_dot", "init_theta", "init_theta_dot"]
DEFAULT_INIT_RAND = {k: {"distrib": "uniform", "low": -0.05, "high": 0.05} for k in LABELS}
DEFAULT_INERTIAL_RAND = {"pole_length": {"distrib": "uniform", "low": -0.25, "high": 0.0},
                         "cart_mass": {"distrib": "uniform", "low": 0.5, "high": 1.5},
                         "pole_mass": {"distrib": "uniform", "low": 0.05, "high": 0.15}}


def cartpole_f(x, F, l, mp, mc, g=9.8):
    s, c = math.sin(x[2]), math.cos(x[2])
    Mm, ml = mc + mp, mp * l
    tmp = (F + ml * x[3] * x[3] * s) / Mm
    thdd = (g * s - c * tmp) / (l * (4.0 / 3.0 - mp * c * c / Mm))
    return np.array([x[1], tmp - ml * thdd * c / Mm, x[3], thdd])


def cartpole_rk4(x, F, dt, l, mp, mc, g=9.8):
    k1 = cartpole_f(x, F, l, mp, mc, g)
    k2 = cartpole_f(x + dt * 0.5 * k1, F, l, mp, mc, g)
    k3 = cartpole_f(x + dt * 0.5 * k2, F, l, mp, mc, g)
    k4 = cartpole_f(x + dt * k3, F, l, mp, mc, g)
    return x + dt / 6.0 * (k1 + 2.0 * k2 + 2.0 * k3 + k4)


class CartPoleOracle:
    NAME = "cartpole"

    def __init__(self, draws=None, seed=None, info_in_reset=False, normalized_rl_action_space=False,
                 task="stabilization", task_info=None, cost="rl_reward", pyb_freq=50, ctrl_freq=50,
                 episode_len_sec=5, init_state=None, randomized_init=True, init_state_randomization_info=None,
                 prior_prop=None, inertial_prop=None, randomized_inertial_prop=False,
                 inertial_prop_randomization_info=None, constraints=None, done_on_violation=False,
                 use_constraint_penalty=False, constraint_penalty=-1, disturbances=None,
                 obs_goal_horizon=0, rew_state_weight=1.0, rew_act_weight=0.0001, rew_exponential=True,
                 done_on_out_of_bound=True, info_mse_metric_state_weight=None, reseed_on_reset=False,
                 dyn_applies_force=False, **_ignored):
        self.GRAVITY_ACC = 9.8
        self.x_threshold, self.theta_threshold_radians = 2.4, 90 * math.pi / 180
        self.action_scale, self.action_threshold = 10.0, 10.0
        self.obs_goal_horizon = obs_goal_horizon
        self.rew_state_weight = np.array(rew_state_weight, ndmin=1, dtype=float)
        self.rew_act_weight = np.array(rew_act_weight, ndmin=1, dtype=float)
        self.rew_exponential, self.done_on_out_of_bound = rew_exponential, done_on_out_of_bound
        self.info_mse_metric_state_weight = np.array([1, 0, 1, 0] if info_mse_metric_state_weight is None
                                                     else info_mse_metric_state_weight, dtype=float)
        ip = dict(pole_length=0.5, pole_mass=0.1, cart_mass=1.0)
        ip.update(inertial_prop or {})
        self.EFFECTIVE_POLE_LENGTH, self.POLE_MASS, self.CART_MASS = ip["pole_length"], ip["pole_mass"], ip["cart_mass"]
        self.TASK, self.COST = task, cost
        self.TASK_INFO = copy.deepcopy(task_info or {"stabilization_goal": [0], "stabilization_goal_tolerance": 0.05,
                                                     "trajectory_type": "circle", "num_cycles": 1, "trajectory_plane": "zx",
                                                     "trajectory_position_offset": [0, 0], "trajectory_scale": 0.2})
        self.CTRL_FREQ, self.PYB_FREQ = ctrl_freq, pyb_freq
        if pyb_freq % ctrl_freq != 0:
            raise ValueError("[ERROR] in BenchmarkEnv.__init__(), pyb_freq is not divisible by env_freq.")
        self.PYB_STEPS_PER_CTRL = int(pyb_freq / ctrl_freq)
        self.CTRL_TIMESTEP, self.PYB_TIMESTEP = 1.0 / ctrl_freq, 1.0 / pyb_freq
        self.EPISODE_LEN_SEC = episode_len_sec
        self.CTRL_STEPS = episode_len_sec * ctrl_freq
        self.RANDOMIZED_INIT = randomized_init
        self.INIT_STATE_RAND_INFO = copy.deepcopy(DEFAULT_INIT_RAND if init_state_randomization_info is None
                                                  else init_state_randomization_info)
        self.RANDOMIZED_INERTIAL_PROP = randomized_inertial_prop
        self.INERTIAL_PROP_RAND_INFO = copy.deepcopy(DEFAULT_INERTIAL_RAND if inertial_prop_randomization_info is None
                                                     else inertial_prop_randomization_info)
        self.NORMALIZED_RL_ACTION_SPACE = normalized_rl_action_space
        lim = 1.0 if normalized_rl_action_space else self.action_threshold
        self.action_low, self.action_high = np.array([-lim], np.float32), np.array([lim], np.float32)
        low = np.array([-2 * self.x_threshold, -F32MAX, -2 * self.theta_threshold_radians, -F32MAX])
        self.state_low, self.state_high = low.astype(np.float32), (-low).astype(np.float32)
        self.action_dim, self.state_dim = 1, 4
        self.obs_dim = 4 * (1 + (obs_goal_horizon if task == "traj_tracking" else 1)) \
            if cost == "rl_reward" and obs_goal_horizon > 0 else 4
        if cost == "quadratic":
            self.Q, self.R = np.eye(4), np.eye(1)
        self.DONE_ON_VIOLATION = done_on_violation
        self.use_constraint_penalty, self.constraint_penalty = use_constraint_penalty, constraint_penalty
        self.constraints = None if constraints is None else [_Constraint(self, c) for c in constraints]
        self.num_constraints = 0 if constraints is None else sum(c.num_constraints for c in self.constraints)
        dims = {"observation": 4, "action": 1, "dynamics": 2}
        self.disturbances = {}
        if disturbances is not None:
            for mode, specs in disturbances.items():
                self.disturbances[mode] = [_Disturbance(self, CHANNELS[mode], k, dims[mode], s) for k, s in enumerate(specs)]
        self.RND_SEED, self.RESEED_ON_RESET = seed, reseed_on_reset
        self._user_draws = draws
        self.draws = draws if draws is not None else SequentialDraws(np.random.RandomState(seed))
        self.INFO_IN_RESET = info_in_reset
        self.initial_reset, self.cnstr_violation, self.episode = False, False, -1
        self.INIT = {k: 0.0 for k in LABELS}
        if init_state is not None:
            for i, k in enumerate(LABELS):
                self.INIT[k] = init_state[i] if isinstance(init_state, np.ndarray) else init_state.get(k, 0.0)
        self.dyn_applies_force = dyn_applies_force
        self.U_GOAL = np.zeros(1)
        ti = self.TASK_INFO
        if task == "stabilization":
            self.X_GOAL = np.hstack([ti["stabilization_goal"][0], 0.0, 0.0, 0.0])
        else:
            pos, vel = generate_trajectory(ti["trajectory_type"], episode_len_sec, ti["num_cycles"], ti["trajectory_plane"],
                                           ti["trajectory_position_offset"], ti["trajectory_scale"], self.CTRL_TIMESTEP)
            z = np.zeros(pos.shape[0])
            self.X_GOAL = np.vstack([pos[:, 0], vel[:, 0], z, z]).T

    def _randomize(self, values, info, stream):
        out, info = copy.deepcopy(values), copy.deepcopy(info)
        for slot, key in enumerate(values):
            if key in info:
                name = info[key].pop("distrib")
                args = info[key].pop("args", [])
                out[key] += self.draws.distrib(name, args, info[key], stream=stream, slot=slot, episode=self.episode)
        return out

    def reset(self):
        self.episode = 0 if self.RESEED_ON_RESET else self.episode + 1
        self.initial_reset = True
        self.pyb_step_counter = self.ctrl_step_counter = 0
        slot = 0
        for mode in self.disturbances:
            for d in self.disturbances[mode]:
                d.reset(self, slot)
                slot += 1
        self.obs_call_index = 0
        prop = {"pole_length": self.EFFECTIVE_POLE_LENGTH, "pole_mass": self.POLE_MASS, "cart_mass": self.CART_MASS}
        if self.RANDOMIZED_INERTIAL_PROP:
            prop = self._randomize(prop, self.INERTIAL_PROP_RAND_INFO, px.STREAM_RESET_INERTIA)
            if any(v < 0 for v in prop.values()):
                raise ValueError("[ERROR] in CartPole.reset(), negative randomized inertial properties.")
        self.OVERRIDDEN = prop
        init = dict(self.INIT)
        if self.RANDOMIZED_INIT:
            init = self._randomize(init, self.INIT_STATE_RAND_INFO, px.STREAM_RESET_INIT)
        self.state = np.array([init[k] for k in LABELS], dtype=float)
        obs = self._get_observation()
        info = {"x_reference": self.X_GOAL, "u_reference": self.U_GOAL}
        if self.constraints is not None:
            vals = [c.value(self) for c in self.constraints if c.var == "state"]
            info["constraint_values"] = np.concatenate(vals) if vals else np.zeros(0)
        return (obs, info) if self.INFO_IN_RESET else obs

    def _get_observation(self):
        obs = self.state.copy()
        for d in self.disturbances.get("observation", []):
            obs = d.apply(obs, self, self.obs_call_index)
        self.obs_call_index += 1
        if self.COST == "rl_reward" and self.obs_goal_horizon > 0:
            if self.TASK == "traj_tracking":
                idx = [min(self.ctrl_step_counter + 1 + i, self.X_GOAL.shape[0] - 1) for i in range(self.obs_goal_horizon)]
                return np.concatenate([obs, self.X_GOAL[idx].flatten()])
            return np.concatenate([obs, self.X_GOAL.flatten()])
        return obs

    def _goal_at(self, idx):
        return self.X_GOAL if self.TASK == "stabilization" else self.X_GOAL[idx]

    def step(self, action):
        if not self.initial_reset:
            raise RuntimeError("[ERROR] You must call env.reset() at least once before using env.step().")
        action = np.atleast_1d(np.asarray(action, dtype=float))
        self.current_raw_input_action = action
        a = np.clip(action, self.action_low.astype(float), self.action_high.astype(float))
        force = a * self.action_scale if self.NORMALIZED_RL_ACTION_SPACE else a
        self.current_preprocessed_action = force
        t = self.ctrl_step_counter
        for d in self.disturbances.get("action", []):
            force = d.apply(force, self, t)
        ext = np.zeros(2)
        for d in self.disturbances.get("dynamics", []):
            ext = d.apply(ext, self, t)
        F = float(force[0]) + (float(ext[0]) if self.dyn_applies_force else 0.0)
        p = self.OVERRIDDEN
        for _ in range(self.PYB_STEPS_PER_CTRL):
            self.state = cartpole_rk4(self.state, F, self.PYB_TIMESTEP, p["pole_length"], p["pole_mass"], p["cart_mass"],
                                      self.GRAVITY_ACC)
        obs = self._get_observation()
        info = {}
        err = (self.state - self._goal_at(min(t, self.X_GOAL.shape[0] - 1))) * self.info_mse_metric_state_weight
        info["mse"] = np.sum(err ** 2)
        done = False
        if self.TASK == "stabilization" and self.COST == "quadratic":
            info["goal_reached"] = bool(np.linalg.norm(self.state - self.X_GOAL) < self.TASK_INFO["stabilization_goal_tolerance"])
            done = info["goal_reached"]
        if self.done_on_out_of_bound:
            x, th = self.state[0], self.state[2]
            if x < -self.x_threshold or x > self.x_threshold or th < -self.theta_threshold_radians or th > self.theta_threshold_radians:
                done = True
        if self.COST == "rl_reward":
            e = self.state - self._goal_at(min(t, self.X_GOAL.shape[0] - 1))
            ea = self.current_preprocessed_action - self.U_GOAL
            rew = -(np.sum(self.rew_state_weight * e * e) + np.sum(self.rew_act_weight * ea * ea))
            rew = np.exp(rew) if self.rew_exponential else rew
        else:
            ex, eu = self.state - self._goal_at(t), self.current_preprocessed_action - self.U_GOAL
            rew = float(-(0.5 * ex @ self.Q @ ex + 0.5 * eu @ self.R @ eu))
        self.pyb_step_counter += self.PYB_STEPS_PER_CTRL
        self.ctrl_step_counter += 1
        if self.constraints is not None:
            cs_ = [c.value(self) for c in self.constraints]
            info["constraint_values"] = np.concatenate(cs_)
            viol = any(c.violated(v) for c, v in zip(self.constraints, cs_))
            self.cnstr_violation = viol
            info["constraint_violation"] = int(viol)
            if viol and self.DONE_ON_VIOLATION:
                done = True
            if self.COST == "rl_reward" and self.use_constraint_penalty and \
                    any(c.almost_active(v) for c, v in zip(self.constraints, cs_)):
                rew += self.constraint_penalty
        if self.ctrl_step_counter >= self.CTRL_STEPS:
            info["TimeLimit.truncated"] = not done
            done = True
        return obs, rew, done, info
