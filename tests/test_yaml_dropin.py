"""Drop-in replay (SURVEY.md section 4 item 4, section 2 "env config must load unchanged"): the reference's own YAML configs
-- competition/level0-3.yaml, getting_started.yaml, tests/scripts/verbose_api.yaml, tests/scripts/tracking.yaml and the registered
default quadrotor.yaml, snapshotted unchanged in tests/golden/ref_yaml_configs.json -- go through ``make('quadrotor', **cfg)`` with
only ``physics: dyn`` (the path this repo implements) and ``gui: False`` (no graphics stack) overridden, and the call sequences
of the reference's scripts are replayed on them:
  tests/scripts/verbose_api.py:16-113   reset -> print every reset-info key -> 2 x (action_space.sample(), step, read info) -> close
  tests/scripts/tracking.py:18-83       env_func = partial(make, 'quadrotor', **cfg) per trajectory type, stepped over the episode
  competition/getting_started.py:60-316 ctrl_freq := 500 (firmware layout), reset info keys, per-step info keys, done -> reset
"""
import json
import os
from functools import partial

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
CONFIGS = json.load(open(os.path.join(HERE, "golden", "ref_yaml_configs.json")))
RESET_INFO_KEYS = ["symbolic_model", "nominal_physical_parameters", "x_reference", "u_reference", "symbolic_constraints", "ctrl_timestep",
                   "ctrl_freq", "episode_len_sec", "quadrotor_kf", "quadrotor_km", "gate_dimensions", "obstacle_dimensions",
                   "nominal_gates_pos_and_type", "nominal_obstacles_pos", "initial_state_randomization", "inertial_prop_randomization",
                   "gates_and_obs_randomization", "disturbances"]   # quadrotor.py:1136-1202, read at getting_started.py:108-137


def test_fixture_is_the_reference_files_unchanged():
    if not os.path.isdir("/root/reference"):
        pytest.skip("reference tree not present on this machine")
    import sys
    sys.path.insert(0, os.path.join(HERE, "golden"))
    from make_yaml_fixtures import read_all
    assert json.loads(json.dumps(read_all())) == CONFIGS


def _cfg(name, **over):
    cfg = json.loads(json.dumps(CONFIGS[name]["config"]))
    cfg.update(physics="dyn", gui=False)
    cfg.update(over)
    return cfg


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["level0", "level1", "level2", "level3", "getting_started"])
def test_competition_yaml_freqs_raise_like_the_reference(name):
    """pyb_freq 500 / ctrl_freq 30 is not divisible: BenchmarkEnv.__init__ raises (benchmark_env.py:154-155);
    getting_started.py:69-76 rewrites ctrl_freq to the firmware's 500 Hz before make()."""
    from safe_control_gym_b200 import make
    with pytest.raises(ValueError, match="pyb_freq is not divisible"):
        make("quadrotor", **_cfg(name))


@pytest.mark.gpu
@pytest.mark.parametrize("num_envs", [1, 1024])
@pytest.mark.parametrize("name,over", [("level0", dict(ctrl_freq=500)), ("level1", dict(ctrl_freq=500)), ("level2", dict(ctrl_freq=500)),
                                       ("level3", dict(ctrl_freq=500)), ("getting_started", dict(ctrl_freq=500)),
                                       ("level2", dict(ctrl_freq=60, pyb_freq=240)), ("verbose_api", {}), ("tracking", {}),
                                       ("quadrotor_default", dict(info_in_reset=True))])
def test_verbose_api_sequence(name, over, num_envs):
    """tests/scripts/verbose_api.py:16-113 on every config; num_envs = 1 with squeeze=True has the reference's scalar signatures."""
    import torch
    from safe_control_gym_b200 import make
    cfg = _cfg(name, **over)
    env = make("quadrotor", num_envs=num_envs, squeeze=True, **cfg)
    initial_obs, initial_info = env.reset()
    nx = {1: 2, 2: 6, 3: 12}[cfg["quad_type"]]
    for key in RESET_INFO_KEYS:
        assert key in initial_info, key
    sm = initial_info["symbolic_model"]
    for attr in ("x_sym", "u_sym", "x_dot", "cost_func"):
        str(getattr(sm, attr))
    for fun in initial_info["symbolic_constraints"]:
        assert callable(fun)
    assert np.shape(initial_info["x_reference"])[-1] == nx and len(initial_info["u_reference"]) == env.action_dim
    assert set(initial_info["nominal_physical_parameters"]) == {"quadrotor_mass", "quadrotor_ixx_inertia", "quadrotor_iyy_inertia",
                                                                "quadrotor_izz_inertia"}
    if num_envs == 1:
        assert isinstance(initial_obs, np.ndarray) and initial_obs.shape == (env.obs_dim,)
    else:
        assert tuple(initial_obs.shape) == (num_envs, env.obs_dim)
    episodes = 1
    for i in range(2):
        action = env.action_space.sample()
        if num_envs > 1:
            action = np.repeat(action[None], num_envs, axis=0)
        obs, reward, done, info = env.step(action)
        if env.constraints is not None:
            cv, viol = info["constraint_values"], info["constraint_violation"]
            assert np.shape(cv)[-1] == env.constraints.num_constraints
        if num_envs == 1:
            assert isinstance(obs, np.ndarray) and isinstance(reward, float) and isinstance(done, bool)
            if env.constraints is not None:
                bool(viol)
            if done:
                episodes += 1
                env.reset()
        else:
            assert tuple(obs.shape) == (num_envs, env.obs_dim) and tuple(reward.shape) == (num_envs,) and done.dtype == torch.bool
    assert env.CTRL_FREQ == cfg["ctrl_freq"] and abs(env.CTRL_TIMESTEP - 1.0 / cfg["ctrl_freq"]) < 1e-15
    env.close()


@pytest.mark.gpu
def test_getting_started_sequence():
    """competition/getting_started.py:60-316 without the firmware wrapper's internals: ctrl_freq := 500, reset, the printed reset info,
    then steps reading every per-step info key the script reads; a violation ends the episode and the env is reset."""
    from safe_control_gym_b200 import make
    cfg = _cfg("getting_started", ctrl_freq=500)
    env = make("quadrotor", num_envs=1, squeeze=True, **cfg)
    obs, info = env.reset()
    vicon_obs = [obs[0], 0, obs[2], 0, obs[4], 0, obs[6], obs[7], obs[8], 0, 0, 0]
    assert len(vicon_obs) == 12 and len(cfg["gates"]) == len(info["nominal_gates_pos_and_type"])
    for key in ("ctrl_timestep", "ctrl_freq", "episode_len_sec", "gate_dimensions", "obstacle_dimensions", "x_reference"):
        info[key]
    lo, hi = env.constraints.input_constraints[0].lower_bounds, env.constraints.input_constraints[0].upper_bounds
    assert len(lo) == 4 and np.all(hi > lo)
    assert len(env.constraints.state_constraints[0].lower_bounds) == len(cfg["constraints"][1]["active_dims"])
    hover = env.MASS * env.GRAVITY_ACC / 4
    resets = 0
    for i in range(40):
        action = np.full(4, hover if i != 20 else 0.0)      # step 20: below the input bound -> constraint violation -> done
        obs, reward, done, info = env.step(action)
        assert info["collision"][1] in (True, False)         # (object id, flag), getting_started.py:209-211
        if "constraint_values" in info and info["constraint_violation"] == True:   # noqa: E712 (the script's own comparison)
            assert i == 20 and done
        for key in ("current_target_gate_id", "current_target_gate_type", "current_target_gate_in_range", "current_target_gate_pos",
                    "at_goal_position", "task_completed", "constraint_values", "constraint_violation", "collision"):
            assert key in info, key
        assert info["current_target_gate_id"] == 0 and not info["task_completed"]
        np.sqrt(action / env.KF)                              # the RPMs the script logs (getting_started.py:243)
        if done:
            resets += 1
            new_initial_obs, _ = env.reset()
            assert new_initial_obs.shape == (12,)
    assert resets == 1
    env.close()


@pytest.mark.gpu
def test_tracking_sequence():
    """tests/scripts/tracking.py:18-83: env_func = partial(make, 'quadrotor', **cfg) for figure8 / circle / square; the controller reads
    env.X_GOAL as its reference and steps the env over the whole episode (here: the on-device DSL PID, scg_rollout_pid, through the
    same partial)."""
    import torch
    from safe_control_gym_b200 import make
    cfg = _cfg("tracking")
    iterations = int(cfg["episode_len_sec"] * cfg["ctrl_freq"])
    for traj in ("figure8", "circle", "square"):
        cfg["task_info"]["trajectory_type"] = traj
        env_func = partial(make, "quadrotor", **cfg)
        env = env_func(num_envs=8)
        reference_traj = env.X_GOAL
        assert reference_traj.shape == (iterations, 6)
        env.reset()
        rew = torch.empty(iterations, 8, device="cuda")
        env.rollout_pid(iterations, reward_out=rew)
        torch.cuda.synchronize()
        assert bool(torch.isfinite(rew).all())
        # the PID tracks the reference: mean reward (= -quadratic loss against X_GOAL[t]) over the second half of the episode
        assert float(rew[iterations // 2:].mean()) > -0.5, (traj, float(rew[iterations // 2:].mean()))
        env.close()


@pytest.mark.gpu
def test_register_with_config_entry_point():
    """envs/__init__.py:6-8: register(id, entry_point, config_entry_point) -- get_config(id) returns the registered default config
    (the keys of the reference's quadrotor.yaml), and make(id, **that) builds the env."""
    from safe_control_gym_b200 import get_config, make, register
    from safe_control_gym_b200.registration import _REGISTRY
    _REGISTRY.pop("quadrotor_yaml", None)
    register(id="quadrotor_yaml", entry_point="safe_control_gym_b200.envs.quadrotor:BatchedQuadrotor",
             config_entry_point="safe_control_gym_b200.envs:quadrotor.yaml")
    with pytest.raises(Exception, match="Cannot re-register"):
        register(id="quadrotor_yaml", entry_point="x:y")
    cfg = get_config("quadrotor_yaml")
    ref = CONFIGS["quadrotor_default"]["config"]
    assert set(cfg) == set(ref)
    assert {k: v for k, v in cfg.items() if k != "physics"} == {k: v for k, v in ref.items() if k != "physics"}
    env = make("quadrotor_yaml", num_envs=4, **cfg)
    env.reset()
    env.step(np.zeros((4, 2)))
    env.close()
