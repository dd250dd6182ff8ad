"""SURVEY.md 8f rows N1 / N2 on the real env: the episode-statistics wrapper agrees with the kernel's own
warp-reduced statistics, and a logged rollout exported in the Logger formats is the stepped trajectory."""
import numpy as np
import pytest
import torch

import bench
from safe_control_gym_b200.trajectory_log import TrajectoryLog
from safe_control_gym_b200.vec_env import VecRecordEpisodeStatistics, make_vec_envs

pytestmark = pytest.mark.gpu


def _make(num_envs, seed=None, **over):
    from safe_control_gym_b200.envs.quadrotor import BatchedQuadrotor
    kw = bench.env_kwargs()
    kw.update(over)
    if seed is not None:
        kw["seed"] = seed
    return BatchedQuadrotor(num_envs=num_envs, attitude="fast", **kw)


def test_wrapper_agrees_with_kernel_statistics():
    n, steps = 2048, 70
    w = VecRecordEpisodeStatistics(make_vec_envs(_make, batch_size=n, seed=11), deque_size=None)
    w.add_tracker("constraint_violation", 0)
    w.reset()
    w.venv.stats(reset=True)
    acts = bench.make_actions(torch, n, steps, torch.device("cuda:0"), 4, 0.005)
    for t in range(steps):
        obs, rew, done, info = w.step(acts[t])
    st = w.venv.stats()
    assert len(w.return_queue) == int(st["episodes"]) and st["episodes"] > 100
    assert sum(w.length_queue) == st["length"]
    assert abs(sum(w.return_queue) - st["return"]) <= 1e-6 * max(1.0, abs(st["return"]))
    open_viol = float(w.episode_stats["constraint_violation"].sum().item())
    assert w.accumulated_stats["constraint_violation"] + open_viol == st["violation_steps"]
    w.close()


def test_logged_rollout_exports_the_stepped_trajectory(tmp_path):
    n, k = 256, 30
    acts = bench.make_actions(torch, n, k, torch.device("cuda:0"), 9, 0.005)
    a, b = _make(n, auto_reset=False, done_on_violation=False), _make(n, auto_reset=False, done_on_violation=False)
    a.reset(); b.reset()
    log = torch.empty(k, 16, n, device="cuda")
    a.rollout(acts, log_out=log, log_stride=1)
    tl = TrajectoryLog(log, logging_freq_hz=a.CTRL_FREQ, envs=[0, 17, 255], output_folder=str(tmp_path))
    for t in range(k):
        b.step(acts[t])
        s = b.state[[0, 17, 255]].double().cpu().numpy()   # [x, xd, y, yd, z, zd, phi, theta, psi, p, q, r]
        assert np.array_equal(tl.states[:, 0:3, t], s[:, [0, 2, 4]].astype(np.float32).astype(np.float64))
        assert np.array_equal(tl.states[:, 3:6, t], s[:, [1, 3, 5]])
        assert np.array_equal(tl.states[:, 6:9, t], s[:, 6:9])
    z = np.load(tl.save())
    assert z["states"].shape == (3, 16, k) and np.all(z["states"][:, 12:16] > 0)   # rpm rows
    a.close(); b.close()


def test_runtime_error_conventions():
    """benchmark_env.py:272-275 (step before reset), :252-254 (cost parameters after the first reset), :256-266
    (adversary control without an adversary), quadrotor.py:464-465 (negative randomised inertia)."""
    env = _make(8)
    with pytest.raises(RuntimeError, match="reset"):
        env.step(torch.zeros(8, 4, device="cuda:0"))
    with pytest.raises(RuntimeError, match="reset"):
        env.rollout(torch.zeros(2, 8, 4, device="cuda:0"))
    env.set_cost_function_param(np.eye(12), np.eye(4))
    env.reset()
    with pytest.raises(RuntimeError, match="after the first reset"):
        env.set_cost_function_param(np.eye(12), np.eye(4))
    with pytest.raises(RuntimeError, match="adversary_disturbance does not exist"):
        env.set_adversary_control(np.zeros(4))
    env.close()
    with pytest.raises(ValueError, match="negative randomized inertial"):
        _make(8, randomized_inertial_prop=True,
              inertial_prop_randomization_info={"M": dict(distrib="uniform", low=-1.0, high=0.0)})


def test_vec_env_attribute_access_and_random_state():
    """vec_env.py:130-163, dummy_vec_env.py:88-143 on the batch: get_attr / set_attr / env_method, and the random state
    (seed, env-id range, episode counters) restores the exact continuation."""
    n = 64
    env = _make(n, seed=5)
    assert env.get_attr("CTRL_FREQ") == [60] * n and env.get_attr("CTRL_FREQ", indices=[3, 9]) == [60, 60]
    assert len(env.get_attr("MASS", indices=7)) == 1
    env.set_attr("my_tag", "a")
    assert env.get_attr("my_tag", 0) == ["a"]
    with pytest.raises(ValueError):
        env.set_attr("my_tag", ["a", "b"], indices=[0, 1])
    seeds = env.env_method("seed", 5, indices=[0, 1])
    assert seeds == [[5], [5]]
    env.reset()
    acts = bench.make_actions(torch, n, 60, torch.device("cuda:0"), 9, 0.02)
    env.rollout(acts[:30])
    rnd, sd = env.get_env_random_state(), env.state_dict()
    env.rollout(acts[30:])
    want = env.state.clone()
    assert int(env.stats()["episodes"]) > 0                 # resets happened, so the draws matter
    env.load_state_dict(sd)
    env.set_env_random_state(rnd)
    env.rollout(acts[30:])
    assert torch.equal(env.state, want)
    other = _make(n, seed=6)
    with pytest.raises(ValueError):
        other.set_env_random_state(rnd)
    other.close()
    env.close()
