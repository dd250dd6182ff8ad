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
