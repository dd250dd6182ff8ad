"""Host-side 'next' rows that need no GPU: the Logger-format exporter on a synthetic record
and the episode-statistics wrapper on a scripted stand-in env, against a literal re-run of the reference's
per-env Python loop (record_episode_statistics.py:140-169)."""
import glob
import os
from collections import deque

import numpy as np
import torch

from safe_control_gym_b200.trajectory_log import TrajectoryLog
from safe_control_gym_b200.vec_env import VecRecordEpisodeStatistics, make_vec_envs


def test_trajectory_log_formats(tmp_path):
    T, N = 7, 5
    rng = np.random.default_rng(0)
    log = rng.normal(size=(T, 16, N))
    tl = TrajectoryLog(torch.as_tensor(log), logging_freq_hz=60, envs=[3, 1], output_folder=str(tmp_path))
    assert tl.states.shape == (2, 16, T) and np.array_equal(tl.states[0], log[:, :, 3].T)
    assert np.allclose(tl.timestamps[1], np.arange(T) / 60)
    z = np.load(tl.save())
    assert set(z.files) == {"timestamps", "states", "controls"} and z["controls"].shape == (2, 12, T)
    assert np.array_equal(z["states"], tl.states)
    d = tl.save_as_csv("unit")
    names = sorted(os.path.basename(p) for p in glob.glob(os.path.join(d, "*.csv")))
    assert len(names) == 2 * 23 and "ya0.csv" in names and "pwm3-1.csv" in names and "yar1.csv" in names
    x0 = np.loadtxt(os.path.join(d, "x0.csv"), delimiter=",")
    assert x0.shape == (T, 2) and np.allclose(x0[:, 1], log[:, 0, 3]) and np.allclose(x0[:, 0], np.arange(T) / 60)
    rr = np.loadtxt(os.path.join(d, "rr1.csv"), delimiter=",")
    assert rr[0, 1] == 0 and np.allclose(rr[1:, 1], np.diff(log[:, 6, 1]) * 60)
    pwm = np.loadtxt(os.path.join(d, "pwm2-0.csv"), delimiter=",")
    assert np.allclose(pwm[:, 1], (log[:, 14, 3] - 4070.3) / 0.2685)


class _Scripted:
    """num_envs lock-step envs with scripted rewards / dones / info (auto-reset semantics are the env's own)."""

    def __init__(self, num_envs, seed=0, steps=40):
        g = np.random.default_rng(seed)
        self.num_envs, self.device = num_envs, torch.device("cpu")
        self.observation_space = self.action_space = None
        self.rew = g.normal(size=(steps, num_envs))
        self.done = g.random((steps, num_envs)) < 0.15
        self.viol = (g.random((steps, num_envs)) < 0.3).astype(np.int32)
        self.mse = g.random((steps, num_envs))
        self.t = 0

    def reset(self):
        self.t = 0
        return torch.zeros(self.num_envs, 2)

    def step(self, a):
        t = self.t
        self.t += 1
        info = {"constraint_violation": torch.as_tensor(self.viol[t]), "mse": torch.as_tensor(self.mse[t])}
        return torch.zeros(self.num_envs, 2), torch.as_tensor(self.rew[t]), torch.as_tensor(self.done[t]), info

    def close(self):
        pass


def test_vec_record_episode_statistics_matches_reference_loop():
    n, steps, dq = 6, 40, 10
    env = make_vec_envs(lambda num_envs, seed: _Scripted(num_envs, seed, steps), batch_size=n, n_processes=4, seed=3)
    w = VecRecordEpisodeStatistics(env, deque_size=dq)
    w.add_tracker("constraint_violation", 0)
    w.add_tracker("mse", 0, mode="queue")
    w.reset()
    # the reference's loop, literally
    ep_ret, ep_len = np.zeros(n), np.zeros(n)
    rq, lq, mq = deque(maxlen=dq), deque(maxlen=dq), deque(maxlen=dq)
    stats = {"constraint_violation": [0] * n, "mse": [0] * n}
    acc = 0
    for t in range(steps):
        obs, rew, done, info = w.step(None)
        for i in range(n):
            ep_ret[i] += env.rew[t, i]
            ep_len[i] += 1
            stats["constraint_violation"][i] += env.viol[t, i]
            stats["mse"][i] += env.mse[t, i]
            if env.done[t, i]:
                rq.append(ep_ret[i]); lq.append(ep_len[i])
                ep_ret[i] = 0; ep_len[i] = 0
                acc += stats["constraint_violation"][i]
                mq.append(stats["mse"][i])
                stats["constraint_violation"][i] = 0; stats["mse"][i] = 0
        if env.done[t].any():
            idx = np.nonzero(env.done[t])[0]
            assert np.array_equal(info["episode"]["env"].numpy(), idx)
        else:
            assert "episode" not in info
    assert np.allclose(list(w.return_queue), list(rq)) and list(w.length_queue) == list(lq)
    assert np.allclose(list(w.queued_stats["mse"]), list(mq))
    assert w.accumulated_stats["constraint_violation"] == acc
    assert np.allclose(w.episode_return.numpy(), ep_ret) and np.allclose(w.episode_length.numpy(), ep_len)
