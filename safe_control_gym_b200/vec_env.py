"""The VecEnv face of the batched envs: what the reference's RL loops put around
the env step -- ``make_vec_envs`` (envs/env_wrappers/vectorized_env/__init__.py:40-65) and
``VecRecordEpisodeStatistics`` (envs/env_wrappers/record_episode_statistics.py:92-169) -- with the per-env Python
loops replaced by tensor operations on the device the env lives on.

A batched env already *is* a vector env (``num_envs`` lock-step copies with DummyVecEnv's auto-reset,
dummy_vec_env.py:34-49), so ``make_vec_envs`` builds ONE env of ``batch_size`` instead of ``batch_size`` processes.
Differences a caller sees, both forced by scale: ``info`` stays a dict of ``[N]`` tensors (no ``info["n"][i]`` dicts),
and a finished episode is reported as ``info["episode"] = {"r", "l", <tracked keys>, "env"}`` tensors over the envs
that ended this step.
"""
from collections import deque

import torch


def make_vec_envs(env_func, env_configs=None, batch_size=1, n_processes=1, seed=None):
    """``env_func(num_envs=batch_size, seed=seed)``: one batched env in place of ``batch_size`` workers.
    ``n_processes`` is accepted and ignored (the batch is one CUDA launch); per-env ``env_configs`` must be identical
    (lock-step envs share their scenario constants)."""
    if env_configs is not None:
        env_configs = list(env_configs)
        if any(c != env_configs[0] for c in env_configs):
            raise ValueError("make_vec_envs: the batched env needs one shared env_config")
    kw = dict(env_configs[0]) if env_configs else {}
    if seed is not None:
        kw["seed"] = seed
    return env_func(num_envs=batch_size, **kw)


class VecEnvWrapper:
    """vec_env.py:139-182 in the reference: forwards everything to the wrapped env."""

    def __init__(self, venv):
        self.venv = venv
        self.num_envs = venv.num_envs
        self.observation_space = venv.observation_space
        self.action_space = venv.action_space

    def step_async(self, actions):
        self._actions = actions

    def step_wait(self):
        return self.venv.step(self._actions)

    def step(self, actions):
        self.step_async(actions)
        return self.step_wait()

    def reset(self, **kwargs):
        return self.venv.reset(**kwargs)

    def close(self):
        return self.venv.close()

    def __getattr__(self, name):
        return getattr(self.venv, name)


class VecRecordEpisodeStatistics(VecEnvWrapper):
    """Episode returns / lengths and tracked info keys (e.g. ``constraint_violation``, ``mse``), same attribute
    names as the reference: ``episode_return, episode_length, return_queue, length_queue, add_tracker,
    episode_stats, accumulated_stats, queued_stats``."""

    def __init__(self, venv, deque_size=None, **kwargs):
        super().__init__(venv)
        self.deque_size = deque_size
        dev = venv.device
        self.episode_return = torch.zeros(self.num_envs, dtype=torch.float64, device=dev)
        self.episode_length = torch.zeros(self.num_envs, dtype=torch.float64, device=dev)
        self.return_queue = deque(maxlen=deque_size)
        self.length_queue = deque(maxlen=deque_size)
        self.episode_stats, self.accumulated_stats, self.queued_stats = {}, {}, {}

    def add_tracker(self, name, init_value, mode="accumulate"):
        self.episode_stats[name] = torch.full((self.num_envs,), float(init_value), dtype=torch.float64, device=self.venv.device)
        if mode == "accumulate":
            self.accumulated_stats[name] = init_value
        elif mode == "queue":
            self.queued_stats[name] = deque(maxlen=self.deque_size)
        else:
            raise Exception("Tracker mode not implemented.")

    def reset(self, **kwargs):
        self.episode_return.zero_()
        self.episode_length.zero_()
        for key in self.episode_stats:
            self.episode_stats[key].zero_()
        return self.venv.reset(**kwargs)

    def _tail(self, values):
        """The last ``deque_size`` finished episodes of this step, in env order (one D2H copy)."""
        if self.deque_size is not None:
            values = values[-self.deque_size:]
        return values.tolist()

    def step_wait(self):
        obs, reward, done, info = self.venv.step(self._actions)
        self.episode_return += reward.to(torch.float64)
        self.episode_length += 1
        for key, acc in self.episode_stats.items():
            if key in info:   # at a done step the env's info is the terminal step's (the reference's "terminal_info")
                acc += info[key].to(torch.float64)
        if bool(done.any()):
            idx = done.nonzero(as_tuple=True)[0]
            ep = {"env": idx, "r": self.episode_return[idx].clone(), "l": self.episode_length[idx].clone()}
            self.return_queue.extend(self._tail(ep["r"]))
            self.length_queue.extend(self._tail(ep["l"]))
            self.episode_return[idx] = 0
            self.episode_length[idx] = 0
            for key, acc in self.episode_stats.items():
                ep[key] = acc[idx].clone()
                if key in self.accumulated_stats:
                    self.accumulated_stats[key] += float(ep[key].sum().item())
                if key in self.queued_stats:
                    self.queued_stats[key].extend(self._tail(ep[key]))
                acc[idx] = 0
            info["episode"] = ep
        return obs, reward, done, info


def info_n(info, done, keys=("constraint_violation", "TimeLimit.truncated", "mse", "constraint_values")):
    """The reference's list-of-dicts form of a step's ``info`` (DummyVecEnv.step_wait, dummy_vec_env.py:38-49): ``out["n"][i]`` is
    env i's dict; an env that finished this step carries ``terminal_observation`` and ``terminal_info`` (its last step's info),
    which is what the PPO rollout loop reads (controllers/ppo/ppo.py:263-273).  One D2H copy per key: for small ``num_envs``."""
    host = {k: info[k].cpu().numpy() for k in keys if k in info}
    d = done.cpu().numpy()
    term = info["terminal_observation"].cpu().numpy() if bool(d.any()) else None
    out = []
    for i in range(len(d)):
        own = {k: v[i] for k, v in host.items()}
        own["TimeLimit.truncated"] = bool(own.get("TimeLimit.truncated", False))
        out.append({"terminal_observation": term[i], "terminal_info": own} if d[i] else own)
    return {"n": out}


class ListInfoVecEnv(VecEnvWrapper):
    """A batched env behind DummyVecEnv's exact return types: numpy ``obs [N, obs_dim]``, ``rew [N]``, ``done [N]`` and
    ``info = {"n": [dict per env]}`` -- for reference code that must run unmodified (e.g. ``PPO.train_step``) on a handful of envs.
    Throughput code keeps the tensors and takes the masked form."""

    def reset(self, **kwargs):
        out = self.venv.reset(**kwargs)
        if isinstance(out, tuple):
            return out[0].cpu().numpy(), {"n": [dict() for _ in range(self.num_envs)]}
        return out.cpu().numpy()

    def step_wait(self):
        obs, rew, done, info = self.venv.step(torch.as_tensor(self._actions))
        return obs.cpu().numpy(), rew.cpu().numpy(), done.cpu().numpy(), info_n(info, done)
