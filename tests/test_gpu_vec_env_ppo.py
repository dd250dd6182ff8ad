"""The PPO rollout loop's truncation handling (controllers/ppo/ppo.py:255-273) on the batched env: the reference's own
access pattern -- ``info["n"][idx]["terminal_info"]["TimeLimit.truncated"]`` and ``["terminal_observation"]`` -- through
``ListInfoVecEnv``, and the masked tensor form INTEGRATION.md section 4 gives, against each other."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _env(n):
    from safe_control_gym_b200.envs.quadrotor import BatchedQuadrotor
    return BatchedQuadrotor(num_envs=n, dtype=torch.float32, auto_reset=True, quad_type=2, ctrl_freq=50, pyb_freq=200,
                            episode_len_sec=0.5, task="stabilization", cost="rl_reward",
                            task_info=dict(stabilization_goal=[0, 1], stabilization_goal_tolerance=0.0),
                            init_state=dict(init_x=0, init_z=1.0), randomized_init=True,
                            init_state_randomization_info={"init_x": dict(distrib="uniform", low=-0.1, high=0.1)},
                            done_on_out_of_bound=True, seed=5)


def test_ppo_truncation_pattern_list_form_equals_masked_form():
    from safe_control_gym_b200.vec_env import ListInfoVecEnv
    n = 6
    a, b = _env(n), _env(n)
    venv = ListInfoVecEnv(a)
    venv.reset(); b.reset()
    critic = lambda o: np.asarray(o, dtype=np.float64).sum(-1)       # stands for agent.ac.critic
    hover = a.MASS * a.GRAVITY_ACC / 2
    truncated_seen = 0
    for t in range(60):
        act = np.full((n, 2), hover, dtype=np.float32)
        act[t % n] *= 1.5 if t > 30 else 1.0                         # some envs leave the box: true termination, not truncation
        next_obs, rew, done, info = venv.step(act)
        assert isinstance(next_obs, np.ndarray) and next_obs.shape == (n, a.obs_dim) and done.shape == (n,)
        # ---- the reference's loop, verbatim (ppo.py:263-273 with the critic call replaced by the stand-in)
        terminal_v = np.zeros(n)
        for idx, inf in enumerate(info["n"]):
            if "terminal_info" not in inf:
                continue
            inff = inf["terminal_info"]
            if "TimeLimit.truncated" in inff and inff["TimeLimit.truncated"]:
                terminal_obs = inf["terminal_observation"]
                terminal_v[idx] = critic(terminal_obs)
                truncated_seen += 1
        # ---- the masked tensor form on a twin env
        obs_t, rew_t, done_t, info_t = b.step(torch.as_tensor(act, device="cuda"))
        trunc = done_t & info_t["TimeLimit.truncated"]
        tv = torch.zeros(n, dtype=torch.float64, device="cuda")
        if bool(trunc.any()):
            tv[trunc] = info_t["terminal_observation"][trunc].double().sum(-1)
        assert np.array_equal(done, done_t.cpu().numpy())
        assert np.allclose(terminal_v, tv.cpu().numpy(), rtol=0, atol=1e-6)
        assert np.array_equal(next_obs, obs_t.cpu().numpy())
    assert truncated_seen >= n     # 25-step episodes: every env was truncated at least once
    a.close(); b.close()
