"""level2-like config (bench.level2_like_kwargs) under the ticket-queue modes: python tools/bench_level2.py"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
from safe_control_gym_b200.envs.quadrotor import BatchedQuadrotor  # noqa: E402
from tools.bench_configs import timed  # noqa: E402

n, k = 65536, 1000
env = BatchedQuadrotor(num_envs=n, device="cuda:0", dtype=torch.float32, attitude="fast_intrinsic", auto_reset=True, **bench.level2_like_kwargs())
env.reset()
a = bench.make_actions(torch, n, k, torch.device("cuda:0"), 5, 0.005)
for mode, chunk in ((-1, 0), (0, 0), (1, 32), (1, 64), (1, 128), (1, 250)):
    env.set_rollout_queue(mode, chunk)
    dt = timed(lambda: env.rollout(a), 8)
    print("level2-like variant", env.lib.scg_kernel_variant(env._h), "queue mode", mode, "chunk", chunk, "queued", env.last_rollout_queued,
          f"{n * k / dt / 1e9:.2f} G env-steps/s")
