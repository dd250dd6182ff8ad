"""Small driver for ncu captures: a few fused-rollout launches of the bench workload.
    python tools/profile_rollout.py [--attitude fast_intrinsic] [--k 200] [--launches 3] [--action-std 0.005]"""
import argparse
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from safe_control_gym_b200.envs.quadrotor import BatchedQuadrotor  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--attitude", default="fast_intrinsic")
ap.add_argument("--k", type=int, default=200)
ap.add_argument("--envs", type=int, default=65536)
ap.add_argument("--launches", type=int, default=3)
ap.add_argument("--action-std", type=float, default=0.005)
ap.add_argument("--steps-api", type=int, default=0, help="also run this many scg_step launches")
a = ap.parse_args()
env = BatchedQuadrotor(num_envs=a.envs, device="cuda:0", dtype=torch.float32, attitude=a.attitude, auto_reset=True, **bench.env_kwargs())
acts = bench.make_actions(torch, a.envs, a.k, torch.device("cuda:0"), 1337, a.action_std)
env.reset()
for _ in range(a.launches):
    env.rollout(acts)
for t in range(a.steps_api):
    env.step(acts[t % a.k])
torch.cuda.synchronize()
print("ok", env.stats()["env_steps"])
