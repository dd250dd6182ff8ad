"""Small driver for ncu captures of the other BASELINE configs (the bench shapes at reduced K).
    python tools/profile_config.py --config 2|3|4|5|5g|2l|pid|step [--k 64] [--launches 3] [--scale 0.25]"""
import argparse
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from safe_control_gym_b200.envs.cartpole import BatchedCartPole  # noqa: E402
from safe_control_gym_b200.envs.quadrotor import BatchedQuadrotor  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--config", default="4")
ap.add_argument("--k", type=int, default=64)
ap.add_argument("--launches", type=int, default=3)
ap.add_argument("--scale", type=float, default=0.25)
a = ap.parse_args()
dev = torch.device("cuda:0")
g = torch.Generator(device=dev); g.manual_seed(1)
if a.config == "3":
    n = int(4 * 1024 * 1024 * a.scale)
    kw = bench.config3_kwargs(); kw.pop("info_in_reset")
    env = BatchedCartPole(num_envs=n, device=dev, dtype=torch.float32, attitude="fast_intrinsic", auto_reset=True, **kw)
    acts = torch.empty(a.k, n, 1, device=dev).normal_(0.0, 3.0, generator=g).clamp_(-10, 10)
    run = lambda: env.rollout(acts)
elif a.config == "4":
    n = int(1024 * 1024 * a.scale)
    env = BatchedQuadrotor(num_envs=n, device=dev, dtype=torch.float32, attitude="fast_intrinsic", auto_reset=True, **bench.config4_kwargs())
    hover = env.MASS * 9.8 / 2
    acts = torch.empty(a.k, n, 2, device=dev).normal_(hover, 0.005, generator=g).clamp_(0, 4 * hover)
    run = lambda: env.rollout(acts)
else:
    n = int((65536 if a.config in ("pid", "step", "2l", "2") else 4 * 1024 * 1024) * (1.0 if a.config in ("pid", "step", "2l", "2") else a.scale))
    kw = bench.config5_kwargs() if a.config == "5g" else (bench.level2_like_kwargs() if a.config == "2l" else bench.env_kwargs())
    env = BatchedQuadrotor(num_envs=n, device=dev, dtype=torch.float32, attitude="fast_intrinsic", auto_reset=True, **kw)
    acts = bench.make_actions(torch, n, a.k, dev, 1, 0.005)
    if a.config in ("5", "5g"):
        log = torch.empty(a.k, 16, n, device=dev)
        run = lambda: env.rollout(acts, log_out=log)
    elif a.config in ("2l", "2"):
        run = lambda: env.rollout(acts)
    elif a.config == "pid":
        run = lambda: env.rollout_pid(a.k)
    else:
        def run():
            for t in range(a.k):
                env.step(acts[t])
env.reset()
for _ in range(a.launches):
    run()
torch.cuda.synchronize()
print("ok", a.config, n, env.stats()["env_steps"], "variant", int(env.lib.scg_kernel_variant(env._h)))
