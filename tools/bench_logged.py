"""Config 5 break-down on one GPU: the same 3D workload at N envs x K steps per launch with and without the trajectory log.
    python tools/bench_logged.py [--envs 4194304] [--ks 20,100]"""
import argparse
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from safe_control_gym_b200.envs.quadrotor import BatchedQuadrotor  # noqa: E402
from tools.bench_configs import timed  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--envs", type=int, default=4 * 1024 * 1024)
ap.add_argument("--ks", default="20,100")
a = ap.parse_args()
dev = torch.device("cuda:0")
for k in [int(x) for x in a.ks.split(",")]:
    env = BatchedQuadrotor(num_envs=a.envs, device=dev, dtype=torch.float32, attitude="fast_intrinsic", auto_reset=True, **bench.env_kwargs())
    env.reset()
    acts = bench.make_actions(torch, a.envs, k, dev, 5, 0.005)
    log = torch.empty(k, 16, a.envs, dtype=torch.float32, device=dev)
    for name, kw in (("no log", {}), ("log every step", dict(log_out=log, log_stride=1)), ("log every 4th step", dict(log_out=log, log_stride=4))):
        dt = timed(lambda: env.rollout(acts, **kw), 5)
        byt = 16 + (64 if "every step" in name else 16 if "4th" in name else 0)
        print(f"N={a.envs} K={k} {name:20s}: {a.envs * k / dt / 1e9:7.2f} G env-steps/s  {byt * a.envs * k / dt / 1e9:7.1f} GB/s algorithmic "
              f"(queued={env.last_rollout_queued})", flush=True)
    env.close()
    del acts, log
