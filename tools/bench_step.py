"""Host-driven step() loop: host-side cost and pipelined time per step.  python tools/bench_step.py [n_envs]"""
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
from safe_control_gym_b200.envs.quadrotor import BatchedQuadrotor  # noqa: E402

n, k = int(sys.argv[1]) if len(sys.argv) > 1 else 65536, 500
env = BatchedQuadrotor(num_envs=n, device="cuda:0", dtype=torch.float32, attitude="fast_intrinsic", auto_reset=True, **bench.env_kwargs())
a = bench.make_actions(torch, n, k, torch.device("cuda:0"), 7, 0.005)
env.reset()
for rep in range(5):
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for t in range(k):
        env.step(a[t])
    t1 = time.perf_counter(); torch.cuda.synchronize(); t2 = time.perf_counter()
    print("SCG_PDL=%s  host side %.2f us/step, pipelined %.2f us/step = %.3e env-steps/s" % (
        os.environ.get("SCG_PDL", "1"), (t1 - t0) / k * 1e6, (t2 - t0) / k * 1e6, n * k / (t2 - t0)))
