"""Quick device-timed sweep of the fused rollout (CUDA events, 65,536 envs): attitude modes x reset pressure.
    python tools/quick_bench.py [--k 200] [--envs 65536]"""
import argparse
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from safe_control_gym_b200.envs.quadrotor import BatchedQuadrotor  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--k", type=int, default=200)
ap.add_argument("--envs", type=int, default=65536)
ap.add_argument("--reps", type=int, default=5)
a = ap.parse_args()
dev = torch.device("cuda:0")
for att in ("fast", "fast_intrinsic", "roundtrip"):
    for std in (0.005, 0.0):
        env = BatchedQuadrotor(num_envs=a.envs, device=dev, dtype=torch.float32, attitude=att, auto_reset=True, **bench.env_kwargs())
        acts = bench.make_actions(torch, a.envs, a.k, dev, 1337, std)
        env.reset()
        for _ in range(2):
            env.rollout(acts)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(a.reps):
            env.rollout(acts)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / a.reps
        st = env.stats()
        print(f"attitude={att:15s} action_std={std:5.3f} K={a.k}: {ms:8.3f} ms/launch  {a.envs * a.k / ms / 1e6:8.2f} G env-steps/s "
              f"({ms * 1e3 / a.k:6.2f} us/step)  episodes/env-step={st['episodes'] / max(st['env_steps'], 1):.4f}", flush=True)
        if std > 0:   # host-driven BenchmarkEnv.step(): one launch per control step
            for t in range(20):
                env.step(acts[t])
            torch.cuda.synchronize()
            e0.record()
            for t in range(a.k):
                env.step(acts[t])
            e1.record()
            torch.cuda.synchronize()
            us = e0.elapsed_time(e1) * 1e3 / a.k
            print(f"    host-driven step(): {us:7.2f} us/step  {a.envs / us / 1e3:7.3f} G env-steps/s", flush=True)
        env.close()
