"""Logged rollouts at 16 M envs: throughput of consecutive phases of a long run (burst -> power-limited steady state).
    python tools/bench_logged_phase.py"""
import sys, os, torch
sys.path.insert(0, '/root/repo')
import bench
from safe_control_gym_b200.envs.quadrotor import BatchedQuadrotor
from tools.bench_configs import timed
n, k = 16*1024*1024, 20
dev = torch.device("cuda:0")
env = BatchedQuadrotor(num_envs=n, device=dev, dtype=torch.float32, attitude="fast_intrinsic", auto_reset=True, **bench.env_kwargs())
env.reset()
a = bench.make_actions(torch, n, k, dev, 5, 0.005)
log = torch.empty(k, 16, n, dtype=torch.float32, device=dev)
if len(sys.argv) > 1:
    for _ in range(int(sys.argv[1])):
        env.rollout(a)
    torch.cuda.synchronize()
    print(f"{sys.argv[1]} launches without log first", flush=True)
for rep in range(int(sys.argv[2]) if len(sys.argv) > 2 else 4):
    dt = timed(lambda: env.rollout(a, log_out=log, log_stride=1), 5, warmup=0)
    st = env.stats()
    print(f"launches {rep*5}..{rep*5+4}: {n*k/dt/1e9:6.2f} G env-steps/s, {80*n*k/dt/1e9:7.1f} GB/s; episodes so far {st['episodes']:.0f}", flush=True)
