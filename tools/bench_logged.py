"""Config 5 break-down on one GPU: the 3D workload at N envs x K steps per launch with and without the trajectory log,
with / without gate geometry; sustained over `--seconds` of back-to-back launches (power and clocks sampled).
    python tools/bench_logged.py [--envs 4194304] [--k 100] [--seconds 2] [--gates 0|1]      (SCG_TLOG=0: plain stores)"""
import argparse
import os
import subprocess
import sys
import threading
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from safe_control_gym_b200.envs.quadrotor import BatchedQuadrotor  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--envs", type=int, default=4 * 1024 * 1024)
ap.add_argument("--k", type=int, default=100)
ap.add_argument("--seconds", type=float, default=2.0)
ap.add_argument("--gates", type=int, default=0)
ap.add_argument("--modes", default="nolog,log")
a = ap.parse_args()
dev = torch.device("cuda:0")


def smi():
    out = subprocess.run(["nvidia-smi", "--query-gpu=clocks.sm,power.draw", "--format=csv,noheader,nounits", "-i", "0"],
                         capture_output=True, text=True).stdout.strip().split(",")
    return float(out[0]), float(out[1])


kw = bench.env_kwargs()
env = BatchedQuadrotor(num_envs=a.envs, device=dev, dtype=torch.float32, attitude="fast_intrinsic", auto_reset=True,
                       gate_geometry=bool(a.gates), **kw)
env.reset()
acts = bench.make_actions(torch, a.envs, a.k, dev, 5, 0.005)
log = torch.empty(a.k, 16, a.envs, dtype=torch.float32, device=dev)
for mode in a.modes.split(","):
    args = dict(log_out=log, log_stride=1) if mode == "log" else {}
    for _ in range(2):
        env.rollout(acts, **args)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    reps, t0, samples, stop = 0, time.perf_counter(), [], threading.Event()

    def sampler():
        while not stop.is_set():
            samples.append(smi())
            stop.wait(0.1)
    th = threading.Thread(target=sampler, daemon=True)
    th.start()
    e0.record()
    marks = []
    while time.perf_counter() - t0 < a.seconds:
        env.rollout(acts, **args)
        reps += 1
        m = torch.cuda.Event()
        m.record()
        marks.append(m)
        if len(marks) > 4:   # keep four launches queued: the GPU never waits for the host, the host never runs far ahead
            marks.pop(0).synchronize()
    e1.record()
    torch.cuda.synchronize()
    stop.set(); th.join()
    dt = e0.elapsed_time(e1) * 1e-3 / reps
    byt = 16 + (64 if mode == "log" else 0)
    mhz = sorted(s[0] for s in samples)[len(samples) // 2]
    watts = max(s[1] for s in samples)
    print(f"N={a.envs} K={a.k} gates={a.gates} {mode:6s} tlog={int(env.last_rollout_tlog)}: {a.envs * a.k / dt / 1e9:7.2f} G env-steps/s  "
          f"{byt * a.envs * a.k / dt / 1e9:7.1f} GB/s algorithmic  ({reps} launches, {dt * 1e3:.2f} ms each, sm {mhz:.0f} MHz, max {watts:.0f} W, "
          f"{byt * a.envs * a.k / dt / 1e9 / max(watts, 1):.2f} GB/s/W)", flush=True)
env.close()
