"""scg_step launches with programmatic stream serialization (the next step's launch overlaps this step's
tail; the kernel waits for its predecessor with griddepcontrol.wait before it touches batch memory).  SCG_PDL=0 restores plain
launches (read once per process): the same 300 host-driven steps, interleaved with torch kernels on the same stream that
write the next action in place, must end in the same bits either way."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SCRIPT = r"""
import hashlib, sys, torch
sys.path.insert(0, %r)
import bench
from safe_control_gym_b200.envs.quadrotor import BatchedQuadrotor
n = 20000
env = BatchedQuadrotor(num_envs=n, dtype=torch.float32, attitude="fast_intrinsic", auto_reset=True, **bench.env_kwargs())
acts = bench.make_actions(torch, n, 300, torch.device("cuda:0"), 3, 0.01)
a = torch.empty(n, 4, device="cuda")
env.reset()
tot = torch.zeros((), device="cuda", dtype=torch.float64)
for t in range(300):
    a.copy_(acts[t])                      # a torch kernel between two steps: rewrites the buffer the previous step read
    obs, rew, done, info = env.step(a)
    tot += rew.double().sum()             # ... and one behind it that reads what the step wrote
torch.cuda.synchronize()
h = hashlib.sha256()
for name in ("state", "ctrl_step", "episode", "flags", "obs", "reward", "done", "c_values"):
    h.update(getattr(env, "buf_" + name).cpu().numpy().tobytes())
print("HASH", h.hexdigest(), float(tot), env.stats()["env_steps"])
""" % ROOT


def _run(pdl):
    env = dict(os.environ, SCG_PDL=pdl)
    out = subprocess.run([sys.executable, "-c", SCRIPT], capture_output=True, text=True, env=env, timeout=300)
    assert out.returncode == 0, out.stderr[-2000:]
    return [l for l in out.stdout.splitlines() if l.startswith("HASH")][0]


def test_programmatic_dependent_launch_changes_no_result():
    assert _run("1") == _run("0")
