"""Single-GPU throughput of the other BASELINE.json configs (3: CartPole 4M envs, 4: Quadrotor 2D 1M envs with
impulse + step disturbances and per-env mass / inertia, 5: Quadrotor 3D 16M envs, logged).  bench.py stays on the
headline config 2; this tool fills BASELINE.md section 4.  One JSON line per config.

    python tools/bench_configs.py [--configs 3,4,5] [--scale 1.0] [--repeats 5]
"""
import argparse
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from safe_control_gym_b200.envs.cartpole import BatchedCartPole  # noqa: E402
from safe_control_gym_b200.envs.quadrotor import BatchedQuadrotor  # noqa: E402

DEV = "cuda:0"


def timed(fn, repeats, warmup=2):
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
    ev[0].record()
    for _ in range(repeats):
        fn()
    ev[1].record()
    torch.cuda.synchronize()
    return ev[0].elapsed_time(ev[1]) * 1e-3 / repeats


def actions(k, n, nu, mean, std, lo, hi, seed=7):
    g = torch.Generator(device=DEV)
    g.manual_seed(seed)
    a = torch.empty(k, n, nu, dtype=torch.float32, device=DEV)
    a.normal_(mean=mean, std=std, generator=g)
    return a.clamp_(lo, hi)


def config3(scale, repeats):
    """CartPole step, 4M envs, white-noise action disturbance, default state + input constraints; K = 50 (one episode
    second at 50 Hz) per launch.  Algorithmic bytes: 4 B action per env-step; flops: RK4 of the cart-pole ODE
    (4 x ~45) + constraints ~ 220 per env-step."""
    n, k = int(4 * 1024 * 1024 * scale), 50
    env = BatchedCartPole(num_envs=n, device=DEV, auto_reset=True, seed=3, randomized_init=True,
                          disturbances={"action": [dict(disturbance_func="white_noise", std=0.2)]},
                          constraints=[dict(constraint_form="default_constraint", constrained_variable="input"),
                                       dict(constraint_form="default_constraint", constrained_variable="state")],
                          done_on_violation=False, done_on_out_of_bound=True)
    env.reset()
    a = actions(k, n, 1, 0.0, 0.5, -1.0, 1.0) if env.NORMALIZED_RL_ACTION_SPACE else actions(k, n, 1, 0.0, 3.0, -10.0, 10.0)
    dt = timed(lambda: env.rollout(a), repeats)
    st = env.stats()
    return dict(config="3 CartPole 4M envs, white-noise action disturbance, default constraints", envs=n, k=k, sec_per_launch=dt,
                env_steps_per_sec=n * k / dt, flops_per_step=220, bytes_per_step=4, episodes=st["episodes"], lean=bool(env.lean))


def config4(scale, repeats):
    """Quadrotor 2D DYN, 1M envs, impulse + step disturbances on the action channel, per-env mass / Iyy drawn at reset
    and used by the integrator (dyn_uses_episode_inertia extension, F9); K = 100 at 60 / 240 Hz.
    Algorithmic: 8 B action per env-step; flops ~ 120 + 40 S = 280."""
    n, k = int(1024 * 1024 * scale), 100
    env = BatchedQuadrotor(num_envs=n, device=DEV, quad_type=2, ctrl_freq=60, pyb_freq=240, episode_len_sec=5, auto_reset=True, seed=4,
                           task="stabilization", cost="rl_reward", randomized_init=True,
                           randomized_inertial_prop=True, dyn_uses_episode_inertia=True,
                           inertial_prop_randomization_info={"M": dict(distrib="uniform", low=-0.005, high=0.005),
                                                             "Iyy": dict(distrib="uniform", low=-1e-6, high=1e-6)},
                           disturbances={"action": [dict(disturbance_func="impulse", magnitude=0.01, step_offset=None, duration=2, decay_rate=0.9),
                                                    dict(disturbance_func="step", magnitude=0.005, step_offset=None)]},
                           constraints=[dict(constraint_form="default_constraint", constrained_variable="input"),
                                        dict(constraint_form="default_constraint", constrained_variable="state")],
                           done_on_violation=False, done_on_out_of_bound=True)
    env.reset()
    hover = 0.027 * 9.8 / 2 if abs(env.MASS - 0.027) < 1e-9 else env.MASS * 9.8 / 2
    a = actions(k, n, 2, hover, 0.005, 0.0, 4 * hover)
    dt = timed(lambda: env.rollout(a), repeats)
    st = env.stats()
    return dict(config="4 Quadrotor 2D 1M envs, impulse+step action disturbances, per-env mass/inertia in the integrator", envs=n, k=k,
                sec_per_launch=dt, env_steps_per_sec=n * k / dt, flops_per_step=280, bytes_per_step=8, episodes=st["episodes"],
                lean=bool(env.lean))


def config5(scale, repeats, gate_geometry):
    """Quadrotor 3D DYN, 16M envs, trajectories logged every control step (16 x 4 B per env-step written), level0's gates
    and obstacles; K = 20 per launch.  Algorithmic bytes: 16 B action + 64 B log record = 80 B per env-step (HBM roofline)."""
    n, k = int(16 * 1024 * 1024 * scale), 20
    kw = bench.env_kwargs()
    env = BatchedQuadrotor(num_envs=n, device=DEV, dtype=torch.float32, attitude="fast_intrinsic", auto_reset=True,
                           gate_geometry=gate_geometry, **kw)
    env.reset()
    a = bench.make_actions(torch, n, k, torch.device(DEV), 5, 0.005)
    log = torch.empty(k, 16, n, dtype=torch.float32, device=DEV)
    dt = timed(lambda: env.rollout(a, log_out=log, log_stride=1), repeats)
    st = env.stats()
    return dict(config="5 Quadrotor 3D 16M envs, every step logged, gates/obstacles" + (" (gate geometry on)" if gate_geometry else ""),
                envs=n, k=k, sec_per_launch=dt, env_steps_per_sec=n * k / dt, flops_per_step=bench.FLOPS_PER_STEP, bytes_per_step=80,
                hbm_gbs=80.0 * n * k / dt / 1e9, episodes=st["episodes"], gates=st.get("gates"), collisions=st.get("collisions"),
                lean=bool(env.lean))


def config_pid(scale, repeats):
    """Closed loop: the headline 3D workload with the action of every step computed by the on-device DSL
    PID (stabilisation at level0's goal) instead of read from a table -- no host or HBM action traffic at all."""
    n, k = int(65536 * scale), 1000
    kw = bench.env_kwargs()
    env = BatchedQuadrotor(num_envs=n, device=DEV, dtype=torch.float32, attitude="fast_intrinsic", auto_reset=True, **kw)
    env.reset()
    dt = timed(lambda: env.rollout_pid(k), repeats)
    st = env.stats()
    return dict(config="closed loop: 3D 65,536 envs x 1,000 steps, on-device PID", envs=n, k=k, sec_per_launch=dt,
                env_steps_per_sec=n * k / dt, flops_per_step=bench.FLOPS_PER_STEP + 330, bytes_per_step=72, episodes=st["episodes"],
                completed=st["completed"], lean=bool(env.lean))


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--configs", default="3,4,5")
    ap.add_argument("--scale", type=float, default=1.0, help="scale the env counts (CPU-side smoke runs use a small value)")
    ap.add_argument("--repeats", type=int, default=5)
    ap.add_argument("--gate-geometry", action="store_true")
    args = ap.parse_args()
    for c in args.configs.split(","):
        r = {"3": lambda: config3(args.scale, args.repeats), "4": lambda: config4(args.scale, args.repeats),
             "5": lambda: config5(args.scale, args.repeats, args.gate_geometry),
             "pid": lambda: config_pid(args.scale, args.repeats)}[c]()
        r["tflops"] = r["flops_per_step"] * r["env_steps_per_sec"] / 1e12
        r["gbs"] = r["bytes_per_step"] * r["env_steps_per_sec"] / 1e9
        print(json.dumps(r), flush=True)
        torch.cuda.empty_cache()
