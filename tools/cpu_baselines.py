"""CPU baselines of BASELINE.md section 3 on this machine's host cores (no GPU needed):
  C1  the single-env numpy oracle classes stepped in a Python loop (the reference's execution style), 1 core;
  C2  the numpy-vectorised 3D step (oracle/batched_numpy.py), 1 core and one process per core (the reference's own
      data-parallel strategy, SubprocVecEnv);
  C3  the C / OpenMP port of the 3D step (oracle/scg_oracle.c), 1 thread and all threads.
One JSON line per measurement.   python tools/cpu_baselines.py [--seconds 5]"""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from oracle.cartpole_oracle import CartPoleOracle  # noqa: E402
from oracle.quad_oracle import QuadOracle  # noqa: E402


def c2_worker(args):
    """One process: `n` envs of BASELINE config 2 stepped for `seconds`; returns env-steps done."""
    n, seconds, seed = args
    os.environ["OMP_NUM_THREADS"] = "1"
    from oracle.batched_numpy import BatchedQuadOracle
    kw = bench.env_kwargs()
    kw.pop("seed")
    env = BatchedQuadOracle(n, kw, seed=seed)
    env.reset()
    rng = np.random.default_rng(seed)
    acts = np.clip(0.03454 * 9.8 / 4 + rng.normal(0, 0.005, size=(16, 4, n)), env.a_lo, env.a_hi)
    steps, t0 = 0, time.perf_counter()
    while time.perf_counter() - t0 < seconds:
        env.step(acts[steps % 16])
        steps += 1
    return steps * n / (time.perf_counter() - t0)


def loop(env, act, seconds):
    env.reset()
    n, t0 = 0, time.perf_counter()
    while time.perf_counter() - t0 < seconds:
        _, _, done, _ = env.step(act)
        n += 1
        if done:
            env.reset()
    return n / (time.perf_counter() - t0)


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--seconds", type=float, default=5.0)
    a = ap.parse_args()
    kw = bench.env_kwargs()
    kw.pop("seed")
    hover = np.full(4, 0.03454 * 9.8 / 4)
    out = []
    out.append(dict(id="C1", config="1 level0 + physics=dyn, 60/240 Hz (S=4), 1 env", cores=1,
                    env_steps_per_sec=loop(QuadOracle(**kw), hover, a.seconds)))
    kw1 = dict(kw, ctrl_freq=240, pyb_freq=240)
    out.append(dict(id="C1", config="1 level0 + physics=dyn, 240/240 Hz (S=1), 1 env", cores=1,
                    env_steps_per_sec=loop(QuadOracle(**kw1), hover, a.seconds)))
    out.append(dict(id="C1", config="3 CartPole, white-noise action disturbance, 1 env", cores=1,
                    env_steps_per_sec=loop(CartPoleOracle(disturbances={"action": [dict(disturbance_func="white_noise", std=0.2)]},
                                                          constraints=[dict(constraint_form="default_constraint", constrained_variable="input"),
                                                                       dict(constraint_form="default_constraint", constrained_variable="state")]),
                                           np.zeros(1), a.seconds)))
    q2 = QuadOracle(quad_type=2, ctrl_freq=60, pyb_freq=240, episode_len_sec=5, randomized_inertial_prop=True, dyn_uses_episode_inertia=True,
                    inertial_prop_randomization_info={"M": dict(distrib="uniform", low=-0.005, high=0.005)},
                    disturbances={"action": [dict(disturbance_func="impulse", magnitude=0.01, step_offset=None, duration=2, decay_rate=0.9),
                                             dict(disturbance_func="step", magnitude=0.005, step_offset=None)]},
                    constraints=[dict(constraint_form="default_constraint", constrained_variable="input"),
                                 dict(constraint_form="default_constraint", constrained_variable="state")])
    out.append(dict(id="C1", config="4 Quadrotor 2D, impulse + step disturbances, 1 env", cores=1,
                    env_steps_per_sec=loop(q2, np.full(2, 0.027 * 9.8 / 2), a.seconds)))
    out.append(dict(id="C2", config="2 Quadrotor 3D, 65,536 envs, numpy-vectorised step", cores=1,
                    env_steps_per_sec=c2_worker((65536, a.seconds, 0))))
    import multiprocessing as mp
    procs = bench.host_threads()
    with mp.get_context("fork").Pool(procs) as pool:
        rates = pool.map(c2_worker, [(65536 // procs, a.seconds, i) for i in range(procs)])
    out.append(dict(id="C2", config="2 Quadrotor 3D, 65,536 envs over one process per core, numpy-vectorised step", cores=procs,
                    env_steps_per_sec=float(sum(rates))))
    for threads in (1, bench.host_threads()):
        v, cores, _ = bench.cpu_reference(4096, 200, 3, threads=threads)
        out.append(dict(id="C3", config="2 Quadrotor 3D, 4096 envs x 200 steps sample, C / OpenMP", cores=cores, env_steps_per_sec=v))
    for r in out:
        print(json.dumps(r), flush=True)
