"""bench.py -- Quadrotor-3D DYN env-steps/sec (BASELINE.json metric) on N B200s.

Workload (BASELINE.json configs[1]): 65,536 envs per GPU x 1,000
control steps, quad_type 3, physics = dyn, 60 / 240 Hz (S = 4 substeps), level0's state/input bound
constraints (14 values), cost = competition, done_on_violation with auto-reset, randomised initial
state (Philox).  One bench "step" = one fused K = 1000 rollout launch over the whole batch
(65.5 M env-steps, 262 M integrator substeps).  Actions are a [1000, N, 4] fp32 table (1.05 GB, larger
than the 126 MB L2) that is re-streamed from HBM every step.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]

Prints ONE JSON line (see the task contract).  `value` = device-timed throughput with inputs resident
in HBM; `e2e` = the same metric through the public API with the action table in pinned HOST memory
(chunked H2D copies inside the timed region, final observations + statistics read back);
`roofline` = the rollout kernel against the live-measured FP32 FMA peak (the kernel is compute-bound:
16 B of HBM traffic per env-step against 632 algorithmic flops) with the HBM view alongside;
`cpu_baseline` = the C/OpenMP oracle (oracle/scg_oracle.c) on the host cores over a bounded sample.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

N_ENVS = 65536
K_CTRL = 1000
SUBSTEPS = 4
FLOPS_PER_STEP = 200 + 108 * SUBSTEPS        # SURVEY.md section 8d accounting (FMA = 2), frozen
BYTES_PER_STEP_FUSED = 16                    # one float4 action per env-step (state amortised over K)
METRIC = "quadrotor3d_dyn_env_steps_per_sec"

INIT_RAND = {"init_x": dict(distrib="uniform", low=-0.1, high=0.1), "init_y": dict(distrib="uniform", low=-0.1, high=0.1),
             "init_z": dict(distrib="uniform", low=0.47, high=1.47), "init_phi": dict(distrib="uniform", low=-0.1, high=0.1),
             "init_theta": dict(distrib="uniform", low=-0.1, high=0.1), "init_psi": dict(distrib="uniform", low=-0.1, high=0.1)}


def env_kwargs():
    """level0.yaml's quadrotor_config (competition/level0.yaml:16-182) with physics=dyn at 60/240 Hz."""
    return dict(quad_type=3, ctrl_freq=60, pyb_freq=240, episode_len_sec=33, cost="competition", task="stabilization",
                task_info=dict(stabilization_goal=[-0.5, 2.9, 0.75], stabilization_goal_tolerance=0.15),
                init_state=dict(init_x=-0.9, init_x_dot=0, init_y=-2.9, init_y_dot=0, init_z=0.03, init_z_dot=0,
                                init_phi=0, init_theta=0, init_psi=0, init_p=0, init_q=0, init_r=0),
                randomized_init=True, init_state_randomization_info=INIT_RAND,
                gates=[[0.5, -2.5, 0, 0, 0, -1.57, 0], [2, -1.5, 0, 0, 0, 0, 1], [0, 0.2, 0, 0, 0, 1.57, 1], [-0.5, 1.5, 0, 0, 0, 0, 0]],
                obstacles=[[1.5, -2.5, 0, 0, 0, 0], [0.5, -1, 0, 0, 0, 0], [1.5, 0, 0, 0, 0, 0], [-1, 0, 0, 0, 0, 0]],
                constraints=[dict(constraint_form="default_constraint", constrained_variable="input"),
                             dict(constraint_form="bounded_constraint", constrained_variable="state", active_dims=[0, 2, 4],
                                  upper_bounds=[3, 3, 2], lower_bounds=[-3, -3, -0.1])],
                done_on_violation=True, done_on_completion=True, done_on_collision=True, done_on_out_of_bound=False,
                normalized_rl_action_space=False, seed=1337)


def level2_like_kwargs():
    """Config 2 with competition/level2.yaml's randomisation (level2.yaml:40-100): inertial properties, gate and obstacle poses
    redrawn per episode besides the initial state.  The DYN integrator keeps the nominal inertia, as the reference's does (F9)."""
    u = lambda lo, hi: dict(distrib="uniform", low=lo, high=hi)
    kw = env_kwargs()
    kw.update(randomized_inertial_prop=True,
              inertial_prop_randomization_info={"M": u(-0.01, 0.01), "Ixx": u(-1e-6, 1e-6), "Iyy": u(-1e-6, 1e-6), "Izz": u(-1e-6, 1e-6)},
              randomized_gates_and_obstacles=True, gates_and_obstacles_randomization_info=dict(gates=u(-0.1, 0.1), obstacles=u(-0.2, 0.2)))
    return kw


def config3_kwargs():
    """BASELINE configs[2]: the cartpole_config of tests/scripts/verbose_api.yaml:1-91 -- 50 / 50 Hz,
    10 s episodes (T = 500), true pole length 1.0 with randomised inertial properties, randomised init, default input +
    state constraints (10 values), cost quadratic, done_on_violation -- with the white-noise disturbance (std 0.1) on the
    ACTION channel, drawn from Philox per env and step."""
    u = lambda lo, hi: dict(distrib="uniform", low=lo, high=hi)
    return dict(ctrl_freq=50, pyb_freq=50, episode_len_sec=10, cost="quadratic", task="traj_tracking",
                task_info=dict(stabilization_goal=[0], stabilization_goal_tolerance=0.01, trajectory_type="circle", num_cycles=1,
                               trajectory_plane="zx", trajectory_position_offset=[0, 0], trajectory_scale=0.2),
                init_state=dict(init_x=0, init_x_dot=0, init_theta=0, init_theta_dot=0), randomized_init=True,
                init_state_randomization_info={k: u(-0.05, 0.05) for k in ["init_x", "init_x_dot", "init_theta", "init_theta_dot"]},
                inertial_prop=dict(pole_length=1.0, pole_mass=0.1, cart_mass=1.0), randomized_inertial_prop=True,
                inertial_prop_randomization_info=dict(pole_length=u(-0.25, 0.0), cart_mass=u(0.5, 1.5), pole_mass=u(0.05, 0.15)),
                constraints=[dict(constraint_form="default_constraint", constrained_variable="input", strict=False),
                             dict(constraint_form="default_constraint", constrained_variable="state")],
                disturbances=dict(action=[dict(disturbance_func="white_noise", std=0.1)]),
                done_on_violation=True, done_on_out_of_bound=True, info_in_reset=True, seed=1337)


def config4_kwargs():
    """BASELINE configs[3]: Quadrotor 2D, 60 / 240 Hz (S = 4), 10 s episodes (T = 600); per-episode
    mass = nominal +- 0.01 kg and Iyy +- 1e-6 that REACH the integrator (dyn_uses_episode_inertia: extension, F9); action
    channel: impulse (0.02 N, duration 10, decay 0.9, random offset) + step (0.01 N, random offset); default input + state
    constraints (16 values); stabilisation, cost rl_reward, out-of-bound termination."""
    u = lambda lo, hi: dict(distrib="uniform", low=lo, high=hi)
    return dict(quad_type=2, ctrl_freq=60, pyb_freq=240, episode_len_sec=10, task="stabilization", cost="rl_reward",
                task_info=dict(stabilization_goal=[0, 1], stabilization_goal_tolerance=0.05),
                init_state=dict(init_x=0, init_x_dot=0, init_z=1.0, init_z_dot=0, init_theta=0, init_theta_dot=0), randomized_init=True,
                init_state_randomization_info={"init_x": u(-0.3, 0.3), "init_z": u(-0.3, 0.3), "init_theta": u(-0.2, 0.2),
                                               "init_theta_dot": u(-0.5, 0.5)},
                randomized_inertial_prop=True, dyn_uses_episode_inertia=True,
                inertial_prop_randomization_info={"M": u(-0.01, 0.01), "Iyy": u(-1e-6, 1e-6)},
                disturbances={"action": [dict(disturbance_func="impulse", magnitude=0.02, step_offset=None, duration=10, decay_rate=0.9),
                                         dict(disturbance_func="step", magnitude=0.01, step_offset=None)]},
                constraints=[dict(constraint_form="default_constraint", constrained_variable="input"),
                             dict(constraint_form="default_constraint", constrained_variable="state")],
                done_on_violation=False, done_on_out_of_bound=True, rew_exponential=True, seed=1337)


def config5_kwargs():
    """BASELINE configs[4]: the headline 3D scenario (level0's gates + obstacles) with the analytic
    gate-pass / contact geometry ON (extension, F5); every env logs the 16-float Logger record every control step."""
    kw = env_kwargs()
    kw["gate_geometry"] = True
    return kw


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md recipe)."""

    def __init__(self, gpu_index):
        self.q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
                  "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
        self.idx, self.rows, self.stop = gpu_index, [], threading.Event()
        self.th = threading.Thread(target=self._run, daemon=True)

    def _run(self):
        while not self.stop.is_set():
            try:
                out = subprocess.run(["nvidia-smi", f"--query-gpu={self.q}", "--format=csv,noheader,nounits", "-i", str(self.idx)],
                                     capture_output=True, text=True, timeout=5).stdout.strip()
                if out:
                    self.rows.append([c.strip() for c in out.split(",")])
            except Exception:
                pass
            self.stop.wait(0.2)

    def __enter__(self):
        self.th.start()
        return self

    def __exit__(self, *a):
        self.stop.set()
        self.th.join(timeout=6)

    def summary(self):
        sm = sorted(int(r[0]) for r in self.rows if r and r[0].isdigit())
        mx = [int(r[1]) for r in self.rows if len(r) > 1 and r[1].isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({names[j] for r in self.rows for j in range(4) if len(r) > 2 + j and r[2 + j].lower().startswith("active")})
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": reasons,
                "samples": len(self.rows)}


def make_actions(torch, n, k, device, seed, std=0.005):
    """hover + N(0, 0.005^2) per motor, clipped to the action bounds."""
    g = torch.Generator(device=device)
    g.manual_seed(seed)
    a = torch.empty(k, n, 4, dtype=torch.float32, device=device)
    a.normal_(mean=0.03454 * 9.8 / 4, std=std, generator=g) if std > 0 else a.fill_(0.03454 * 9.8 / 4)
    return a.clamp_(0.0281616878, 0.1483414471)


def host_threads():
    """All host threads this process may use (torchrun exports OMP_NUM_THREADS=1 for its workers: the CPU arm
    asks for its thread count explicitly instead of inheriting that)."""
    try:
        return max(1, len(os.sched_getaffinity(0)))
    except AttributeError:
        return max(1, os.cpu_count() or 1)


def cpu_reference(sample_envs, sample_steps, repeats, threads=0):
    """Times the C/OpenMP oracle on the host cores: env-steps/s over `repeats` rollouts of a bounded
    sample (sample_envs envs x sample_steps control steps) of the same workload."""
    import numpy as np
    threads = threads or host_threads()
    from oracle.c_oracle import COracle3D
    from oracle.quad_oracle import QuadOracle
    kw = env_kwargs()
    kw.pop("seed")
    co = COracle3D(QuadOracle(**kw), n=sample_envs, seed=1337, init_rand=INIT_RAND, auto_reset=True, threads=threads)
    rng = np.random.default_rng(1337)
    acts = np.clip(0.03454 * 9.8 / 4 + rng.normal(0, 0.005, size=(sample_steps, sample_envs, 4)), 0.0281616878, 0.1483414471)
    co.reset()
    co.rollout(acts[:max(1, sample_steps // 10)])   # warm-up
    times = []
    for _ in range(repeats):
        t0 = time.perf_counter()
        co.rollout(acts)
        times.append(time.perf_counter() - t0)
    return sample_envs * sample_steps / (sum(times) / len(times)), co.num_threads(), times


# --------------------------------------------------------------------------------------------------------------------
# The other BASELINE configs and the two other modes of config 2 (host-driven step(), closed
# loop), each measured the same way: warm-up launch, then `reps` back-to-back launches between CUDA events on the launching
# stream, barrier + synchronize on both sides, MAX over ranks.  Inputs are larger than L2 (config 3: 8.4 GB, 4: 5 GB,
# 5: 27 GB of actions) except where noted.
def cpu_port(kind, sample_envs, sample_steps, threads=0):
    """env-steps/s of the C/OpenMP port of config `kind` (oracle/scg_oracle.c) on the host cores, one rollout of the sample."""
    import numpy as np
    from oracle import c_oracle as co
    threads = threads or host_threads()
    rng = np.random.default_rng(1337)
    if kind == "3":
        from oracle.cartpole_oracle import CartPoleOracle
        kw = config3_kwargs(); seed = kw.pop("seed")
        port = co.COracleCartPole(CartPoleOracle(**kw), n=sample_envs, seed=seed, threads=threads)
        acts, extra = rng.normal(0, 3.0, size=(sample_steps, sample_envs, 1)), {}
    else:
        from oracle.quad_oracle import QuadOracle
        kw = config4_kwargs() if kind == "4" else config5_kwargs(); seed = kw.pop("seed")
        port = co.COracleQuadExt(QuadOracle(**kw), n=sample_envs, seed=seed, threads=threads)
        if kind == "4":
            hover = 0.03454 * 9.8 / 2
            acts, extra = np.clip(hover + rng.normal(0, 0.005, size=(sample_steps, sample_envs, 2)), 0, 4 * hover), {}
        else:
            acts = np.clip(0.03454 * 9.8 / 4 + rng.normal(0, 0.005, size=(sample_steps, sample_envs, 4)), 0.0281616878, 0.1483414471)
            extra = dict(log_out=np.zeros((sample_steps, 16, sample_envs)))
    port.reset()
    port.rollout(acts[:max(1, sample_steps // 10)])
    t0 = time.perf_counter()
    port.rollout(acts, **extra)
    dt = time.perf_counter() - t0
    return {"value": sample_envs * sample_steps / dt, "unit": "env-steps/s", "cores": port.num_threads(), "kind": "port",
            "sample": f"{sample_envs} envs x {sample_steps} ctrl steps, C/OpenMP port (oracle/scg_oracle.c), fp64"}


def timed_launches(torch, dist, world, dev, fn, reps, sampler=None):
    """Average seconds per call of fn over `reps` back-to-back calls, device-timed, max over ranks."""
    fn()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    marks = []
    for _ in range(reps):
        fn()
        if sampler is not None:   # long runs: keep a few launches queued so that neither side waits for the other
            m = torch.cuda.Event(); m.record(); marks.append(m)
            if len(marks) > 4:
                marks.pop(0).synchronize()
    e1.record()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    t = torch.tensor([e0.elapsed_time(e1) * 1e-3 / reps], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def run_configs(torch, dist, world, rank, dev, args, fp32_peak, hbm_peak):
    """BASELINE configs[2..4] + the host-driven and closed-loop modes of config 2 -> the `configs` array of the JSON line."""
    from safe_control_gym_b200.envs.cartpole import BatchedCartPole
    from safe_control_gym_b200.envs.quadrotor import BatchedQuadrotor
    out = []
    sc = args.config_scale

    def entry(cid, workload, n, k, dt, bound, per_step, cpu, parity, **more):
        value = world * n * k / dt
        per_gpu = n * k / dt
        ach = per_gpu * per_step / (1e12 if bound == "fp32" else 1e9)
        peak = fp32_peak if bound == "fp32" else hbm_peak
        e = {"id": cid, "workload": workload, "envs_per_gpu": n, "ctrl_steps_per_launch": k, "value": value, "unit": "env-steps/s",
             "ms_per_launch": dt * 1e3, "scaling": "weak",
             "roofline": {"bound": bound, "achieved": ach, "peak": peak, "unit": "TFLOP/s" if bound == "fp32" else "GB/s", "frac": ach / peak,
                          ("flops_per_env_step" if bound == "fp32" else "bytes_per_env_step"): per_step, "traffic": None},
             "cpu_baseline": cpu, "parity_test": parity}
        e.update(more)
        return e

    def cpu(kind, envs, steps):
        return cpu_port(kind, envs, steps) if (rank == 0 and world == 1 and not args.no_cpu) else None

    if "3" in args.configs:
        n, k = max(64, int(4 * 1024 * 1024 * sc)), 500
        kw = config3_kwargs(); kw.pop("info_in_reset")
        env = BatchedCartPole(num_envs=n, device=dev, dtype=torch.float32, attitude=args.attitude, auto_reset=True, seed_offset=rank * n, **kw)
        g = torch.Generator(device=dev); g.manual_seed(3 + rank)
        a = torch.empty(k, n, 1, dtype=torch.float32, device=dev).normal_(0.0, 3.0, generator=g).clamp_(-10.0, 10.0)
        env.reset()
        dt = timed_launches(torch, dist, world, dev, lambda: env.rollout(a), 8)
        out.append(entry("3", "cartpole_4Mx500_white_noise_action_default_constraints", n, k, dt, "fp32", 240, cpu("3", 262144, 200),
                         "tests/test_gpu_teacher_forced.py::test_config3_cartpole_teacher_forced (1e-5 per step vs the fp64 oracle; parity "
                         "unpinned: the env is absent from the reference fork, SURVEY.md F2)",
                         kernel_variant=int(env.lib.scg_kernel_variant(env._h)), episodes=env.stats()["episodes"]))
        env.close(); del env, a
        torch.cuda.empty_cache()
    if "4" in args.configs:
        n, k = max(64, int(1024 * 1024 * sc)), 600
        kw = config4_kwargs()
        env = BatchedQuadrotor(num_envs=n, device=dev, dtype=torch.float32, attitude=args.attitude, auto_reset=True, seed_offset=rank * n, **kw)
        hover = env.MASS * 9.8 / 2
        g = torch.Generator(device=dev); g.manual_seed(4 + rank)
        a = torch.empty(k, n, 2, dtype=torch.float32, device=dev).normal_(hover, 0.005, generator=g).clamp_(0.0, 4 * hover)
        env.reset()
        dt = timed_launches(torch, dist, world, dev, lambda: env.rollout(a), 16)
        out.append(entry("4", "quadrotor2d_1Mx600_impulse_step_per_env_inertia", n, k, dt, "fp32", 120 + 60 * SUBSTEPS, cpu("4", 65536, 200),
                         "tests/test_gpu_teacher_forced.py::test_config4_quad2d_teacher_forced (1e-5 per step vs the fp64 oracle; per-env "
                         "inertia in the integrator is an extension over the reference's DYN behaviour, SURVEY.md F9)",
                         kernel_variant=int(env.lib.scg_kernel_variant(env._h)), episodes=env.stats()["episodes"]))
        env.close(); del env, a
        torch.cuda.empty_cache()
    if "5" in args.configs:
        # weak: 16 M envs on one GPU, 2 M per GPU on N > 1 ("N fixed per GPU"); strong: 16 M in total
        k = 100
        shapes = [("5", max(64, int((16 if world == 1 else 2) * 1024 * 1024 * sc)), "weak")]
        if world > 1:
            shapes.append(("5-strong", max(64, int(16 * 1024 * 1024 * sc) // world), "strong"))
        for cid, n, scaling in shapes:
            for gates in (True, False):
                kw = config5_kwargs(); kw["gate_geometry"] = gates
                env = BatchedQuadrotor(num_envs=n, device=dev, dtype=torch.float32, attitude=args.attitude, auto_reset=True,
                                       seed_offset=rank * n, **kw)
                a = make_actions(torch, n, k, dev, 5 + rank, args.action_std)
                log = torch.empty(k, 16, n, dtype=torch.float32, device=dev)
                env.reset()
                dt = timed_launches(torch, dist, world, dev, lambda: env.rollout(a, log_out=log, log_stride=1), 8 if gates else 24)
                st = env.stats()
                e = entry(cid if gates else cid + "-gates-off", "quadrotor3d_%dx100_logged_every_step_gate_geometry_%s" % (n * world, "on" if gates else "off"),
                          n, k, dt, "hbm", 80, cpu("5", 32768, 100) if (gates and cid == "5") else None,
                          "tests/test_gpu_gates.py (gate / contact geometry vs oracle/gate_geometry.py: an extension, parity unpinned -- PyBullet's "
                          "queries are dead in DYN, SURVEY.md F5); tests/test_gpu_batch.py::test_logged_rollout_record, "
                          "::test_tma_staged_log_equals_plain_log (the Logger record, bit-exact)",
                          log_store="tma" if env.last_rollout_tlog else "plain", gates_passed=st["gates"], collisions=st["collisions"])
                e["scaling"] = scaling
                out.append(e)
                env.close(); del env, a, log
                torch.cuda.empty_cache()
    if "2h" in args.configs:
        n, k = args.envs, 500
        env = BatchedQuadrotor(num_envs=n, device=dev, dtype=torch.float32, attitude=args.attitude, auto_reset=True, seed_offset=rank * n, **env_kwargs())
        a = make_actions(torch, n, k, dev, 7 + rank, args.action_std)
        env.reset()

        def loop():
            for t in range(k):
                env.step(a[t])
        dt = timed_launches(torch, dist, world, dev, loop, 2)
        out.append(entry("2-host-driven", "quadrotor3d_dyn_65536_host_driven_step", n, k, dt, "hbm", 180, None,
                         "tests/test_gpu_batch.py::test_fused_rollout_equals_host_driven_steps; tests/test_gpu_golden.py (every golden scenario steps "
                         "through scg_step)", us_per_step=dt / k * 1e6,
                         note="one scg_step launch per control step from Python (facade + launch latency included); state is L2-resident"))
        env.close(); del env, a
    if "2l" in args.configs:
        # competition/level2.yaml's randomisation set on config 2: per-episode inertial properties, initial state, gate / obstacle poses
        n, k = args.envs, 1000
        kw = level2_like_kwargs()
        env = BatchedQuadrotor(num_envs=n, device=dev, dtype=torch.float32, attitude=args.attitude, auto_reset=True, seed_offset=rank * n, **kw)
        a = make_actions(torch, n, k, dev, 9 + rank, args.action_std)
        env.reset()
        dt = timed_launches(torch, dist, world, dev, lambda: env.rollout(a), 8)
        out.append(entry("2-level2-like", "quadrotor3d_dyn_65536x1000_randomised_inertia_init_gates_obstacles", n, k, dt, "fp32", FLOPS_PER_STEP, None,
                         "tests/test_gpu_golden.py (level2_dyn fixture from the reference's own Quadrotor); tests/test_gpu_config4.py::"
                         "test_ds_kernel_3d_level2_like_matches_lean_kernel; tests/test_gpu_batch.py::test_spare_reset_with_inertia_and_offsets_equals_stepping",
                         kernel_variant=int(env.lib.scg_kernel_variant(env._h)), episodes=env.stats()["episodes"]))
        env.close(); del env, a
    if "2c" in args.configs:
        n, k = args.envs, 1000
        env = BatchedQuadrotor(num_envs=n, device=dev, dtype=torch.float32, attitude=args.attitude, auto_reset=True, seed_offset=rank * n, **env_kwargs())
        env.reset()
        dt = timed_launches(torch, dist, world, dev, lambda: env.rollout_pid(k), 8)
        out.append(entry("2-closed-loop", "quadrotor3d_dyn_65536x1000_on_device_pid", n, k, dt, "fp32", FLOPS_PER_STEP, None,
                         "tests/test_gpu_pid.py (actions vs oracle/pid_oracle.py, itself pinned on the reference's PIDController)",
                         note="actions computed on the device by the DSL PID (quadrotor_utils.py:70-278): no action table, no H2D traffic; the PID's own "
                              "~330 flop per step are not counted"))
        env.close(); del env
    torch.cuda.empty_cache()
    return out


def run_reference(args):
    """--impl reference: the reference's CPU implementation of the path.  The reference itself cannot be
    imported on this image (pybullet / gym / casadi absent) and its DYN path does not run as
    shipped (F3), so the arm times the oracle port (C + OpenMP, all host threads) on a bounded sample."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    sample_envs, sample_steps = 4096, 200
    t0 = time.perf_counter()
    vals = []
    for _ in range(args.warmup):
        cpu_reference(sample_envs, sample_steps, 1)
    v, cores, times = cpu_reference(sample_envs, sample_steps, max(args.steps, 1))
    line = {"metric": METRIC, "value": v, "unit": "env-steps/s", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": 1e3 * sum(times) / len(times), "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f64", "data": "synthetic", "impl": "reference",
            "config": {"workload": "quadrotor3d_dyn_65536x1000_level0_constraints", "substeps": SUBSTEPS,
                       "sample": f"{sample_envs} envs x {sample_steps} ctrl steps per step"},
            "cpu_baseline": {"value": v, "unit": "env-steps/s", "cores": cores, "kind": "port",
                             "sample": f"{sample_envs} envs x {sample_steps} ctrl steps, C/OpenMP oracle (oracle/scg_oracle.c), fp64"},
            "e2e": {"value": v, "unit": "env-steps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0, "wall_s": time.perf_counter() - t0}
    emit(line)


def run_gpu(args):
    import torch
    import torch.distributed as dist
    from safe_control_gym_b200 import _lib as L
    from safe_control_gym_b200.envs.quadrotor import BatchedQuadrotor
    import ctypes as C
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    n, k = args.envs, args.ctrl_steps
    kw = env_kwargs()
    env = BatchedQuadrotor(num_envs=n, device=dev, dtype=torch.float32, attitude=args.attitude, auto_reset=True,
                           seed_offset=rank * n, **kw)   # Philox keyed by GLOBAL env id: shard-independent results
    actions = make_actions(torch, n, k, dev, 1337 + rank, args.action_std)
    from safe_control_gym_b200.sharding import StatsMerger
    merger = StatsMerger(dev)

    def one_step():
        env.rollout(actions)
        if world > 1:
            # The only collective of the path: merge the episode statistics.  It runs on NCCL's
            # stream behind a snapshot of the 16 doubles, two buffers in turn, so the next launch does not wait for it;
            # drain() makes the timed region end only after the last merge has completed.
            merger.submit(env.stats_vector())

    def drain():
        merger.drain()

    env.reset()
    env.buf_stats.zero_()
    for _ in range(args.warmup):
        one_step()
    drain()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    launches0 = env.launch_count
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(args.steps + 1)]
    with ClockSampler(local) as clk:
        torch.cuda.synchronize()
        ev[0].record()
        for s in range(args.steps):
            one_step()
            ev[s + 1].record()
        drain()
        ev_end = torch.cuda.Event(enable_timing=True)
        ev_end.record()
        torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    total_ms = ev[0].elapsed_time(ev_end)   # includes the last statistics merge
    per = [ev[s].elapsed_time(ev[s + 1]) for s in range(args.steps)]
    launches = env.launch_count - launches0
    queued = bool(env.last_rollout_queued)   # the launches of the timed region went through the time-chunk ticket queue
    t = torch.tensor([total_ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    total_ms = float(t.item())
    value = world * n * k * args.steps / (total_ms * 1e-3)
    stats = env.stats()

    # ---- end to end through the public API: host action table -> rollout_host -> obs + stats on the host
    host_actions = torch.empty(k, n, 4, dtype=torch.float32, pin_memory=True)
    host_actions.copy_(actions)
    obs_host = torch.empty(12, n, dtype=torch.float32, pin_memory=True)
    e2e_steps = max(1, min(args.steps, 5))

    def one_e2e():
        env.rollout_host(host_actions, chunk_steps=args.chunk)
        obs_host.copy_(env.buf_obs, non_blocking=True)
        st = env.stats_vector().cpu()
        torch.cuda.synchronize()
        return st
    one_e2e()
    if world > 1:
        dist.barrier()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        one_e2e()
    e2e_s = time.perf_counter() - t0
    t = torch.tensor([e2e_s], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    e2e_value = world * n * k * e2e_steps / float(t.item())
    h2d_bytes = int(host_actions.numel() * 4)

    # ---- sustained: >= 2 s of back-to-back headline launches with the clocks / power sampled inside that window
    class _PowerSampler(ClockSampler):
        def __init__(self, idx):
            super().__init__(idx)
            self.q += ",power.draw"

        def power_max(self):
            p = [float(r[6]) for r in self.rows if len(r) > 6 and r[6].replace(".", "", 1).isdigit()]
            return max(p) if p else None
    sus_reps = max(10, int(args.sustain_seconds / (total_ms * 1e-3 / args.steps)))
    with _PowerSampler(local) as sclk:
        sus_dt = timed_launches(torch, dist, world, dev, one_step, sus_reps, sampler=sclk)
        drain()
        torch.cuda.synchronize()
    value_sustained = world * n * k / sus_dt
    sustained = {"value": value_sustained, "unit": "env-steps/s", "seconds": sus_dt * sus_reps, "launches": sus_reps,
                 "clocks": sclk.summary(), "power_w_max": sclk.power_max()}

    # ---- end to end, closed loop: what the reference's own loops do (competition/getting_started.py:142-316, tests/scripts/tracking.py)
    # -- a controller turns observations into actions.  Here the DSL PID runs on the device (scg_rollout_pid): nothing goes to the GPU
    # per step, the final observations and the statistics come back.
    env.reset()

    def one_e2e_pid():
        env.rollout_pid(k)
        obs_host.copy_(env.buf_obs, non_blocking=True)
        st = env.stats_vector().cpu()
        torch.cuda.synchronize()
        return st
    one_e2e_pid()
    if world > 1:
        dist.barrier()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        one_e2e_pid()
    t = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    e2e_pid_value = world * n * k * e2e_steps / float(t.item())
    variant_id = int(env.lib.scg_kernel_variant(env._h))
    del actions, host_actions
    env.close()
    torch.cuda.empty_cache()
    peak = C.c_double()
    L.check(L.load().scg_measure_fp32_peak(C.byref(peak)))
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        hbm_peak, hbm_src = float(peaks["hbm_gbs"]), "MEASURED_PEAKS.json"
    except Exception:
        hbm_peak, hbm_src = 6650.0, "fallback (B200_PROFILING.md)"
    configs = run_configs(torch, dist, world, rank, dev, args, peak.value, hbm_peak) if args.configs else []

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    # ---- roofline of the dominant kernel (k_rollout<3, float, ATT, ...>): one launch per step
    variant = {0: "GENERAL", 1: "LEAN", 2: "LEAN,PLAIN", 3: "LEAN,PLAIN,SPEC"}[variant_id]
    if queued:
        variant += ",QUEUE"
    kernel_ms = sum(per) / len(per)   # one launch per step; stats all-reduce excluded at N = 1
    flops = float(n) * k * FLOPS_PER_STEP
    achieved_tf = flops / (kernel_ms * 1e-3) / 1e12
    hbm_bytes = float(n) * k * BYTES_PER_STEP_FUSED + n * (12 * 4 * 3 + 24)
    traffic, fma_pipe = None, None
    tp = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tp):
        try:
            tj = json.load(open(tp))
            traffic, fma_pipe = tj.get("k_env_rollout_dram_bytes_per_launch"), tj.get("k_env_rollout_fma_pipe_frac")
        except Exception:
            traffic = None
    cpu_v, cores, _ = cpu_reference(4096, 200, 3) if not args.no_cpu else (None, 0, None)
    line = {
        "metric": METRIC, "value": value, "unit": "env-steps/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": total_ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic",
        "config": {"workload": "quadrotor3d_dyn_65536x1000_level0_constraints", "envs_per_gpu": n, "ctrl_steps_per_step": k,
                   "substeps": SUBSTEPS, "env_substeps_per_sec": value * SUBSTEPS, "mode": "fused rollout, auto-reset, Philox init",
                   "schedule": "ticket queue over (64-step chunk, 64-env block) units, one persistent launch" if queued else "one block per 64 envs",
                   "attitude": args.attitude, "action_std": args.action_std,
                   "l2_policy": "inputs larger than L2 (1.05 GB action table re-streamed each step)"},
        "value_sustained": sustained,
        "roofline": {"bound": "fp32", "achieved": achieved_tf, "peak": peak.value, "unit": "TFLOP/s", "frac": achieved_tf / peak.value,
                     "frac_meaning": "ALGORITHMIC flops of the literal reference sequence (632 per env-step, SURVEY.md 8d) over the measured FFMA peak; the "
                                     "fast attitude path does not execute the quaternion round trip it is credited for.  The hardware view (ncu "
                                     "sm__inst_executed_pipe_fma of the same kernel) is `fma_pipe_frac`.",
                     "fma_pipe_frac": fma_pipe,
                     "traffic": traffic, "kernel": "scg::k_rollout<3,float,%s,%s>" % (args.attitude, variant), "kernel_ms": kernel_ms,
                     "flops_per_env_step": FLOPS_PER_STEP,
                     "peak_source": "measured live: register-only FFMA kernel (scg_measure_fp32_peak); MEASURED_PEAKS.json has no non-tensor fp32 figure",
                     "hbm": {"bound": "hbm", "achieved": hbm_bytes / (kernel_ms * 1e-3) / 1e9, "peak": hbm_peak, "unit": "GB/s",
                             "frac": hbm_bytes / (kernel_ms * 1e-3) / 1e9 / hbm_peak, "peak_source": hbm_src,
                             "bytes_per_env_step": BYTES_PER_STEP_FUSED}},
        "cpu_baseline": {"value": cpu_v, "unit": "env-steps/s", "cores": cores, "kind": "port",
                         "sample": "4096 envs x 200 ctrl steps x 3, C/OpenMP oracle (oracle/scg_oracle.c), fp64"},
        "e2e": {"value": e2e_value, "unit": "env-steps/s", "h2d_bytes_per_step": h2d_bytes,
                "d2h_bytes_per_step": int(obs_host.numel() * 4 + L.STATS_LEN * 8), "steps": e2e_steps, "chunk_ctrl_steps": args.chunk,
                "mode": "open loop: a [1000, N, 4] fp32 action table in pinned host memory is streamed to the device every step (16 B per env-step "
                        "over PCIe: %.1f GB/s per GPU)" % (h2d_bytes * e2e_value / (world * n * k) / 1e9),
                "closed_loop": {"value": e2e_pid_value, "unit": "env-steps/s", "h2d_bytes_per_step": 0,
                                "d2h_bytes_per_step": int(obs_host.numel() * 4 + L.STATS_LEN * 8),
                                "mode": "actions from the on-device DSL PID (scg_rollout_pid), final observations + statistics read back: the shape of "
                                        "the reference's own loops (controller in the loop, competition/getting_started.py:142-316)"}},
        "configs": configs,
        "gpu_launches": launches, "clocks": clk.summary(),
        "episode_stats": {kk: stats[kk] for kk in ("episodes", "return", "length", "violation_steps", "env_steps")},
    }
    emit(line)
    if world > 1:
        dist.destroy_process_group()


_REAL_STDOUT = None


def emit(line):
    text = (json.dumps(line) + "\n").encode()
    if _REAL_STDOUT is None:
        sys.stdout.write(text.decode()); sys.stdout.flush()
    else:
        sys.stdout.flush()
        os.write(_REAL_STDOUT, text)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=300)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200")
    ap.add_argument("--envs", type=int, default=N_ENVS)
    ap.add_argument("--ctrl-steps", type=int, default=K_CTRL)
    ap.add_argument("--chunk", type=int, default=50)
    ap.add_argument("--attitude", default="fast_intrinsic", help="fast_intrinsic (MUFU sin/cos) | fast (polynomial sin/cos) | roundtrip (literal)")
    ap.add_argument("--action-std", type=float, default=0.005)
    ap.add_argument("--configs", default="3,4,5,2h,2l,2c", type=lambda v: [c for c in v.split(",") if c],
                    help="extra BASELINE configs measured after the headline: 3, 4, 5, 2h (host-driven step), 2l (level2-like randomisation), 2c (closed loop); '' = none")
    ap.add_argument("--config-scale", type=float, default=1.0, help="scale the env counts of the extra configs (smoke runs)")
    ap.add_argument("--sustain-seconds", type=float, default=2.0)
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline legs")
    args = ap.parse_args()
    # stdout carries exactly one JSON line: anything a library prints there (NCCL's version banner under torchrun) goes
    # to stderr instead -- fd 1 is pointed at fd 2 for the run and the line is written to the saved descriptor
    global _REAL_STDOUT
    sys.stdout.flush()
    _REAL_STDOUT = os.dup(1)
    os.dup2(2, 1)
    if args.impl == "reference":
        run_reference(args)
    else:
        run_gpu(args)


if __name__ == "__main__":
    main()
