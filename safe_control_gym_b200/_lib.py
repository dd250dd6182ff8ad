"""ctypes binding of libscg_b200.so -- mirrors include/scg_b200.h field by field.

This is the stub a reference maintainer would add.  There is no
fallback: if the shared library is missing or does not load, importing the batched
envs raises.
"""
import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("SCG_LIB", os.path.join(HERE, "libscg_b200.so"))   # SCG_LIB: an alternative build of the same library (A/B measurements)

MAX_NX, MAX_NU, MAX_CONSTRAINTS, MAX_CON_ROWS, MAX_ROWS = 12, 4, 8, 24, 64
MAX_DIST, MAX_GATES, MAX_OBSTACLES, STATS_LEN = 4, 8, 8, 16
STAT_REPLICAS = 64   # SCG_STAT_REPLICAS: the device accumulators are [STAT_REPLICAS][STATS_LEN], totals = sum over axis 0
NOALLOC_QUEUE_WORDS = 131072

QUAD_1D, QUAD_2D, QUAD_3D, CARTPOLE = 1, 2, 3, 10
TASK_STABILIZATION, TASK_TRAJ_TRACKING = 0, 1
COST_RL_REWARD, COST_QUADRATIC, COST_COMPETITION = 0, 1, 2
CON_BOUNDED, CON_LINEAR, CON_QUADRATIC, CON_SYMMETRIC = 0, 1, 2, 3
VAR_STATE, VAR_INPUT = 0, 1
DIST_IMPULSE, DIST_STEP, DIST_UNIFORM, DIST_WHITE, DIST_PERIODIC = 0, 1, 2, 3, 4
CH_OBSERVATION, CH_ACTION, CH_DYNAMICS = 0, 1, 2
RAND_NONE, RAND_UNIFORM, RAND_NORMAL = 0, 1, 2
ATT_ROUNDTRIP, ATT_FAST, ATT_FAST_INTRINSIC = 0, 1, 2
SYM_FC, SYM_FD_RK4, SYM_DF, SYM_LOSS, SYM_FD_EULER = 0, 1, 2, 3, 4

INFO_VIOLATION, INFO_TRUNCATED, INFO_COLLISION, INFO_GATE_PASSED = 0x01, 0x02, 0x04, 0x08
INFO_AT_GOAL, INFO_COMPLETED, INFO_GOAL_REACHED, INFO_OUT_OF_BOUND = 0x10, 0x20, 0x40, 0x80
FLAG_CNSTR_VIOLATION, FLAG_AT_GOAL, FLAG_FLIPPED, FLAG_COMPLETED = 0x1, 0x2, 0x4, 0x8
FLAG_GATE_SHIFT, FLAG_HOLD_SHIFT = 8, 16
STAT_NAMES = ["episodes", "return", "length", "violation_steps", "mse", "gates", "collisions", "truncated",
              "completed", "env_steps", "violated_episodes"]

EXPORTS = ["scg_create", "scg_destroy", "scg_bind_buffers", "scg_set_injection", "scg_reset", "scg_step",
           "scg_rollout", "scg_rollout_host", "scg_rollout_pid", "scg_stats_reset", "scg_stats_read", "scg_config_word", "scg_last_rollout_spec", "scg_set_spec_enabled", "scg_constraints_eval",
           "scg_symbolic_eval", "scg_symbolic_loss", "scg_gate_info", "scg_measure_fp32_peak", "scg_launch_count", "scg_kernel_variant",
           "scg_set_rollout_queue", "scg_last_rollout_queue", "scg_set_seed", "scg_rollout_noalloc", "scg_last_rollout_tlog",
           "scg_last_error", "scg_version"]


class Constraint(C.Structure):
    _fields_ = [("form", C.c_int32), ("var", C.c_int32), ("strict", C.c_int32), ("n_rows", C.c_int32),
                ("n_active", C.c_int32), ("active_dims", C.c_int32 * MAX_NX),
                ("lower", C.c_double * MAX_NX), ("upper", C.c_double * MAX_NX),
                ("A", C.c_double * (MAX_NX * MAX_NX)), ("b", C.c_double * MAX_CON_ROWS),
                ("has_tolerance", C.c_int32), ("tolerance", C.c_double * MAX_CON_ROWS)]


class Disturbance(C.Structure):
    _fields_ = [("kind", C.c_int32), ("slot", C.c_int32), ("has_mask", C.c_int32), ("mask", C.c_double * MAX_NX),
                ("magnitude", C.c_double), ("step_offset", C.c_int32), ("max_step", C.c_int32),
                ("duration", C.c_double), ("decay_rate", C.c_double),
                ("low", C.c_double * MAX_NX), ("high", C.c_double * MAX_NX), ("std", C.c_double * MAX_NX),
                ("scale", C.c_double), ("frequency", C.c_double), ("inject", C.c_void_p)]


class Rand(C.Structure):
    _fields_ = [("distrib", C.c_int32), ("a", C.c_double), ("b", C.c_double)]


class Config(C.Structure):
    _fields_ = [
        ("env_kind", C.c_int32), ("precision", C.c_int32), ("attitude_mode", C.c_int32),
        ("substeps", C.c_int32), ("ctrl_steps", C.c_int32), ("ctrl_freq", C.c_int32), ("pyb_freq", C.c_int32),
        ("pyb_dt", C.c_double),
        ("mass", C.c_double), ("gravity_acc", C.c_double), ("gravity_force", C.c_double),
        ("J", C.c_double * 3), ("J_inv", C.c_double * 3), ("kf", C.c_double), ("km", C.c_double), ("arm", C.c_double),
        ("pwm2rpm_scale", C.c_double), ("pwm2rpm_const", C.c_double), ("pwm_min", C.c_double), ("pwm_max", C.c_double),
        ("cp_length", C.c_double), ("cp_pole_mass", C.c_double), ("cp_cart_mass", C.c_double),
        ("normalized_action", C.c_int32), ("norm_act_scale", C.c_double), ("hover_thrust", C.c_double),
        ("act_low", C.c_double * MAX_NU), ("act_high", C.c_double * MAX_NU),
        ("state_low", C.c_double * MAX_NX), ("state_high", C.c_double * MAX_NX),
        ("done_on_out_of_bound", C.c_int32),
        ("task", C.c_int32), ("cost", C.c_int32), ("rew_exponential", C.c_int32), ("obs_goal_horizon", C.c_int32),
        ("rew_state_weight", C.c_double * MAX_NX), ("rew_act_weight", C.c_double * MAX_NU),
        ("Q", C.c_double * (MAX_NX * MAX_NX)), ("R", C.c_double * (MAX_NU * MAX_NU)),
        ("u_goal", C.c_double * MAX_NU), ("mse_weight", C.c_double * MAX_NX), ("goal_tolerance", C.c_double),
        ("n_goal_rows", C.c_int32), ("x_goal", C.POINTER(C.c_double)),
        ("done_on_violation", C.c_int32), ("done_on_collision", C.c_int32), ("done_on_completion", C.c_int32),
        ("use_constraint_penalty", C.c_int32), ("constraint_penalty", C.c_double),
        ("n_constraints", C.c_int32), ("constraints", Constraint * MAX_CONSTRAINTS),
        ("n_dist", C.c_int32 * 3), ("dist", (Disturbance * MAX_DIST) * 3),
        ("adversary_channel", C.c_int32),
        ("init_state", C.c_double * (MAX_NX + 1)), ("init_rand", Rand * MAX_NX),
        ("inertia_nominal", C.c_double * 4), ("inertia_rand", Rand * 4),
        ("reseed_on_reset", C.c_int32),
        ("n_gates", C.c_int32), ("n_obstacles", C.c_int32),
        ("gates", (C.c_double * 7) * MAX_GATES), ("obstacles", (C.c_double * 6) * MAX_OBSTACLES),
        ("gate_rand", Rand), ("obstacle_rand", Rand),
        ("seed", C.c_uint32), ("env_id_base", C.c_int64),
        ("dyn_uses_episode_inertia", C.c_int32), ("dyn_applies_force", C.c_int32), ("gate_geometry", C.c_int32),
        ("auto_reset", C.c_int32),
        ("dyn_aero", C.c_int32), ("aero_drag_coeff", C.c_double * 3), ("aero_gnd_eff_coeff", C.c_double),
        ("aero_prop_radius", C.c_double), ("aero_gnd_eff_h_clip", C.c_double), ("aero_prop_xy", (C.c_double * 2) * 4)]


class Buffers(C.Structure):
    _fields_ = [("n", C.c_int64), ("state", C.c_void_p), ("ctrl_step", C.c_void_p), ("episode", C.c_void_p),
                ("flags", C.c_void_p), ("dist_offset", C.c_void_p), ("inertia", C.c_void_p),
                ("obs", C.c_void_p), ("env_state", C.c_void_p), ("reward", C.c_void_p), ("done", C.c_void_p),
                ("info", C.c_void_p), ("mse", C.c_void_p), ("c_values", C.c_void_p), ("terminal_obs", C.c_void_p),
                ("ep_return", C.c_void_p), ("ep_length", C.c_void_p), ("ep_violations", C.c_void_p),
                ("stats", C.c_void_p), ("adv_action", C.c_void_p), ("aero_mem", C.c_void_p)]


class RolloutArgs(C.Structure):
    _fields_ = [("actions", C.c_void_p), ("k_steps", C.c_int32), ("reward_out", C.c_void_p), ("done_out", C.c_void_p),
                ("log_out", C.c_void_p), ("log_stride", C.c_int32)]


class PidConfig(C.Structure):
    """scg_pid_config: defaults = PIDController.__init__ (quadrotor_utils.py:76-90)."""
    _fields_ = [("g", C.c_double), ("m", C.c_double), ("kf", C.c_double), ("km", C.c_double),
                ("pwm2rpm_scale", C.c_double), ("pwm2rpm_const", C.c_double), ("min_pwm", C.c_double), ("max_pwm", C.c_double),
                ("p_coeff_for", C.c_double * 3), ("i_coeff_for", C.c_double * 3), ("d_coeff_for", C.c_double * 3),
                ("p_coeff_tor", C.c_double * 3), ("i_coeff_tor", C.c_double * 3), ("d_coeff_tor", C.c_double * 3)]

    def __init__(self, g=9.8, m=0.027, kf=3.16e-10, km=7.94e-12, pwm2rpm_scale=0.2685, pwm2rpm_const=4070.3, min_pwm=20000,
                 max_pwm=65535, p_coeff_for=(.4, .4, 1.25), i_coeff_for=(.05, .05, .05), d_coeff_for=(.2, .2, .5),
                 p_coeff_tor=(70000., 70000., 60000.), i_coeff_tor=(.0, .0, 500.), d_coeff_tor=(20000., 20000., 12000.)):
        super().__init__()
        self.g, self.m, self.kf, self.km = g, m, kf, km
        self.pwm2rpm_scale, self.pwm2rpm_const, self.min_pwm, self.max_pwm = pwm2rpm_scale, pwm2rpm_const, min_pwm, max_pwm
        for name, v in (("p_coeff_for", p_coeff_for), ("i_coeff_for", i_coeff_for), ("d_coeff_for", d_coeff_for),
                        ("p_coeff_tor", p_coeff_tor), ("i_coeff_tor", i_coeff_tor), ("d_coeff_tor", d_coeff_tor)):
            setattr(self, name, (C.c_double * 3)(*[float(x) for x in v]))


_lib = None


def load():
    """Load libscg_b200.so (built in-tree by ``python -m safe_control_gym_b200.build``)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(f"{LIB_PATH} is missing: build it with `python -m safe_control_gym_b200.build` "
                          "(there is no CPU fallback for the batched env step)")
    lib = C.CDLL(LIB_PATH)
    vp, i32, i64 = C.c_void_p, C.c_int32, C.c_int64
    lib.scg_create.argtypes = [C.POINTER(Config), C.POINTER(vp)]
    lib.scg_destroy.argtypes = [vp]
    lib.scg_bind_buffers.argtypes = [vp, C.POINTER(Buffers)]
    lib.scg_set_injection.argtypes = [vp, C.c_int, C.c_int, vp]
    lib.scg_reset.argtypes = [vp, vp, vp, vp, vp, vp]
    lib.scg_step.argtypes = [vp, vp, vp]
    lib.scg_rollout.argtypes = [vp, C.POINTER(RolloutArgs), vp]
    lib.scg_rollout_host.argtypes = [vp, vp, i32, i32, vp]
    lib.scg_rollout_noalloc.argtypes = [vp, C.POINTER(RolloutArgs), vp]
    lib.scg_rollout_pid.argtypes = [vp, C.POINTER(PidConfig), vp, C.POINTER(RolloutArgs), vp, vp]
    lib.scg_last_rollout_spec.argtypes = [vp]
    lib.scg_set_spec_enabled.argtypes = [vp, C.c_int]
    lib.scg_config_word.argtypes = [vp]
    lib.scg_config_word.restype = C.c_uint32
    lib.scg_stats_reset.argtypes = [vp, vp]
    lib.scg_stats_read.argtypes = [vp, C.POINTER(C.c_double), vp]
    lib.scg_constraints_eval.argtypes = [vp, vp, vp, i64, vp, vp, vp]
    lib.scg_symbolic_eval.argtypes = [vp, C.c_int, vp, vp, i64, C.c_double, vp, vp, vp]
    lib.scg_symbolic_loss.argtypes = [vp, vp, vp, vp, vp, C.POINTER(C.c_double), C.POINTER(C.c_double), i64, vp, vp, vp, vp]
    lib.scg_gate_info.argtypes = [vp, vp, vp, vp]
    lib.scg_measure_fp32_peak.argtypes = [C.POINTER(C.c_double)]
    lib.scg_launch_count.argtypes = [vp]
    lib.scg_launch_count.restype = i64
    lib.scg_kernel_variant.argtypes = [vp]
    lib.scg_set_rollout_queue.argtypes = [vp, C.c_int32, C.c_int32]
    lib.scg_last_rollout_queue.argtypes = [vp]
    lib.scg_set_seed.argtypes = [vp, C.c_uint32]
    lib.scg_last_rollout_tlog.argtypes = [vp]
    lib.scg_last_error.restype = C.c_char_p
    lib.scg_version.restype = C.c_char_p
    for name in EXPORTS:
        fn = getattr(lib, name)
        if fn.restype is C.c_int:
            fn.restype = C.c_int
    _lib = lib
    return lib


class ScgError(RuntimeError):
    pass


def check(rc):
    if rc != 0:
        msg = load().scg_last_error().decode("utf8", "replace")
        if "pyb_freq is not divisible" in msg:
            raise ValueError(msg)
        raise ScgError(f"scg error {rc}: {msg}")
