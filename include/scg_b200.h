/*
 * scg_b200.h -- C ABI of the B200-native batched env-step library (libscg_b200.so).
 *
 * Drop-in boundary for ONE hot path of ekumenlabs/safe-control-gym: the per-control-step
 * work of Quadrotor.step()/reset() with physics=DYN (and the CartPole symbolic step), run
 * in lock-step over N independent environments on one GPU.  The reference has no FFI for
 * this path (it is pure Python); each entry point below names the reference interface it
 * replaces (paths under safe_control_gym/).  The Python binding a maintainer would add is
 * a ctypes stub; safe_control_gym_b200/_lib.py is that stub.
 *
 * Conventions
 *  - plain C types only; every function returns an int status (SCG_OK = 0, negative =
 *    error; scg_last_error() gives the text); nothing throws, nothing aborts.
 *  - all batch memory is owned by the CALLER (device pointers, e.g. torch tensors);
 *    the library owns only a small per-handle constant block.
 *  - work is enqueued on the given CUDA stream (cudaStream_t passed as void*), no hidden
 *    synchronisation.  One handle is used from one host thread at a time.
 *  - batch tensors are struct-of-arrays "[field][n]" of the handle's real type
 *    (float for precision 32, double for 64) unless stated; actions are env-major
 *    "[n][nu]" (one 16-byte vector load per 3D env).
 */
#ifndef SCG_B200_H
#define SCG_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define SCG_OK 0
#define SCG_ERR_BAD_CONFIG (-1)
#define SCG_ERR_NULL (-2)
#define SCG_ERR_SHAPE (-3)
#define SCG_ERR_STATE (-4)   /* e.g. step before reset (benchmark_env.py:268-275) */
#define SCG_ERR_CUDA (-100)  /* -100 - cudaError_t */

#define SCG_MAX_NX 12
#define SCG_MAX_NU 4
#define SCG_MAX_CONSTRAINTS 8   /* constraint objects (YAML entries) */
#define SCG_MAX_CON_ROWS 24     /* values per object: bounded = 2 * 12 */
#define SCG_MAX_ROWS 64         /* total constraint values */
#define SCG_MAX_DIST 4          /* disturbances per channel */
#define SCG_MAX_GATES 8
#define SCG_MAX_OBSTACLES 8
#define SCG_STATS_LEN 16
/* The device accumulators are SCG_STAT_REPLICAS copies of the statistics vector; a warp adds to the copy its index selects and the
 * totals are the sums over the copies (scg_stats_read returns them).  One copy made every warp of a launch queue on the same
 * few L2 words: 6 us of a 15 us host-driven step at 65,536 envs. */
#define SCG_STAT_REPLICAS 64

/* env_kind: QuadType (quadrotor_utils.py:11-18) or the cart-pole */
enum { SCG_QUAD_1D = 1, SCG_QUAD_2D = 2, SCG_QUAD_3D = 3, SCG_CARTPOLE = 10 };
/* Task / Cost (benchmark_env.py:19-33) */
enum { SCG_TASK_STABILIZATION = 0, SCG_TASK_TRAJ_TRACKING = 1 };
enum { SCG_COST_RL_REWARD = 0, SCG_COST_QUADRATIC = 1, SCG_COST_COMPETITION = 2 };
/* constraint_form (constraints.py:586-591, 371-415) */
enum { SCG_CON_BOUNDED = 0, SCG_CON_LINEAR = 1, SCG_CON_QUADRATIC = 2, SCG_CON_SYMMETRIC = 3 };
enum { SCG_VAR_STATE = 0, SCG_VAR_INPUT = 1 };
/* disturbance_func (disturbances.py:307-312) and channel (quadrotor.py:808-813) */
enum { SCG_DIST_IMPULSE = 0, SCG_DIST_STEP = 1, SCG_DIST_UNIFORM = 2, SCG_DIST_WHITE = 3, SCG_DIST_PERIODIC = 4 };
enum { SCG_CH_OBSERVATION = 0, SCG_CH_ACTION = 1, SCG_CH_DYNAMICS = 2 };
/* distribution of a randomised reset value (benchmark_env.py:277-308) */
enum { SCG_RAND_NONE = 0, SCG_RAND_UNIFORM = 1, SCG_RAND_NORMAL = 2 };
/* attitude update: literal PyBullet quaternion round trip, or the equivalent wrap with
 * the round trip kept only outside the identity region */
enum { SCG_ATT_ROUNDTRIP = 0, SCG_ATT_FAST = 1, SCG_ATT_FAST_INTRINSIC = 2 /* fp32: sin/cos on the SFU */ };
/* scg_symbolic_eval ops (symbolic_systems.py:63-123) */
enum { SCG_SYM_FC = 0, SCG_SYM_FD_RK4 = 1, SCG_SYM_DF = 2, SCG_SYM_LOSS = 3, SCG_SYM_FD_EULER = 4 };

/* info bits written per env by scg_step (quadrotor.py:1004-1134, benchmark_env.py:437-461) */
#define SCG_INFO_VIOLATION   0x01u  /* info["constraint_violation"] */
#define SCG_INFO_TRUNCATED   0x02u  /* info["TimeLimit.truncated"]  */
#define SCG_INFO_COLLISION   0x04u  /* info["collision"][1]          */
#define SCG_INFO_GATE_PASSED 0x08u  /* stepped_through_gate          */
#define SCG_INFO_AT_GOAL     0x10u  /* info["at_goal_position"]      */
#define SCG_INFO_COMPLETED   0x20u  /* info["task_completed"]        */
#define SCG_INFO_GOAL_REACHED 0x40u /* info["goal_reached"]: the PREVIOUS step's flag (the reference's _get_info runs before _get_done) */
#define SCG_INFO_OUT_OF_BOUND 0x80u /* out-of-bound termination      */

/* layout of the per-env packed "flags" word kept between steps */
#define SCG_FLAG_CNSTR_VIOLATION 0x1u  /* previous step's violation (benchmark_env.py:438-449, F8) */
#define SCG_FLAG_AT_GOAL         0x2u
#define SCG_FLAG_FLIPPED         0x4u  /* 2D: Euler fold left roll = yaw = pi */
#define SCG_FLAG_COMPLETED       0x8u
#define SCG_FLAG_GOAL_REACHED    0x10u /* goal_reached of the previous _get_done (quadrotor.py:965, read at :1013 one step later) */
#define SCG_FLAG_GATE_SHIFT 8          /* bits 8..15  current_gate       */
#define SCG_FLAG_HOLD_SHIFT 16         /* bits 16..31 steps_at_goal_pos  */

/* indices into the float64[SCG_STATS_LEN] episode-statistics vector (record_episode_statistics.py:142-169) */
enum { SCG_STAT_EPISODES = 0, SCG_STAT_RETURN = 1, SCG_STAT_LENGTH = 2, SCG_STAT_VIOLATION_STEPS = 3,
       SCG_STAT_MSE = 4, SCG_STAT_GATES = 5, SCG_STAT_COLLISIONS = 6, SCG_STAT_TRUNCATED = 7,
       SCG_STAT_COMPLETED = 8, SCG_STAT_ENV_STEPS = 9, SCG_STAT_VIOLATED_EPISODES = 10 };

/* one YAML constraint entry (constraints.py:21-415) */
typedef struct {
    int32_t form;                          /* SCG_CON_*  */
    int32_t var;                           /* SCG_VAR_*  */
    int32_t strict;                        /* constraints.py:137-140 */
    int32_t n_rows;                        /* num_constraints */
    int32_t n_active;                      /* filtered input dim */
    int32_t active_dims[SCG_MAX_NX];       /* rows of the filter eye(dim)[active_dims] (:77) */
    double lower[SCG_MAX_NX];              /* bounded / default / symmetric(-bound) */
    double upper[SCG_MAX_NX];
    double A[SCG_MAX_NX * SCG_MAX_NX];     /* linear: n_rows x n_active row-major; quadratic: P */
    double b[SCG_MAX_CON_ROWS];            /* linear rows / quadratic b[0] */
    int32_t has_tolerance;                 /* is_almost_active (:143-158) */
    double tolerance[SCG_MAX_CON_ROWS];
} scg_constraint;

/* one YAML disturbance entry (disturbances.py:89-288) */
typedef struct {
    int32_t kind;                          /* SCG_DIST_* */
    int32_t slot;                          /* row in dist_offset / Philox randint slot: running index
                                              over all disturbances in YAML order (DisturbanceList.reset) */
    int32_t has_mask;
    double mask[SCG_MAX_NX];
    double magnitude;                      /* impulse / step */
    int32_t step_offset;                   /* >= 0 fixed, -1 = randint(max_step) at reset */
    int32_t max_step;                      /* int(EPISODE_LEN_SEC / CTRL_TIMESTEP) */
    double duration, decay_rate;           /* impulse */
    double low[SCG_MAX_NX], high[SCG_MAX_NX]; /* uniform */
    double std[SCG_MAX_NX];                /* white noise */
    double scale, frequency;               /* periodic */
    /* optional injected draws, device pointer [dim][n] for the NEXT step/reset call (raw
     * values as numpy would return them: U(low,high), N(0,std), phase); NULL = Philox */
    const void* inject;
} scg_disturbance;

typedef struct {
    int32_t distrib;                       /* SCG_RAND_* */
    double a, b;                           /* uniform: low, high; normal: loc, scale */
} scg_rand;

/* everything Quadrotor.__init__ / BenchmarkEnv.__init__ derive from the YAML kwargs */
typedef struct {
    int32_t env_kind, precision, attitude_mode;
    int32_t substeps;                      /* PYB_STEPS_PER_CTRL (benchmark_env.py:156) */
    int32_t ctrl_steps;                    /* CTRL_STEPS (:161) */
    int32_t ctrl_freq, pyb_freq;
    double pyb_dt;                         /* PYB_TIMESTEP == the DYN integrator step (F3) */
    /* physical constants (base_aviary.py:104-147, cf2x.urdf) */
    double mass, gravity_acc, gravity_force, J[3], J_inv[3], kf, km, arm;
    double pwm2rpm_scale, pwm2rpm_const, pwm_min, pwm_max;
    /* cart-pole (A13): effective pole half-length, pole mass, cart mass */
    double cp_length, cp_pole_mass, cp_cart_mass;
    /* action map (quadrotor.py:699-733, 815-842) */
    int32_t normalized_action;
    double norm_act_scale, hover_thrust;
    double act_low[SCG_MAX_NU], act_high[SCG_MAX_NU];
    /* state space (quadrotor.py:735-806) */
    double state_low[SCG_MAX_NX], state_high[SCG_MAX_NX];
    int32_t done_on_out_of_bound;
    /* task / cost */
    int32_t task, cost, rew_exponential, obs_goal_horizon;
    double rew_state_weight[SCG_MAX_NX], rew_act_weight[SCG_MAX_NU];
    double Q[SCG_MAX_NX * SCG_MAX_NX], R[SCG_MAX_NU * SCG_MAX_NU];
    double u_goal[SCG_MAX_NU], mse_weight[SCG_MAX_NX], goal_tolerance;
    int32_t n_goal_rows;                   /* 1 (stabilisation) or T (trajectory) */
    const double* x_goal;                  /* HOST pointer, n_goal_rows x nx row-major (copied) */
    /* termination (benchmark_env.py:422-463, quadrotor.py:956-1002) */
    int32_t done_on_violation, done_on_collision, done_on_completion;
    int32_t use_constraint_penalty;
    double constraint_penalty;
    /* constraints */
    int32_t n_constraints;
    scg_constraint constraints[SCG_MAX_CONSTRAINTS];
    /* disturbances: [channel][k] */
    int32_t n_dist[3];
    scg_disturbance dist[3][SCG_MAX_DIST];
    int32_t adversary_channel;             /* -1 none, SCG_CH_ACTION, SCG_CH_DYNAMICS */
    /* reset (quadrotor.py:356-504) */
    double init_state[SCG_MAX_NX + 1];     /* nominal INIT_*, state order; 3D/1D unused tail = 0 */
    scg_rand init_rand[SCG_MAX_NX];
    double inertia_nominal[4];             /* M, Ixx, Iyy, Izz after inertial_prop override */
    scg_rand inertia_rand[4];
    int32_t reseed_on_reset;               /* every episode replays episode 0's draws */
    /* gates / obstacles (quadrotor.py:371-434, 1060-1130) */
    int32_t n_gates, n_obstacles;
    double gates[SCG_MAX_GATES][7];
    double obstacles[SCG_MAX_OBSTACLES][6];
    scg_rand gate_rand, obstacle_rand;
    /* RNG contract (oracle/philox.py): key = (seed, env_id_base + i) */
    uint32_t seed;
    int64_t env_id_base;
    /* extensions beyond the reference's DYN behaviour (off = parity mode) */
    int32_t dyn_uses_episode_inertia;      /* F9 */
    int32_t dyn_applies_force;             /* F5 */
    int32_t gate_geometry;                 /* analytic gate-pass / collision tests */
    int32_t auto_reset;                    /* DummyVecEnv.step_wait semantics (dummy_vec_env.py:38-49) */
    /* extension (quad_type 3): the aerodynamic terms of the reference's PYB_* modes (base_aviary.py:437-497) folded
     * into DYN as one more external force / body torque per substep.  Bit 0: _drag, bit 1: _ground_effect (_downwash is zero for
     * one drone per env).  Coefficients as _parse_urdf_parameters reads them (base_aviary.py:637-644, :145-147). */
    int32_t dyn_aero;
    double aero_drag_coeff[3];             /* DRAG_COEFF = (xy, xy, z) */
    double aero_gnd_eff_coeff;             /* GND_EFF_COEFF */
    double aero_prop_radius;               /* PROP_RADIUS */
    double aero_gnd_eff_h_clip;            /* GND_EFF_H_CLIP */
    double aero_prop_xy[4][2];             /* propeller link positions in the body frame (cf2x.urdf) */
} scg_config;

/* caller-owned batch memory (device pointers; NULL = not wanted where marked optional) */
typedef struct {
    int64_t n;
    /* persistent per-env state */
    void* state;              /* [nx][n] real, reference state order; angles are the pre-round-trip Euler
                                 angles, rate slots hold rpy_rates (base_aviary.py:576) */
    int32_t* ctrl_step;       /* [n] ctrl_step_counter */
    int32_t* episode;         /* [n] episodes started (Philox counter word) */
    uint32_t* flags;          /* [n] SCG_FLAG_* */
    int32_t* dist_offset;     /* [sum n_dist][n] impulse / step offsets, optional if none */
    void* inertia;            /* [4][n] real episode M, Ixx, Iyy, Izz, optional */
    /* per-step outputs */
    void* obs;                /* [obs_dim][n] */
    void* env_state;          /* [nx][n] env.state (optional) */
    void* reward;             /* [n] */
    uint8_t* done;            /* [n] */
    uint8_t* info;            /* [n] SCG_INFO_* */
    void* mse;                /* [n] optional */
    void* c_values;           /* [rows][n] optional, info["constraint_values"] */
    void* terminal_obs;       /* [obs_dim][n] optional (auto_reset) */
    /* episode statistics (optional) */
    void* ep_return;          /* [n] real running return */
    int32_t* ep_length;       /* [n] */
    int32_t* ep_violations;   /* [n] */
    double* stats;            /* [SCG_STAT_REPLICAS][SCG_STATS_LEN] device accumulators (totals: sum over the replicas) */
    /* adversary action for the next step, [n][dim] env-major (optional) */
    const void* adv_action;
    /* dyn_aero drag: sum(2 pi rpm / 60) of the previous substep's action (base_aviary.py:224, :297, :320), [n] real; required then */
    void* aero_mem;
} scg_buffers;

/* K-step fused rollout (state stays in registers for all K * substeps integrator steps) */
typedef struct {
    const void* actions;      /* [K][n][nu] real, device */
    int32_t k_steps;
    void* reward_out;         /* [K][n] optional */
    uint8_t* done_out;        /* [K][n] optional */
    void* log_out;            /* [K][16][n] optional: pos3 vel3 rpy3 ang_vel3 rpm4 (Logger.py:50-66) */
    int32_t log_stride;       /* log every log_stride-th ctrl step (>= 1) */
} scg_rollout_args;

typedef struct scg_env scg_env;

/* Quadrotor.__init__ (quadrotor.py:148-354) / make() (utils/registration.py:145-149) */
int scg_create(const scg_config* cfg, scg_env** out);
/* close() (base_aviary.py:198-206) */
int scg_destroy(scg_env* env);
/* (re)bind the batch tensors; validates n and required pointers */
int scg_bind_buffers(scg_env* env, const scg_buffers* buf);
/* update injected-draw pointers / adversary pointer between calls (cheap, host only) */
int scg_set_injection(scg_env* env, int channel, int k, const void* dev_ptr);
/* BenchmarkEnv.seed(seed) (benchmark_env.py:220-235) after construction: every later reset / disturbance draw is keyed
 * by (seed, global env id).  Host only, takes effect with the next launch. */
int scg_set_seed(scg_env* env, uint32_t seed);
/* Quadrotor.reset() (quadrotor.py:356-504): mask = NULL resets every env, else envs with
 * mask[i] != 0.  inject_* optional device tensors of pre-drawn additive offsets:
 * init [nx][n], inertia [4][n], dist offsets [slots][n]. */
int scg_reset(scg_env* env, const uint8_t* mask, const void* inject_init, const void* inject_inertia,
              const int32_t* inject_offsets, void* stream);
/* Quadrotor.step(action) (quadrotor.py:507-558) for all n envs; action [n][nu] device */
int scg_step(scg_env* env, const void* action, void* stream);
/* K fused ctrl steps with auto-reset; the batched analogue of the PPO rollout loop
 * (controllers/ppo/ppo.py:247-273) driving DummyVecEnv.step_wait */
int scg_rollout(scg_env* env, const scg_rollout_args* args, void* stream);
/* same, actions in pinned HOST memory [K][n][nu]; copies are chunked and overlapped
 * with compute on an internal copy stream; final obs / reward sums land in the bound buffers */
int scg_rollout_host(scg_env* env, const void* host_actions, int32_t k_steps, int32_t chunk_steps, void* stream);
/* zero the accumulators / read their totals, float64[SCG_STATS_LEN] (summed over the replicas on the device, async copy to host_out) */
int scg_stats_reset(scg_env* env, void* stream);
int scg_stats_read(scg_env* env, double* host_out, void* stream);
/* ConstraintList.get_values / is_violated (constraints.py:503-563) on caller-given
 * state [nx][n] and raw action [n][nu]; out_values [rows][n], out_flags [n] */
int scg_constraints_eval(scg_env* env, const void* state, const void* action, int64_t n,
                         void* out_values, uint8_t* out_flags, void* stream);
/* SymbolicModel.fc_func / fd (RK4, mpc_utils.py:79-101) / df_func / loss
 * (symbolic_systems.py:63-123): x [nx][n], u [nu][n].  out0 / out1 per op:
 *   SCG_SYM_FC f [nx][n]; SCG_SYM_FD_RK4 / _EULER x_next [nx][n] over dt;
 *   SCG_SYM_DF df/dx [nx*nx][n] (entry r*nx+c), df/du [nx*nu][n]; SCG_SYM_LOSS l [n], l_x [nx][n]
 *   (handle's Q, R, U_GOAL, Xr = 0; scg_symbolic_loss takes them as arguments) */
int scg_symbolic_eval(scg_env* env, int op, const void* x, const void* u, int64_t n, double dt,
                      void* out0, void* out1, void* stream);
/* SymbolicModel.loss(x, u, Xr, Ur, Q, R) (symbolic_systems.py:114-123): xr [nx][n] / ur [nu][n] device or
 * NULL (0 / U_GOAL); Q [nx*nx], R [nu*nu] HOST row-major or NULL (the handle's); out_l [n],
 * out_lx [nx][n] and out_lu [nu][n] optional.  l_xx = (Q + Q')/2, l_uu = (R + R')/2, l_xu = 0 are
 * constants the host forms itself. */
int scg_symbolic_loss(scg_env* env, const void* x, const void* u, const void* xr, const void* ur, const double* Q,
                      const double* R, int64_t n, void* out_l, void* out_lx, void* out_lu, void* stream);
/* FP32 FMA microbenchmark (measurement aid, not on the path): runs a register-only FFMA kernel on the
 * current device and returns the achieved TFLOP/s (FMA = 2 flop) -- the roofline denominator for the
 * compute-bound fused rollout, since MEASURED_PEAKS.json carries no non-tensor fp32 figure. */
int scg_measure_fp32_peak(double* tflops_out);
/* number of kernels this handle has launched (bench.py "gpu_launches") */
/* Gains and motor constants of the DSL CF2.x PID (quadrotor_utils.py:76-122, `PIDController.__init__`). */
typedef struct scg_pid_config {
    double g, m, kf, km, pwm2rpm_scale, pwm2rpm_const, min_pwm, max_pwm;
    double p_coeff_for[3], i_coeff_for[3], d_coeff_for[3], p_coeff_tor[3], i_coeff_tor[3], d_coeff_tor[3];
} scg_pid_config;

/* Closed-loop fused rollout: like scg_rollout, but the action of every control step is computed on the device by the
 * PID from the env's state before the step and the reference point X_GOAL (stabilisation) / X_GOAL[ctrl_step]
 * (trajectory tracking), as the reference's loops do on the host (competition_utils.py:338-356 `thrusts` for
 * quad_type 3, controllers/pid/pid.py:97-141 for quad_type 2).  args->actions is ignored.  pid_mem [9][n] (precision
 * of the env) is the controller memory (integral position error, last rpy, integral rotation error); zero it for a
 * fresh controller; it is zeroed for an env whenever that env auto-resets.  actions_out (optional) [k][n][nu]. */
int scg_rollout_pid(scg_env* env, const scg_pid_config* pid, void* pid_mem, const scg_rollout_args* args, void* actions_out,
                    void* stream);

/* info["current_target_gate_in_range"] / ["current_target_gate_pos"] of Quadrotor._get_info (quadrotor.py:1093-1112)
 * for every env, from the bound state / flags / episode buffers: out_pose [6][n] (precision of the env) = the
 * current gate's true pose (x, y, z + height, r, p, yaw incl. this episode's offsets) when the drone is within the
 * 0.45 m visibility range of the gate body, else its nominal row; zeros once all gates are passed.
 * out_in_range [n] u8.  Either output may be NULL.  quad_type 3 only. */
int scg_gate_info(scg_env* env, void* out_pose, uint8_t* out_in_range, void* stream);
int64_t scg_launch_count(const scg_env* env);
/* Which instantiation of the step kernel the config selected.  1 = LEAN: the auto-reset takes a spare episode start
 * (initial state, inertial properties, disturbance offsets) drawn ahead of time at a warp-convergent cadence;
 * 2 = LEAN + PLAIN (fp32): in addition reset() draws nothing but uniform initial-state offsets and no per-step extra
 * is configured (disturbances, adversary, external force, dense / tolerance / symmetric constraints, off-diagonal Q
 * or R, constraint penalty, gate geometry), so all of that is compiled out; 3 = PLAIN + SPEC: a Quadrotor 3D config of
 * the competition family (stabilisation task, cost "competition", done on violation and completion, position box +
 * input bounds, positions-only mse, auto-reset) whose fused launches without per-step outputs run with those switches
 * as compile-time constants (scg_step and launches with per-step outputs take variant 2); 4 = PLAIN + DS (fp32; quad_type 2, 3 and
 * the cart-pole): like 2, but the config's extras are the cheap ones -- action-channel impulse / step / white-noise disturbances
 * evaluated inline, per-episode inertial properties and offsets taken from the spare (quad_type 3: with the nominal inertia in
 * the integrator, the reference's own DYN behaviour), gate geometry behind its inline broad phase -- so the rest of the LEAN
 * kernel is compiled out (BASELINE configs 3, 4, 5 and the level1-3 randomisation sets; competition-family configs run with the
 * switches of variant 3 compiled in as well; closed-loop launches, scg_rollout_pid, of a variant 2 / 3 config take this loop
 * too); 0 = GENERAL: SCG_ATT_ROUNDTRIP only
 * (validation: literal quaternion round trip, reset one lane at a time).  Same results in every variant. */
int scg_kernel_variant(const scg_env* env);
/* Diagnostic: the packed word of uniform scenario switches the kernels branch on (csrc/scg_params.cuh: CFG_*); the SPEC
 * kernels are the ones compiled for particular values of it.  No reference counterpart. */
uint32_t scg_config_word(const scg_env* env);
/* Diagnostics: whether the last fused launch ran a SPEC instantiation; keep every launch of this handle on the non-SPEC kernels
 * (enabled = 0) -- same results, used by the parity tests and for A/B measurements.  No reference counterpart. */
int scg_last_rollout_spec(const scg_env* env);
int scg_set_spec_enabled(scg_env* env, int enabled);
/* Scheduling of the fused rollout launches (scg_rollout / scg_rollout_host / scg_rollout_pid; no reference counterpart:
 * the reference steps one env at a time).  mode -1 (default) = automatic: launches of >= 64 control steps whose env
 * blocks fill the resident block slots of the GPU unevenly (between half a wave and six waves) are cut into chunks of
 * control steps that the resident blocks take from a ticket counter, so every SM sub-partition stays equally loaded;
 * 0 = always one block per 64 envs for the whole launch; 1 = always through the queue.  chunk_steps 0 = default (64).
 * The environment variables SCG_QUEUE / SCG_QUEUE_CHUNK give the initial values.  Results do not depend on it (a K-step
 * rollout equals any split of it into launches); the episode statistics agree up to the order of their float sums. */
int scg_set_rollout_queue(scg_env* env, int32_t mode, int32_t chunk_steps);
/* scg_rollout without any allocation.  No counterpart in the reference; added for
 * callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * Same arguments and results as scg_rollout, always through the time-chunk ticket queue, whose scratch (pop counter, push
 * counter, ready list) lives in a file-scope static device array of SCG_NOALLOC_QUEUE_WORDS int32 (512 KiB, held by every
 * GPU that loads the library until exit) instead of the handle's cudaMalloc'ed list.  Limits: precision 32 with a fast
 * attitude mode, k_steps >= 64, and 2 + ceil(k_steps / chunk) * ceil(n / 64) <= SCG_NOALLOC_QUEUE_WORDS (chunk = 64 unless
 * scg_set_rollout_queue changed it), e.g. 65,536 envs x 8,000 steps or 1 M envs x 448 steps per launch; anything beyond
 * returns SCG_ERR_SHAPE with the numbers in scg_last_error().  Calls no cudaMalloc / cudaFree / allocating library routine;
 * the words it uses are cleared on `stream` before the kernel reads them.  ALL handles and calls on one GPU share the
 * array: calls that overlap on the GPU (different streams, no ordering between them) race -- keep them on one stream or
 * order them with events.  The kernel's first launch in a process loads its module; warm up once before capturing. */
#define SCG_NOALLOC_QUEUE_WORDS 131072
int scg_rollout_noalloc(scg_env* env, const scg_rollout_args* args, void* stream);
/* 1 when the last fused rollout launch went through the ticket queue, else 0 */
int scg_last_rollout_queue(const scg_env* env);
/* 1 when the last fused rollout launch wrote its trajectory log (scg_rollout_args.log_out: the Logger record,
 * envs/gym_pybullet_drones/Logger.py:50-66) through the TMA unit -- records staged in shared memory, one bulk tensor store
 * per warp and pair of logged steps -- else 0 (plain strided stores).  Taken for quad_type 3, precision 32, a fast attitude
 * mode, n % 4 == 0 and a 16-byte aligned log tensor, unless the queue is forced on or SCG_TLOG=0.  Same bytes either way. */
int scg_last_rollout_tlog(const scg_env* env);
const char* scg_last_error(void);
const char* scg_version(void);

#ifdef __cplusplus
}
#endif
#endif /* SCG_B200_H */
