// scg_params.cuh -- device-side parameter blocks built once by scg_create() from scg_config.
//
// Hot<real>  : passed BY VALUE as a __grid_constant__ kernel parameter, so every field is a
//              constant-bank operand (no load instruction, no register) in the substep loop.
// Cold<real> : library-owned device memory; read with uniform loads only when a feature
//              that needs it is enabled (dense constraint rows, disturbances, goal table, gates).
#pragma once
#include <cuda.h>   // CUtensorMap (types only: the encode entry point is fetched through the runtime, no -lcuda)

#include "scg_math.cuh"
#include "../../include/scg_b200.h"

namespace scg {

constexpr int MAX_BOUNDED = 4;   // bounded/default constraint objects kept in the hot block
constexpr int MAX_DENSE_ROWS = 24;

template <typename real> struct BoundedCon {
    real lo[SCG_MAX_NX], hi[SCG_MAX_NX];       // per FULL dim; inactive dims: -inf / +inf
    int8_t row_lo[SCG_MAX_NX], row_hi[SCG_MAX_NX];  // output row (YAML order) or -1
    int8_t var, strict, has_tol, symmetric;
    uint32_t mask;                             // bit d set: FULL dim d is constrained
};

template <typename real> struct RandP { int32_t distrib; real a, b; };

template <typename real> struct Hot {
    // timing
    int32_t substeps, ctrl_steps, ctrl_freq, pyb_freq;
    real dt;
    // physics (base_aviary.py:522-576)
    real kf, km, arm_s2;            // arm_s2 = L / sqrt(2)
    real inv_mass, gravity_force;   // nominal (parity mode)
    real J[3], J_inv[3];
    real dJ[3], kJ[3];             // fast substep (scg_step.cuh): (Izz-Iyy, Ixx-Izz, Iyy-Ixx) and dt / J
    // per-step scalars of the fast 3D path in two 16-byte rows: [dt, arm_s2, km_over_kf, inv_mass] [-dJ[1], kJ[1], -g / m * m = -GRAVITY_ACC, 0]
    alignas(16) real step_row[8];
    alignas(16) real ndJ_xz[2];    // (-dJ[0], -dJ[2]) and (kJ[0], kJ[2]): 64-bit constant operands of the packed substep, one 16-byte row
    real kJ_xz[2];
    real pwm_scale, pwm_const, pwm_min, pwm_max, inv_pwm_scale;
    real inv_nkf;                   // 1 / (n_motor * KF)
    real f_min, f_max, inv_n_motor, km_over_kf, inv_kf;   // per-rotor force at PWM_min / PWM_max, 1 / n_motor, KM / KF, 1 / KF (fast motor map)
    real gravity_acc, cp_length, cp_pole_mass, cp_cart_mass;   // cart-pole (A13): effective (half) pole length, masses
    // action map
    int32_t normalized_action;
    real norm_act_scale, hover_thrust;
    alignas(16) real act_low[SCG_MAX_NU];
    alignas(16) real act_high[SCG_MAX_NU];
    // state space / termination
    real state_low[SCG_MAX_NX], state_high[SCG_MAX_NX];
    int32_t done_on_out_of_bound, done_on_violation, done_on_collision, done_on_completion;
    // task / cost
    int32_t task, cost, rew_exponential, obs_goal_horizon, n_goal_rows;
    real rew_state_weight[SCG_MAX_NX], rew_act_weight[SCG_MAX_NU], u_goal[SCG_MAX_NU];
    real x_goal0[SCG_MAX_NX];       // stabilisation goal (row 0)
    real mse_weight[SCG_MAX_NX];
    uint32_t mse_mask;              // dims with a non-zero mse weight
    int32_t need_full_state;        // some per-step consumer reads the body rates p, q, r (else only the final obs does)
    real goal_tol;
    int32_t quadratic_diag;         // Q, R diagonal -> hot path, else dense from Cold
    real q_diag[SCG_MAX_NX], r_diag[SCG_MAX_NU];
    int32_t use_constraint_penalty;
    real constraint_penalty;
    // constraints
    int32_t n_bounded, n_dense_rows, n_rows, any_tolerance;
    int32_t need_cold_constraints;  // tolerances, symmetric or dense rows: the flags alone need constraint_values_cold
    BoundedCon<real> bounded[MAX_BOUNDED];
    // the rows of the non-symmetric bounded constraints as a flat table indexed by output row (Cold::crow_*; per-step constraint
    // VALUES of configs without tolerances / dense rows: constraint_values_flat); n_crows = 0 when some row is not of that kind
    int32_t n_crows, crow_strict;
    real viol_thr;                  // fp32: 5e-9 (== round(c, 8) > 0 in exact arithmetic)
    // LEAN configs: all bounded constraints merged into one box per state / input dim (violation flag only)
    real box_lo[SCG_MAX_NX], box_hi[SCG_MAX_NX];
    alignas(16) real ubox_lo[SCG_MAX_NU];   // 16-byte rows: one LDCU.128 each
    alignas(16) real ubox_hi[SCG_MAX_NU];
    uint32_t box_mask, ubox_mask;   // dims that carry a bound
    real box_thr;                   // violated <=> max_k(c_k) * 1e8 > box_thr (0.5, or just below -0.5 when strict)
    // the per-step position constants of a 3D env in four 16-byte rows (one LDCU.128 each instead of a dozen scalar loads):
    // [box_lo x y z, box_thr] [box_hi x y z, goal_tol^2] [x_goal0 x y z, goal_tol] [mse_weight x y z, 0]
    alignas(16) real pos_pack[16];
    // feature switches (uniform branches)
    int32_t n_dist_obs, n_dist_act, n_dist_dyn, adversary_channel;
    int32_t n_gates, gate_geometry, dyn_uses_episode_inertia, dyn_applies_force, auto_reset;
    int32_t reseed_on_reset, any_init_rand, any_inertia_rand, any_gate_rand;   // *_rand: bit 0 uniform, bit 1 normal draws needed
    real init_state[SCG_MAX_NX + 1];
    real inertia_nominal[4];
    RandP<real> init_rand[SCG_MAX_NX], inertia_rand[4];   // distrib NONE: a = b = 0
    uint32_t init_block_mask;       // Philox blocks (4 labels each) holding a randomised init label
    uint32_t seed;
    uint32_t env_id_base_lo;
    // CFG_DIST_SIMPLE: the action-channel disturbances the control-step loop evaluates INLINE -- impulse / step / white noise
    // without mask or injected draws, no adversary, no external force (BASELINE configs 3 and 4, competition/level1-3.yaml):
    // parameters as constant-bank operands instead of generic loads from Cold::dist behind an out-of-line call
    int32_t ds_n, ds_kind[SCG_MAX_DIST], ds_off[SCG_MAX_DIST] /* >= 0 fixed, -1: this episode's draw */, ds_slot[SCG_MAX_DIST];
    real ds_mag[SCG_MAX_DIST], ds_half[SCG_MAX_DIST], ds_decay[SCG_MAX_DIST], ds_l2decay[SCG_MAX_DIST], ds_std[SCG_MAX_DIST][SCG_MAX_NU];
    // gate_geometry: the first broad phase of the contact scan, inline (step_env): nominal axis (x, y) and widened squared radius of every
    // gate, then every obstacle.  A drone outside all of them -- almost every env-step -- never enters gate_scan (scg_env.cuh).
    // dyn_aero (extension): see aero_terms (scg_env.cuh)
    int32_t aero;
    real aero_drag[3], aero_gnd_coeff, aero_prop_radius, aero_h_clip, aero_prop_xy[4][2];
    int32_t geo_n;
    // two objects per packed operation: (-x_2j, -x_2j+1), (-y_2j, -y_2j+1), (r2_2j, r2_2j+1); an odd count is padded with an object at infinity
    alignas(8) float geo_negx[(SCG_MAX_GATES + SCG_MAX_OBSTACLES) / 2][2], geo_negy[(SCG_MAX_GATES + SCG_MAX_OBSTACLES) / 2][2];
    float geo_r2[(SCG_MAX_GATES + SCG_MAX_OBSTACLES) / 2][2];
    // the uniform switches the control-step loop tests, packed (CFG_*): one constant-bank word feeds all of
    // them, where a separate field per switch costs a dependent load -> compare -> branch chain each
    uint32_t cfg;
};

enum : uint32_t {
    CFG_STAB = 1u << 0,             // task == stabilization
    CFG_COST_RL = 1u << 1, CFG_COST_QUAD = 1u << 2, CFG_REW_EXP = 1u << 3,
    CFG_OOB = 1u << 4,              // done_on_out_of_bound
    CFG_DONE_COMPLETION = 1u << 5, CFG_DONE_VIOLATION = 1u << 6,
    CFG_HAS_ROWS = 1u << 7,         // some constraint row exists
    CFG_AUTO_RESET = 1u << 8,
    CFG_FULL_STATE = 1u << 9,       // a per-step consumer reads p, q, r (or the launch writes per-step outputs)
    CFG_RESEED = 1u << 10,
    CFG_UBOX = 1u << 11,            // LEAN: an input dim carries a bound
    CFG_MSE_RATES = 1u << 12,       // 3D: info["mse"] weighs p, q, r
    CFG_STEP_OUTPUTS = 1u << 13,    // set per launch
    CFG_GATE_GEOMETRY = 1u << 14,   // 3D: gate rays + contact scan on the integrated pose (extension)
    CFG_DISTURBED = 1u << 15,       // an action / dynamics disturbance, an adversary or an external force acts per step
    CFG_COLD_CONSTRAINTS = 1u << 16,// dense rows, tolerances, symmetric or mixed-strictness bounds: no merged box
    CFG_QR_DENSE = 1u << 17,        // quadratic cost with off-diagonal Q / R
    CFG_PENALTY = 1u << 18,         // rl_reward with use_constraint_penalty
    CFG_MSE_POS = 1u << 20,         // 3D: info["mse"] weighs nothing but the positions x, y, z (the default)
    CFG_BOX_POS = 1u << 21,         // 3D: the merged state box bounds nothing but x, y, z
    CFG_CLIP_ONCE = 1u << 22,       // 3D: the action bounds lie inside the rotor force bounds, the motor map needs no second clip
    CFG_DIST_SIMPLE = 1u << 23,     // (with CFG_DISTURBED) every per-step disturbance is one of Hot::ds_*: evaluated inline
    CFG_SPARE_SIMPLE = 1u << 24,    // no impulse / step disturbance with a per-episode offset outside the action channel: apply_spare_simple
    CFG_AERO = 1u << 25,            // 3D: drag / ground-effect terms per substep (extension; LEAN / GENERAL kernels only)
    CFG_RESET_EXTRAS = 1u << 19,    // reset() draws more than uniform init offsets: inertial properties, disturbance offsets, normal draws
};
// SPEC kernels (scg_step.cuh): the switches of the competition scenario family (competition/level0-3.yaml: stabilisation
// task, cost "competition", done on violation / completion / collision, bounded position + default input constraints,
// positions-only mse, auto-reset, no per-step outputs) as compile-time constants.  A launch takes the SPEC kernel only when
// its runtime word agrees with CFG_SPEC_VAL on every bit of CFG_SPEC_MASK.
constexpr uint32_t CFG_SPEC_MASK = CFG_STAB | CFG_COST_RL | CFG_COST_QUAD | CFG_OOB | CFG_DONE_COMPLETION | CFG_DONE_VIOLATION |
                                   CFG_HAS_ROWS | CFG_AUTO_RESET | CFG_FULL_STATE | CFG_RESEED | CFG_UBOX | CFG_MSE_RATES |
                                   CFG_STEP_OUTPUTS | CFG_MSE_POS | CFG_BOX_POS | CFG_CLIP_ONCE;
constexpr uint32_t CFG_SPEC_VAL = CFG_STAB | CFG_DONE_COMPLETION | CFG_DONE_VIOLATION | CFG_HAS_ROWS | CFG_AUTO_RESET | CFG_UBOX |
                                  CFG_MSE_POS | CFG_BOX_POS | CFG_CLIP_ONCE;

// Per env kind: which value of the switch word the SPEC instantiation of that kind is compiled for.  quad_type 3: the competition
// family above.  quad_type 2 (PLAIN + DS loop only): the RL family of BASELINE config 4 -- stabilisation, cost rl_reward with the
// exponential, out-of-bound termination, default constraints, auto-reset.  Cart-pole (PLAIN + DS only): BASELINE config 3 -- trajectory
// tracking, cost quadratic, done on violation and out of bound.  The launcher compares and falls back, as for quad_type 3.
template <int KIND> struct SpecCfg { static constexpr uint32_t MASK = CFG_SPEC_MASK, VAL = CFG_SPEC_VAL; };
template <> struct SpecCfg<SCG_QUAD_2D> {
    static constexpr uint32_t MASK = CFG_SPEC_MASK | CFG_REW_EXP;
    static constexpr uint32_t VAL = CFG_STAB | CFG_COST_RL | CFG_REW_EXP | CFG_OOB | CFG_HAS_ROWS | CFG_AUTO_RESET | CFG_FULL_STATE | CFG_UBOX;
};
template <> struct SpecCfg<SCG_CARTPOLE> {
    static constexpr uint32_t MASK = CFG_SPEC_MASK;
    static constexpr uint32_t VAL = CFG_COST_QUAD | CFG_OOB | CFG_DONE_VIOLATION | CFG_HAS_ROWS | CFG_AUTO_RESET | CFG_FULL_STATE | CFG_UBOX;
};


template <typename real> struct DenseRow {     // linear: c = a . x - b ; quadratic: c = x' P x - b
    real a[SCG_MAX_NX];
    real b, tol;
    int8_t var, strict, has_tol, quadratic;    // quadratic rows index Cold::P[p_index]
    int32_t out_row, p_index;
};

template <typename real> struct DistP {
    int32_t kind, has_mask, step_offset, max_step, slot, dim;
    real mask[SCG_MAX_NX];
    real magnitude, duration, half_duration, decay_rate;
    real low[SCG_MAX_NX], span[SCG_MAX_NX], std[SCG_MAX_NX];
    real scale, omega;              // omega = 2 pi frequency
    const real* inject;             // [dim][n] or nullptr
};

template <typename real> struct Cold {
    DenseRow<real> dense[MAX_DENSE_ROWS];
    real P[2][SCG_MAX_NX * SCG_MAX_NX];
    real bounded_tol[MAX_BOUNDED][2 * SCG_MAX_NX];   // per (object, lo/hi dim)
    real Q[SCG_MAX_NX * SCG_MAX_NX], R[SCG_MAX_NU * SCG_MAX_NU];
    DistP<real> dist[3][SCG_MAX_DIST];
    RandP<real> gate_rand, obstacle_rand;
    real gates[SCG_MAX_GATES][9], obstacles[SCG_MAX_OBSTACLES][8];   // reference rows + (cos, sin) of the nominal yaw
    real gate_broad2, obstacle_broad2;   // squared horizontal distance from the NOMINAL axis beyond which nothing can touch (inf: unbounded draws)
    int32_t n_obstacles;
    const real* x_goal;             // [n_goal_rows][nx] device
    // flat constraint rows (Hot::n_crows): source (< nx: state dim, else nx + input dim), sign (+1: x - bound, -1: bound - x), bound.
    // In the cold block: only launches with per-step outputs (scg_step) read them, through the L1.
    // one 16-byte entry per row: {source index, sign, bound, 0}, the index stored as a real (exact: < 32)
    alignas(16) real crow[SCG_MAX_ROWS][4];
};

// caller-owned batch memory, typed
template <typename real> struct Buf {
    int64_t n;
    real* state; int32_t* ctrl_step; int32_t* episode; uint32_t* flags; int32_t* dist_offset; real* inertia;
    real* obs; real* env_state; real* reward; uint8_t* done; uint8_t* info; real* mse; real* c_values;
    real* terminal_obs; real* ep_return; int32_t* ep_length; int32_t* ep_violations; double* stats;
    const real* adv_action;
    real* aero_mem;
};

// the DSL PID's gains and motor constants (quadrotor_utils.py:76-122), for closed-loop rollouts (scg_pid.cuh)
template <typename real> struct PidP {
    real p_for[3], i_for[3], d_for[3], p_tor[3], i_tor[3], d_tor[3];
    real gravity, inv_4kf, kf, pwm_scale, pwm_const, inv_pwm_scale, min_pwm, max_pwm, dt, inv_dt;
};

template <typename real> struct Roll {
    const real* actions; int32_t k_steps; real* reward_out; uint8_t* done_out; real* log_out; int32_t log_stride;
    int32_t step_outputs;           // 1: write the per-step outputs of Buf (scg_step), 0: only final obs/state
    int32_t use_pid;                // closed loop: actions come from the on-device PID instead of `actions`
    real* pid_mem;                  // [9][n] controller memory
    real* actions_out;              // optional [K][n][nu]: the actions the PID commanded
    PidP<real> pid;
    // time-chunk queue (k_rollout<..., QUEUE>): q_mem[0] = pop counter, q_mem[1] = push counter, q_mem[2 + s] = ready list;
    // q_tickets = chunks * q_env_blocks ranges of q_chunk control steps each
    int32_t* q_mem; int32_t q_chunk; int32_t q_env_blocks; int32_t q_tickets; uint32_t q_sleep_ns;
    int32_t pdl;                    // host-driven single steps: launched with programmatic stream serialization (see k_rollout)
};

// Staged trajectory log (k_rollout_tlog, scg_step.cuh): two TMA descriptors over the caller's log tensor viewed as the 2-D
// array [rows = logged steps x 16][n] of floats -- `full` stores a box of LOG_T logged steps x 16 fields x 32 envs (what one
// warp stages), `one` a single logged step (the tail of a launch).  Encoded on the host per launch (scg_k_3d_f_tlog.cu).
struct alignas(64) LogMaps { CUtensorMap full; CUtensorMap one; };
#ifndef SCG_LOG_T
#define SCG_LOG_T 2
#endif
constexpr int LOG_T = SCG_LOG_T;             // logged steps per TMA store
#ifndef SCG_LOG_STAGES
#define SCG_LOG_STAGES 1
#endif
constexpr int LOG_STAGES = SCG_LOG_STAGES;   // stages per warp.  One is enough: a stage is rewritten a whole control step after it was handed to the TMA
                                             // unit (the wait before the rewrite never blocks), and 8 KB per block keep nine blocks per SM resident --
                                             // measured 53.0 G env-steps/s against 49.9 G with two stages / eight blocks (profiles/r2_summary.md)
constexpr int LOG_STAGE = LOG_T * 16 * 32;   // floats of one stage of one warp (two warps per block)

template <typename real> struct ResetInject { const real* init; const real* inertia; const int32_t* offsets; };

template <int KIND> struct Dims;
template <> struct Dims<SCG_QUAD_1D> { static constexpr int NX = 2, NU = 1, NF = 1; };
template <> struct Dims<SCG_QUAD_2D> { static constexpr int NX = 6, NU = 2, NF = 2; };
template <> struct Dims<SCG_QUAD_3D> { static constexpr int NX = 12, NU = 4, NF = 3; };
template <> struct Dims<SCG_CARTPOLE> { static constexpr int NX = 4, NU = 1, NF = 2; };

}  // namespace scg
