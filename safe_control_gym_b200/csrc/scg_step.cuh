// scg_step.cuh -- the fused reset / step / K-step rollout kernel for the quadrotor and cart-pole envs.
// One thread = one env.  Everything Quadrotor.step() does per control step (rows A1-A11 of
// SURVEY.md section 8a) plus DummyVecEnv's auto-reset (A14) happens between one load and one
// store of the env's state; the substep loop touches registers and constant-bank operands only.
//
// Code-size discipline (profiles/r1a_summary.md): the control-step loop must fit the 32 KB L1.5
// instruction cache, because 65,536 envs are only 3.5 warps per scheduler and nothing hides an
// instruction-fetch miss.  So (a) everything rare -- reset with its Philox draws, the literal
// quaternion round trip, disturbance kinds, dense constraint rows, the extended observation -- lives
// in __noinline__ functions outside the loop, and (b) the kernel is instantiated twice: LEAN (no
// disturbances, no adversary, no external force, bounded constraints without tolerances, nominal
// inertia, diagonal Q/R, plain observation: what scg_create() finds in e.g. level0.yaml) compiles
// those features out, GENERAL keeps them behind uniform branches.
#pragma once
#include <type_traits>

#include "scg_env.cuh"
#include "scg_pid.cuh"

namespace scg {

// 65,536 envs -> 1,024 blocks = 6.9 per SM.  The register file is split per scheduler (16,384 registers each), so
// the 3.5 warps per scheduler of that batch are resident together only at <= 128 registers per thread: the LEAN
// fp32 kernel is bounded to 8 blocks per SM (one wave); at 255 registers half of the blocks wait for a second wave.
constexpr int ENV_BLOCK = 64;

template <int KIND, typename real> struct EnvRegs {
    static constexpr int NX = Dims<KIND>::NX;
    real x[NX];          // internal state, reference order; 3D angles pre-round-trip, rates = rpy_rates
    int32_t t;           // ctrl_step_counter
    int32_t resets;      // number of resets so far (episode index = resets - 1)
    // SCG_FLAG_* of the flags buffer, unpacked while the env lives in registers (2D's fold bit is a2.flipped)
    bool viol, at_goal, completed;
    bool goal_prev;      // goal_reached of the previous _get_done (info["goal_reached"] lags one step, F10); survives reset
    uint32_t gate, hold;
    Inertia<real> in;
    real* offc;          // CFG_DIST_SIMPLE: this episode's impulse / step offsets (Hot::ds_off < 0), mirrored from Buf::dist_offset into a
                         // per-thread shared-memory column [k * ENV_BLOCK] (exact small integers; registers are what the loop is short of)
    Att3<real> a3;       // 3D
    Att2<real> a2;       // 2D
    // PAIRED loops (3D fp32 PLAIN, see step_env): the state of the control-step loop in packed pairs -- (x, y), (vx, vy),
    // (roll rate, yaw rate), (z, pitch), (vz, pitch rate), (roll, yaw), thrust direction (x, y) and z.  While they are
    // live, x[0..5], x[9..11], a3.rpy and a3.R[2 / 5 / 8] are not (x[6..8] and a3.sc stay scalar).
    f32x2 Pxy, Vxy, Wxz, ZA, ZB, Axz, Rxy;
    real R8;
    real aero_prev;      // dyn_aero drag: sum(2 pi rpm / 60) of the previous substep's action (LEAN / GENERAL kernels only)
};

#ifndef SCG_NO_PAIRS
#define SCG_NO_PAIRS 0
#endif

template <int KIND, typename real, int ATT, bool LEAN, bool PLAIN> struct Paired {
    static constexpr bool value = KIND == SCG_QUAD_3D && sizeof(real) == 4 && ATT != SCG_ATT_ROUNDTRIP && LEAN && PLAIN && !SCG_NO_PAIRS;
};
SCG_DEV void pairs_from_scalars(EnvRegs<SCG_QUAD_3D, float>& e) {
    e.Pxy = pk2(e.x[0], e.x[2]); e.Vxy = pk2(e.x[1], e.x[3]); e.Wxz = pk2(e.x[9], e.x[11]); e.ZB = pk2(e.x[5], e.x[10]);
    e.ZA = pk2(e.x[4], e.a3.rpy[1]); e.Axz = pk2(e.a3.rpy[0], e.a3.rpy[2]); e.Rxy = pk2(e.a3.R[2], e.a3.R[5]); e.R8 = e.a3.R[8];
}
SCG_DEV void scalars_from_pairs(EnvRegs<SCG_QUAD_3D, float>& e) {
    e.x[0] = lo2(e.Pxy); e.x[2] = hi2(e.Pxy); e.x[1] = lo2(e.Vxy); e.x[3] = hi2(e.Vxy); e.x[4] = lo2(e.ZA); e.x[5] = lo2(e.ZB);
    e.x[9] = lo2(e.Wxz); e.x[10] = hi2(e.ZB); e.x[11] = hi2(e.Wxz);
    e.a3.rpy[0] = lo2(e.Axz); e.a3.rpy[1] = hi2(e.ZA); e.a3.rpy[2] = hi2(e.Axz);
    e.a3.R[2] = lo2(e.Rxy); e.a3.R[5] = hi2(e.Rxy); e.a3.R[8] = e.R8;
    e.a3.form = 1;
}
template <int KIND, typename real> SCG_DEV void pairs_from_scalars(EnvRegs<KIND, real>&) {}
template <int KIND, typename real> SCG_DEV void scalars_from_pairs(EnvRegs<KIND, real>&) {}

template <int KIND, typename real> SCG_DEV void unpack_flags(EnvRegs<KIND, real>& e, uint32_t f) {
    e.viol = (f & SCG_FLAG_CNSTR_VIOLATION) != 0u;
    e.at_goal = (f & SCG_FLAG_AT_GOAL) != 0u;
    e.completed = (f & SCG_FLAG_COMPLETED) != 0u;
    e.goal_prev = (f & SCG_FLAG_GOAL_REACHED) != 0u;
    e.a2.flipped = (f & SCG_FLAG_FLIPPED) ? 1u : 0u;
    e.gate = (f >> SCG_FLAG_GATE_SHIFT) & 0xFFu;
    e.hold = f >> SCG_FLAG_HOLD_SHIFT;
}

template <int KIND, typename real> SCG_DEV uint32_t pack_flags(const EnvRegs<KIND, real>& e) {
    uint32_t f = (e.viol ? SCG_FLAG_CNSTR_VIOLATION : 0u) | (e.at_goal ? SCG_FLAG_AT_GOAL : 0u) |
                 (e.completed ? SCG_FLAG_COMPLETED : 0u) | (e.goal_prev ? SCG_FLAG_GOAL_REACHED : 0u) | (e.gate << SCG_FLAG_GATE_SHIFT) | (e.hold << SCG_FLAG_HOLD_SHIFT);
    if (KIND == SCG_QUAD_2D && e.a2.flipped) f |= SCG_FLAG_FLIPPED;
    return f;
}

// the integrator's per-episode constants from this episode's (M, Ixx, Iyy, Izz) (extension F9; KIND != cart-pole)
template <int KIND, typename real> SCG_DEV void inertia_from_values(const Hot<real>& h, const real (&v)[4], Inertia<real>& in) {
    in.inv_mass = real(1) / v[0];
    in.g_over_m = h.gravity_force * h.inv_mass;   // g * m_env / m_env == GRAVITY_ACC
#pragma unroll
    for (int j = 0; j < 3; ++j) { in.J[j] = v[j + 1]; in.J_inv[j] = real(1) / in.J[j]; }
#pragma unroll
    for (int j = 0; j < 3; ++j) { in.dJ[j] = in.J[(j + 2) % 3] - in.J[(j + 1) % 3]; in.kJ[j] = h.dt * in.J_inv[j]; }
}

// CFG_DIST_SIMPLE: the episode's impulse / step offsets into registers
template <int KIND, typename real> SCG_DEV void load_offsets(const Hot<real>& h, const int32_t* __restrict__ dist_offset, int64_t n, int64_t i,
                                                             EnvRegs<KIND, real>& e) {
#pragma unroll
    for (int k = 0; k < SCG_MAX_DIST; ++k) {
        if (k < h.ds_n && h.ds_off[k] < 0 && h.ds_kind[k] != SCG_DIST_WHITE) e.offc[k * ENV_BLOCK] = real(dist_offset[int64_t(h.ds_slot[k]) * n + i]);
    }
}

template <int KIND, typename real> SCG_DEV void load_inertia(const Hot<real>& h, const real* __restrict__ inertia, int64_t n, int64_t i,
                                                   Inertia<real>& in) {
    in.cp_l = h.cp_length; in.cp_mp = h.cp_pole_mass; in.cp_mc = h.cp_cart_mass;
    if constexpr (KIND == SCG_CARTPOLE) {   // per-episode pole length / masses always reach the cart-pole ODE
        if (inertia && h.any_inertia_rand) { in.cp_l = inertia[i]; in.cp_mp = inertia[n + i]; in.cp_mc = inertia[2 * n + i]; }
    }
    if (KIND != SCG_CARTPOLE && h.dyn_uses_episode_inertia && inertia) {   // extension (F9)
        const real v[4] = {inertia[i], inertia[n + i], inertia[2 * n + i], inertia[3 * n + i]};
        inertia_from_values<KIND, real>(h, v, in);
    } else {
        in.inv_mass = h.inv_mass;
        in.g_over_m = h.gravity_force * h.inv_mass;
#pragma unroll
        for (int j = 0; j < 3; ++j) { in.J[j] = h.J[j]; in.J_inv[j] = h.J_inv[j]; in.dJ[j] = h.dJ[j]; in.kJ[j] = h.kJ[j]; }
    }
}

template <int KIND, typename real, int ATT> SCG_DEV void derive_attitude(EnvRegs<KIND, real>& e) {
    if constexpr (KIND == SCG_QUAD_3D) {
        const real pre[3] = {e.x[6], e.x[7], e.x[8]};
        attitude3<real, ATT>(pre, e.a3);
    } else if constexpr (KIND == SCG_QUAD_2D) {
        attitude2<real, ATT>(e.x[4], e.a2.flipped, e.a2);
        e.x[4] = e.a2.theta;
    }
}

// env.state (quadrotor.py:853-869) from the registers; ang = the angular-velocity vector
// PyBullet holds (rpy_rates after a step, INIT_ANG_VEL right after reset).  FULL = false leaves
// the 3D body rates p, q, r (slots 9-11) untouched: they need the whole rotation matrix.
template <int KIND, typename real, bool FULL = true, bool PAIRED = false>
SCG_DEV void make_state(EnvRegs<KIND, real>& e, const real (&ang)[3], real (&st)[Dims<KIND>::NX]) {
    if constexpr (KIND == SCG_QUAD_3D) {
        if constexpr (PAIRED) {
            static_assert(!FULL, "the paired state carries no full rotation matrix");
            st[0] = lo2(e.Pxy); st[1] = lo2(e.Vxy); st[2] = hi2(e.Pxy); st[3] = hi2(e.Vxy); st[4] = lo2(e.ZA); st[5] = lo2(e.ZB);
            st[6] = lo2(e.Axz); st[7] = hi2(e.ZA); st[8] = hi2(e.Axz);
        } else {
#pragma unroll
            for (int d = 0; d < 6; ++d) st[d] = e.x[d];
            st[6] = e.a3.rpy[0]; st[7] = e.a3.rpy[1]; st[8] = e.a3.rpy[2];
        }
        if (FULL) {
            attitude3_complete<real>(e.a3);
            const real* R = e.a3.R;   // pqr = R^T ang_v (quadrotor.py:862-864)
            st[9] = (R[0] * ang[0] + R[3] * ang[1]) + R[6] * ang[2];
            st[10] = (R[1] * ang[0] + R[4] * ang[1]) + R[7] * ang[2];
            st[11] = (R[2] * ang[0] + R[5] * ang[1]) + R[8] * ang[2];
        }
    } else if constexpr (KIND == SCG_QUAD_2D) {
#pragma unroll
        for (int d = 0; d < 4; ++d) st[d] = e.x[d];
        st[4] = e.a2.theta; st[5] = ang[1];
    } else if constexpr (KIND == SCG_CARTPOLE) {
#pragma unroll
        for (int d = 0; d < 4; ++d) st[d] = e.x[d];
    } else {
        st[0] = e.x[0]; st[1] = e.x[1];
    }
}

// goal row for step index idx (X_GOAL or X_GOAL[idx])
template <int KIND, typename real>
SCG_DEV void goal_row(const Hot<real>& h, const Cold<real>* __restrict__ cold, int32_t idx, real (&g)[Dims<KIND>::NX]) {
    constexpr int NX = Dims<KIND>::NX;
    if (h.task == SCG_TASK_STABILIZATION) {
#pragma unroll
        for (int d = 0; d < NX; ++d) g[d] = h.x_goal0[d];
    } else {
        const real* row = cold->x_goal + int64_t(idx) * NX;
#pragma unroll
        for (int d = 0; d < NX; ++d) g[d] = row[d];
    }
}

// obs = state + observation disturbances, then extend_obs (quadrotor.py:878-884, benchmark_env.py:406-420).
// write_obs_body inline / write_obs out of line, chosen per instantiation by measurement (profiles/r2_summary.md): inline for the
// kernels that write per-step outputs (the host-driven step saves the ABI call and reads the scenario constants from the constant
// bank: 10.16 -> 9.75 us per step) and for the cart-pole (config 3 +10 %); out of line for the 3D SPEC loops, whose register
// allocation the inlined epilogue disturbs (level2-like -16 %, headline -1.4 %).  Observation disturbances stay out of line.
template <int KIND, typename real>
SCG_DEV void write_obs_body(const Hot<real>& h, const Cold<real>* __restrict__ cold, const int32_t* __restrict__ dist_offset,
                                       real* __restrict__ dst, const real* __restrict__ st, int32_t t_pre, uint32_t obs_index,
                                       RngKey key, uint32_t episode, int64_t n, int64_t i) {
    constexpr int NX = Dims<KIND>::NX;
    real o[NX];
#pragma unroll
    for (int d = 0; d < NX; ++d) o[d] = st[d];
    if (h.n_dist_obs)
        apply_channel<real, NX>(cold, SCG_CH_OBSERVATION, h.n_dist_obs, o, t_pre, real(t_pre * h.substeps) * h.dt, key,
                                obs_index, episode, dist_offset, n, i);
#pragma unroll
    for (int d = 0; d < NX; ++d) dst[int64_t(d) * n + i] = o[d];
    if (h.cost == SCG_COST_RL_REWARD && h.obs_goal_horizon > 0) {
        const int rows = h.task == SCG_TASK_TRAJ_TRACKING ? h.obs_goal_horizon : 1;
        for (int r = 0; r < rows; ++r) {
            int32_t idx = t_pre + 1 + r;
            idx = idx < h.n_goal_rows - 1 ? idx : h.n_goal_rows - 1;
            real g[NX];
            goal_row<KIND, real>(h, cold, idx, g);
#pragma unroll
            for (int d = 0; d < NX; ++d) dst[int64_t((r + 1) * NX + d) * n + i] = g[d];
        }
    }
}

template <int KIND, typename real>
__device__ __noinline__ void write_obs(const Hot<real>& h, const Cold<real>* __restrict__ cold, const int32_t* __restrict__ dist_offset,
                                       real* __restrict__ dst, const real* __restrict__ st, int32_t t_pre, uint32_t obs_index,
                                       RngKey key, uint32_t episode, int64_t n, int64_t i) {
    write_obs_body<KIND, real>(h, cold, dist_offset, dst, st, t_pre, obs_index, key, episode, n, i);
}
template <int KIND, typename real, bool INL>
SCG_DEV void write_obs_sel(const Hot<real>& h, const Cold<real>* __restrict__ cold, const int32_t* __restrict__ dist_offset,
                           real* __restrict__ dst, const real* __restrict__ st, int32_t t_pre, uint32_t obs_index,
                           RngKey key, uint32_t episode, int64_t n, int64_t i) {
    if constexpr (INL) write_obs_body<KIND, real>(h, cold, dist_offset, dst, st, t_pre, obs_index, key, episode, n, i);
    else write_obs<KIND, real>(h, cold, dist_offset, dst, st, t_pre, obs_index, key, episode, n, i);
}

template <typename real> SCG_DEV real rand_value(const RandP<real>& rp, real u, real z) {
    if (rp.distrib == SCG_RAND_UNIFORM) return rp.a + (rp.b - rp.a) * u;
    if (rp.distrib == SCG_RAND_NORMAL) return rp.a + rp.b * z;
    return real(0);
}

// Quadrotor.reset() (quadrotor.py:356-504) for one env.  Leaves e ready for the next step and
// returns env.state of the fresh episode in st.
template <int KIND, typename real, int ATT>
SCG_DEV void reset_env_body(const Hot<real>& h, const Cold<real>* __restrict__ cold, const Buf<real>& b,
                            const ResetInject<real>& inj, const RngKey& key, int64_t n, int64_t i,
                            EnvRegs<KIND, real>& e, real (&st)[Dims<KIND>::NX], bool write = true) {
    constexpr int NX = Dims<KIND>::NX;
    const uint32_t ep = h.reseed_on_reset ? 0u : uint32_t(e.resets);
    e.resets += 1;
    e.t = 0;
    e.at_goal = e.completed = false; e.gate = e.hold = 0u; e.a2.flipped = 0u;   // cnstr_violation survives reset (benchmark_env.py:218)
    // disturbance offsets (DisturbanceList.reset -> randint(max_step), disturbances.py:122-126,169-172)
    for (int ch = 0; ch < 3; ++ch) {
        const int nd = ch == 0 ? h.n_dist_obs : ch == 1 ? h.n_dist_act : h.n_dist_dyn;
        for (int k = 0; k < nd; ++k) {
            const DistP<real>& D = cold->dist[ch][k];
            if ((D.kind == SCG_DIST_IMPULSE || D.kind == SCG_DIST_STEP) && D.step_offset < 0) {
                int32_t off;
                if (inj.offsets) off = inj.offsets[int64_t(D.slot) * n + i];
                else {
                    const U4 w = philox4x32_10(0u, ep, STREAM_RESET_OFFSETS, uint32_t(D.slot >> 2), key.seed, key.env);
                    off = randint(word_of(w, D.slot & 3), D.max_step);
                }
                if (write) b.dist_offset[int64_t(D.slot) * n + i] = off;
            }
        }
    }
    // inertial properties (quadrotor.py:447-467): value += distrib(...)
    if (b.inertia) {
        real u[4] = {real(0), real(0), real(0), real(0)}, z[4] = {real(0), real(0), real(0), real(0)};
        if (!inj.inertia) {
            if (h.any_inertia_rand & 1) draw_uniforms<real, 4>(key, 0u, ep, STREAM_RESET_INERTIA, u);
            if (h.any_inertia_rand & 2) draw_normals<real, 4>(key, 0u, ep, STREAM_RESET_INERTIA, z);
        }
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            real v = h.inertia_nominal[j];
            if (inj.inertia) v += inj.inertia[int64_t(j) * n + i];
            else if (h.any_inertia_rand) v += rand_value(h.inertia_rand[j], u[j], z[j]);
            if (write) b.inertia[int64_t(j) * n + i] = v;
        }
    }
    load_inertia<KIND, real>(h, b.inertia, n, i, e.in);
    // initial state (quadrotor.py:477-493), label order == state order
    real init[NX];
#pragma unroll
    for (int d = 0; d < NX; ++d) init[d] = h.init_state[d];
    if (inj.init) {
#pragma unroll
        for (int d = 0; d < NX; ++d) init[d] += inj.init[int64_t(d) * n + i];
    } else if (h.any_init_rand) {
        real u[NX], z[NX];
#pragma unroll
        for (int d = 0; d < NX; ++d) { u[d] = real(0); z[d] = real(0); }
        if (h.any_init_rand & 1) draw_uniforms<real, NX>(key, 0u, ep, STREAM_RESET_INIT, u);
        if (h.any_init_rand & 2) draw_normals<real, NX>(key, 0u, ep, STREAM_RESET_INIT, z);
#pragma unroll
        for (int d = 0; d < NX; ++d) init[d] += rand_value(h.init_rand[d], u[d], z[d]);
    }
    real ang[3] = {real(0), real(0), real(0)};
    if constexpr (KIND == SCG_QUAD_3D) {
#pragma unroll
        for (int d = 0; d < 9; ++d) e.x[d] = init[d];
        e.x[9] = e.x[10] = e.x[11] = real(0);          // rpy_rates restart at 0 (base_aviary.py:232-233)
        ang[0] = init[9]; ang[1] = init[10]; ang[2] = init[11];   // INIT_ANG_VEL lives in PyBullet only
    } else if constexpr (KIND == SCG_QUAD_2D) {
#pragma unroll
        for (int d = 0; d < 5; ++d) e.x[d] = init[d];
        e.x[5] = real(0);
        ang[1] = init[5];
    } else if constexpr (KIND == SCG_CARTPOLE) {
#pragma unroll
        for (int d = 0; d < 4; ++d) e.x[d] = init[d];
    } else {
        // 1D labels are init_x / init_x_dot (quadrotor.py:208): they set the unobserved x axis;
        // z and z_dot always start at 0.
        e.x[0] = real(0); e.x[1] = real(0);
    }
    derive_attitude<KIND, real, ATT>(e);
    make_state<KIND, real>(e, ang, st);
}

// Auto-reset of a LEAN config (no disturbance offsets, no inertia draws, uniform-only initial-state
// randomisation: host-checked in scg_create) is split in two so that the Philox work does not run
// one lane at a time.  With short episodes a warp sees a reset on a large share of its steps (48 %
// at 2 % resets per env-step):
//   draw_init_lean    the initial state of episode `ep` (counter-based: it can be drawn at any time),
//                     executed by all lanes that miss a spare at a fixed cadence of the loop;
//   reset_from_spare  what the finished lanes do on the spot: take the spare, restart counters,
//                     evaluate the attitude (env.state is rebuilt later, only if it is observed).
template <int KIND, typename real>
SCG_DEV void draw_init_lean(const Hot<real>& h, const RngKey& key, uint32_t ep, real (&init)[Dims<KIND>::NX]) {
    constexpr int NX = Dims<KIND>::NX;
#pragma unroll
    for (int d = 0; d < NX; ++d) init[d] = h.init_state[d];
#pragma unroll
    for (int blk = 0; blk < (NX + 3) / 4; ++blk) {
        if (!((h.init_block_mask >> blk) & 1u)) continue;   // uniform: no randomised label in this block
        const U4 w = philox4x32_10(0u, ep, STREAM_RESET_INIT, uint32_t(blk), key.seed, key.env);
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int d = 4 * blk + j;
            if (d < NX) init[d] += h.init_rand[d].a + (h.init_rand[d].b - h.init_rand[d].a) * u01<real>(word_of(w, j));
        }
    }
}

// The spare lives in shared memory, one column per thread ([slot][ENV_BLOCK], conflict-free, never
// shared between threads), together with everything reset() derives from it -- the wrapped angles,
// their sin / cos and the thrust direction -- so that the lanes that restart only move registers.
template <int KIND, typename real, int ATT> struct Spare {
    static constexpr int NX = Dims<KIND>::NX;
    // fp64 evaluates the gimbal band with the literal sequence (full R): it derives the attitude on the spot
    static constexpr bool PRE3 = KIND == SCG_QUAD_3D && sizeof(real) == 4 && ATT != SCG_ATT_ROUNDTRIP;
    static constexpr int N = NX + (PRE3 ? 12 : KIND == SCG_QUAD_2D ? 4 : 0);
};

template <int KIND, typename real, int ATT>
SCG_DEV void spare_store(const real (&init)[Dims<KIND>::NX], real* __restrict__ sp) {
    constexpr int NX = Dims<KIND>::NX;
#pragma unroll
    for (int d = 0; d < NX; ++d) sp[d * ENV_BLOCK] = init[d];
    if constexpr (Spare<KIND, real, ATT>::PRE3) {
        Att3<real> a;
        const real pre[3] = {init[6], init[7], init[8]};
        attitude3<real, ATT>(pre, a);
#pragma unroll
        for (int j = 0; j < 3; ++j) sp[(NX + j) * ENV_BLOCK] = a.rpy[j];
#pragma unroll
        for (int j = 0; j < 6; ++j) sp[(NX + 3 + j) * ENV_BLOCK] = a.sc[j];
        sp[(NX + 9) * ENV_BLOCK] = a.R[2]; sp[(NX + 10) * ENV_BLOCK] = a.R[5]; sp[(NX + 11) * ENV_BLOCK] = a.R[8];
    } else if constexpr (KIND == SCG_QUAD_2D) {
        Att2<real> a;
        attitude2<real, ATT>(init[4], 0u, a);
        sp[(NX + 0) * ENV_BLOCK] = a.theta; sp[(NX + 1) * ENV_BLOCK] = a.r3x; sp[(NX + 2) * ENV_BLOCK] = a.r3z;
        sp[(NX + 3) * ENV_BLOCK] = a.flipped ? real(1) : real(0);
    }
}

template <int KIND, typename real, int ATT, bool PAIRED = false>
SCG_DEV void reset_from_spare(const real* __restrict__ sp, EnvRegs<KIND, real>& e) {
    constexpr int NX = Dims<KIND>::NX;
    e.resets += 1;
    e.t = 0;
    e.at_goal = e.completed = false; e.gate = e.hold = 0u; e.a2.flipped = 0u;   // cnstr_violation survives reset (benchmark_env.py:218)
    if constexpr (PAIRED) {
        static_assert(!PAIRED || Spare<KIND, real, ATT>::PRE3, "the paired loop takes the precomputed attitude from the spare");
        e.Pxy = pk2(sp[0 * ENV_BLOCK], sp[2 * ENV_BLOCK]); e.Vxy = pk2(sp[1 * ENV_BLOCK], sp[3 * ENV_BLOCK]);
        e.ZA = pk2(sp[4 * ENV_BLOCK], sp[(NX + 1) * ENV_BLOCK]); e.ZB = pk2(sp[5 * ENV_BLOCK], 0.f);
        e.Wxz = pk2(0.f, 0.f);                                    // rpy_rates restart at 0 (base_aviary.py:232-233)
        e.Axz = pk2(sp[(NX + 0) * ENV_BLOCK], sp[(NX + 2) * ENV_BLOCK]);
#pragma unroll
        for (int d = 6; d < 9; ++d) e.x[d] = sp[d * ENV_BLOCK];
#pragma unroll
        for (int j = 0; j < 6; ++j) e.a3.sc[j] = sp[(NX + 3 + j) * ENV_BLOCK];
        e.Rxy = pk2(sp[(NX + 9) * ENV_BLOCK], sp[(NX + 10) * ENV_BLOCK]); e.R8 = sp[(NX + 11) * ENV_BLOCK];
    } else if constexpr (KIND == SCG_QUAD_3D) {
#pragma unroll
        for (int d = 0; d < 9; ++d) e.x[d] = sp[d * ENV_BLOCK];
        e.x[9] = e.x[10] = e.x[11] = real(0);          // rpy_rates restart at 0 (base_aviary.py:232-233)
        if constexpr (Spare<KIND, real, ATT>::PRE3) {
#pragma unroll
            for (int j = 0; j < 3; ++j) e.a3.rpy[j] = sp[(NX + j) * ENV_BLOCK];
#pragma unroll
            for (int j = 0; j < 6; ++j) e.a3.sc[j] = sp[(NX + 3 + j) * ENV_BLOCK];
            e.a3.R[2] = sp[(NX + 9) * ENV_BLOCK]; e.a3.R[5] = sp[(NX + 10) * ENV_BLOCK]; e.a3.R[8] = sp[(NX + 11) * ENV_BLOCK];
            e.a3.form = 1;
        } else derive_attitude<KIND, real, ATT>(e);
    } else if constexpr (KIND == SCG_QUAD_2D) {
#pragma unroll
        for (int d = 0; d < 4; ++d) e.x[d] = sp[d * ENV_BLOCK];
        e.x[5] = real(0);
        e.a2.theta = sp[(NX + 0) * ENV_BLOCK]; e.a2.r3x = sp[(NX + 1) * ENV_BLOCK]; e.a2.r3z = sp[(NX + 2) * ENV_BLOCK];
        e.a2.flipped = sp[(NX + 3) * ENV_BLOCK] != real(0) ? 1u : 0u;
        e.x[4] = e.a2.theta;
    } else if constexpr (KIND == SCG_CARTPOLE) {
#pragma unroll
        for (int d = 0; d < 4; ++d) e.x[d] = sp[d * ENV_BLOCK];
    } else {
        e.x[0] = real(0); e.x[1] = real(0);
    }
}

// The general spare (CFG_RESET_EXTRAS): everything Quadrotor.reset() draws for episode `ep` -- initial state with
// uniform or normal offsets, inertial properties, impulse / step disturbance offsets -- with the same Philox
// counters as reset_env_body.  Slots after the state / attitude block: 4 inertial properties, then one per
// disturbance slot.  Out of line: it runs for all lanes together at the refill cadence.
constexpr int SPARE_EXTRA = 4 + 3 * SCG_MAX_DIST + SCG_MAX_DIST;   // (+ the CURRENT episode's action-channel offsets, EnvRegs::offc)

template <int KIND, typename real, int ATT>
SCG_DEV void draw_spare_general_body(const Hot<real>& h, const Cold<real>* __restrict__ cold, RngKey key, uint32_t ep,
                                     real* __restrict__ sp) {
    constexpr int NX = Dims<KIND>::NX, N0 = Spare<KIND, real, ATT>::N;
    real init[NX];
#pragma unroll
    for (int d = 0; d < NX; ++d) init[d] = h.init_state[d];
    if (h.any_init_rand) {
        real u[NX], z[NX];
#pragma unroll
        for (int d = 0; d < NX; ++d) { u[d] = real(0); z[d] = real(0); }
        if (h.any_init_rand & 1) draw_uniforms<real, NX>(key, 0u, ep, STREAM_RESET_INIT, u);
        if (h.any_init_rand & 2) draw_normals<real, NX>(key, 0u, ep, STREAM_RESET_INIT, z);
#pragma unroll
        for (int d = 0; d < NX; ++d) init[d] += rand_value(h.init_rand[d], u[d], z[d]);
    }
    spare_store<KIND, real, ATT>(init, sp);
    if (h.any_inertia_rand) {   // quadrotor.py:447-467
        real u[4] = {real(0), real(0), real(0), real(0)}, z[4] = {real(0), real(0), real(0), real(0)};
        if (h.any_inertia_rand & 1) draw_uniforms<real, 4>(key, 0u, ep, STREAM_RESET_INERTIA, u);
        if (h.any_inertia_rand & 2) draw_normals<real, 4>(key, 0u, ep, STREAM_RESET_INERTIA, z);
#pragma unroll
        for (int j = 0; j < 4; ++j) sp[(N0 + j) * ENV_BLOCK] = h.inertia_nominal[j] + rand_value(h.inertia_rand[j], u[j], z[j]);
    }
    for (int ch = 0; ch < 3; ++ch) {   // disturbances.py:122-126, 169-172
        const int nd = ch == 0 ? h.n_dist_obs : ch == 1 ? h.n_dist_act : h.n_dist_dyn;
        for (int k = 0; k < nd; ++k) {
            const DistP<real>& D = cold->dist[ch][k];
            if ((D.kind == SCG_DIST_IMPULSE || D.kind == SCG_DIST_STEP) && D.step_offset < 0 && D.slot < 3 * SCG_MAX_DIST) {
                const U4 w = philox4x32_10(0u, ep, STREAM_RESET_OFFSETS, uint32_t(D.slot >> 2), key.seed, key.env);
                sp[(N0 + 4 + D.slot) * ENV_BLOCK] = real(randint(word_of(w, D.slot & 3), D.max_step));   // exact: offsets are small integers
            }
        }
    }
}

// The LEAN kernels call it out of line (rare code, kept out of their loop); the PLAIN + DS loop inlines it: a call inside the packed
// loop pushes the loop-carried pairs through the ABI's few callee-saved registers.
template <int KIND, typename real, int ATT>
__device__ __noinline__ void draw_spare_general(const Hot<real>& h, const Cold<real>* __restrict__ cold, RngKey key, uint32_t ep,
                                                real* __restrict__ sp) {
    draw_spare_general_body<KIND, real, ATT>(h, cold, key, ep, sp);
}

// ... and what the restarting lane does with it beyond reset_from_spare: publish the episode's inertial properties
// and disturbance offsets to the batch buffers the step reads them from.
template <int KIND, typename real, int ATT>
__device__ __noinline__ void apply_spare_extras(const Hot<real>& h, const Cold<real>* __restrict__ cold, real* __restrict__ inertia,
                                                int32_t* __restrict__ dist_offset, const real* __restrict__ sp, int64_t n, int64_t i,
                                                Inertia<real>* in_out, bool write) {
    constexpr int N0 = Spare<KIND, real, ATT>::N;
    if (!write) {   // an out-of-range lane shadowing env n - 1: the buffers belong to the owner lane
        Inertia<real> in;
        load_inertia<KIND, real>(h, inertia, n, i, in);
        *in_out = in;
        return;
    }
    if (inertia) {
#pragma unroll
        for (int j = 0; j < 4; ++j)
            inertia[int64_t(j) * n + i] = h.any_inertia_rand ? sp[(N0 + j) * ENV_BLOCK] : h.inertia_nominal[j];
    }
    Inertia<real> in;
    load_inertia<KIND, real>(h, inertia, n, i, in);
    *in_out = in;
    for (int ch = 0; ch < 3; ++ch) {
        const int nd = ch == 0 ? h.n_dist_obs : ch == 1 ? h.n_dist_act : h.n_dist_dyn;
        for (int k = 0; k < nd; ++k) {
            const DistP<real>& D = cold->dist[ch][k];
            if ((D.kind == SCG_DIST_IMPULSE || D.kind == SCG_DIST_STEP) && D.step_offset < 0 && D.slot < 3 * SCG_MAX_DIST)
                dist_offset[int64_t(D.slot) * n + i] = int32_t(sp[(N0 + 4 + D.slot) * ENV_BLOCK]);
        }
    }
}

// The same for configs whose per-step disturbances are all inline (CFG_DIST_SIMPLE) or absent: no table walks, no read-back --
// the restarting lane stores the episode's inertial properties and offsets (owner lane only) and forms its constants from the
// spare's values directly.  Inline: with ~1 % of env-steps ending an episode a third of all warp-steps pass through here.
template <int KIND, typename real, int ATT>
SCG_DEV void apply_spare_simple(const Hot<real>& h, real* __restrict__ inertia, int32_t* __restrict__ dist_offset, const real* __restrict__ sp,
                                int64_t n, int64_t i, EnvRegs<KIND, real>& e, bool write) {
    constexpr int N0 = Spare<KIND, real, ATT>::N;
    if (inertia) {
        real v[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) v[j] = h.any_inertia_rand ? sp[(N0 + j) * ENV_BLOCK] : h.inertia_nominal[j];
        if (write) {
#pragma unroll
            for (int j = 0; j < 4; ++j) inertia[int64_t(j) * n + i] = v[j];
        }
        if constexpr (KIND == SCG_CARTPOLE) {
            if (h.any_inertia_rand) { e.in.cp_l = v[0]; e.in.cp_mp = v[1]; e.in.cp_mc = v[2]; }
        } else if (h.dyn_uses_episode_inertia) inertia_from_values<KIND, real>(h, v, e.in);
    }
#pragma unroll
    for (int k = 0; k < SCG_MAX_DIST; ++k) {
        if (k < h.ds_n && h.ds_off[k] < 0 && h.ds_kind[k] != SCG_DIST_WHITE) {
            const real o = sp[(N0 + 4 + h.ds_slot[k]) * ENV_BLOCK];
            e.offc[k * ENV_BLOCK] = o;
            if (write) dist_offset[int64_t(h.ds_slot[k]) * n + i] = int32_t(o);
        }
    }
}

// PyBullet-side INIT_ANG_VEL of the episode the spare started (valid until the next draw)
template <int KIND, typename real>
SCG_DEV void spare_ang0(const real* __restrict__ sp, real (&ang0)[3]) {
    ang0[0] = ang0[1] = ang0[2] = real(0);
    if constexpr (KIND == SCG_QUAD_3D) { ang0[0] = sp[9 * ENV_BLOCK]; ang0[1] = sp[10 * ENV_BLOCK]; ang0[2] = sp[11 * ENV_BLOCK]; }
    else if constexpr (KIND == SCG_QUAD_2D) { ang0[1] = sp[5 * ENV_BLOCK]; }
}

// The auto-reset inside the rollout loop: one out-of-line copy, entered by the lanes whose episode
// ended this step (dummy_vec_env.py:38-49).  Works on private copies so that the loop's registers
// are not pinned to local memory by the call.
template <int KIND, typename real, int ATT>
__device__ __noinline__ void reset_env_cold(const Hot<real>& h, const Cold<real>* __restrict__ cold, const Buf<real>& b,
                                            RngKey key, int64_t n, int64_t i, EnvRegs<KIND, real>* e_io, real* __restrict__ st_out,
                                            bool write) {
    constexpr int NX = Dims<KIND>::NX;
    EnvRegs<KIND, real> e = *e_io;
    real st[NX];
    const ResetInject<real> none{nullptr, nullptr, nullptr};
    reset_env_body<KIND, real, ATT>(h, cold, b, none, key, n, i, e, st, write);
    *e_io = e;
#pragma unroll
    for (int d = 0; d < NX; ++d) st_out[d] = st[d];
}

// Action-channel disturbances + adversary on the commanded thrusts, and the external force of the dynamics channel
// (extension, F5: the reference never integrates it in DYN).  Out of line: the undisturbed loop must not carry it.
template <int KIND, typename real>
__device__ __noinline__ void disturb_inputs(const Hot<real>& h, const Cold<real>* __restrict__ cold, const int32_t* __restrict__ dist_offset,
                                            const real* __restrict__ adv_action, RngKey key, int32_t t, uint32_t ep, int64_t n, int64_t i,
                                            real* __restrict__ thr_io, real* __restrict__ f_out) {
    constexpr int NU = Dims<KIND>::NU, NF = Dims<KIND>::NF;
    const real t_sim = real(t * h.substeps) * h.dt;
    real thr[NU];
#pragma unroll
    for (int j = 0; j < NU; ++j) thr[j] = thr_io[j];
    if (h.n_dist_act) apply_channel<real, NU>(cold, SCG_CH_ACTION, h.n_dist_act, thr, t, t_sim, key, uint32_t(t), ep, dist_offset, n, i);
    if (h.adversary_channel == SCG_CH_ACTION && adv_action) {
#pragma unroll
        for (int j = 0; j < NU; ++j) thr[j] += adv_action[i * NU + j];
    }
#pragma unroll
    for (int j = 0; j < NU; ++j) thr_io[j] = thr[j];
    f_out[0] = f_out[1] = f_out[2] = real(0);
    if (h.dyn_applies_force && (h.n_dist_dyn || h.adversary_channel == SCG_CH_DYNAMICS)) {
        real f[NF];
#pragma unroll
        for (int d = 0; d < NF; ++d) f[d] = real(0);
        if (h.n_dist_dyn) apply_channel<real, NF>(cold, SCG_CH_DYNAMICS, h.n_dist_dyn, f, t, t_sim, key, uint32_t(t), ep, dist_offset, n, i);
        if (h.adversary_channel == SCG_CH_DYNAMICS && adv_action) {
#pragma unroll
            for (int d = 0; d < NF; ++d) f[d] += adv_action[i * NF + d];
        }
        if constexpr (KIND == SCG_QUAD_1D) { f_out[2] = f[0]; }
        else if constexpr (KIND == SCG_QUAD_2D || KIND == SCG_CARTPOLE) { f_out[0] = f[0]; f_out[2] = f[1]; }
        else { f_out[0] = f[0]; f_out[1] = f[1]; f_out[2] = f[2]; }
    }
}

// e' M e for a dense weight matrix (quadratic cost with off-diagonal Q / R); out of line
template <typename real, int DIM, int LD>
__device__ __noinline__ real dense_form(const real* __restrict__ M, const real* __restrict__ e) {
    real acc = real(0);
    for (int a = 0; a < DIM; ++a) {
        real row = real(0);
        for (int c = 0; c < DIM; ++c) row += M[a * LD + c] * e[c];
        acc += e[a] * row;
    }
    return acc;
}

// One Quadrotor.step(): returns reward / done / info bits, leaves env.state in st
// (3D: p, q, r only when `full`).
// DS (PLAIN kernels only): the config's extras are exactly the ones that stay cheap inline -- CFG_DIST_SIMPLE disturbances and / or
// per-episode inertia / offsets through apply_spare_simple -- so everything else a LEAN kernel carries (out-of-line disturbance and
// constraint evaluation, dense costs, gate geometry, the PID) is compiled out: a third of the code size of the LEAN kernel.
// (quad_type 3 takes a DS kernel only with nominal inertia in the integrator -- the reference's own DYN behaviour, F9 -- so that the
// packed substep with its constant-bank inertia rows stays valid; per-episode inertia in the 3D integrator takes the LEAN kernel.)
template <int KIND, typename real, int ATT, bool LEAN, bool PLAIN, bool DS = false>
SCG_DEV void step_env(const Hot<real>& h, const Cold<real>* __restrict__ cold, const Buf<real>& b,
                      const RngKey& key, int64_t n, int64_t i, EnvRegs<KIND, real>& e,
                      const real (&raw)[Dims<KIND>::NU], uint32_t cfg, real (&st)[Dims<KIND>::NX], real& reward, bool& done,
                      uint32_t& info, real& mse, Motors<real>& mot, real* __restrict__ c_out) {
    constexpr int NX = Dims<KIND>::NX, NU = Dims<KIND>::NU;
    constexpr bool FASTMATH = sizeof(real) == 4 && ATT != SCG_ATT_ROUNDTRIP;
    const uint32_t ep = (cfg & CFG_RESEED) ? 0u : uint32_t(e.resets - 1);
    const int32_t t = e.t;
    const bool full = (cfg & CFG_FULL_STATE) != 0u;   // uniform
    info = 0u;
    // ---- before_step / _preprocess_control (quadrotor.py:815-842)
    real thr[NU], pre[NU];
#pragma unroll
    for (int j = 0; j < NU; ++j) {
        const real c = m_clip(raw[j], h.act_low[j], h.act_high[j]);
        if constexpr (KIND == SCG_CARTPOLE) pre[j] = h.normalized_action ? c * h.hover_thrust : c;   // force = action_scale * a
        else if constexpr (PLAIN) pre[j] = c;   // a normalised action space takes the LEAN kernel
        else pre[j] = h.normalized_action ? (real(1) + h.norm_act_scale * c) * h.hover_thrust : c;
        thr[j] = pre[j];
    }
    real fx = real(0), fy = real(0), fz = real(0);
    if ((!PLAIN || DS) && (cfg & CFG_DIST_SIMPLE)) {   // uniform: impulse / step / white noise on the action channel, inline (Hot::ds_*)
#pragma unroll
        for (int k = 0; k < SCG_MAX_DIST; ++k) {
            if (k >= h.ds_n) break;
            if (h.ds_kind[k] == SCG_DIST_WHITE) {   // disturbances.py:237-248; draw = (index t, episode, stream of this disturbance)
                real z[NU];
                draw_normals_f<real, NU, FASTMATH>(key.seed, key.env, uint32_t(t), ep, STREAM_DISTURB_BASE + 8u * SCG_CH_ACTION + uint32_t(k), z);
#pragma unroll
                for (int j = 0; j < NU; ++j) thr[j] = m_fma(z[j], h.ds_std[k][j], thr[j]);
            } else {                                // disturbances.py:128-143, 174-184
                const real off = h.ds_off[k] >= 0 ? real(h.ds_off[k]) : e.offc[k * ENV_BLOCK];
                real sn = real(0);
                if (real(t) >= off) {
                    if (h.ds_kind[k] == SCG_DIST_STEP) sn = h.ds_mag[k];
                    else {
                        const int32_t peak = static_cast<int32_t>(off + h.ds_half[k]);
                        const int32_t kk = t >= peak ? t - peak : peak - t;
                        if (real(kk) < h.ds_half[k]) sn = h.ds_mag[k] * pow_int(h.ds_decay[k], h.ds_l2decay[k], kk);
                    }
                }
#pragma unroll
                for (int j = 0; j < NU; ++j) thr[j] += sn;
            }
        }
    } else if (!PLAIN && (cfg & CFG_DISTURBED)) {   // uniform; one out-of-line call on private copies
        real tl[NU], fl[3];
#pragma unroll
        for (int j = 0; j < NU; ++j) tl[j] = thr[j];
        disturb_inputs<KIND, real>(h, cold, b.dist_offset, b.adv_action, key, t, ep, n, i, tl, fl);
#pragma unroll
        for (int j = 0; j < NU; ++j) thr[j] = tl[j];
        fx = fl[0]; fy = fl[1]; fz = fl[2];
    }
    if constexpr (KIND != SCG_CARTPOLE) motors_from_thrust<KIND, real, FASTMATH>(h, thr, mot, !(PLAIN && !DS && (cfg & CFG_CLIP_ONCE)));
    else { mot.T = mot.tx = mot.ty = mot.tz = real(0); mot.rpm[0] = mot.rpm[1] = mot.rpm[2] = mot.rpm[3] = real(0); }
    // ---- _advance_simulation: S explicit substeps (base_aviary.py:266-322, 522-576)
    constexpr bool ROWS = PLAIN && KIND == SCG_QUAD_3D && sizeof(real) == 4 && ATT != SCG_ATT_ROUNDTRIP;   // nominal inertia: constants from Hot::step_row
    const real Tm = mot.T * (ROWS ? ldc4(h.step_row).w : e.in.inv_mass);
    const real dt = ROWS ? ldc4(h.step_row).x : h.dt;
    if constexpr (KIND == SCG_QUAD_3D) {
        const real axf = fx * e.in.inv_mass, ayf = fy * e.in.inv_mass, azf = fz * e.in.inv_mass - e.in.g_over_m;
        const real Tdt = Tm * dt, gdt = e.in.g_over_m * dt;
        (void)Tdt; (void)gdt;
        // dyn_aero (extension, uniform): the rotor speed sum _drag reads -- the previous substep's, then this step's
        real om_sub = real(0), om_now = real(0);
        if constexpr (!PLAIN) {
            if (cfg & CFG_AERO) {
                om_sub = e.aero_prev;
#pragma unroll
                for (int j = 0; j < 4; ++j) om_now += real(2) * Consts<real>::pi * motor_rpm<real, FASTMATH>(h, mot, j) / real(60);
            }
        }
        (void)om_sub; (void)om_now;
        // One substep.  LAST = false (all but the final substep of a control step, LEAN kernels): the pre-round-trip
        // angles are not copied into e.x[6..8] -- only the final ones are ever read (derive_attitude, the next launch).
        auto substep = [&](auto last_tag) {
            constexpr bool LAST = decltype(last_tag)::value;
            const real w0 = e.x[9], w1 = e.x[10], w2 = e.x[11];
            real fax = real(0), fay = real(0), faz = real(0), tax = real(0), tay = real(0);
            if constexpr (!PLAIN) {
                if (cfg & CFG_AERO) {
                    real o[5];
                    aero_terms<real, FASTMATH>(h, e.x[4], e.x[1], e.x[3], e.x[5], e.a3, mot.rpm[0], mot.rpm[1], mot.rpm[2], mot.rpm[3], om_sub, o);
                    fax = o[0] * e.in.inv_mass; fay = o[1] * e.in.inv_mass; faz = o[2] * e.in.inv_mass; tax = o[3]; tay = o[4];
                    om_sub = om_now;   // base_aviary.py:320
                }
            }
            (void)fax; (void)fay; (void)faz; (void)tax; (void)tay;
            if constexpr (FASTMATH) {
                // same update with the constants folded (fp32 fast modes; a few ulp from the literal order below):
                // w x Jw = (w1 w2 (Izz - Iyy), w2 w0 (Ixx - Izz), w0 w1 (Iyy - Ixx)) for a diagonal J; dt / J and dt T / m premultiplied
                if constexpr (PLAIN) {
                    e.x[1] = m_fma(e.a3.R[2], Tdt, e.x[1]); e.x[3] = m_fma(e.a3.R[5], Tdt, e.x[3]);
                    e.x[5] = m_fma(e.a3.R[8], Tdt, e.x[5] - gdt);
                } else {
                    e.x[1] = m_fma(e.a3.R[2], Tdt, e.x[1] + (axf + fax) * dt); e.x[3] = m_fma(e.a3.R[5], Tdt, e.x[3] + (ayf + fay) * dt);
                    e.x[5] = m_fma(e.a3.R[8], Tdt, e.x[5] + (azf + faz) * dt);
                }
                e.x[9] = m_fma(m_fma(-(w1 * w2), e.in.dJ[0], mot.tx + tax), e.in.kJ[0], e.x[9]);
                e.x[10] = m_fma(m_fma(-(w2 * w0), e.in.dJ[1], mot.ty + tay), e.in.kJ[1], e.x[10]);
                e.x[11] = m_fma(m_fma(-(w0 * w1), e.in.dJ[2], mot.tz), e.in.kJ[2], e.x[11]);
            } else {
                const real ax = e.a3.R[2] * Tm + axf + fax, ay = e.a3.R[5] * Tm + ayf + fay, az = e.a3.R[8] * Tm + azf + faz;
                const real jw0 = e.in.J[0] * w0, jw1 = e.in.J[1] * w1, jw2 = e.in.J[2] * w2;
                const real d0 = e.in.J_inv[0] * ((mot.tx + tax) - (w1 * jw2 - w2 * jw1));
                const real d1 = e.in.J_inv[1] * ((mot.ty + tay) - (w2 * jw0 - w0 * jw2));
                const real d2 = e.in.J_inv[2] * (mot.tz - (w0 * jw1 - w1 * jw0));
                e.x[1] += dt * ax; e.x[3] += dt * ay; e.x[5] += dt * az;          // vel first
                e.x[9] += dt * d0; e.x[10] += dt * d1; e.x[11] += dt * d2;         // then rates
            }
            e.x[0] += dt * e.x[1]; e.x[2] += dt * e.x[3]; e.x[4] += dt * e.x[5];   // pos uses NEW vel
            const real pre3[3] = {e.a3.rpy[0] + dt * e.x[9], e.a3.rpy[1] + dt * e.x[10], e.a3.rpy[2] + dt * e.x[11]};
            if constexpr (LAST) { e.x[6] = pre3[0]; e.x[7] = pre3[1]; e.x[8] = pre3[2]; }
            attitude3<real, ATT>(pre3, e.a3);
        };
        if constexpr (Paired<KIND, real, ATT, LEAN, PLAIN>::value) {
            // The same substep on packed pairs (FFMA2 / FADD2, scg_math.cuh): every half is the scalar operation of
            // `substep` above, bit for bit; the loop is issue-bound and a pair takes one issue slot instead of two.
            // Pairing: (x, y) of position and velocity; (roll, yaw) of angles and rates, so that the two wraps that
            // follow the pitch fold run as one; the z lane rides with the pitch lane: (z, pitch) and (vz, pitch rate).
            // The pairs are the loop-carried state of the whole control-step loop (EnvRegs: Pxy ... R8; rollout_range packs
            // them once before the loop), so no step pays for moving values between the two layouts.
            f32x2 &Pxy = e.Pxy, &Vxy = e.Vxy, &Wxz = e.Wxz, &ZB = e.ZB, &ZA = e.ZA, &Axz = e.Axz, &Rxy = e.Rxy;
            float& R8 = e.R8;
            // PLAIN: nominal inertia, so the constant pairs come straight from the constant bank as 64-bit operands
            const f32x2 ndJxz = *reinterpret_cast<const f32x2*>(h.ndJ_xz), kJxz = *reinterpret_cast<const f32x2*>(h.kJ_xz);
            const f32x2 Txz = pk2(mot.tx, mot.tz);
            const Row4<float> sb = ldc4(h.step_row + 4);   // [-dJ[1], kJ[1], -g, 0]
            const float ndJ1 = sb.x, kJ1 = sb.y, ngm = sb.z;
            float sph, cph, sth, cth, sps, cps;
            auto substep2 = [&](auto last_tag) {
                constexpr bool LAST = decltype(last_tag)::value;
                const float w0 = lo2(Wxz), w2 = hi2(Wxz), w1 = hi2(ZB);
                const float m0 = w1 * w2, m1 = w2 * w0, m2 = w0 * w1;
                Vxy = fma2(Rxy, bc2(Tdt), Vxy);
                const float vz0 = fmaf(ngm, dt, lo2(ZB));
                const f32x2 in_xz = fma2(pk2(m0, m2), ndJxz, Txz);
                const float in_y = fmaf(m1, ndJ1, mot.ty);
                Wxz = fma2(in_xz, kJxz, Wxz);
                ZB = pk2(fmaf(R8, Tdt, vz0), fmaf(in_y, kJ1, w1));   // scalar: the two products share no operand pair
                Pxy = fma2(bc2(dt), Vxy, Pxy);
                const f32x2 ZA2 = fma2(bc2(dt), ZB, ZA);      // (z, pre-round-trip pitch)
                const f32x2 PRE = fma2(bc2(dt), Wxz, Axz);    // pre-round-trip (roll, yaw)
                if constexpr (LAST) { e.x[6] = lo2(PRE); e.x[7] = hi2(ZA2); e.x[8] = hi2(PRE); }
                // attitude3<float, ATT> on the pairs
                const float thw = att_wrap<ATT>(hi2(ZA2));
                const bool folded = fabsf(thw) > Consts<float>::half_pi;
                float th = thw;
                if (folded) th = copysignf(Consts<float>::pi - fabsf(thw), thw);   // == copysign(pi, thw) - thw, one constant less
                f32x2 PH = add2(PRE, bc2(folded ? Consts<float>::pi : 0.f));
                {
                    const f32x2 K = add2(fma2(PH, bc2(0.15915494309189535f), bc2(12582912.f)), bc2(-12582912.f));
                    if (ATT == SCG_ATT_FAST_INTRINSIC) PH = fma2(K, bc2(-6.2831855f), PH);
                    else PH = fma2(K, bc2(1.7484555e-07f), fma2(K, bc2(-6.2831855f), PH));
                }
                att_sincos<ATT>(th, &sth, &cth);
                const bool band = !(fabsf(sth) < 0.99999f);
                float r0 = lo2(PH), r1 = th, r2 = hi2(PH);
                att_sincos<ATT>(r0, &sph, &cph);
                att_sincos<ATT>(r2, &sps, &cps);
                Rxy = pk2(cps * sth * cph + sps * sph, sps * sth * cph - cps * sph);
                R8 = cth * cph;
                if (band) {   // see attitude3
                    const bool up = sth > 0.f;
                    const float hp = copysignf(Consts<float>::half_pi, sth);   // the snapped pitch
                    const float r = 0.5f * (thw - hp);
                    const float sr = folded ? -sph : sph, cr = folded ? -cph : cph;
                    const float u = up ? r * cr : -(r * cr);
                    const float corr = 2.f * r * sr * (1.f - u + u * u);
                    float yaw = (up ? hi2(PRE) - lo2(PRE) : hi2(PRE) + lo2(PRE)) + corr;
                    const float k = fast_rint(yaw * 0.07957747154594767f);
                    yaw = fmaf(-k, -3.4969110629390343e-07f, fmaf(-k, 12.566370964050293f, yaw));
                    r0 = 0.f; r1 = hp; r2 = yaw;
                }
                Axz = pk2(r0, r2);
                ZA = pk2(lo2(ZA2), r1);
            };
            if (h.substeps == 4) {   // PYB_FREQ / CTRL_FREQ of every shipped config: straight-line code, no loop-carried copies
                substep2(std::false_type{}); substep2(std::false_type{}); substep2(std::false_type{}); substep2(std::true_type{});
            } else {
#pragma unroll 1
                for (int s = 1; s < h.substeps; ++s) substep2(std::false_type{});
                substep2(std::true_type{});
            }
            e.a3.sc[0] = sph; e.a3.sc[1] = cph; e.a3.sc[2] = sth; e.a3.sc[3] = cth; e.a3.sc[4] = sps; e.a3.sc[5] = cps;
            e.a3.form = 1;
        } else if constexpr (LEAN) {
            if (FASTMATH && h.substeps == 4) {   // straight-line code for the usual PYB_FREQ / CTRL_FREQ
                substep(std::false_type{}); substep(std::false_type{}); substep(std::false_type{}); substep(std::true_type{});
            } else {
#pragma unroll 1
                for (int s = 1; s < h.substeps; ++s) substep(std::false_type{});
                substep(std::true_type{});
            }
        } else {
#pragma unroll 1
            for (int s = 0; s < h.substeps; ++s) substep(std::true_type{});
        }
        if constexpr (!PLAIN) { if (cfg & CFG_AERO) e.aero_prev = om_now; }
    } else if constexpr (KIND == SCG_QUAD_2D) {
        const real axf = fx * e.in.inv_mass, azf = fz * e.in.inv_mass - e.in.g_over_m;
        auto substep2d = [&]() {
            const real ax = e.a2.r3x * Tm + axf, az = e.a2.r3z * Tm + azf;
            const real dw = e.in.J_inv[1] * mot.ty;
            e.x[1] += dt * ax; e.x[3] += dt * az;
            e.x[5] += dt * dw;
            e.x[0] += dt * e.x[1]; e.x[2] += dt * e.x[3];
            attitude2<real, ATT>(e.a2.theta + dt * e.x[5], e.a2.flipped, e.a2);
            e.x[4] = e.a2.theta;
        };
        if (LEAN && h.substeps == 4) { substep2d(); substep2d(); substep2d(); substep2d(); }   // the usual PYB_FREQ / CTRL_FREQ: straight-line
        else {
#pragma unroll 1
            for (int s = 0; s < h.substeps; ++s) substep2d();
        }
    } else if constexpr (KIND == SCG_CARTPOLE) {
        // RK4 over the cart-pole ODE at the simulation rate; a dynamics-channel disturbance acts as an
        // extra horizontal force on the cart when dyn_applies_force is on.
        const real F = thr[0] + fx;
        (void)fy; (void)fz; (void)Tm;
#pragma unroll 1
        for (int s = 0; s < h.substeps; ++s) cartpole_rk4<real, ATT>(e.in, h.gravity_acc, e.x, F, dt);
    } else {
        const real az = Tm + (fz * e.in.inv_mass - e.in.g_over_m);
#pragma unroll 1
        for (int s = 0; s < h.substeps; ++s) { e.x[1] += dt * az; e.x[0] += dt * e.x[1]; }
    }
    // ---- _get_observation: env.state (obs itself is written by the caller, it may be replaced on reset)
    constexpr bool PAIRED = Paired<KIND, real, ATT, LEAN, PLAIN>::value;
    real ang[3];
    if constexpr (PAIRED) { ang[0] = lo2(e.Wxz); ang[1] = hi2(e.ZB); ang[2] = hi2(e.Wxz); }
    else if constexpr (KIND == SCG_QUAD_3D) { ang[0] = e.x[9]; ang[1] = e.x[10]; ang[2] = e.x[11]; }
    else if constexpr (KIND == SCG_QUAD_2D) { ang[0] = real(0); ang[1] = e.x[5]; ang[2] = real(0); }
    else { ang[0] = ang[1] = ang[2] = real(0); }
    if constexpr (KIND == SCG_QUAD_3D) {
        make_state<KIND, real, false, PAIRED>(e, ang, st);
        st[9] = st[10] = st[11] = real(0);
        if (full) {   // uniform: h.need_full_state or per-step outputs
            Att3<real> c = e.a3;   // a copy: columns 0 and 1 must not stay live across control steps
            if constexpr (PAIRED) { c.R[2] = lo2(e.Rxy); c.R[5] = hi2(e.Rxy); c.R[8] = e.R8; }
            attitude3_complete<real>(c);
            const real* R = c.R;
            st[9] = (R[0] * ang[0] + R[3] * ang[1]) + R[6] * ang[2];
            st[10] = (R[1] * ang[0] + R[4] * ang[1]) + R[7] * ang[2];
            st[11] = (R[2] * ang[0] + R[5] * ang[1]) + R[8] * ang[2];
        }
    } else make_state<KIND, real>(e, ang, st);
    // ---- _get_info (quadrotor.py:1004-1134): goal hold (collision scan is dead in DYN, F5)
    bool gate_passed = false, collided = false;
    if constexpr (KIND == SCG_QUAD_3D && (!PLAIN || DS)) {
        if (cfg & CFG_GATE_GEOMETRY) {   // extension: the contact scan and the gate rays on the integrated pose
            real r8;
            if constexpr (PAIRED) r8 = e.R8; else r8 = e.a3.R[8];
            collided = ground_contact<real, FASTMATH>(st[4], r8);
            // first broad phase, inline on constant-bank operands (Hot::geo): is the drone near ANY gate / obstacle axis?
            bool near = false;
            const f32x2 PX = bc2(float(st[0])), PY = bc2(float(st[2]));
#pragma unroll
            for (int j = 0; j < (SCG_MAX_GATES + SCG_MAX_OBSTACLES) / 2; ++j) {   // two objects per packed operation
                if (2 * j >= h.geo_n) break;   // uniform
                const f32x2 DX = add2(PX, *reinterpret_cast<const f32x2*>(h.geo_negx[j])), DY = add2(PY, *reinterpret_cast<const f32x2*>(h.geo_negy[j]));
                const f32x2 S = fma2(DX, DX, mul2(DY, DY));
                near |= (lo2(S) <= h.geo_r2[j][0]) | (hi2(S) <= h.geo_r2[j][1]);
            }
            if (near) {   // rare: the whole rotation matrix (the gate rays work in the body frame) and the exact scan, out of line
                Att3<real> c = e.a3;
                if constexpr (PAIRED) { c.R[2] = lo2(e.Rxy); c.R[5] = hi2(e.Rxy); c.R[8] = e.R8; }
                attitude3_complete<real>(c);
                // pyb_step_counter > 0.5 PYB_FREQ, the counter still at t * S (it advances in after_step, benchmark_env.py:427)
                //   <=>  2 t > CTRL_FREQ (quadrotor.py:1060)
                const bool check_gate = 2 * t > h.ctrl_freq && e.gate < uint32_t(h.n_gates);
                uint32_t bits;
                if constexpr (DS) bits = gate_scan_body<real>(h, cold, key, ep, st[0], st[2], st[4], c.R[0], c.R[1], c.R[2], c.R[3], c.R[4], c.R[5],
                                                              c.R[6], c.R[7], c.R[8], e.gate, check_gate);
                else bits = gate_scan<real>(h, cold, key, ep, st[0], st[2], st[4], c.R[0], c.R[1], c.R[2], c.R[3], c.R[4], c.R[5],
                                            c.R[6], c.R[7], c.R[8], e.gate, check_gate);
                if (bits & 1u) { e.gate += 1u; gate_passed = true; }
                collided |= (bits & 2u) != 0u;
            }
        }
    }
    e.at_goal = false;
    if constexpr (KIND == SCG_QUAD_3D) {
        if (e.gate == uint32_t(h.n_gates) && (cfg & CFG_STAB)) {   // quadrotor.py:1117-1130
            const Row4<real> g = ldc4(h.pos_pack + 8);   // x_goal0[0, 2, 4]
            const real dx = st[0] - g.x, dy = st[2] - g.y, dz = st[4] - g.z;
            // sqrt(d2) < tol  <=>  d2 < tol^2 up to one rounding of the square; the literal modes keep the sqrt
            const real d2 = dx * dx + dy * dy + dz * dz;
            const bool inside = FASTMATH ? (d2 < ldc4(h.pos_pack + 4).w) : (m_sqrt(d2) < h.goal_tol);   // goal_tol^2 rides in the upper-bound row
            e.at_goal = inside;
            e.hold = inside ? (e.hold < 0xFFFFu ? e.hold + 1u : e.hold) : 0u;
            if (e.hold > uint32_t(2 * h.ctrl_freq)) e.completed = true;
        }
    }
    info = (e.at_goal ? SCG_INFO_AT_GOAL : 0u) | (e.completed ? SCG_INFO_COMPLETED : 0u) | (gate_passed ? SCG_INFO_GATE_PASSED : 0u) |
           (collided ? SCG_INFO_COLLISION : 0u);
    // ---- _get_done (quadrotor.py:956-1002)
    done = false;
    if ((cfg & (CFG_STAB | CFG_COST_QUAD)) == (CFG_STAB | CFG_COST_QUAD)) {
        // _get_info ran BEFORE this _get_done (quadrotor.py:553-555): info["goal_reached"] is the previous step's flag (F10)
        // (the cart-pole, absent from this fork, follows the upstream order: _get_done first, the flag is this step's)
        if (KIND != SCG_CARTPOLE && e.goal_prev) info |= SCG_INFO_GOAL_REACHED;
        real nn = real(0);
#pragma unroll
        for (int d = 0; d < NX; ++d) { const real er = st[d] - h.x_goal0[d]; nn += er * er; }
        e.goal_prev = m_sqrt(nn) < h.goal_tol;
        if (e.goal_prev) done = true;
        if (KIND == SCG_CARTPOLE && e.goal_prev) info |= SCG_INFO_GOAL_REACHED;
    }
    if (cfg & CFG_OOB) {
        bool oob = false;
#pragma unroll
        for (int d = 0; d < NX; ++d) {
            const bool masked = KIND == SCG_QUAD_3D ? (d == 0 || d == 2 || d == 4 || d == 6 || d == 7 || d == 8)
                                                   : ((d & 1) == 0);
            if (masked) oob |= (st[d] < h.state_low[d]) | (st[d] > h.state_high[d]);
        }
        if (oob) { done = true; info |= SCG_INFO_OUT_OF_BOUND; }
    }
    if (collided && h.done_on_collision) done = true;
    if ((cfg & CFG_DONE_COMPLETION) && e.completed) done = true;
    // ---- info["mse"] (quadrotor.py:1016-1021) and _get_reward (quadrotor.py:886-954); t is still the
    // pre-increment counter.  Written once against a goal accessor so that a stabilisation goal stays
    // a constant-bank operand and only trajectory tracking loads a row into registers.
    auto costs = [&](auto goal_at, auto goal_q_at, auto goal_pos_at) {
        // a zero weight contributes exactly 0: every dim is evaluated, branch-free (p, q, r only where they are weighed)
        mse = real(0);
        constexpr int NM = KIND == SCG_QUAD_3D ? 9 : NX;
        if (KIND == SCG_QUAD_3D && (cfg & CFG_MSE_POS)) {   // same sum: the skipped terms are exact zeros
            const Row4<real> wt = ldc4(h.pos_pack + 12);   // mse_weight[0, 2, 4]
            { const real er = (st[0] - goal_pos_at(0)) * wt.x; mse += er * er; }
            { const real er = (st[2] - goal_pos_at(2)) * wt.y; mse += er * er; }
            { const real er = (st[4] - goal_pos_at(4)) * wt.z; mse += er * er; }
        } else {
#pragma unroll
            for (int d = 0; d < NM; ++d) { const real er = (st[d] - goal_at(d)) * h.mse_weight[d]; mse += er * er; }
        }
        if (NM < NX && (cfg & CFG_MSE_RATES)) {
#pragma unroll
            for (int d = NM; d < NX; ++d) { const real er = (st[d] - goal_at(d)) * h.mse_weight[d]; mse += er * er; }
        }
        if (cfg & CFG_COST_RL) {
            real dist = real(0), dact = real(0);
#pragma unroll
            for (int d = 0; d < NX; ++d) { const real er = st[d] - goal_at(d); dist += h.rew_state_weight[d] * er * er; }
#pragma unroll
            for (int j = 0; j < NU; ++j) { const real er = pre[j] - h.u_goal[j]; dact += h.rew_act_weight[j] * er * er; }
            reward = -(dist + dact);
            if (cfg & CFG_REW_EXP) {   // fp32 fast modes: MUFU.EX2 (|err| ~ 2 ulp of a value in (0, 1]); literal modes: expf / exp
                if constexpr (FASTMATH) reward = __expf(float(reward)); else reward = m_exp(reward);
            }
        } else if (cfg & CFG_COST_QUAD) {
            real lx = real(0), lu = real(0);
            if (PLAIN || !(cfg & CFG_QR_DENSE)) {
#pragma unroll
                for (int d = 0; d < NX; ++d) { const real er = st[d] - goal_q_at(d); lx += h.q_diag[d] * er * er; }
#pragma unroll
                for (int j = 0; j < NU; ++j) { const real er = pre[j] - h.u_goal[j]; lu += h.r_diag[j] * er * er; }
            } else {
                real ex[NX], eu[NU];
#pragma unroll
                for (int c = 0; c < NX; ++c) ex[c] = st[c] - goal_q_at(c);
#pragma unroll
                for (int c = 0; c < NU; ++c) eu[c] = pre[c] - h.u_goal[c];
                lx = dense_form<real, NX, SCG_MAX_NX>(cold->Q, ex);
                lu = dense_form<real, NU, SCG_MAX_NU>(cold->R, eu);
            }
            reward = -(real(0.5) * lx + real(0.5) * lu);
        } else {
            reward = real(0);
            if (gate_passed) reward += real(100);
            if (e.at_goal) reward += real(100);
            if (collided) reward -= real(1000);
            if (e.viol) reward -= real(100);   // previous step's flag (F8)
        }
    };
    if (cfg & CFG_STAB) {
        auto g0 = [&](int d) { return h.x_goal0[d]; };
        const Row4<real> g4 = ldc4(h.pos_pack + 8);
        auto gp = [&](int d) { return d == 0 ? g4.x : (d == 2 ? g4.y : g4.z); };   // x_goal0[d] for d = 0, 2, 4
        costs(g0, g0, gp);
    } else {
        // X_GOAL[t]; the index is clamped to the last row where the reference would raise an IndexError
        // (t >= rows never happens inside an episode: rows == ctrl_steps, quadrotor.py:924-926, 1016-1021)
        const int32_t gi = t < h.n_goal_rows - 1 ? t : h.n_goal_rows - 1;
        real ga[NX];
#pragma unroll
        for (int d = 0; d < NX; ++d) ga[d] = cold->x_goal[int64_t(gi) * NX + d];
        auto gt = [&](int d) { return ga[d]; };
        costs(gt, gt, gt);
    }
    // ---- after_step (benchmark_env.py:422-463)
    e.t = t + 1;
    if (cfg & CFG_HAS_ROWS) {
        bool violated, almost;
        eval_constraints<KIND, real, PLAIN>(h, cold, st, raw, false, c_out, n, i, violated, almost, cfg);
        e.viol = violated;
        if (violated) { info |= SCG_INFO_VIOLATION; if (cfg & CFG_DONE_VIOLATION) done = true; }
        if (!PLAIN && (cfg & CFG_PENALTY) && almost) reward += h.constraint_penalty;
    }
    if (e.t >= h.ctrl_steps) {
        if (!done) info |= SCG_INFO_TRUNCATED;
        done = true;
    }
}

SCG_DEV float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// the Logger record (Logger.py:50-66): pos3 vel3 rpy3 ang_vel3 rpm4, [16][n] per logged step
template <int KIND, typename real, bool FASTMATH>
SCG_DEV void log_record(const Hot<real>& h, const EnvRegs<KIND, real>& e, const Motors<real>& mot, real (&rec)[16]) {
    if constexpr (KIND == SCG_QUAD_3D) {
        rec[0] = e.x[0]; rec[1] = e.x[2]; rec[2] = e.x[4]; rec[3] = e.x[1]; rec[4] = e.x[3]; rec[5] = e.x[5];
        rec[6] = e.a3.rpy[0]; rec[7] = e.a3.rpy[1]; rec[8] = e.a3.rpy[2];
        rec[9] = e.x[9]; rec[10] = e.x[10]; rec[11] = e.x[11];
    } else if constexpr (KIND == SCG_QUAD_2D) {
        rec[0] = e.x[0]; rec[1] = 0; rec[2] = e.x[2]; rec[3] = e.x[1]; rec[4] = 0; rec[5] = e.x[3];
        rec[6] = 0; rec[7] = e.a2.theta; rec[8] = 0; rec[9] = 0; rec[10] = e.x[5]; rec[11] = 0;
    } else if constexpr (KIND == SCG_CARTPOLE) {
        rec[0] = e.x[0]; rec[1] = 0; rec[2] = 0; rec[3] = e.x[1]; rec[4] = 0; rec[5] = 0;
        rec[6] = 0; rec[7] = e.x[2]; rec[8] = 0; rec[9] = 0; rec[10] = e.x[3]; rec[11] = 0;
    } else {
        rec[0] = 0; rec[1] = 0; rec[2] = e.x[0]; rec[3] = 0; rec[4] = 0; rec[5] = e.x[1];
        rec[6] = rec[7] = rec[8] = rec[9] = rec[10] = rec[11] = 0;
    }
#pragma unroll
    for (int j = 0; j < 4; ++j) rec[12 + j] = motor_rpm<real, FASTMATH>(h, mot, j);
}
template <int KIND, typename real, bool FASTMATH>
SCG_DEV void write_log(const Hot<real>& h, const EnvRegs<KIND, real>& e, const Motors<real>& mot, real* __restrict__ L, int64_t n) {
    real rec[16];
    log_record<KIND, real, FASTMATH>(h, e, mot, rec);
#pragma unroll
    for (int j = 0; j < 16; ++j) L[int64_t(j) * n] = rec[j];
}

// ---- staged log store (TLOG): shared memory -> global through the TMA unit (cp.async.bulk.tensor, SASS UTMASTG).
// Every lane parks its 16-float record in its warp's stage ([LOG_T][16][32] floats, conflict-free); after LOG_T logged steps
// ONE lane issues ONE bulk tensor store of the [LOG_T * 16][32] box -- 16 STS per env-step instead of 16 strided STG with
// their 64-bit address arithmetic.  The hardware clips the box at the tensor's edges (ragged last warp: columns >= n).
SCG_DEV void tma_fence_smem_writes() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
SCG_DEV void tma_store_2d(const CUtensorMap* map, const float* smem, int32_t col, int32_t row) {
    asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%1, %2}], [%3];"
                 :: "l"(reinterpret_cast<uint64_t>(map)), "r"(col), "r"(row), "r"(uint32_t(__cvta_generic_to_shared(smem))) : "memory");
}
SCG_DEV void tma_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
template <int N> SCG_DEV void tma_wait_read() { asm volatile("cp.async.bulk.wait_group.read %0;" :: "n"(N) : "memory"); }
SCG_DEV void tma_wait_all() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }

template <int NU, typename real> struct ActVec;
template <> struct ActVec<4, float> { using type = float4; };
template <> struct ActVec<2, float> { using type = float2; };
template <> struct ActVec<1, float> { using type = float; };
template <> struct ActVec<4, double> { using type = double4; };
template <> struct ActVec<2, double> { using type = double2; };
template <> struct ActVec<1, double> { using type = double; };

// one vector load per env-step: the action row of env i ([n][nu] env-major, 16 B for 3D fp32)
template <int NU, typename real> SCG_DEV void load_action(const real* __restrict__ p, real (&a)[NU]) {
    using V = typename ActVec<NU, real>::type;
    if constexpr (NU == 4 && sizeof(real) == 8) {   // 32-byte rows: two 16-byte loads
        const double2 lo = __ldg(reinterpret_cast<const double2*>(p)), hi = __ldg(reinterpret_cast<const double2*>(p) + 1);
        a[0] = lo.x; a[1] = lo.y; a[2] = hi.x; a[3] = hi.y;
    } else {
        const V v = __ldg(reinterpret_cast<const V*>(p));
        const real* q = reinterpret_cast<const real*>(&v);
#pragma unroll
        for (int j = 0; j < NU; ++j) a[j] = q[j];
    }
}


// MODE 0: reset (mask optional)
template <int KIND, typename real, int ATT>
__global__ void __launch_bounds__(ENV_BLOCK)
k_reset(const __grid_constant__ Hot<real> h, const Cold<real>* __restrict__ cold, const Buf<real> b,
        const ResetInject<real> inj, const uint8_t* __restrict__ mask) {
    constexpr int NX = Dims<KIND>::NX, NU = Dims<KIND>::NU;
    const int64_t n = b.n;
    const int64_t i = int64_t(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i >= n || (mask != nullptr && mask[i] == 0)) return;
    const RngKey key{h.seed, h.env_id_base_lo + uint32_t(i)};
    EnvRegs<KIND, real> e;
#pragma unroll
    for (int d = 0; d < NX; ++d) e.x[d] = b.state[int64_t(d) * n + i];
    e.t = b.ctrl_step[i];
    e.resets = b.episode[i];
    unpack_flags<KIND, real>(e, b.flags[i]);
    real st[NX];
    reset_env_body<KIND, real, ATT>(h, cold, b, inj, key, n, i, e, st);
#pragma unroll
    for (int d = 0; d < NX; ++d) b.state[int64_t(d) * n + i] = e.x[d];
    b.ctrl_step[i] = e.t; b.episode[i] = e.resets; b.flags[i] = pack_flags<KIND, real>(e);
    if (b.env_state) {
#pragma unroll
        for (int d = 0; d < NX; ++d) b.env_state[int64_t(d) * n + i] = st[d];
    }
    write_obs<KIND, real>(h, cold, b.dist_offset, b.obs, st, 0, 0u, key, h.reseed_on_reset ? 0u : uint32_t(e.resets - 1), n, i);
    if (b.ep_return) b.ep_return[i] = real(0);
    if (b.aero_mem) b.aero_mem[i] = real(0);   // last_clipped_action = 0 (base_aviary.py:224)
    if (b.ep_length) b.ep_length[i] = 0;
    if (b.ep_violations) b.ep_violations[i] = 0;
    if (h.n_rows > 0 && b.c_values) {   // after_reset: state constraints only (benchmark_env.py:376-383)
        bool v, a;
        real raw0[NU];
#pragma unroll
        for (int j = 0; j < NU; ++j) raw0[j] = real(0);
        eval_constraints<KIND, real>(h, cold, st, raw0, true, b.c_values, n, i, v, a, h.cfg | CFG_FULL_STATE);
    }
}

// K fused control steps (scg_step is K = 1 with the per-step outputs)
// LEAN: the auto-reset takes a spare initial state drawn ahead (reset() draws nothing but uniform init offsets).
// PLAIN (LEAN only): none of the per-step extras is configured -- disturbances, adversary, external force, dense /
// tolerance / symmetric constraints, off-diagonal Q / R, constraint penalty, gate geometry -- and their uniform
// branches are compiled out (the headline workload: the loop must stay inside the instruction cache).
// SPEC (PLAIN only): the scenario switches of CFG_SPEC_MASK are compile-time constants (scg_params.cuh).
//
// rollout_range runs control steps [k0, k1) of env i: state from / to Buf, exactly what a launch of k1 - k0 steps does
// (a K-step rollout equals any split of it into launches).  Statistics go to the caller's shared-memory columns
// acc[j * ACC_LD], j < N_ACC, which the kernel reduces once at its end.
#ifndef SCG_BLOCKS_PER_SM
#define SCG_BLOCKS_PER_SM 9   // resident 64-thread blocks per SM the SPEC rollout kernels are bounded for (96 registers, no spill); the other fp32
                              // kernels: 8 (128 registers) -- at 96 the PLAIN kernel with per-step outputs spills and the single step goes 21 -> 29 us
#endif
#ifndef SCG_DS_BLOCKS
#define SCG_DS_BLOCKS 8       // resident blocks per SM the PLAIN + DS kernels are bounded for
#endif
#ifndef SCG_SPARE_CADENCE
#define SCG_SPARE_CADENCE 16
#endif
constexpr int SPARE_CADENCE = SCG_SPARE_CADENCE;   // control steps between the warp-wide spare refills (power of two)
constexpr int ACC_LD = ENV_BLOCK + 1;   // column ENV_BLOCK is the scratch column of out-of-range lanes
constexpr int N_ACC = 11;   // episodes, return, length, truncated, completed, violated episodes, violation steps, (mse: double column), env steps, gates, collisions
// SPEC: 0 none; 1 the kind's family (SpecCfg) for fused launches without per-step outputs; 2 (quad_type 3 PLAIN only) the same family
// for launches WITH per-step outputs -- the host-driven scg_step (9.73 -> 9.12 us per step at 65,536 envs).
template <int KIND, typename real, int ATT, bool LEAN, bool PLAIN, int SPEC, bool TLOG = false, bool DS = false>
SCG_DEV void rollout_range(const Hot<real>& h, const Cold<real>* __restrict__ cold, const Buf<real>& b, const Roll<real>& ro,
                           const int64_t i, const int k0, const int k1, float* const ep_acc, double* const mse_acc, real* const sp,
                           const LogMaps* const lg_maps = nullptr, float* const lg_warp = nullptr, real* const pm = nullptr) {
    constexpr int NX = Dims<KIND>::NX, NU = Dims<KIND>::NU;
    const int64_t n = b.n;
    // Out-of-range lanes of the last env block READ env n - 1 and compute along with the warp, but every global side
    // effect of the loop -- state, PID memory, disturbance offsets / inertia, per-step outputs, the log -- is predicated on
    // `valid`: each word of batch memory has exactly one writer.  (Leaving the function early instead makes the whole loop
    // body potentially divergent code, where the compiler cannot use the uniform datapath for the scenario constants:
    // measured 6.11e10 -> 5.78e10 env-steps/s on the headline workload, profiles/r2_summary.md.)  What a shadow lane
    // computes may drift from its owner (it can observe the owner's PID memory / offsets at another step); nothing of it
    // is stored, and its statistics go to the scratch column.
    const bool valid = i < n;
    ep_acc[1 * ACC_LD] = 0.f;   // return of the episodes that end in this range: summed in fp32 per range, in fp64 across ranges
    const int64_t ii = valid ? i : n - 1;
    const RngKey key{h.seed, h.env_id_base_lo + uint32_t(ii)};
    EnvRegs<KIND, real> e;
#pragma unroll
    for (int d = 0; d < NX; ++d) e.x[d] = b.state[int64_t(d) * n + ii];
    e.t = b.ctrl_step[ii];
    e.resets = b.episode[ii];
    unpack_flags<KIND, real>(e, b.flags[ii]);
    e.offc = (PLAIN && !DS) ? nullptr : (LEAN ? sp + (Spare<KIND, real, ATT>::N + 4 + 3 * SCG_MAX_DIST) * ENV_BLOCK : sp);
    // SPEC: cost "competition" is a compile-time fact, the goal_reached bit (cost "quadratic" only) cannot change: it is
    // not carried through the loop (one register less at the 96-register bound) but re-read where the flags are stored
    if constexpr (SPEC && KIND == SCG_QUAD_3D) e.goal_prev = false;
    real st[NX];   // env.state after the last step (LEAN: rebuilt from the registers after the loop)
    load_inertia<KIND, real>(h, (PLAIN && !DS) ? nullptr : b.inertia, n, ii, e.in);
    if constexpr (!PLAIN || DS) { if (h.cfg & CFG_DIST_SIMPLE) load_offsets<KIND, real>(h, b.dist_offset, n, ii, e); }
    derive_attitude<KIND, real, ATT>(e);
    e.aero_prev = real(0);
    if constexpr (!PLAIN && KIND == SCG_QUAD_3D) { if ((h.cfg & CFG_AERO) && b.aero_mem) e.aero_prev = b.aero_mem[ii]; }
    constexpr bool PAIRED = Paired<KIND, real, ATT, LEAN, PLAIN>::value;
    if constexpr (PAIRED) pairs_from_scalars(e);
    real ep_ret = b.ep_return ? b.ep_return[ii] : real(0);
    int32_t ep_len = b.ep_length ? b.ep_length[ii] : 0;
    int32_t ep_viol = b.ep_violations ? b.ep_violations[ii] : 0;
    // per-thread statistics of this range (added to the shared-memory columns at its end)
    float s_viol = 0.f, s_mse = 0.f;
    // the statistics that only change when an episode ends live in a per-thread column of shared memory (ep_acc[0..5]):
    // six values that would otherwise be held (or spilled) across every control step for an update on 2 % of them
    float s_gates = 0.f, s_coll = 0.f;   // gate_geometry only
    real reward = real(0), mse = real(0);
    bool done = false;
    uint32_t info = 0u;
    Motors<real> mot;
    // the uniform switches of the loop in one word (see CFG_*)
    const uint32_t cfg_rt = h.cfg | (ro.step_outputs ? (CFG_FULL_STATE | CFG_STEP_OUTPUTS) : 0u);
    const uint32_t cfg = SPEC == 1 ? ((cfg_rt & ~SpecCfg<KIND>::MASK) | SpecCfg<KIND>::VAL)
                       : SPEC == 2 ? ((cfg_rt & ~SpecCfg<KIND>::MASK) | SpecCfg<KIND>::VAL | CFG_FULL_STATE | CFG_STEP_OUTPUTS) : cfg_rt;
    const bool full = (cfg & CFG_FULL_STATE) != 0u;
    real* const c_out = SPEC == 1 ? nullptr : ((ro.step_outputs && valid) ? b.c_values : nullptr);
    // software prefetch of the next action (one 16-byte load per 3D env per step)
    real raw[NU], nxt[NU];
    // uniform: actions from the on-device PID (scg_pid.cuh); its memory rides in the shared-memory column pm[j * ENV_BLOCK]
    // (which loops carry the controller: the LEAN / GENERAL ones of the quadrotors and the 3D PLAIN + DS loop)
    constexpr bool PIDK = (!PLAIN || (DS && KIND == SCG_QUAD_3D)) && !TLOG && (KIND == SCG_QUAD_3D || KIND == SCG_QUAD_2D);
    const bool closed_loop = PIDK && ro.use_pid != 0;
    if constexpr (PIDK) {
        if (closed_loop) {
#pragma unroll
            for (int j = 0; j < 9; ++j) pm[j * ENV_BLOCK] = ro.pid_mem[int64_t(j) * n + ii];
        }
    }
    const real* ap = ro.actions + (int64_t(k0) * n + ii) * NU;   // action row of this env at the step being prefetched
    if (!closed_loop) load_action<NU, real>(ap, nxt);
    else {
#pragma unroll
        for (int j = 0; j < NU; ++j) nxt[j] = real(0);
    }
    int lg_slot = 0, lg_stage = 0, lg_row0 = 0;   // TLOG (warp-uniform): records in the open stage, which stage, its first logged step
    bool fresh = false;   // obs must come from a reset this step
    // LEAN: spare initial state for the next episode (drawn ahead, see draw_init_lean)
    bool spare_ok = false;
    const bool ahead = k1 - k0 >= 16;   // uniform: host-driven single steps draw on demand only
    const bool any_out = ro.reward_out != nullptr || ro.done_out != nullptr || ro.log_out != nullptr;   // uniform
#pragma unroll 1
    for (int k = k0; k < k1; ++k) {
#pragma unroll
        for (int j = 0; j < NU; ++j) raw[j] = nxt[j];
        if constexpr (PIDK) {
            if (closed_loop) {   // action = PID(state before the step, reference point of this step)
                real pos[3], vel[3], rpy[3], tp[3], tv[3], act[NU];
                const int32_t gi = e.t < h.n_goal_rows - 1 ? e.t : h.n_goal_rows - 1;
                const bool stab = (cfg & CFG_STAB) != 0u;
                auto goal_at = [&](int d) { return stab ? h.x_goal0[d] : cold->x_goal[int64_t(gi) * NX + d]; };
                if constexpr (KIND == SCG_QUAD_3D) {
                    if constexpr (PAIRED) {   // the loop-carried pairs (EnvRegs): no conversion, the halves are registers
                        pos[0] = lo2(e.Pxy); pos[1] = hi2(e.Pxy); pos[2] = lo2(e.ZA); vel[0] = lo2(e.Vxy); vel[1] = hi2(e.Vxy); vel[2] = lo2(e.ZB);
                        rpy[0] = lo2(e.Axz); rpy[1] = hi2(e.ZA); rpy[2] = hi2(e.Axz);
                    } else {
                        pos[0] = e.x[0]; pos[1] = e.x[2]; pos[2] = e.x[4]; vel[0] = e.x[1]; vel[1] = e.x[3]; vel[2] = e.x[5];
                        rpy[0] = e.a3.rpy[0]; rpy[1] = e.a3.rpy[1]; rpy[2] = e.a3.rpy[2];
                    }
                    tp[0] = goal_at(0); tp[1] = goal_at(2); tp[2] = goal_at(4);
                    tv[0] = stab ? real(0) : goal_at(1); tv[1] = stab ? real(0) : goal_at(3); tv[2] = stab ? real(0) : goal_at(5);
                } else {
                    pos[0] = e.x[0]; pos[1] = real(0); pos[2] = e.x[2]; vel[0] = e.x[1]; vel[1] = real(0); vel[2] = e.x[3];
                    rpy[0] = real(0); rpy[1] = e.a2.theta; rpy[2] = real(0);
                    tp[0] = goal_at(0); tp[1] = real(0); tp[2] = goal_at(2);
                    tv[0] = stab ? real(0) : goal_at(1); tv[1] = real(0); tv[2] = stab ? real(0) : goal_at(3);
                }
                pid_action<KIND, real, ATT>(ro.pid, pos, vel, rpy, tp, tv, pm, ENV_BLOCK, act);
#pragma unroll
                for (int j = 0; j < NU; ++j) raw[j] = act[j];
                if (ro.actions_out && valid) {
#pragma unroll
                    for (int j = 0; j < NU; ++j) ro.actions_out[(int64_t(k) * n + ii) * NU + j] = act[j];
                }
            }
        }
        ap += n * NU;
        if (!closed_loop && k + 1 < k1) load_action<NU, real>(ap, nxt);
        // LEAN: env.state of this step is loop-local (it aliases the registers of e); what is observed
        // after the loop is rebuilt from e there
        real sl[LEAN ? NX : 1];
        real (&sk)[NX] = *reinterpret_cast<real (*)[NX]>(LEAN ? sl : st);
        step_env<KIND, real, ATT, LEAN, PLAIN, DS>(h, cold, b, key, n, ii, e, raw, cfg, sk, reward, done, info, mse, mot, c_out);
        ep_ret += reward; ep_len += 1;
        const bool viol = (info & SCG_INFO_VIOLATION) != 0u;
        ep_viol += viol ? 1 : 0;
        s_viol += viol ? 1.f : 0.f;
        s_mse += float(mse);
        if ((!PLAIN || DS) && (cfg & CFG_GATE_GEOMETRY)) {
            s_gates += (info & SCG_INFO_GATE_PASSED) ? 1.f : 0.f;
            s_coll += (info & SCG_INFO_COLLISION) ? 1.f : 0.f;
        }
        if (any_out) {
            if (ro.reward_out && valid) ro.reward_out[int64_t(k) * n + ii] = reward;
            if (ro.done_out && valid) ro.done_out[int64_t(k) * n + ii] = done ? 1 : 0;
            if constexpr (TLOG) {
                if (k % ro.log_stride == 0) {   // uniform: the whole warp (shadow lanes included -- their columns are clipped) stages
                    float rec[16];
                    if constexpr (PAIRED) {
                        EnvRegs<KIND, real> el = e;
                        scalars_from_pairs(el);
                        log_record<KIND, real, true>(h, el, mot, rec);
                    } else log_record<KIND, real, true>(h, e, mot, rec);
                    const int lane = threadIdx.x & 31;
                    if (lg_slot == 0) {   // this stage's previous store (two stores ago) must have finished reading it
                        if (lane == 0) tma_wait_read<LOG_STAGES - 1>();
                        __syncwarp();
                        lg_row0 = k / ro.log_stride;
                    }
                    float* const stg = lg_warp + lg_stage * LOG_STAGE;
#pragma unroll
                    for (int j = 0; j < 16; ++j) stg[(lg_slot * 16 + j) * 32 + lane] = rec[j];
                    if (++lg_slot == LOG_T) {
                        tma_fence_smem_writes();
                        __syncwarp();
                        if (lane == 0) { tma_store_2d(&lg_maps->full, stg, int32_t(i) - lane, lg_row0 * 16); tma_commit(); }
                        lg_slot = 0; lg_stage = lg_stage + 1 == LOG_STAGES ? 0 : lg_stage + 1;
                    }
                }
            } else
            if (ro.log_out && (k % ro.log_stride) == 0 && valid) {
                if constexpr (PAIRED) {
                    EnvRegs<KIND, real> el = e;   // the record reads the scalar layout
                    scalars_from_pairs(el);
                    write_log<KIND, real, true>(h, el, mot, ro.log_out + int64_t(k / ro.log_stride) * 16 * n + ii, n);
                } else
                write_log<KIND, real, (sizeof(real) == 4 && ATT != SCG_ATT_ROUNDTRIP)>(h, e, mot, ro.log_out + int64_t(k / ro.log_stride) * 16 * n + ii, n);
            }
        }
        fresh = false;
        if constexpr (LEAN) {
            // The spare is made ready BEFORE the episode-end block so that statistics and restart share one divergent
            // region.  One inline copy of the draw: at the cadence every lane that used its spare refills it together;
            // a lane that ends two episodes inside one period draws on the spot.
            const bool restart = done && (cfg & CFG_AUTO_RESET) != 0u;
            const bool cadence = ahead && ((k - k0) & (SPARE_CADENCE - 1)) == 0;   // uniform
            if (bool(int(!spare_ok) & (int(restart) | int(cadence)))) {
                const uint32_t next_ep = (cfg & CFG_RESEED) ? 0u : uint32_t(e.resets);
                if ((!PLAIN || DS) && (cfg & CFG_RESET_EXTRAS)) {
                    if constexpr (DS) draw_spare_general_body<KIND, real, ATT>(h, cold, key, next_ep, sp);
                    else draw_spare_general<KIND, real, ATT>(h, cold, key, next_ep, sp);
                }
                else {
                    real init[NX];
                    draw_init_lean<KIND, real>(h, key, next_ep, init);
                    spare_store<KIND, real, ATT>(init, sp);
                }
                spare_ok = true;
            }
        }
        if (done) {
            ep_acc[0 * ACC_LD] += 1.f; ep_acc[1 * ACC_LD] += float(ep_ret); ep_acc[2 * ACC_LD] += float(ep_len);
            if (info & SCG_INFO_TRUNCATED) ep_acc[3 * ACC_LD] += 1.f;
            if (info & SCG_INFO_COMPLETED) ep_acc[4 * ACC_LD] += 1.f;
            if (ep_viol > 0) ep_acc[5 * ACC_LD] += 1.f;
            if (cfg & CFG_AUTO_RESET) {   // DummyVecEnv.step_wait (dummy_vec_env.py:38-49)
                if ((cfg & CFG_STEP_OUTPUTS) && b.terminal_obs && valid) {
                    real tl[NX];
#pragma unroll
                    for (int d = 0; d < NX; ++d) tl[d] = sk[d];
                    write_obs_sel<KIND, real, (SPEC != 1 || KIND == SCG_CARTPOLE)>(h, cold, b.dist_offset, b.terminal_obs, tl, e.t - 1, uint32_t(e.t), key,
                                                                              (cfg & CFG_RESEED) ? 0u : uint32_t(e.resets - 1), n, ii);
                }
                if constexpr (!LEAN) {
                    EnvRegs<KIND, real> ec = e;
                    real sn[NX];
                    reset_env_cold<KIND, real, ATT>(h, cold, b, key, n, ii, &ec, sn, valid);
                    e = ec;
                    if (cfg & CFG_DIST_SIMPLE) load_offsets<KIND, real>(h, b.dist_offset, n, ii, e);
#pragma unroll
                    for (int d = 0; d < NX; ++d) st[d] = sn[d];
                }
                ep_ret = real(0); ep_len = 0; ep_viol = 0;
                fresh = true;
                if constexpr (!PLAIN) e.aero_prev = real(0);   // (after reset_env_cold copied the registers back; the LEAN restart below keeps it)
                if constexpr (PIDK) {
                    if (closed_loop) {   // PIDController.reset() with the env (quadrotor_utils.py:124-135)
#pragma unroll
                        for (int j = 0; j < 9; ++j) pm[j * ENV_BLOCK] = real(0);
                    }
                }
                if constexpr (LEAN) {
                    reset_from_spare<KIND, real, ATT, PAIRED>(sp, e);
                    if constexpr (DS) {
                        if (cfg & CFG_RESET_EXTRAS) apply_spare_simple<KIND, real, ATT>(h, b.inertia, b.dist_offset, sp, n, ii, e, valid);
                    } else if (!PLAIN && (cfg & CFG_RESET_EXTRAS)) {
                        if ((cfg & CFG_SPARE_SIMPLE) && ((cfg & CFG_DIST_SIMPLE) || h.n_dist_act == 0)) apply_spare_simple<KIND, real, ATT>(h, b.inertia, b.dist_offset, sp, n, ii, e, valid);
                        else {
                            Inertia<real> in;
                            apply_spare_extras<KIND, real, ATT>(h, cold, b.inertia, b.dist_offset, sp, n, ii, &in, valid);
                            e.in = in;
                            if (cfg & CFG_DIST_SIMPLE) {   // (offsets of this episode as apply_spare_extras published them)
#pragma unroll
                                for (int q = 0; q < SCG_MAX_DIST; ++q)
                                    if (q < h.ds_n && h.ds_off[q] < 0 && h.ds_kind[q] != SCG_DIST_WHITE)
                                        e.offc[q * ENV_BLOCK] = sp[(Spare<KIND, real, ATT>::N + 4 + h.ds_slot[q]) * ENV_BLOCK];
                            }
                        }
                    }
                    spare_ok = false;
                }
            }
        }
    }
    if constexpr (TLOG) {   // the records of an incomplete stage go out one logged step at a time; then drain
        const int lane = threadIdx.x & 31;
        if (lg_slot > 0) { tma_fence_smem_writes(); __syncwarp(); }
        if (lane == 0) {
            for (int s = 0; s < lg_slot; ++s)
                tma_store_2d(&lg_maps->one, lg_warp + lg_stage * LOG_STAGE + s * 16 * 32, int32_t(i) - lane, (lg_row0 + s) * 16);
            tma_commit();
            tma_wait_all();
        }
        __syncwarp();
    }
    if constexpr (PIDK) {
        if (closed_loop && valid) {   // (a shadow lane's copy is dropped: the memory belongs to the owner lane)
#pragma unroll
            for (int j = 0; j < 9; ++j) ro.pid_mem[int64_t(j) * n + ii] = pm[j * ENV_BLOCK];
        }
    }
    if constexpr (PAIRED) scalars_from_pairs(e);
    if constexpr (!PLAIN && KIND == SCG_QUAD_3D) { if ((h.cfg & CFG_AERO) && b.aero_mem && valid) b.aero_mem[ii] = e.aero_prev; }
    // a range that ends inside the launch (ticket queue) hands over the state only: observation, env.state and the
    // per-step scalars are rewritten by the range that ends the launch
    const bool last_range = k1 >= ro.k_steps;   // uniform
    if (!valid) {
        // nothing of a shadow lane is stored
    } else if (!last_range) {
#pragma unroll
        for (int d = 0; d < NX; ++d) b.state[int64_t(d) * n + ii] = e.x[d];
        b.ctrl_step[ii] = e.t; b.episode[ii] = e.resets;
        b.flags[ii] = pack_flags<KIND, real>(e) | ((SPEC && KIND == SCG_QUAD_3D) ? (b.flags[ii] & SCG_FLAG_GOAL_REACHED) : 0u);
    } else {
    if constexpr (LEAN) {
        // env.state as observed after the loop: the fresh episode's (PyBullet still holds INIT_ANG_VEL)
        // or the last step's (ang_v = rpy_rates)
        real ang[3] = {real(0), real(0), real(0)};
        if (fresh) spare_ang0<KIND, real>(sp, ang);
        else if constexpr (KIND == SCG_QUAD_3D) { ang[0] = e.x[9]; ang[1] = e.x[10]; ang[2] = e.x[11]; }
        else if constexpr (KIND == SCG_QUAD_2D) { ang[1] = e.x[5]; }
        // sin / cos of the angles are not carried through the loop: the same arithmetic on the stored
        // pre-round-trip angles gives them back bit for bit
        if constexpr (KIND == SCG_QUAD_3D) derive_attitude<KIND, real, ATT>(e);
        make_state<KIND, real>(e, ang, st);
    } else if constexpr (KIND == SCG_QUAD_3D) {
        if (!full && !fresh) {   // the final observation needs p, q, r
            attitude3_complete<real>(e.a3);
            const real* R = e.a3.R;
            st[9] = (R[0] * e.x[9] + R[3] * e.x[10]) + R[6] * e.x[11];
            st[10] = (R[1] * e.x[9] + R[4] * e.x[10]) + R[7] * e.x[11];
            st[11] = (R[2] * e.x[9] + R[5] * e.x[10]) + R[8] * e.x[11];
        }
    }
#pragma unroll
    for (int d = 0; d < NX; ++d) b.state[int64_t(d) * n + ii] = e.x[d];
    b.ctrl_step[ii] = e.t; b.episode[ii] = e.resets;
    b.flags[ii] = pack_flags<KIND, real>(e) | ((SPEC && KIND == SCG_QUAD_3D) ? (b.flags[ii] & SCG_FLAG_GOAL_REACHED) : 0u);
    if (b.env_state) {
#pragma unroll
        for (int d = 0; d < NX; ++d) b.env_state[int64_t(d) * n + ii] = st[d];
    }
    {
        const uint32_t ep = h.reseed_on_reset ? 0u : uint32_t(e.resets - 1);
        real fl[NX];
#pragma unroll
        for (int d = 0; d < NX; ++d) fl[d] = st[d];
        write_obs_sel<KIND, real, (SPEC != 1 || KIND == SCG_CARTPOLE)>(h, cold, b.dist_offset, b.obs, fl, fresh ? 0 : e.t - 1, fresh ? 0u : uint32_t(e.t), key, ep, n, ii);
    }
    b.reward[ii] = reward; b.done[ii] = done ? 1 : 0; b.info[ii] = uint8_t(info);
    if (b.mse) b.mse[ii] = mse;
    }
    if (b.ep_return && valid) b.ep_return[ii] = ep_ret;
    if (b.ep_length && valid) b.ep_length[ii] = ep_len;
    if (b.ep_violations && valid) b.ep_violations[ii] = ep_viol;
    // statistics of this range into the caller's columns (an out-of-range lane was handed the scratch column)
    // (the mse sum goes into a double column: a sum of fp32 terms in fp64 is exact, so the total does not depend on
    // which resident thread ran which chunk; the other statistics are small integers, exact in fp32)
    ep_acc[6 * ACC_LD] += s_viol; ep_acc[8 * ACC_LD] += float(k1 - k0);
    mse_acc[0] += double(s_mse); mse_acc[ACC_LD] += double(ep_acc[1 * ACC_LD]);
    if (!PLAIN || DS) { ep_acc[9 * ACC_LD] += s_gates; ep_acc[10 * ACC_LD] += s_coll; }
}

// episode statistics of a block: warp-shuffle reduction, one atomic per warp per statistic
template <bool PLAIN>
SCG_DEV void reduce_stats(double* __restrict__ stats, const float* __restrict__ acc_buf, const double* __restrict__ mse_buf) {
    if (stats) {
        stats += ((blockIdx.x * (ENV_BLOCK / 32) + (threadIdx.x >> 5)) & (SCG_STAT_REPLICAS - 1)) * SCG_STATS_LEN;   // this warp's replica
        constexpr int NS = PLAIN ? 9 : 11;
        const float* const a = acc_buf + threadIdx.x;
        float v[11] = {a[0 * ACC_LD], a[1 * ACC_LD], a[2 * ACC_LD], a[6 * ACC_LD], a[7 * ACC_LD], a[3 * ACC_LD],
                       a[4 * ACC_LD], a[5 * ACC_LD], a[8 * ACC_LD], a[9 * ACC_LD], a[10 * ACC_LD]};
        const int slot[11] = {SCG_STAT_EPISODES, SCG_STAT_RETURN, SCG_STAT_LENGTH, SCG_STAT_VIOLATION_STEPS,
                              SCG_STAT_MSE, SCG_STAT_TRUNCATED, SCG_STAT_COMPLETED, SCG_STAT_VIOLATED_EPISODES, SCG_STAT_ENV_STEPS,
                              SCG_STAT_GATES, SCG_STAT_COLLISIONS};
        // per-thread counts are exact in fp32 (<= 2^24 control steps per launch and thread); their sum over a warp is
        // formed in fp64 so that it stays exact for any launch length
#pragma unroll
        for (int j = 0; j < NS; ++j) {
            if (slot[j] == SCG_STAT_MSE || slot[j] == SCG_STAT_RETURN) continue;
            // (counts: integers below 2^24 per thread, so the warp sum is one exact integer reduction)
            const int r = __reduce_add_sync(0xffffffffu, __float2int_rn(v[j]));
            if ((threadIdx.x & 31) == 0 && r != 0) atomicAdd(stats + slot[j], double(r));
        }
        double m = mse_buf[threadIdx.x], r = mse_buf[ACC_LD + threadIdx.x];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) { m += __shfl_xor_sync(0xffffffffu, m, o); r += __shfl_xor_sync(0xffffffffu, r, o); }
        if ((threadIdx.x & 31) == 0 && m != 0.0) atomicAdd(stats + SCG_STAT_MSE, m);
        if ((threadIdx.x & 31) == 0 && r != 0.0) atomicAdd(stats + SCG_STAT_RETURN, r);
    }
}

SCG_DEV int32_t ld_acquire_gpu(const int32_t* p) {
    int32_t v;
    asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
SCG_DEV void st_release_gpu(int32_t* p, int32_t v) { asm volatile("st.release.gpu.global.s32 [%0], %1;" ::"l"(p), "r"(v) : "memory"); }

// QUEUE = false: block b owns envs [64 b, 64 b + 64) for all K steps.
// QUEUE = true (fused launches whose envs fit in a few waves): the K steps are cut into chunks of ro.q_chunk; the
// resident blocks pop (chunk, env block) ranges from a ready list in global memory and a block that finishes a chunk
// pushes the env block's next one (release / acquire on the list slot), the state travelling through Buf.  With 65,536 envs
// the static split leaves 272 of the 592 schedulers with four warps and 320 with three for the whole launch, and the
// four-warp ones set the makespan; through the queue every scheduler hosts four warps, the ~14 % that wait for a range
// sleep, and a scheduler that runs ahead simply takes more ranges (same result: the split into ranges is invisible, see
// rollout_range).
template <int KIND, typename real, int ATT, bool LEAN, bool PLAIN = false, int SPEC = 0, bool QUEUE = false, bool DS = false>
__global__ void __launch_bounds__(ENV_BLOCK, sizeof(real) == 4 ? (DS ? SCG_DS_BLOCKS : (SPEC == 1 ? SCG_BLOCKS_PER_SM : 8)) : 1)
k_rollout(const __grid_constant__ Hot<real> h, const Cold<real>* __restrict__ cold, const Buf<real> b, const Roll<real> ro) {
    if constexpr (!QUEUE) {
        if (ro.pdl) {   // uniform; scg_inst.cuh: launch_rollout_v.  Nothing above this line reads batch memory.
            asm volatile("griddepcontrol.launch_dependents;");
            asm volatile("griddepcontrol.wait;" ::: "memory");
        }
    }
    __shared__ float acc_buf[N_ACC * ACC_LD];
    __shared__ double mse_buf[2 * ACC_LD];   // fp64 columns: mse sum, return sum
    mse_buf[threadIdx.x] = 0.0; mse_buf[ACC_LD + threadIdx.x] = 0.0;
    __shared__ real sp_buf[LEAN ? (Spare<KIND, real, ATT>::N + ((PLAIN && !DS) ? 0 : SPARE_EXTRA)) * ENV_BLOCK : SCG_MAX_DIST * ENV_BLOCK];
    real* const sp = sp_buf + threadIdx.x;
    __shared__ real pid_buf[((PLAIN && !(DS && KIND == SCG_QUAD_3D)) || KIND == SCG_CARTPOLE || KIND == SCG_QUAD_1D) ? 1 : 9 * ENV_BLOCK];   // the closed loop's controller memory, one column per thread (scg_pid.cuh)
#pragma unroll
    for (int j = 0; j < N_ACC; ++j) acc_buf[j * ACC_LD + threadIdx.x] = 0.f;
    if constexpr (!QUEUE) {
        const int64_t i = int64_t(blockIdx.x) * blockDim.x + threadIdx.x;
        rollout_range<KIND, real, ATT, LEAN, PLAIN, SPEC, false, DS>(h, cold, b, ro, i, 0, ro.k_steps, acc_buf + (i < b.n ? threadIdx.x : ENV_BLOCK),
                                                          mse_buf + (i < b.n ? threadIdx.x : ENV_BLOCK), sp, nullptr, nullptr, pid_buf + threadIdx.x);
    } else {
        // q_mem: [0] pop counter, [1] push counter, [2 + s] the ready list: item s = 1 + env block + env_blocks * chunk.
        // Items 0 .. env_blocks - 1 are implicit (chunk 0 of every env block is ready at launch); every finished chunk
        // but the last of its env block pushes the next one, so the list fills in COMPLETION order and the block that
        // has waited longest gets whichever env block is handed over first (no head-of-line blocking behind a slow one).
        __shared__ int32_t s_item;
        int32_t* const items = ro.q_mem + 2;
        int32_t push = 0;   // thread 0: the item to publish for the range this block has just finished (0: none)
        for (;;) {
            __syncthreads();   // the previous range is complete: its stores precede thread 0's release below (and everybody has read s_item)
            if (threadIdx.x == 0) {
                // hand-over and pop in one go: the two counter atomics are independent and overlap; st.release.gpu carries
                // the fence that makes the block's state stores visible before the item
                const int32_t slot = atomicAdd(ro.q_mem, 1);
                if (push) st_release_gpu(items + ro.q_env_blocks + atomicAdd(ro.q_mem + 1, 1), push);
                int32_t item = slot < ro.q_env_blocks ? slot + 1 : 0;
                if (slot >= ro.q_env_blocks && slot < ro.q_tickets) {
                    while ((item = ld_acquire_gpu(items + slot)) == 0) __nanosleep(ro.q_sleep_ns);
                }
                s_item = slot < ro.q_tickets ? item : 0;
            }
            __syncthreads();
            const int32_t item = s_item;
            if (item == 0) break;   // every range of the launch has been handed out
            const int32_t chunk = (item - 1) / ro.q_env_blocks, eb = (item - 1) - chunk * ro.q_env_blocks;
            const int64_t i = int64_t(eb) * ENV_BLOCK + threadIdx.x;
            const int k0 = chunk * ro.q_chunk, k1 = k0 + ro.q_chunk < ro.k_steps ? k0 + ro.q_chunk : ro.k_steps;
            rollout_range<KIND, real, ATT, LEAN, PLAIN, SPEC, false, DS>(h, cold, b, ro, i, k0, k1, acc_buf + (i < b.n ? threadIdx.x : ENV_BLOCK),
                                                              mse_buf + (i < b.n ? threadIdx.x : ENV_BLOCK), sp, nullptr, nullptr, pid_buf + threadIdx.x);
            push = k1 < ro.k_steps ? item + ro.q_env_blocks : 0;
        }
    }
    reduce_stats<(PLAIN && !DS)>(b.stats, acc_buf, mse_buf);
}

// The same launch shape as k_rollout<..., QUEUE = false> with the trajectory log staged in shared memory and written by the
// TMA unit (rollout_range<..., TLOG = true>; 3D fp32, launched from scg_k_3d_f_tlog.cu when the log tensor qualifies).
// DS: the PLAIN + DS loop (inline disturbances, gate geometry behind its inline broad phase, cheap restart) -- BASELINE config 5.
template <int ATT, bool DS, bool SPEC>
#ifndef SCG_TLOG_BLOCKS
#define SCG_TLOG_BLOCKS 9   // 96 registers, like the SPEC rollout kernels (SCG_BLOCKS_PER_SM)
#endif
__global__ void __launch_bounds__(ENV_BLOCK, (SPEC && !DS) ? SCG_TLOG_BLOCKS : 8)   // (the other kernels spill at 96 registers)
k_rollout_tlog(const __grid_constant__ Hot<float> h, const Cold<float>* __restrict__ cold, const Buf<float> b, const Roll<float> ro,
               const __grid_constant__ LogMaps maps) {
    constexpr int KIND = SCG_QUAD_3D;
    __shared__ float acc_buf[N_ACC * ACC_LD];
    __shared__ double mse_buf[2 * ACC_LD];
    __shared__ float sp_buf[(Spare<KIND, float, ATT>::N + (DS ? SPARE_EXTRA : 0)) * ENV_BLOCK];
    __shared__ __align__(128) float lg_buf[2 * LOG_STAGES * LOG_STAGE];   // [warp][stage][LOG_T][16][32]
    mse_buf[threadIdx.x] = 0.0; mse_buf[ACC_LD + threadIdx.x] = 0.0;
#pragma unroll
    for (int j = 0; j < N_ACC; ++j) acc_buf[j * ACC_LD + threadIdx.x] = 0.f;
    const int64_t i = int64_t(blockIdx.x) * blockDim.x + threadIdx.x;
    rollout_range<KIND, float, ATT, true, true, SPEC, true, DS>(h, cold, b, ro, i, 0, ro.k_steps, acc_buf + (i < b.n ? threadIdx.x : ENV_BLOCK),
                                                               mse_buf + (i < b.n ? threadIdx.x : ENV_BLOCK), sp_buf + threadIdx.x, &maps,
                                                               lg_buf + (threadIdx.x >> 5) * LOG_STAGES * LOG_STAGE);
    reduce_stats<!DS>(b.stats, acc_buf, mse_buf);
}


}  // namespace scg
