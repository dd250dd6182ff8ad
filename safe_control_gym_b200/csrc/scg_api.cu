// scg_api.cu -- host side of libscg_b200.so: the extern "C" ABI declared in include/scg_b200.h.
// Builds the device parameter blocks from scg_config, validates shapes, and launches the fused
// kernels of scg_step.cuh / scg_symbolic.cuh on the caller's stream.
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <limits>
#include <new>
#include <string>

#include "scg_host.cuh"
#include "scg_symbolic.cuh"

using namespace scg;

static thread_local std::string g_err;
int scg_fail(int code, const std::string& msg) { g_err = msg; return code; }
int scg_cuda_fail(cudaError_t e, const char* what) {
    g_err = std::string(what) + ": " + cudaGetErrorString(e);
    return SCG_ERR_CUDA - int(e);
}
static int fail(int code, const std::string& msg) { return scg_fail(code, msg); }
static int cuda_fail(cudaError_t e, const char* what) { return scg_cuda_fail(e, what); }
#define CU(x) SCG_CU(x)

static int kind_dims(int kind, int* nx, int* nu) {
    switch (kind) {
        case SCG_QUAD_1D: *nx = 2; *nu = 1; return 0;
        case SCG_QUAD_2D: *nx = 6; *nu = 2; return 0;
        case SCG_QUAD_3D: *nx = 12; *nu = 4; return 0;
        case SCG_CARTPOLE: *nx = 4; *nu = 1; return 0;
    }
    return -1;
}

template <typename real>
static int build_params(const scg_config& c, int nx, int nu, Hot<real>& h, Cold<real>& k, int* n_rows_out, int* n_slots_out, bool* lean_out) {
    std::memset(&h, 0, sizeof(h));
    std::memset(&k, 0, sizeof(k));
    const real INF = std::numeric_limits<real>::infinity();
    h.substeps = c.substeps; h.ctrl_steps = c.ctrl_steps; h.ctrl_freq = c.ctrl_freq; h.pyb_freq = c.pyb_freq;
    h.dt = real(c.pyb_dt);
    h.kf = real(c.kf); h.km = real(c.km); h.arm_s2 = real(c.arm / std::sqrt(2.0));
    h.inv_mass = real(1.0 / c.mass); h.gravity_force = real(c.gravity_force);
    for (int j = 0; j < 3; ++j) { h.J[j] = real(c.J[j]); h.J_inv[j] = real(c.J_inv[j]); }
    for (int j = 0; j < 3; ++j) { h.dJ[j] = h.J[(j + 2) % 3] - h.J[(j + 1) % 3]; h.kJ[j] = h.dt * h.J_inv[j]; }
    h.ndJ_xz[0] = -h.dJ[0]; h.ndJ_xz[1] = -h.dJ[2]; h.kJ_xz[0] = h.kJ[0]; h.kJ_xz[1] = h.kJ[2];
    h.pwm_scale = real(c.pwm2rpm_scale); h.pwm_const = real(c.pwm2rpm_const);
    h.pwm_min = real(c.pwm_min); h.pwm_max = real(c.pwm_max); h.inv_pwm_scale = real(1.0 / c.pwm2rpm_scale);
    if (c.env_kind != SCG_CARTPOLE) {
        h.inv_nkf = real(1.0 / ((4 / nu) * c.kf));
        const double rmin = c.pwm2rpm_scale * c.pwm_min + c.pwm2rpm_const, rmax = c.pwm2rpm_scale * c.pwm_max + c.pwm2rpm_const;
        h.f_min = real(c.kf * rmin * rmin); h.f_max = real(c.kf * rmax * rmax);
        h.inv_n_motor = real(1.0 / (4 / nu)); h.km_over_kf = real(c.km / c.kf); h.inv_kf = real(1.0 / c.kf);
    }
    h.gravity_acc = real(c.gravity_acc); h.cp_length = real(c.cp_length); h.cp_pole_mass = real(c.cp_pole_mass); h.cp_cart_mass = real(c.cp_cart_mass);
    h.normalized_action = c.normalized_action; h.norm_act_scale = real(c.norm_act_scale); h.hover_thrust = real(c.hover_thrust);
    for (int j = 0; j < nu; ++j) { h.act_low[j] = real(c.act_low[j]); h.act_high[j] = real(c.act_high[j]); }
    for (int d = 0; d < nx; ++d) {
        h.state_low[d] = real(c.state_low[d]); h.state_high[d] = real(c.state_high[d]);
        h.rew_state_weight[d] = real(c.rew_state_weight[d]); h.mse_weight[d] = real(c.mse_weight[d]);
        if (c.mse_weight[d] != 0.0) h.mse_mask |= 1u << d;
        h.x_goal0[d] = c.x_goal ? real(c.x_goal[d]) : real(0);
        h.init_state[d] = real(c.init_state[d]);
    }
    for (int j = 0; j < nu; ++j) { h.rew_act_weight[j] = real(c.rew_act_weight[j]); h.u_goal[j] = real(c.u_goal[j]); }
    h.done_on_out_of_bound = c.done_on_out_of_bound; h.done_on_violation = c.done_on_violation;
    h.done_on_collision = c.done_on_collision; h.done_on_completion = c.done_on_completion;
    h.task = c.task; h.cost = c.cost; h.rew_exponential = c.rew_exponential; h.obs_goal_horizon = c.obs_goal_horizon;
    h.n_goal_rows = c.n_goal_rows; h.goal_tol = real(c.goal_tolerance);
    h.use_constraint_penalty = c.use_constraint_penalty; h.constraint_penalty = real(c.constraint_penalty);
    // quadratic cost: diagonal fast path
    bool diag = true;
    for (int a = 0; a < nx; ++a) for (int b = 0; b < nx; ++b) {
        k.Q[a * SCG_MAX_NX + b] = real(c.Q[a * SCG_MAX_NX + b]);
        if (a != b && c.Q[a * SCG_MAX_NX + b] != 0.0) diag = false;
    }
    for (int a = 0; a < nu; ++a) for (int b = 0; b < nu; ++b) {
        k.R[a * SCG_MAX_NU + b] = real(c.R[a * SCG_MAX_NU + b]);
        if (a != b && c.R[a * SCG_MAX_NU + b] != 0.0) diag = false;
    }
    h.quadratic_diag = diag ? 1 : 0;
    for (int a = 0; a < nx; ++a) h.q_diag[a] = real(c.Q[a * SCG_MAX_NX + a]);
    for (int a = 0; a < nu; ++a) h.r_diag[a] = real(c.R[a * SCG_MAX_NU + a]);
    // ---- constraints: bounded objects -> per-dim tables (hot), everything else -> dense rows (cold)
    int row = 0, nb = 0, nd = 0, np = 0;
    if (c.n_constraints < 0 || c.n_constraints > SCG_MAX_CONSTRAINTS) return fail(SCG_ERR_BAD_CONFIG, "n_constraints out of range");
    for (int ci = 0; ci < c.n_constraints; ++ci) {
        const scg_constraint& s = c.constraints[ci];
        const int full = s.var == SCG_VAR_STATE ? nx : nu;
        if (s.var != SCG_VAR_STATE && s.var != SCG_VAR_INPUT) return fail(SCG_ERR_BAD_CONFIG, "constraint var must be state or input");
        if (s.n_active < 1 || s.n_active > full) return fail(SCG_ERR_BAD_CONFIG, "constraint n_active out of range");
        for (int j = 0; j < s.n_active; ++j)
            if (s.active_dims[j] < 0 || s.active_dims[j] >= full) return fail(SCG_ERR_BAD_CONFIG, "active_dim not strictly smaller than dim");
        if (s.form == SCG_CON_BOUNDED || s.form == SCG_CON_SYMMETRIC) {
            if (nb >= MAX_BOUNDED) return fail(SCG_ERR_BAD_CONFIG, "too many bounded constraints");
            BoundedCon<real>& bc = h.bounded[nb];
            for (int d = 0; d < SCG_MAX_NX; ++d) { bc.lo[d] = -INF; bc.hi[d] = INF; bc.row_lo[d] = bc.row_hi[d] = -1; }
            for (int d = 0; d < 2 * SCG_MAX_NX; ++d) k.bounded_tol[nb][d] = -INF;
            bc.var = int8_t(s.var); bc.strict = int8_t(s.strict); bc.has_tol = int8_t(s.has_tolerance);
            bc.symmetric = s.form == SCG_CON_SYMMETRIC;
            bc.mask = 0u;
            const int nrows = bc.symmetric ? s.n_active : 2 * s.n_active;
            if (s.n_rows != nrows) return fail(SCG_ERR_BAD_CONFIG, "bounded constraint n_rows mismatch");
            for (int j = 0; j < s.n_active; ++j) {
                const int d = s.active_dims[j];
                if (bc.row_lo[d] >= 0) return fail(SCG_ERR_BAD_CONFIG, "duplicates in active_dims");
                bc.lo[d] = real(s.lower[j]); bc.hi[d] = real(s.upper[j]);
                bc.mask |= 1u << d;
                bc.row_lo[d] = int8_t(row + j);
                if (!bc.symmetric) bc.row_hi[d] = int8_t(row + s.n_active + j);
                if (s.has_tolerance) {
                    k.bounded_tol[nb][d] = real(s.tolerance[j]);
                    if (!bc.symmetric) k.bounded_tol[nb][SCG_MAX_NX + d] = real(s.tolerance[s.n_active + j]);
                }
            }
            if (s.has_tolerance) h.any_tolerance = 1;
            row += nrows; ++nb;
        } else if (s.form == SCG_CON_LINEAR) {
            if (s.n_rows < 1 || s.n_rows > SCG_MAX_CON_ROWS) return fail(SCG_ERR_BAD_CONFIG, "linear constraint n_rows out of range");
            for (int r = 0; r < s.n_rows; ++r) {
                if (nd >= MAX_DENSE_ROWS) return fail(SCG_ERR_BAD_CONFIG, "too many dense constraint rows");
                DenseRow<real>& dr = k.dense[nd++];
                for (int j = 0; j < s.n_active; ++j) dr.a[s.active_dims[j]] += real(s.A[r * s.n_active + j]);
                dr.b = real(s.b[r]); dr.var = int8_t(s.var); dr.strict = int8_t(s.strict);
                dr.has_tol = int8_t(s.has_tolerance); dr.tol = s.has_tolerance ? real(s.tolerance[r]) : real(0);
                dr.quadratic = 0; dr.out_row = row + r;
            }
            if (s.has_tolerance) h.any_tolerance = 1;
            row += s.n_rows;
        } else if (s.form == SCG_CON_QUADRATIC) {
            if (nd >= MAX_DENSE_ROWS || np >= 2) return fail(SCG_ERR_BAD_CONFIG, "too many quadratic constraints");
            DenseRow<real>& dr = k.dense[nd++];
            for (int a = 0; a < s.n_active; ++a) for (int b = 0; b < s.n_active; ++b)   // P_full = F' P F
                k.P[np][s.active_dims[a] * SCG_MAX_NX + s.active_dims[b]] += real(s.A[a * s.n_active + b]);
            dr.b = real(s.b[0]); dr.var = int8_t(s.var); dr.strict = int8_t(s.strict);
            dr.has_tol = int8_t(s.has_tolerance); dr.tol = s.has_tolerance ? real(s.tolerance[0]) : real(0);
            dr.quadratic = 1; dr.p_index = np++; dr.out_row = row;
            if (s.has_tolerance) h.any_tolerance = 1;
            row += 1;
        } else return fail(SCG_ERR_BAD_CONFIG, "unknown constraint form");
    }
    if (row > SCG_MAX_ROWS) return fail(SCG_ERR_BAD_CONFIG, "too many constraint rows");
    h.n_bounded = nb; h.n_dense_rows = nd; h.n_rows = row;
    *n_rows_out = row;
    // ---- disturbances
    int slots = 0;
    const int dims[3] = {nx, nu, c.env_kind == SCG_CARTPOLE ? 2 : c.env_kind};
    for (int ch = 0; ch < 3; ++ch) {
        if (c.n_dist[ch] < 0 || c.n_dist[ch] > SCG_MAX_DIST) return fail(SCG_ERR_BAD_CONFIG, "n_dist out of range");
        for (int q = 0; q < c.n_dist[ch]; ++q) {
            const scg_disturbance& s = c.dist[ch][q];
            DistP<real>& D = k.dist[ch][q];
            if (s.kind < SCG_DIST_IMPULSE || s.kind > SCG_DIST_PERIODIC) return fail(SCG_ERR_BAD_CONFIG, "unknown disturbance kind");
            D.kind = s.kind; D.has_mask = s.has_mask; D.step_offset = s.step_offset; D.max_step = s.max_step;
            D.slot = s.slot; D.dim = dims[ch];
            if (s.slot + 1 > slots) slots = s.slot + 1;
            for (int d = 0; d < SCG_MAX_NX; ++d) {
                D.mask[d] = real(s.mask[d]); D.low[d] = real(s.low[d]); D.span[d] = real(s.high[d] - s.low[d]); D.std[d] = real(s.std[d]);
            }
            D.magnitude = real(s.magnitude); D.duration = real(s.duration); D.half_duration = real(s.duration / 2.0);
            D.decay_rate = real(s.decay_rate); D.scale = real(s.scale); D.omega = real(2.0 * M_PI * s.frequency);
            D.inject = static_cast<const real*>(s.inject);
        }
    }
    *n_slots_out = slots;
    h.n_dist_obs = c.n_dist[0]; h.n_dist_act = c.n_dist[1]; h.n_dist_dyn = c.n_dist[2];
    h.adversary_channel = c.adversary_channel;
    h.n_gates = c.n_gates; h.gate_geometry = c.gate_geometry;
    h.dyn_uses_episode_inertia = c.dyn_uses_episode_inertia; h.dyn_applies_force = c.dyn_applies_force;
    h.auto_reset = c.auto_reset; h.reseed_on_reset = c.reseed_on_reset;
    for (int d = 0; d < nx; ++d) {
        h.init_rand[d].distrib = c.init_rand[d].distrib; h.init_rand[d].a = real(c.init_rand[d].a); h.init_rand[d].b = real(c.init_rand[d].b);
        if (c.init_rand[d].distrib == SCG_RAND_NONE) h.init_rand[d].a = h.init_rand[d].b = real(0);
        else h.init_block_mask |= 1u << (d / 4);
        if (c.init_rand[d].distrib == SCG_RAND_UNIFORM) h.any_init_rand |= 1;
        if (c.init_rand[d].distrib == SCG_RAND_NORMAL) h.any_init_rand |= 2;
    }
    for (int j = 0; j < 4; ++j) {
        h.inertia_nominal[j] = real(c.inertia_nominal[j]);
        h.inertia_rand[j].distrib = c.inertia_rand[j].distrib; h.inertia_rand[j].a = real(c.inertia_rand[j].a); h.inertia_rand[j].b = real(c.inertia_rand[j].b);
        if (c.inertia_rand[j].distrib == SCG_RAND_UNIFORM) h.any_inertia_rand |= 1;
        if (c.inertia_rand[j].distrib == SCG_RAND_NORMAL) h.any_inertia_rand |= 2;
    }
    k.gate_rand.distrib = c.gate_rand.distrib; k.gate_rand.a = real(c.gate_rand.a); k.gate_rand.b = real(c.gate_rand.b);
    k.obstacle_rand.distrib = c.obstacle_rand.distrib; k.obstacle_rand.a = real(c.obstacle_rand.a); k.obstacle_rand.b = real(c.obstacle_rand.b);
    h.any_gate_rand = (c.gate_rand.distrib != SCG_RAND_NONE || c.obstacle_rand.distrib != SCG_RAND_NONE) ? 1 : 0;
    if (c.n_gates < 0 || c.n_gates > SCG_MAX_GATES || c.n_obstacles < 0 || c.n_obstacles > SCG_MAX_OBSTACLES)
        return fail(SCG_ERR_BAD_CONFIG, "too many gates / obstacles");
    for (int g = 0; g < c.n_gates; ++g) {
        for (int j = 0; j < 7; ++j) k.gates[g][j] = real(c.gates[g][j]);
        k.gates[g][7] = real(std::cos(c.gates[g][5])); k.gates[g][8] = real(std::sin(c.gates[g][5]));
    }
    for (int g = 0; g < c.n_obstacles; ++g) {
        for (int j = 0; j < 6; ++j) k.obstacles[g][j] = real(c.obstacles[g][j]);
        k.obstacles[g][6] = real(std::cos(c.obstacles[g][5])); k.obstacles[g][7] = real(std::sin(c.obstacles[g][5]));
    }
    {   // broad phase of the contact scan: 0.35 m covers every part of a gate / obstacle plus the drone; a uniform
        // pose offset widens it by sqrt(2) max|offset|, a normal one is unbounded
        auto broad2 = [](const scg_rand& r) -> double {
            if (r.distrib == SCG_RAND_NONE) return 0.35 * 0.35;
            if (r.distrib == SCG_RAND_UNIFORM) { const double m = std::max(std::fabs(r.a), std::fabs(r.b)); const double d = 0.35 + 1.4143 * m; return d * d; }
            return double(INFINITY);
        };
        k.gate_broad2 = real(broad2(c.gate_rand)); k.obstacle_broad2 = real(broad2(c.obstacle_rand));
        // the inline copy of the same test (Hot::geo_*): fp32, radius widened by 2e-4 relative so that it never rejects what the scan accepts
        h.geo_n = 0;
        for (int q = 0; q < (SCG_MAX_GATES + SCG_MAX_OBSTACLES) / 2; ++q)
            for (int w = 0; w < 2; ++w) { h.geo_negx[q][w] = -1e18f; h.geo_negy[q][w] = -1e18f; h.geo_r2[q][w] = 0.f; }   // padding: never near
        auto put = [&](double x, double y, double r2) {
            h.geo_negx[h.geo_n >> 1][h.geo_n & 1] = -float(x); h.geo_negy[h.geo_n >> 1][h.geo_n & 1] = -float(y);
            h.geo_r2[h.geo_n >> 1][h.geo_n & 1] = float(r2 * 1.0002 + 1e-6);
            ++h.geo_n;
        };
        for (int g = 0; g < c.n_gates; ++g) put(c.gates[g][0], c.gates[g][1], broad2(c.gate_rand));
        for (int g = 0; g < c.n_obstacles; ++g) put(c.obstacles[g][0], c.obstacles[g][1], broad2(c.obstacle_rand));
    }
    k.n_obstacles = c.n_obstacles;
    h.seed = c.seed; h.env_id_base_lo = uint32_t(c.env_id_base & 0xFFFFFFFFll);
    h.viol_thr = real(5e-9);
    // ---- which per-step consumers read the 3D body rates p, q, r (they need the whole rotation matrix)
    bool rates = c.env_kind != SCG_QUAD_3D;
    if (c.env_kind == SCG_QUAD_3D) {
        constexpr uint32_t RATE_BITS = 0xE00u;
        if (h.mse_mask & RATE_BITS) rates = true;
        for (int q = 0; q < nb; ++q) if (h.bounded[q].var == SCG_VAR_STATE && (h.bounded[q].mask & RATE_BITS)) rates = true;
        for (int q = 0; q < nd; ++q) if (k.dense[q].var == SCG_VAR_STATE) rates = true;
        if (c.cost == SCG_COST_RL_REWARD) for (int d = 9; d < 12; ++d) if (c.rew_state_weight[d] != 0.0) rates = true;
        if (c.cost == SCG_COST_QUADRATIC) rates = true;   // Q weighs the rates (identity by default); goal_reached norms the full state
    }
    h.need_full_state = rates ? 1 : 0;
    h.cfg = (c.task == SCG_TASK_STABILIZATION ? CFG_STAB : 0u) | (c.cost == SCG_COST_RL_REWARD ? CFG_COST_RL : 0u) |
            (c.cost == SCG_COST_QUADRATIC ? CFG_COST_QUAD : 0u) | (c.rew_exponential ? CFG_REW_EXP : 0u) |
            (c.done_on_out_of_bound ? CFG_OOB : 0u) | (c.done_on_completion ? CFG_DONE_COMPLETION : 0u) |
            (c.done_on_violation ? CFG_DONE_VIOLATION : 0u) | (c.auto_reset ? CFG_AUTO_RESET : 0u) |
            (rates ? CFG_FULL_STATE : 0u) | (c.reseed_on_reset ? CFG_RESEED : 0u) | ((h.mse_mask & 0xE00u) ? CFG_MSE_RATES : 0u) |
            ((c.gate_geometry && c.env_kind == SCG_QUAD_3D) ? CFG_GATE_GEOMETRY : 0u) |
            ((c.dyn_aero && c.env_kind == SCG_QUAD_3D) ? CFG_AERO : 0u);
    h.aero = c.env_kind == SCG_QUAD_3D ? c.dyn_aero : 0;
    for (int j = 0; j < 3; ++j) h.aero_drag[j] = real(c.aero_drag_coeff[j]);
    h.aero_gnd_coeff = real(c.aero_gnd_eff_coeff); h.aero_prop_radius = real(c.aero_prop_radius); h.aero_h_clip = real(c.aero_gnd_eff_h_clip);
    for (int j = 0; j < 4; ++j) { h.aero_prop_xy[j][0] = real(c.aero_prop_xy[j][0]); h.aero_prop_xy[j][1] = real(c.aero_prop_xy[j][1]); }
    // ---- LEAN kernel specialisation (scg_step.cuh): none of the rare features is configured
    bool symmetric = false;
    for (int q = 0; q < nb; ++q) symmetric |= h.bounded[q].symmetric != 0;
    h.need_cold_constraints = (nd > 0 || h.any_tolerance || symmetric) ? 1 : 0;
    // merged violation box: any(c_k > thr) == (max_k c_k > thr) when every bounded constraint shares one strictness
    bool same_strict = true;
    for (int d = 0; d < SCG_MAX_NX; ++d) { h.box_lo[d] = -INF; h.box_hi[d] = INF; }
    for (int d = 0; d < SCG_MAX_NU; ++d) { h.ubox_lo[d] = -INF; h.ubox_hi[d] = INF; }
    for (int q = 0; q < nb; ++q) {
        const BoundedCon<real>& bc = h.bounded[q];
        if (bc.strict != h.bounded[0].strict) same_strict = false;
        if (bc.symmetric) continue;
        const int full = bc.var == SCG_VAR_STATE ? nx : nu;
        for (int d = 0; d < full; ++d) {
            if (!((bc.mask >> d) & 1u)) continue;
            if (bc.var == SCG_VAR_STATE) {
                h.box_mask |= 1u << d;
                if (bc.lo[d] > h.box_lo[d]) h.box_lo[d] = bc.lo[d];
                if (bc.hi[d] < h.box_hi[d]) h.box_hi[d] = bc.hi[d];
            } else {
                h.ubox_mask |= 1u << d;
                if (bc.lo[d] > h.ubox_lo[d]) h.ubox_lo[d] = bc.lo[d];
                if (bc.hi[d] < h.ubox_hi[d]) h.ubox_hi[d] = bc.hi[d];
            }
        }
    }
    h.box_thr = (nb > 0 && h.bounded[0].strict) ? std::nextafter(real(-0.5), -INF) : real(0.5);
    // flat value table (every row from a non-symmetric bounded object without tolerance, one strictness)
    h.n_crows = 0; h.crow_strict = nb > 0 ? h.bounded[0].strict : 0;
    if (nb > 0 && nd == 0 && !h.any_tolerance && !symmetric && same_strict && h.n_rows <= SCG_MAX_ROWS) {
        int filled = 0;
        for (int r = 0; r < SCG_MAX_ROWS; ++r) { k.crow[r][0] = real(0); k.crow[r][1] = real(0); k.crow[r][2] = real(0); k.crow[r][3] = real(0); }
        for (int q = 0; q < nb; ++q) {
            const BoundedCon<real>& bc = h.bounded[q];
            const int full = bc.var == SCG_VAR_STATE ? nx : nu, off = bc.var == SCG_VAR_STATE ? 0 : nx;
            for (int d = 0; d < full; ++d) {
                if (!((bc.mask >> d) & 1u)) continue;
                const int rl = bc.row_lo[d], rh = bc.row_hi[d];
                k.crow[rl][0] = real(off + d); k.crow[rl][1] = real(-1); k.crow[rl][2] = bc.lo[d];
                k.crow[rh][0] = real(off + d); k.crow[rh][1] = real(1); k.crow[rh][2] = bc.hi[d];
                filled += 2;
            }
        }
        if (filled == h.n_rows) h.n_crows = h.n_rows;
    }
    if (h.n_rows > 0) h.cfg |= CFG_HAS_ROWS;
    if (h.ubox_mask) h.cfg |= CFG_UBOX;
    if (c.env_kind == SCG_QUAD_3D && !(h.mse_mask & ~0x15u)) h.cfg |= CFG_MSE_POS;
    if (c.env_kind == SCG_QUAD_3D && !(h.box_mask & ~0x15u)) h.cfg |= CFG_BOX_POS;
    if (c.env_kind == SCG_QUAD_3D && !c.normalized_action) {
        bool inside = true;
        for (int j = 0; j < nu; ++j) inside = inside && h.act_low[j] >= h.f_min && h.act_high[j] <= h.f_max;
        if (inside) h.cfg |= CFG_CLIP_ONCE;
    }
    if (h.need_cold_constraints || !same_strict) h.cfg |= CFG_COLD_CONSTRAINTS;
    if (c.n_dist[1] > 0 || c.n_dist[2] > 0 || c.adversary_channel == SCG_CH_ACTION || c.adversary_channel == SCG_CH_DYNAMICS)
        h.cfg |= CFG_DISTURBED;
    {   // inline evaluation of the action-channel disturbances (Hot::ds_*)
        bool simple = c.n_dist[1] > 0 && c.adversary_channel < 0 && !(c.n_dist[2] > 0 && c.dyn_applies_force);
        for (int q = 0; q < c.n_dist[1] && simple; ++q) {
            const scg_disturbance& s = c.dist[1][q];
            simple = (s.kind == SCG_DIST_IMPULSE || s.kind == SCG_DIST_STEP || s.kind == SCG_DIST_WHITE) && !s.has_mask && !s.inject;
        }
        h.ds_n = 0;
        if (simple) {
            h.ds_n = c.n_dist[1];
            for (int q = 0; q < c.n_dist[1]; ++q) {
                const scg_disturbance& s = c.dist[1][q];
                h.ds_kind[q] = s.kind; h.ds_off[q] = s.step_offset; h.ds_slot[q] = s.slot;
                h.ds_mag[q] = real(s.magnitude); h.ds_half[q] = real(s.duration / 2.0); h.ds_decay[q] = real(s.decay_rate);
                h.ds_l2decay[q] = real(std::log2(s.decay_rate > 0.0 ? s.decay_rate : 1.0));
                for (int j = 0; j < SCG_MAX_NU; ++j) h.ds_std[q][j] = real(s.std[j]);
            }
            h.cfg |= CFG_DIST_SIMPLE;
        }
    }
    if (!h.quadratic_diag) h.cfg |= CFG_QR_DENSE;
    if (c.cost == SCG_COST_RL_REWARD && c.use_constraint_penalty) h.cfg |= CFG_PENALTY;
    // ---- LEAN kernel (scg_step.cuh): the auto-reset can take a spare initial state drawn ahead of time, i.e. reset()
    // draws nothing but uniform initial-state offsets -- no disturbance offsets, no inertial properties, no normal
    // draws.  Per-step features (disturbances, adversary, dense constraints, ...) are uniform branches in both kernels.
    bool reset_draws = (h.any_init_rand & 2) != 0 || h.any_inertia_rand != 0 || c.dyn_uses_episode_inertia != 0;
    for (int ch = 0; ch < 3; ++ch)
        for (int q = 0; q < c.n_dist[ch]; ++q) {
            const scg_disturbance& s = c.dist[ch][q];
            if ((s.kind == SCG_DIST_IMPULSE || s.kind == SCG_DIST_STEP) && s.step_offset < 0) reset_draws = true;
        }
    if (reset_draws) h.cfg |= CFG_RESET_EXTRAS;
    {
        bool outside = false;
        for (int ch = 0; ch < 3; ch += 2)
            for (int q = 0; q < c.n_dist[ch]; ++q) {
                const scg_disturbance& s = c.dist[ch][q];
                if ((s.kind == SCG_DIST_IMPULSE || s.kind == SCG_DIST_STEP) && s.step_offset < 0) outside = true;
            }
        if (!outside) h.cfg |= CFG_SPARE_SIMPLE;
    }
    h.step_row[0] = h.dt; h.step_row[1] = h.arm_s2; h.step_row[2] = h.km_over_kf; h.step_row[3] = h.inv_mass;
    h.step_row[4] = -h.dJ[1]; h.step_row[5] = h.kJ[1]; h.step_row[6] = -(h.gravity_force * h.inv_mass); h.step_row[7] = real(0);
    for (int j = 0; j < 3; ++j) {
        h.pos_pack[j] = h.box_lo[2 * j]; h.pos_pack[4 + j] = h.box_hi[2 * j];
        h.pos_pack[8 + j] = h.x_goal0[2 * j]; h.pos_pack[12 + j] = h.mse_weight[2 * j];
    }
    h.pos_pack[3] = h.box_thr; h.pos_pack[7] = h.goal_tol * h.goal_tol; h.pos_pack[11] = h.goal_tol; h.pos_pack[15] = real(0);
    *lean_out = true;   // every config takes the spare reset; SCG_ATT_ROUNDTRIP (validation) is routed to the GENERAL kernel by the launcher
    return SCG_OK;
}

extern "C" int scg_set_rollout_queue(scg_env* env, int32_t mode, int32_t chunk_steps) {
    if (!env) return scg_fail(SCG_ERR_NULL, "scg_set_rollout_queue: null env");
    if (mode < -1 || mode > 1 || chunk_steps < 0) return scg_fail(SCG_ERR_BAD_CONFIG, "scg_set_rollout_queue: mode in {-1, 0, 1}, chunk_steps >= 0");
    env->queue_mode = mode; env->queue_chunk = chunk_steps;
    return SCG_OK;
}
extern "C" int scg_last_rollout_queue(const scg_env* env) { return env ? env->last_queue : 0; }
extern "C" int scg_last_rollout_tlog(const scg_env* env) { return env ? env->last_tlog : 0; }
extern "C" int scg_last_rollout_spec(const scg_env* env) { return env ? env->last_spec : 0; }
extern "C" int scg_set_spec_enabled(scg_env* env, int enabled) {
    if (!env) return fail(SCG_ERR_NULL, "scg_set_spec_enabled: null handle");
    env->spec_enabled = enabled ? 1 : 0;
    return SCG_OK;
}
// BenchmarkEnv.seed (benchmark_env.py:220-235): the Philox key of every later draw.  The constant block travels by value
// with each launch, so the host copy is all there is to patch.
extern "C" int scg_set_seed(scg_env* env, uint32_t seed) {
    if (!env) return scg_fail(SCG_ERR_NULL, "scg_set_seed: null env");
    env->cfg.seed = seed; env->hf.seed = seed; env->hd.seed = seed;
    return SCG_OK;
}
extern "C" const char* scg_last_error(void) { return g_err.c_str(); }
extern "C" const char* scg_version(void) { return "scg_b200 0.1 (sm_100a)"; }
extern "C" int64_t scg_launch_count(const scg_env* env) { return env ? env->launches : 0; }
extern "C" uint32_t scg_config_word(const scg_env* env) { return env ? (env->cfg.precision == 32 ? env->hf.cfg : env->hd.cfg) : 0u; }
extern "C" int scg_kernel_variant(const scg_env* env) {
    if (!env || !env->lean) return 0;
    const bool quad_att = env->cfg.env_kind == SCG_QUAD_3D || env->cfg.env_kind == SCG_QUAD_2D;
    if (quad_att && env->cfg.attitude_mode == SCG_ATT_ROUNDTRIP) return 0;   // validation mode: GENERAL kernel (scg_inst.cuh)
    if (!env->plain && env->cfg.precision == 32 && env->cfg.env_kind != SCG_QUAD_1D &&
        !(env->cfg.env_kind == SCG_QUAD_3D && env->cfg.dyn_uses_episode_inertia)) {
        // PLAIN + DS (scg_inst.cuh: ds_kernel_ok; launches with injected draws still take variant 1; closed-loop launches of a PLAIN config take the PLAIN + DS loop too)
        const uint32_t c = env->hf.cfg;
        const bool general = (c & (CFG_COLD_CONSTRAINTS | CFG_QR_DENSE | CFG_PENALTY | CFG_AERO)) || env->hf.normalized_action;
        const bool dist_ok = !(c & CFG_DISTURBED) || (c & CFG_DIST_SIMPLE);
        const bool reset_ok = !(c & CFG_RESET_EXTRAS) || ((c & CFG_SPARE_SIMPLE) && ((c & CFG_DIST_SIMPLE) || env->hf.n_dist_act == 0));
        if (!general && dist_ok && reset_ok) return 4;
    }
    if (!(env->plain && env->cfg.precision == 32)) return 1;
    return env->cfg.env_kind == SCG_QUAD_3D && (env->hf.cfg & CFG_SPEC_MASK) == CFG_SPEC_VAL ? 3 : 2;
}

template <typename real> static int upload_params(scg_env* e) {
    Cold<real>* k = new (std::nothrow) Cold<real>;
    if (!k) return fail(SCG_ERR_BAD_CONFIG, "out of host memory");
    Hot<real>& h = *reinterpret_cast<Hot<real>*>(sizeof(real) == 4 ? (void*)&e->hf : (void*)&e->hd);
    int rc = build_params<real>(e->cfg, e->nx, e->nu, h, *k, &e->n_rows, &e->n_slots, &e->lean);
    e->plain = e->lean && !(h.cfg & (CFG_DISTURBED | CFG_COLD_CONSTRAINTS | CFG_QR_DENSE | CFG_PENALTY | CFG_GATE_GEOMETRY | CFG_RESET_EXTRAS | CFG_AERO)) &&
               !h.normalized_action;
    if (rc != SCG_OK) { delete k; return rc; }
    const int rows = e->cfg.n_goal_rows > 0 ? e->cfg.n_goal_rows : 1;
    real* goal_host = new real[size_t(rows) * e->nx];
    for (int r = 0; r < rows; ++r) for (int d = 0; d < e->nx; ++d)
        goal_host[size_t(r) * e->nx + d] = e->cfg.x_goal ? real(e->cfg.x_goal[size_t(r) * e->nx + d]) : real(0);
    cudaError_t ce = cudaMalloc(&e->x_goal_dev, sizeof(real) * rows * e->nx);
    if (ce == cudaSuccess) ce = cudaMemcpy(e->x_goal_dev, goal_host, sizeof(real) * rows * e->nx, cudaMemcpyHostToDevice);
    delete[] goal_host;
    if (ce != cudaSuccess) { delete k; return cuda_fail(ce, "x_goal upload"); }
    k->x_goal = static_cast<const real*>(e->x_goal_dev);
    ce = cudaMalloc(&e->cold_dev, sizeof(Cold<real>));
    if (ce == cudaSuccess) ce = cudaMemcpy(e->cold_dev, k, sizeof(Cold<real>), cudaMemcpyHostToDevice);
    if (ce != cudaSuccess) { delete k; return cuda_fail(ce, "cold block upload"); }
    e->cold_host = k;
    return SCG_OK;
}

extern "C" int scg_create(const scg_config* cfg, scg_env** out) {
    if (!cfg || !out) return fail(SCG_ERR_NULL, "scg_create: null argument");
    int nx, nu;
    if (kind_dims(cfg->env_kind, &nx, &nu)) return fail(SCG_ERR_BAD_CONFIG, "unknown env_kind");
    if (cfg->precision != 32 && cfg->precision != 64) return fail(SCG_ERR_BAD_CONFIG, "precision must be 32 or 64");
    if (cfg->attitude_mode < SCG_ATT_ROUNDTRIP || cfg->attitude_mode > SCG_ATT_FAST_INTRINSIC) return fail(SCG_ERR_BAD_CONFIG, "bad attitude_mode");
    if (cfg->attitude_mode == SCG_ATT_FAST_INTRINSIC && cfg->precision != 32) return fail(SCG_ERR_BAD_CONFIG, "SCG_ATT_FAST_INTRINSIC is fp32 only");
    if (cfg->substeps < 1 || cfg->pyb_freq < 1 || cfg->ctrl_freq < 1 || cfg->pyb_freq % cfg->ctrl_freq != 0)
        return fail(SCG_ERR_BAD_CONFIG, "[ERROR] in BenchmarkEnv.__init__(), pyb_freq is not divisible by env_freq.");
    if (cfg->substeps != cfg->pyb_freq / cfg->ctrl_freq) return fail(SCG_ERR_BAD_CONFIG, "substeps != pyb_freq / ctrl_freq");
    if (cfg->n_goal_rows < 1 || !cfg->x_goal) return fail(SCG_ERR_BAD_CONFIG, "x_goal table missing");
    if (cfg->obs_goal_horizon < 0) return fail(SCG_ERR_BAD_CONFIG, "negative obs_goal_horizon");
    if (!(cfg->mass > 0.0)) return fail(SCG_ERR_BAD_CONFIG, "mass must be positive");
    scg_env* e = new (std::nothrow) scg_env;
    if (!e) return fail(SCG_ERR_BAD_CONFIG, "out of host memory");
    e->cfg = *cfg; e->nx = nx; e->nu = nu;
    int mul = 1;
    if (cfg->cost == SCG_COST_RL_REWARD && cfg->obs_goal_horizon > 0)
        mul = cfg->task == SCG_TASK_TRAJ_TRACKING ? 1 + cfg->obs_goal_horizon : 2;
    e->obs_dim = nx * mul;
    std::memset(&e->buf, 0, sizeof(e->buf));
    int rc = cfg->precision == 32 ? upload_params<float>(e) : upload_params<double>(e);
    if (rc != SCG_OK) { scg_destroy(e); return rc; }
    e->cfg.x_goal = nullptr;   // host table was copied; do not keep the caller's pointer
    if (const char* q = getenv("SCG_QUEUE")) { const int m = atoi(q); e->queue_mode = m < 0 ? -1 : (m > 0 ? 1 : 0); }
    if (const char* q = getenv("SCG_QUEUE_SLEEP")) { const int c = atoi(q); if (c >= 0) e->queue_sleep_ns = unsigned(c); }
    if (const char* q = getenv("SCG_QUEUE_BLOCKS_PER_SM")) { const int c = atoi(q); if (c >= 0) e->queue_blocks_per_sm = c; }
    if (const char* q = getenv("SCG_QUEUE_CHUNK")) { const int c = atoi(q); e->queue_chunk = c > 0 ? c : 0; }
    *out = e;
    return SCG_OK;
}

extern "C" int scg_destroy(scg_env* e) {
    if (!e) return SCG_OK;
    if (e->cold_dev) cudaFree(e->cold_dev);
    if (e->stats_total_dev) cudaFree(e->stats_total_dev);
    if (e->q_mem) cudaFree(e->q_mem);
    if (e->x_goal_dev) cudaFree(e->x_goal_dev);
    for (int j = 0; j < 2; ++j) {
        if (e->stage[j]) cudaFree(e->stage[j]);
        if (e->ev_copied[j]) cudaEventDestroy(e->ev_copied[j]);
        if (e->ev_consumed[j]) cudaEventDestroy(e->ev_consumed[j]);
    }
    if (e->copy_stream) cudaStreamDestroy(e->copy_stream);
    if (e->cold_host) { if (e->cfg.precision == 32) delete static_cast<Cold<float>*>(e->cold_host); else delete static_cast<Cold<double>*>(e->cold_host); }
    delete e;
    return SCG_OK;
}

extern "C" int scg_bind_buffers(scg_env* e, const scg_buffers* b) {
    if (!e || !b) return fail(SCG_ERR_NULL, "scg_bind_buffers: null argument");
    if (b->n < 1 || b->n > (int64_t(1) << 31) - 1024) return fail(SCG_ERR_SHAPE, "n out of range");
    if (!b->state || !b->ctrl_step || !b->episode || !b->flags || !b->obs || !b->reward || !b->done || !b->info)
        return fail(SCG_ERR_NULL, "scg_bind_buffers: state, ctrl_step, episode, flags, obs, reward, done, info are required");
    if (e->n_slots > 0 && !b->dist_offset) return fail(SCG_ERR_NULL, "dist_offset required when disturbances carry step offsets");
    if (e->cfg.dyn_uses_episode_inertia && !b->inertia) return fail(SCG_ERR_NULL, "inertia buffer required for dyn_uses_episode_inertia");
    if (e->cfg.env_kind == SCG_QUAD_3D && (e->cfg.dyn_aero & 1) && !b->aero_mem) return fail(SCG_ERR_NULL, "aero_mem buffer required for dyn_aero drag");
    if (e->cfg.adversary_channel >= 0 && e->cfg.adversary_channel != SCG_CH_ACTION && e->cfg.adversary_channel != SCG_CH_DYNAMICS)
        return fail(SCG_ERR_BAD_CONFIG, "adversary channel must be action or dynamics");
    e->buf = *b;
    e->bound = true;
    return SCG_OK;
}

extern "C" int scg_set_injection(scg_env* e, int channel, int k, const void* dev_ptr) {
    if (!e) return fail(SCG_ERR_NULL, "scg_set_injection: null env");
    if (channel == -1) { e->buf.adv_action = dev_ptr; return SCG_OK; }   // adversary action for the next step
    if (channel < 0 || channel > 2 || k < 0 || k >= e->cfg.n_dist[channel]) return fail(SCG_ERR_SHAPE, "no such disturbance");
    if (channel == SCG_CH_ACTION) {
        // injected draws are read by the general (out-of-line) evaluation only: the inline path is taken while no
        // action-channel disturbance has a pointer set.  The constant block travels by value with every launch.
        e->cfg.dist[channel][k].inject = dev_ptr;
        bool any = false;
        for (int q = 0; q < e->cfg.n_dist[SCG_CH_ACTION]; ++q) any = any || e->cfg.dist[SCG_CH_ACTION][q].inject != nullptr;
        const bool simple = !any && e->hf.ds_n + e->hd.ds_n > 0;
        if (simple) { e->hf.cfg |= CFG_DIST_SIMPLE; e->hd.cfg |= CFG_DIST_SIMPLE; }
        else { e->hf.cfg &= ~CFG_DIST_SIMPLE; e->hd.cfg &= ~CFG_DIST_SIMPLE; }
    }
    if (e->cfg.precision == 32) {
        Cold<float>* c = static_cast<Cold<float>*>(e->cold_host);
        c->dist[channel][k].inject = static_cast<const float*>(dev_ptr);
        const size_t off = reinterpret_cast<char*>(&c->dist[channel][k].inject) - reinterpret_cast<char*>(c);
        CU(cudaMemcpy(static_cast<char*>(e->cold_dev) + off, &c->dist[channel][k].inject, sizeof(void*), cudaMemcpyHostToDevice));
    } else {
        Cold<double>* c = static_cast<Cold<double>*>(e->cold_host);
        c->dist[channel][k].inject = static_cast<const double*>(dev_ptr);
        const size_t off = reinterpret_cast<char*>(&c->dist[channel][k].inject) - reinterpret_cast<char*>(c);
        CU(cudaMemcpy(static_cast<char*>(e->cold_dev) + off, &c->dist[channel][k].inject, sizeof(void*), cudaMemcpyHostToDevice));
    }
    return SCG_OK;
}

template <typename real>
static int dispatch_reset(scg_env* e, const ResetInject<real>& inj, const uint8_t* mask, cudaStream_t st) {
    switch (e->cfg.env_kind) {
        case SCG_QUAD_1D: return launch_reset<SCG_QUAD_1D, real>(e, inj, mask, st);
        case SCG_QUAD_2D: return launch_reset<SCG_QUAD_2D, real>(e, inj, mask, st);
        case SCG_QUAD_3D: return launch_reset<SCG_QUAD_3D, real>(e, inj, mask, st);
        case SCG_CARTPOLE: return launch_reset<SCG_CARTPOLE, real>(e, inj, mask, st);
    }
    return fail(SCG_ERR_BAD_CONFIG, "unknown env_kind");
}

template <typename real>
static int dispatch_rollout(scg_env* e, const Roll<real>& ro, cudaStream_t st) {
    switch (e->cfg.env_kind) {
        case SCG_QUAD_1D: return launch_rollout<SCG_QUAD_1D, real>(e, ro, st);
        case SCG_QUAD_2D: return launch_rollout<SCG_QUAD_2D, real>(e, ro, st);
        case SCG_QUAD_3D: return launch_rollout<SCG_QUAD_3D, real>(e, ro, st);
        case SCG_CARTPOLE: return launch_rollout<SCG_CARTPOLE, real>(e, ro, st);
    }
    return fail(SCG_ERR_BAD_CONFIG, "unknown env_kind");
}

extern "C" int scg_reset(scg_env* e, const uint8_t* mask, const void* inject_init, const void* inject_inertia,
                         const int32_t* inject_offsets, void* stream) {
    if (!e) return fail(SCG_ERR_NULL, "scg_reset: null env");
    if (!e->bound) return fail(SCG_ERR_STATE, "scg_reset: buffers not bound");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    int rc;
    if (e->cfg.precision == 32) {
        ResetInject<float> inj{(const float*)inject_init, (const float*)inject_inertia, inject_offsets};
        rc = dispatch_reset<float>(e, inj, mask, st);
    } else {
        ResetInject<double> inj{(const double*)inject_init, (const double*)inject_inertia, inject_offsets};
        rc = dispatch_reset<double>(e, inj, mask, st);
    }
    if (rc == SCG_OK) e->did_reset = true;
    return rc;
}

template <typename real>
static int run_steps_t(scg_env* e, const void* actions, int k_steps, void* reward_out, uint8_t* done_out, void* log_out,
                       int log_stride, int step_outputs, const scg_pid_config* pid, void* pid_mem, void* actions_out, cudaStream_t st) {
    Roll<real> ro{};
    ro.actions = (const real*)actions; ro.k_steps = k_steps; ro.reward_out = (real*)reward_out; ro.done_out = done_out;
    ro.log_out = (real*)log_out; ro.log_stride = log_stride; ro.step_outputs = step_outputs;
    static const bool pdl = [] { const char* v = getenv("SCG_PDL"); return !(v && atoi(v) == 0); }();
    ro.pdl = (pdl && k_steps == 1 && step_outputs) ? 1 : 0;   // scg_step: the next step's launch overlaps this one's tail
    if (pid) {
        ro.use_pid = 1; ro.pid_mem = (real*)pid_mem; ro.actions_out = (real*)actions_out;
        PidP<real>& P = ro.pid;
        for (int j = 0; j < 3; ++j) {
            P.p_for[j] = real(pid->p_coeff_for[j]); P.i_for[j] = real(pid->i_coeff_for[j]); P.d_for[j] = real(pid->d_coeff_for[j]);
            P.p_tor[j] = real(pid->p_coeff_tor[j]); P.i_tor[j] = real(pid->i_coeff_tor[j]); P.d_tor[j] = real(pid->d_coeff_tor[j]);
        }
        P.gravity = real(pid->g * pid->m); P.inv_4kf = real(1.0 / (4.0 * pid->kf)); P.kf = real(pid->kf);
        P.pwm_scale = real(pid->pwm2rpm_scale); P.pwm_const = real(pid->pwm2rpm_const); P.inv_pwm_scale = real(1.0 / pid->pwm2rpm_scale);
        P.min_pwm = real(pid->min_pwm); P.max_pwm = real(pid->max_pwm);
        const double dt = 1.0 / double(e->cfg.ctrl_freq);
        P.dt = real(dt); P.inv_dt = real(double(e->cfg.ctrl_freq));
    }
    return dispatch_rollout<real>(e, ro, st);
}

static int run_steps(scg_env* e, const void* actions, int k_steps, void* reward_out, uint8_t* done_out, void* log_out,
                     int log_stride, int step_outputs, cudaStream_t st, const scg_pid_config* pid = nullptr, void* pid_mem = nullptr,
                     void* actions_out = nullptr) {
    if (e->cfg.precision == 32)
        return run_steps_t<float>(e, actions, k_steps, reward_out, done_out, log_out, log_stride, step_outputs, pid, pid_mem, actions_out, st);
    return run_steps_t<double>(e, actions, k_steps, reward_out, done_out, log_out, log_stride, step_outputs, pid, pid_mem, actions_out, st);
}

extern "C" int scg_step(scg_env* e, const void* action, void* stream) {
    if (!e || !action) return fail(SCG_ERR_NULL, "scg_step: null argument");
    if (!e->bound) return fail(SCG_ERR_STATE, "scg_step: buffers not bound");
    if (!e->did_reset) return fail(SCG_ERR_STATE, "[ERROR] You must call env.reset() at least once before using env.step().");
    return run_steps(e, action, 1, nullptr, nullptr, nullptr, 1, 1, static_cast<cudaStream_t>(stream));
}

extern "C" int scg_rollout(scg_env* e, const scg_rollout_args* a, void* stream) {
    if (!e || !a || !a->actions) return fail(SCG_ERR_NULL, "scg_rollout: null argument");
    if (!e->bound) return fail(SCG_ERR_STATE, "scg_rollout: buffers not bound");
    if (!e->did_reset) return fail(SCG_ERR_STATE, "[ERROR] You must call env.reset() at least once before using env.step().");
    if (a->k_steps < 1) return fail(SCG_ERR_SHAPE, "k_steps must be >= 1");
    if (a->log_out && a->log_stride < 1) return fail(SCG_ERR_SHAPE, "log_stride must be >= 1");
    return run_steps(e, a->actions, a->k_steps, a->reward_out, a->done_out, a->log_out, a->log_stride > 0 ? a->log_stride : 1,
                     0, static_cast<cudaStream_t>(stream));
}

// ---------------------------------------------------------------- scg_rollout_noalloc
// The scratch of the time-chunk ticket queue (pop counter, push counter, ready list) for callers that must not allocate:
// one file-scope device array per GPU that loads the library, held until exit.  g_queue_scratch is a host-side stand-in
// in host code -- its device address comes from cudaGetSymbolAddress, once per device, with that device current.
static __device__ __align__(256) int32_t g_queue_scratch[SCG_NOALLOC_QUEUE_WORDS];
static int32_t* g_queue_scratch_ptr[64] = {nullptr};

extern "C" int scg_rollout_noalloc(scg_env* e, const scg_rollout_args* a, void* stream) {
    if (!e || !a || !a->actions) return fail(SCG_ERR_NULL, "scg_rollout_noalloc: null argument");
    if (!e->bound) return fail(SCG_ERR_STATE, "scg_rollout_noalloc: buffers not bound");
    if (!e->did_reset) return fail(SCG_ERR_STATE, "[ERROR] You must call env.reset() at least once before using env.step().");
    if (a->log_out && a->log_stride < 1) return fail(SCG_ERR_SHAPE, "log_stride must be >= 1");
    const bool quad_att = e->cfg.env_kind == SCG_QUAD_3D || e->cfg.env_kind == SCG_QUAD_2D;
    if (e->cfg.precision != 32 || (quad_att && e->cfg.attitude_mode == SCG_ATT_ROUNDTRIP))
        return fail(SCG_ERR_BAD_CONFIG, "scg_rollout_noalloc: the queued rollout kernel exists for precision 32 with a fast attitude mode only");
    if (a->k_steps < 64) return fail(SCG_ERR_SHAPE, "scg_rollout_noalloc: k_steps must be >= 64 (shorter launches take the static grid, which needs no scratch: use scg_rollout)");
    // the ranges of this launch, as launch_queued (scg_inst.cuh) will cut them
    const int64_t env_blocks = (e->buf.n + 63) / 64;
    int chunk = e->queue_chunk > 0 ? e->queue_chunk : 64;
    if (chunk > a->k_steps) chunk = a->k_steps;
    const int64_t tickets = int64_t((a->k_steps + chunk - 1) / chunk) * env_blocks;
    if (2 + tickets > SCG_NOALLOC_QUEUE_WORDS) {
        char msg[256];
        snprintf(msg, sizeof(msg), "scg_rollout_noalloc: %lld (64-env block, %d-step chunk) ranges need %lld words of queue scratch, the static array holds "
                 "%d: use fewer envs / steps per launch, a longer chunk (scg_set_rollout_queue), or scg_rollout", (long long)tickets, chunk,
                 (long long)(2 + tickets), SCG_NOALLOC_QUEUE_WORDS);
        return fail(SCG_ERR_SHAPE, msg);
    }
    int dev = 0;
    CU(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64) return fail(SCG_ERR_BAD_CONFIG, "scg_rollout_noalloc: device ordinal out of range");
    if (!g_queue_scratch_ptr[dev]) {
        int32_t* p = nullptr;
        CU(cudaGetSymbolAddress((void**)&p, g_queue_scratch));
        g_queue_scratch_ptr[dev] = p;
    }
    // For this call the handle's queue scratch IS the static array, large enough by the check above, so the launcher neither
    // frees nor allocates; it clears the words it uses (cudaMemsetAsync on `stream`) before the kernel reads them.
    int32_t* const keep_mem = e->q_mem; const int64_t keep_cap = e->q_cap; const int keep_mode = e->queue_mode;
    e->q_mem = g_queue_scratch_ptr[dev]; e->q_cap = SCG_NOALLOC_QUEUE_WORDS; e->queue_mode = 1;
    const int rc = run_steps(e, a->actions, a->k_steps, a->reward_out, a->done_out, a->log_out, a->log_stride > 0 ? a->log_stride : 1, 0,
                             static_cast<cudaStream_t>(stream));
    const bool queued = e->last_queue == 1 && e->q_mem == g_queue_scratch_ptr[dev];
    e->q_mem = keep_mem; e->q_cap = keep_cap; e->queue_mode = keep_mode;
    if (rc != SCG_OK) return rc;
    if (!queued) return fail(SCG_ERR_STATE, "scg_rollout_noalloc: the launcher did not take the queued kernel (the rollout ran on the static grid, which uses no scratch)");
    return SCG_OK;
}

// Closed loop: K fused control steps whose actions come from the on-device DSL PID tracking the env's X_GOAL
// (quadrotor_utils.py:70-278; competition_utils.py:338-356 / controllers/pid/pid.py:97-141).
extern "C" int scg_rollout_pid(scg_env* e, const scg_pid_config* pid, void* pid_mem, const scg_rollout_args* a, void* actions_out,
                               void* stream) {
    if (!e || !pid || !pid_mem || !a) return fail(SCG_ERR_NULL, "scg_rollout_pid: null argument");
    if (!e->bound) return fail(SCG_ERR_STATE, "scg_rollout_pid: buffers not bound");
    if (!e->did_reset) return fail(SCG_ERR_STATE, "[ERROR] You must call env.reset() at least once before using env.step().");
    if (e->cfg.env_kind != SCG_QUAD_3D && e->cfg.env_kind != SCG_QUAD_2D)
        return fail(SCG_ERR_BAD_CONFIG, "scg_rollout_pid: the PID drives quad_type 2 and 3");
    if (a->k_steps < 1) return fail(SCG_ERR_SHAPE, "k_steps must be >= 1");
    if (a->log_out && a->log_stride < 1) return fail(SCG_ERR_SHAPE, "log_stride must be >= 1");
    return run_steps(e, nullptr, a->k_steps, a->reward_out, a->done_out, a->log_out, a->log_stride > 0 ? a->log_stride : 1, 0,
                     static_cast<cudaStream_t>(stream), pid, pid_mem, actions_out);
}

// Actions in pinned host memory: chunked H2D copies on a private stream, double-buffered against
// the rollout launches on the caller's stream.
extern "C" int scg_rollout_host(scg_env* e, const void* host_actions, int32_t k_steps, int32_t chunk_steps, void* stream) {
    if (!e || !host_actions) return fail(SCG_ERR_NULL, "scg_rollout_host: null argument");
    if (!e->bound) return fail(SCG_ERR_STATE, "scg_rollout_host: buffers not bound");
    if (!e->did_reset) return fail(SCG_ERR_STATE, "[ERROR] You must call env.reset() at least once before using env.step().");
    if (k_steps < 1 || chunk_steps < 1) return fail(SCG_ERR_SHAPE, "k_steps and chunk_steps must be >= 1");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const size_t real_bytes = e->cfg.precision == 32 ? 4 : 8;
    const size_t step_bytes = size_t(e->buf.n) * e->nu * real_bytes;
    const size_t need = step_bytes * chunk_steps;
    if (!e->copy_stream) {
        CU(cudaStreamCreateWithFlags(&e->copy_stream, cudaStreamNonBlocking));
        for (int j = 0; j < 2; ++j) {
            CU(cudaEventCreateWithFlags(&e->ev_copied[j], cudaEventDisableTiming));
            CU(cudaEventCreateWithFlags(&e->ev_consumed[j], cudaEventDisableTiming));
        }
    }
    if (need > e->stage_bytes) {
        CU(cudaStreamSynchronize(e->copy_stream));
        for (int j = 0; j < 2; ++j) {
            if (e->stage_used[j]) CU(cudaEventSynchronize(e->ev_consumed[j]));   // the old buffer may still be read
            if (e->stage[j]) CU(cudaFree(e->stage[j]));
            CU(cudaMalloc(&e->stage[j], need));
            e->stage_used[j] = false;
        }
        e->stage_bytes = need;
    }
    const char* src = static_cast<const char*>(host_actions);
    int chunk = 0;
    for (int k0 = 0; k0 < k_steps; k0 += chunk_steps, ++chunk) {
        const int ks = k_steps - k0 < chunk_steps ? k_steps - k0 : chunk_steps;
        const int s = chunk & 1;
        // a slot is rewritten only after its last consumer -- of this call or of an earlier one -- has run
        if (e->stage_used[s]) CU(cudaStreamWaitEvent(e->copy_stream, e->ev_consumed[s], 0));
        CU(cudaMemcpyAsync(e->stage[s], src + size_t(k0) * step_bytes, step_bytes * ks, cudaMemcpyHostToDevice, e->copy_stream));
        CU(cudaEventRecord(e->ev_copied[s], e->copy_stream));
        CU(cudaStreamWaitEvent(st, e->ev_copied[s], 0));
        int rc = run_steps(e, e->stage[s], ks, nullptr, nullptr, nullptr, 1, 0, st);
        // recorded on failure too: whatever was enqueued on `st` so far orders the next copy into this slot
        const cudaError_t ce = cudaEventRecord(e->ev_consumed[s], st);
        e->stage_used[s] = true;
        if (rc != SCG_OK) return rc;
        CU(ce);
    }
    return SCG_OK;
}

extern "C" int scg_stats_reset(scg_env* e, void* stream) {
    if (!e || !e->bound || !e->buf.stats) return fail(SCG_ERR_NULL, "scg_stats_reset: no stats buffer bound");
    CU(cudaMemsetAsync(e->buf.stats, 0, sizeof(double) * SCG_STAT_REPLICAS * SCG_STATS_LEN, static_cast<cudaStream_t>(stream)));
    return SCG_OK;
}

static __global__ void k_stats_fold(const double* __restrict__ rep, double* __restrict__ total) {
    double s = 0.0;
    for (int r = 0; r < SCG_STAT_REPLICAS; ++r) s += rep[r * SCG_STATS_LEN + threadIdx.x];
    total[threadIdx.x] = s;
}

extern "C" int scg_stats_read(scg_env* e, double* host_out, void* stream) {
    if (!e || !host_out || !e->bound || !e->buf.stats) return fail(SCG_ERR_NULL, "scg_stats_read: no stats buffer bound");
    if (!e->stats_total_dev) CU(cudaMalloc(&e->stats_total_dev, sizeof(double) * SCG_STATS_LEN));
    k_stats_fold<<<1, SCG_STATS_LEN, 0, static_cast<cudaStream_t>(stream)>>>(e->buf.stats, e->stats_total_dev);
    CU(cudaGetLastError());
    CU(cudaMemcpyAsync(host_out, e->stats_total_dev, sizeof(double) * SCG_STATS_LEN, cudaMemcpyDeviceToHost, static_cast<cudaStream_t>(stream)));
    return SCG_OK;
}

template <typename real> static int constraints_eval_t(scg_env* e, const void* state, const void* action, int64_t n,
                                                       void* out_values, uint8_t* out_flags, cudaStream_t st) {
    const real* s_ = (const real*)state; const real* a_ = (const real*)action; real* o_ = (real*)out_values;
    switch (e->cfg.env_kind) {
        case SCG_QUAD_1D: return launch_constraints<SCG_QUAD_1D, real>(e, s_, a_, n, o_, out_flags, st);
        case SCG_QUAD_2D: return launch_constraints<SCG_QUAD_2D, real>(e, s_, a_, n, o_, out_flags, st);
        case SCG_QUAD_3D: return launch_constraints<SCG_QUAD_3D, real>(e, s_, a_, n, o_, out_flags, st);
        case SCG_CARTPOLE: return launch_constraints<SCG_CARTPOLE, real>(e, s_, a_, n, o_, out_flags, st);
    }
    return fail(SCG_ERR_BAD_CONFIG, "unknown env_kind");
}

extern "C" int scg_constraints_eval(scg_env* e, const void* state, const void* action, int64_t n, void* out_values,
                                    uint8_t* out_flags, void* stream) {
    if (!e || !state) return fail(SCG_ERR_NULL, "scg_constraints_eval: null argument");
    if (n < 1) return fail(SCG_ERR_SHAPE, "n must be >= 1");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    return e->cfg.precision == 32 ? constraints_eval_t<float>(e, state, action, n, out_values, out_flags, st)
                                  : constraints_eval_t<double>(e, state, action, n, out_values, out_flags, st);
}

static int symbolic_dispatch(scg_env* e, int op, const void* x, const void* u, const void* xr, const void* ur,
                             const double* Q, const double* R, int64_t n, double dt, void* out0, void* out1, void* out2,
                             cudaStream_t st) {
    int rc = e->cfg.precision == 32
                 ? symbolic_launch<float>(e->cfg, op, (const float*)x, (const float*)u, (const float*)xr, (const float*)ur, Q, R, n,
                                          float(dt), (float*)out0, (float*)out1, (float*)out2, st)
                 : symbolic_launch<double>(e->cfg, op, (const double*)x, (const double*)u, (const double*)xr, (const double*)ur, Q, R,
                                           n, dt, (double*)out0, (double*)out1, (double*)out2, st);
    if (rc == -2) return fail(SCG_ERR_NULL, "scg_symbolic_eval: SCG_SYM_DF needs out1 (df/du)");
    if (rc != 0) return fail(SCG_ERR_BAD_CONFIG, "symbolic op not available for this env kind");
    e->launches += 1;
    CU(cudaGetLastError());
    return SCG_OK;
}

extern "C" int scg_symbolic_eval(scg_env* e, int op, const void* x, const void* u, int64_t n, double dt, void* out0,
                                 void* out1, void* stream) {
    if (!e || !x || !u || !out0) return fail(SCG_ERR_NULL, "scg_symbolic_eval: null argument");
    if (n < 1) return fail(SCG_ERR_SHAPE, "n must be >= 1");
    if (op < SCG_SYM_FC || op > SCG_SYM_FD_EULER) return fail(SCG_ERR_BAD_CONFIG, "unknown symbolic op");
    // SCG_SYM_LOSS through this entry point: the handle's Q, R and U_GOAL, Xr = 0, out1 = l_x, no l_u
    return symbolic_dispatch(e, op, x, u, nullptr, nullptr, nullptr, nullptr, n, dt, out0, out1, nullptr,
                             static_cast<cudaStream_t>(stream));
}

extern "C" int scg_symbolic_loss(scg_env* e, const void* x, const void* u, const void* xr, const void* ur, const double* Q,
                                 const double* R, int64_t n, void* out_l, void* out_lx, void* out_lu, void* stream) {
    if (!e || !x || !u || !out_l) return fail(SCG_ERR_NULL, "scg_symbolic_loss: null argument");
    if (n < 1) return fail(SCG_ERR_SHAPE, "n must be >= 1");
    return symbolic_dispatch(e, SCG_SYM_LOSS, x, u, xr, ur, Q, R, n, 0.0, out_l, out_lx, out_lu, static_cast<cudaStream_t>(stream));
}

// ---------------------------------------------------------------- fp32 peak microbenchmark
__global__ void __launch_bounds__(256) k_fma_peak(float* out, int iters, float a, float b) {
    float v[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) v[j] = float(threadIdx.x + j) * 1e-3f;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int r = 0; r < 16; ++r) {
#pragma unroll
            for (int j = 0; j < 8; ++j) v[j] = fmaf(v[j], a, b);
        }
    }
    float s = 0.f;
#pragma unroll
    for (int j = 0; j < 8; ++j) s += v[j];
    if (s == 12345.678f) out[0] = s;   // never true; keeps the chain alive
}

extern "C" int scg_measure_fp32_peak(double* tflops_out) {
    if (!tflops_out) return fail(SCG_ERR_NULL, "scg_measure_fp32_peak: null argument");
    int dev = 0, sms = 0;
    CU(cudaGetDevice(&dev));
    CU(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    float* out = nullptr;
    CU(cudaMalloc(&out, sizeof(float)));
    cudaEvent_t e0, e1;
    CU(cudaEventCreate(&e0)); CU(cudaEventCreate(&e1));
    const int blocks = sms * 8, threads = 256, iters = 4096;
    double best = 0.0;
    for (int rep = 0; rep < 5; ++rep) {
        CU(cudaEventRecord(e0));
        k_fma_peak<<<blocks, threads>>>(out, iters, 0.999f, 1e-3f);
        CU(cudaEventRecord(e1));
        CU(cudaEventSynchronize(e1));
        float ms = 0.f;
        CU(cudaEventElapsedTime(&ms, e0, e1));
        const double flops = 2.0 * 8 * 16 * double(iters) * blocks * threads;
        const double tf = flops / (ms * 1e-3) / 1e12;
        if (rep > 0 && tf > best) best = tf;
    }
    cudaEventDestroy(e0); cudaEventDestroy(e1); cudaFree(out);
    *tflops_out = best;
    return SCG_OK;
}
