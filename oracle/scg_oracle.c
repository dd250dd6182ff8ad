/*
 * scg_oracle.c -- plain-C (fp64) restatement of the reference Quadrotor-3D physics=dyn control
 * step, batched over envs with OpenMP.  TEST INFRASTRUCTURE ONLY (see oracle/__init__.py): it
 * is the native CPU baseline of bench.py (`cpu_baseline`, `--impl reference`) and the large-N
 * checker of tests/; the product never links or calls it.
 *
 * Follows, in /root/reference/safe_control_gym/:
 *   envs/gym_pybullet_drones/quadrotor.py:815-842   _preprocess_control
 *   envs/gym_pybullet_drones/quadrotor_utils.py:21-67  cmd2pwm / pwm2rpm
 *   envs/gym_pybullet_drones/base_aviary.py:266-322,522-576  _advance_simulation / _dynamics
 *       (TIMESTEP := PYB_TIMESTEP)
 *   envs/gym_pybullet_drones/quadrotor.py:844-1002,1117-1130  observation, reward (competition),
 *       done, goal hold
 *   envs/benchmark_env.py:422-463  after_step (bounded constraints, violation flag, time limit)
 *   envs/constraints.py:106-141,273-304  value rounding and violation test
 *   envs/env_wrappers/vectorized_env/dummy_vec_env.py:38-49  auto-reset
 * PyBullet's quaternion helpers are restated from bullet3 (oracle/bullet_math.py).
 * Parity: pinned against the numpy oracle (itself pinned on the reference's own code) in
 * tests/test_oracle_c.py.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

typedef struct {
    int32_t substeps, ctrl_steps, ctrl_freq;
    double dt, mass, gravity_force, J[3], J_inv[3], kf, km, arm;
    double pwm_scale, pwm_const, pwm_min, pwm_max;
    double act_low[4], act_high[4];
    int32_t n_bounded;              /* bounded constraint objects */
    int32_t var[4], strict[4];      /* 0 state / 1 input */
    double lo[4][12], hi[4][12];    /* per full dim, +-inf when inactive */
    double goal[12], goal_tol;
    int32_t n_gates, done_on_violation, done_on_completion, auto_reset;
    double init_state[12], init_lo[12], init_hi[12];   /* uniform additive randomisation (lo == hi == 0: none) */
    uint32_t seed;
    int64_t env_id_base;
} oracle_cfg;

/* ---- Philox4x32-10 (oracle/philox.py) */
static void philox(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t k0, uint32_t k1, uint32_t out[4]) {
    for (int r = 0; r < 10; ++r) {
        uint64_t p0 = (uint64_t)0xD2511F53u * c0, p1 = (uint64_t)0xCD9E8D57u * c2;
        uint32_t n0 = (uint32_t)(p1 >> 32) ^ c1 ^ k0, n2 = (uint32_t)(p0 >> 32) ^ c3 ^ k1;
        c1 = (uint32_t)p1; c3 = (uint32_t)p0; c0 = n0; c2 = n2;
        k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
    }
    out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}
static double u01(uint32_t w) { return ((double)(w >> 8) + 0.5) * 5.9604644775390625e-08; }

/* ---- bullet3 helpers */
static void quat_from_euler(const double rpy[3], double q[4]) {
    double sph = sin(rpy[0] / 2.0), cph = cos(rpy[0] / 2.0), sth = sin(rpy[1] / 2.0), cth = cos(rpy[1] / 2.0);
    double sps = sin(rpy[2] / 2.0), cps = cos(rpy[2] / 2.0);
    q[0] = sph * cth * cps - cph * sth * sps;
    q[1] = cph * sth * cps + sph * cth * sps;
    q[2] = cph * cth * sps - sph * sth * cps;
    q[3] = cph * cth * cps + sph * sth * sps;
    double len = sqrt(q[0] * q[0] + q[1] * q[1] + q[2] * q[2] + q[3] * q[3]);
    q[0] /= len; q[1] /= len; q[2] /= len; q[3] /= len;
}
static void euler_from_quat(const double q[4], double rpy[3]) {
    double x = q[0], y = q[1], z = q[2], w = q[3];
    double sarg = -2.0 * (x * z - w * y);
    if (sarg <= -0.99999) { rpy[0] = 0; rpy[1] = -0.5 * M_PI; rpy[2] = 2 * atan2(x, -y); }
    else if (sarg >= 0.99999) { rpy[0] = 0; rpy[1] = 0.5 * M_PI; rpy[2] = 2 * atan2(-x, y); }
    else {
        rpy[0] = atan2(2 * (y * z + w * x), w * w - x * x - y * y + z * z);
        rpy[1] = asin(sarg);
        rpy[2] = atan2(2 * (x * y + w * z), w * w + x * x - y * y - z * z);
    }
}
static void matrix_from_quat(const double q[4], double R[9]) {
    double x = q[0], y = q[1], z = q[2], w = q[3];
    double s = 2.0 / (x * x + y * y + z * z + w * w);
    double xs = x * s, ys = y * s, zs = z * s;
    double wx = w * xs, wy = w * ys, wz = w * zs, xx = x * xs, xy = x * ys, xz = x * zs, yy = y * ys, yz = y * zs, zz = z * zs;
    R[0] = 1.0 - (yy + zz); R[1] = xy - wz; R[2] = xz + wy;
    R[3] = xy + wz; R[4] = 1.0 - (xx + zz); R[5] = yz - wx;
    R[6] = xz - wy; R[7] = yz + wx; R[8] = 1.0 - (xx + yy);
}

typedef struct { double pos[3], vel[3], pre[3], rates[3], rpy[3], q[4]; } env_t;

static void derive(env_t* e) { quat_from_euler(e->pre, e->q); euler_from_quat(e->q, e->rpy); }

static void make_state(const env_t* e, const double ang[3], double st[12]) {
    double R[9];
    matrix_from_quat(e->q, R);
    st[0] = e->pos[0]; st[1] = e->vel[0]; st[2] = e->pos[1]; st[3] = e->vel[1]; st[4] = e->pos[2]; st[5] = e->vel[2];
    st[6] = e->rpy[0]; st[7] = e->rpy[1]; st[8] = e->rpy[2];
    st[9] = (R[0] * ang[0] + R[3] * ang[1]) + R[6] * ang[2];
    st[10] = (R[1] * ang[0] + R[4] * ang[1]) + R[7] * ang[2];
    st[11] = (R[2] * ang[0] + R[5] * ang[1]) + R[8] * ang[2];
}

static void reset_env(const oracle_cfg* c, env_t* e, uint32_t env_id, int32_t* resets, int32_t* t, uint32_t* flags, double st[12]) {
    uint32_t ep = (uint32_t)(*resets);
    *resets += 1; *t = 0; *flags &= 1u;
    double init[12];
    for (int b = 0; b < 3; ++b) {
        uint32_t w[4];
        philox(0u, ep, 0u, (uint32_t)b, c->seed, env_id, w);
        for (int j = 0; j < 4; ++j) {
            int d = 4 * b + j;
            init[d] = c->init_state[d];
            if (c->init_lo[d] != 0.0 || c->init_hi[d] != 0.0) init[d] += c->init_lo[d] + (c->init_hi[d] - c->init_lo[d]) * u01(w[j]);
        }
    }
    e->pos[0] = init[0]; e->vel[0] = init[1]; e->pos[1] = init[2]; e->vel[1] = init[3]; e->pos[2] = init[4]; e->vel[2] = init[5];
    e->pre[0] = init[6]; e->pre[1] = init[7]; e->pre[2] = init[8];
    e->rates[0] = e->rates[1] = e->rates[2] = 0.0;
    derive(e);
    double ang[3] = {init[9], init[10], init[11]};
    make_state(e, ang, st);
}

static double round8(double c) { return rint(c * 1e8) / 1e8; }

/* One Quadrotor.step(); returns done, writes reward / violation. */
static int step_env(const oracle_cfg* c, env_t* e, const double raw[4], int32_t* t, uint32_t* flags, double st[12],
                    double* reward, int* violated, int* truncated) {
    double rpm[4], f[4], zt[4];
    for (int j = 0; j < 4; ++j) {
        double thr = fmin(fmax(raw[j], c->act_low[j]), c->act_high[j]);
        thr = fmax(thr, 0.0);
        double pwm = (sqrt(thr / 1 / c->kf) - c->pwm_const) / c->pwm_scale;
        pwm = fmin(fmax(pwm, c->pwm_min), c->pwm_max);
        rpm[j] = c->pwm_scale * pwm + c->pwm_const;
        f[j] = rpm[j] * rpm[j] * c->kf; zt[j] = rpm[j] * rpm[j] * c->km;
    }
    const double T = ((f[0] + f[1]) + f[2]) + f[3];
    const double tz = ((zt[0] - zt[1]) + zt[2]) - zt[3];
    const double tx = (((f[0] + f[1]) - f[2]) - f[3]) * (c->arm / sqrt(2.0));
    const double ty = (((-f[0] + f[1]) + f[2]) - f[3]) * (c->arm / sqrt(2.0));
    for (int s = 0; s < c->substeps; ++s) {
        double R[9];
        matrix_from_quat(e->q, R);
        const double fw[3] = {R[2] * T, R[5] * T, R[8] * T - c->gravity_force};
        const double* w = e->rates;
        const double jw[3] = {c->J[0] * w[0], c->J[1] * w[1], c->J[2] * w[2]};
        const double tq[3] = {tx - (w[1] * jw[2] - w[2] * jw[1]), ty - (w[2] * jw[0] - w[0] * jw[2]), tz - (w[0] * jw[1] - w[1] * jw[0])};
        for (int a = 0; a < 3; ++a) e->vel[a] += c->dt * (fw[a] / c->mass);
        for (int a = 0; a < 3; ++a) e->rates[a] += c->dt * (c->J_inv[a] * tq[a]);
        for (int a = 0; a < 3; ++a) e->pos[a] += c->dt * e->vel[a];
        for (int a = 0; a < 3; ++a) e->pre[a] = e->rpy[a] + c->dt * e->rates[a];
        derive(e);
    }
    make_state(e, e->rates, st);
    /* goal hold (all gates passed only; with gates present and no gate geometry it never is) */
    int at_goal = 0, completed = (*flags & 8u) != 0;
    if (c->n_gates == 0) {
        uint32_t hold = *flags >> 16;
        double dx = st[0] - c->goal[0], dy = st[2] - c->goal[2], dz = st[4] - c->goal[4];
        if (sqrt(dx * dx + dy * dy + dz * dz) < c->goal_tol) { at_goal = 1; if (hold < 0xFFFFu) hold += 1; } else hold = 0;
        if (hold > (uint32_t)(2 * c->ctrl_freq)) completed = 1;
        *flags = (*flags & 0xFFFFu) | (hold << 16);
    }
    *flags = (*flags & ~(2u | 8u)) | (at_goal ? 2u : 0u) | (completed ? 8u : 0u);
    int done = 0;
    if (c->done_on_completion && completed) done = 1;
    *reward = 0.0;
    if (at_goal) *reward += 100.0;
    if (*flags & 1u) *reward -= 100.0;   /* previous step's violation (F8) */
    *t += 1;
    int v = 0;
    for (int k = 0; k < c->n_bounded; ++k) {
        const int nd = c->var[k] == 0 ? 12 : 4;
        for (int d = 0; d < nd; ++d) {
            const double x = c->var[k] == 0 ? st[d] : raw[d];
            const double clo = round8(c->lo[k][d] - x), chi = round8(x - c->hi[k][d]);
            v |= c->strict[k] ? (clo >= 0.0 || chi >= 0.0) : (clo > 0.0 || chi > 0.0);
        }
    }
    *violated = v;
    if (v) { *flags |= 1u; if (c->done_on_violation) done = 1; } else *flags &= ~1u;
    *truncated = 0;
    if (*t >= c->ctrl_steps) { *truncated = !done; done = 1; }
    return done;
}

/* state layout [12][n] identical to the CUDA library's (angles pre-round-trip, rates = rpy_rates).
 * stats: [0] episodes [1] return [2] length [3] violation steps [9] env steps. */
int oracle_quad3d_rollout(const oracle_cfg* c, int64_t n, int32_t k_steps, const double* actions, double* state,
                          int32_t* ctrl_step, int32_t* episode, uint32_t* flags, double* obs_out, double* reward_out,
                          uint8_t* done_out, double* stats, int32_t threads) {
    double s_eps = 0, s_ret = 0, s_len = 0, s_viol = 0;
#ifdef _OPENMP
    if (threads > 0) omp_set_num_threads(threads);
#endif
#pragma omp parallel for schedule(static) reduction(+ : s_eps, s_ret, s_len, s_viol)
    for (int64_t i = 0; i < n; ++i) {
        env_t e;
        e.pos[0] = state[0 * n + i]; e.vel[0] = state[1 * n + i]; e.pos[1] = state[2 * n + i]; e.vel[1] = state[3 * n + i];
        e.pos[2] = state[4 * n + i]; e.vel[2] = state[5 * n + i];
        for (int a = 0; a < 3; ++a) { e.pre[a] = state[(6 + a) * n + i]; e.rates[a] = state[(9 + a) * n + i]; }
        derive(&e);
        int32_t t = ctrl_step[i], resets = episode[i];
        uint32_t fl = flags[i];
        double st[12], ep_ret = 0.0, reward = 0.0;
        int32_t ep_len = 0;
        int done = 0;
        for (int k = 0; k < k_steps; ++k) {
            int viol, trunc;
            done = step_env(c, &e, actions + ((int64_t)k * n + i) * 4, &t, &fl, st, &reward, &viol, &trunc);
            ep_ret += reward; ep_len += 1; s_viol += viol;
            if (reward_out) reward_out[(int64_t)k * n + i] = reward;
            if (done_out) done_out[(int64_t)k * n + i] = (uint8_t)done;
            if (done) {
                s_eps += 1; s_ret += ep_ret; s_len += ep_len;
                if (c->auto_reset) { reset_env(c, &e, (uint32_t)(c->env_id_base + i), &resets, &t, &fl, st); ep_ret = 0; ep_len = 0; }
            }
        }
        state[0 * n + i] = e.pos[0]; state[1 * n + i] = e.vel[0]; state[2 * n + i] = e.pos[1]; state[3 * n + i] = e.vel[1];
        state[4 * n + i] = e.pos[2]; state[5 * n + i] = e.vel[2];
        for (int a = 0; a < 3; ++a) { state[(6 + a) * n + i] = e.pre[a]; state[(9 + a) * n + i] = e.rates[a]; }
        ctrl_step[i] = t; episode[i] = resets; flags[i] = fl;
        if (obs_out) for (int d = 0; d < 12; ++d) obs_out[d * n + i] = st[d];
    }
    if (stats) { stats[0] += s_eps; stats[1] += s_ret; stats[2] += s_len; stats[3] += s_viol; stats[9] += (double)n * k_steps; }
    return 0;
}

int oracle_quad3d_reset(const oracle_cfg* c, int64_t n, double* state, int32_t* ctrl_step, int32_t* episode,
                        uint32_t* flags, double* obs_out) {
#pragma omp parallel for schedule(static)
    for (int64_t i = 0; i < n; ++i) {
        env_t e;
        double st[12];
        int32_t t = 0, resets = episode[i];
        uint32_t fl = flags[i];
        reset_env(c, &e, (uint32_t)(c->env_id_base + i), &resets, &t, &fl, st);
        state[0 * n + i] = e.pos[0]; state[1 * n + i] = e.vel[0]; state[2 * n + i] = e.pos[1]; state[3 * n + i] = e.vel[1];
        state[4 * n + i] = e.pos[2]; state[5 * n + i] = e.vel[2];
        for (int a = 0; a < 3; ++a) { state[(6 + a) * n + i] = e.pre[a]; state[(9 + a) * n + i] = e.rates[a]; }
        ctrl_step[i] = t; episode[i] = resets; flags[i] = fl;
        if (obs_out) for (int d = 0; d < 12; ++d) obs_out[d * n + i] = st[d];
    }
    return 0;
}

int oracle_num_threads(void) {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}

/* ======================================================================================================
 * BASELINE config 3: the cart-pole step (restates oracle/cartpole_oracle.py, the contract inferred from
 * tests/scripts/verbose_api.yaml:1-91 -- the env is absent from the reference fork, SURVEY.md F2: parity
 * unpinned).  cost "quadratic" with Q = R = I, one white-noise action disturbance (disturbances.py:215-248) drawn
 * from Philox, default input + state constraints, RK4 at the simulation rate (mpc_utils.py:79-101), out-of-bound
 * and violation termination, auto-reset with uniform initial-state / inertial-property draws.
 * ====================================================================================================== */
typedef struct {
    int32_t substeps, ctrl_steps, task_tracking, n_goal_rows;
    double dt, gravity;
    double act_low, act_high;            /* action space == default input constraint */
    double state_low[4], state_high[4];  /* default state constraint (float32 Box bounds promoted) */
    double x_threshold, theta_threshold; /* done_on_out_of_bound */
    double noise_std;                    /* white noise on the action channel (0: none) */
    double goal_tol;
    int32_t done_on_violation, done_on_out_of_bound, auto_reset, randomized_inertia;
    double init_state[4], init_lo[4], init_hi[4];
    double prop[3], prop_lo[3], prop_hi[3];   /* pole_length, pole_mass, cart_mass: nominal + uniform additive offsets */
    const double* x_goal;                /* [n_goal_rows][4] */
    uint32_t seed;
    int64_t env_id_base;
} cartpole_cfg;

static void cp_f(const double x[4], double F, double l, double mp, double mc, double g, double dx[4]) {
    const double s = sin(x[2]), c = cos(x[2]);
    const double Mm = mc + mp, ml = mp * l;
    const double tmp = (F + ml * x[3] * x[3] * s) / Mm;
    const double thdd = (g * s - c * tmp) / (l * (4.0 / 3.0 - mp * c * c / Mm));
    dx[0] = x[1]; dx[1] = tmp - ml * thdd * c / Mm; dx[2] = x[3]; dx[3] = thdd;
}
static void cp_rk4(double x[4], double F, double dt, const double p[3], double g) {
    double k1[4], k2[4], k3[4], k4[4], y[4];
    cp_f(x, F, p[0], p[1], p[2], g, k1);
    for (int d = 0; d < 4; ++d) y[d] = x[d] + dt * 0.5 * k1[d];
    cp_f(y, F, p[0], p[1], p[2], g, k2);
    for (int d = 0; d < 4; ++d) y[d] = x[d] + dt * 0.5 * k2[d];
    cp_f(y, F, p[0], p[1], p[2], g, k3);
    for (int d = 0; d < 4; ++d) y[d] = x[d] + dt * k3[d];
    cp_f(y, F, p[0], p[1], p[2], g, k4);
    for (int d = 0; d < 4; ++d) x[d] = x[d] + dt / 6.0 * (k1[d] + 2.0 * k2[d] + 2.0 * k3[d] + k4[d]);
}
static void cp_reset(const cartpole_cfg* c, uint32_t env_id, int32_t* resets, int32_t* t, double x[4], double p[3]) {
    const uint32_t ep = (uint32_t)(*resets);
    *resets += 1; *t = 0;
    uint32_t w[4];
    philox(0u, ep, 1u, 0u, c->seed, env_id, w);   /* STREAM_RESET_INERTIA: slot j = word j */
    for (int j = 0; j < 3; ++j) {
        p[j] = c->prop[j];
        if (c->randomized_inertia) p[j] += c->prop_lo[j] + (c->prop_hi[j] - c->prop_lo[j]) * u01(w[j]);
    }
    philox(0u, ep, 0u, 0u, c->seed, env_id, w);   /* STREAM_RESET_INIT */
    for (int d = 0; d < 4; ++d) {
        x[d] = c->init_state[d];
        if (c->init_lo[d] != 0.0 || c->init_hi[d] != 0.0) x[d] += c->init_lo[d] + (c->init_hi[d] - c->init_lo[d]) * u01(w[d]);
    }
}

/* state [4][n], inertia [3][n] (this episode's pole_length, pole_mass, cart_mass), actions [k][n].
 * stats: [0] episodes [1] return [2] length [3] violation steps [9] env steps. */
int oracle_cartpole_rollout(const cartpole_cfg* c, int64_t n, int32_t k_steps, const double* actions, double* state, double* inertia,
                            int32_t* ctrl_step, int32_t* episode, uint32_t* flags, double* reward_out, uint8_t* done_out,
                            double* stats, int32_t threads) {
    double s_eps = 0, s_ret = 0, s_len = 0, s_viol = 0;
#ifdef _OPENMP
    if (threads > 0) omp_set_num_threads(threads);
#endif
#pragma omp parallel for schedule(static) reduction(+ : s_eps, s_ret, s_len, s_viol)
    for (int64_t i = 0; i < n; ++i) {
        const uint32_t env_id = (uint32_t)(c->env_id_base + i);
        double x[4], p[3], ep_ret = 0.0;
        for (int d = 0; d < 4; ++d) x[d] = state[d * n + i];
        for (int j = 0; j < 3; ++j) p[j] = inertia[j * n + i];
        int32_t t = ctrl_step[i], resets = episode[i], ep_len = 0;
        uint32_t fl = flags[i];
        for (int k = 0; k < k_steps; ++k) {
            const double raw = actions[(int64_t)k * n + i];
            const double pre = fmin(fmax(raw, c->act_low), c->act_high);
            double F = pre;
            if (c->noise_std > 0.0) {   /* normals(1, index = t, episode, STREAM_DISTURB_BASE + 8 * CHANNEL_ACTION + 0)[0] */
                uint32_t w[4];
                philox((uint32_t)t, (uint32_t)(resets - 1), 16u + 8u * 1u + 0u, 0u, c->seed, env_id, w);
                F += c->noise_std * (sqrt(-2.0 * log(u01(w[0]))) * cos(2.0 * M_PI * u01(w[1])));
            }
            for (int s = 0; s < c->substeps; ++s) cp_rk4(x, F, c->dt, p, c->gravity);
            int done = 0;
            const double* goal = c->x_goal + 4 * (c->task_tracking ? t : 0);   /* X_GOAL[ctrl_step_counter], un-clamped (pre-increment) */
            double ex2 = 0.0;
            for (int d = 0; d < 4; ++d) ex2 += (x[d] - goal[d]) * (x[d] - goal[d]);
            if (!c->task_tracking && sqrt(ex2) < c->goal_tol) done = 1;
            if (c->done_on_out_of_bound && (x[0] < -c->x_threshold || x[0] > c->x_threshold || x[2] < -c->theta_threshold || x[2] > c->theta_threshold)) done = 1;
            const double reward = -(0.5 * ex2 + 0.5 * pre * pre);
            t += 1;
            int v = round8(c->act_low - raw) > 0.0 || round8(raw - c->act_high) > 0.0;
            for (int d = 0; d < 4; ++d) v |= round8(c->state_low[d] - x[d]) > 0.0 || round8(x[d] - c->state_high[d]) > 0.0;
            if (v) { fl |= 1u; if (c->done_on_violation) done = 1; } else fl &= ~1u;
            if (t >= c->ctrl_steps) done = 1;
            ep_ret += reward; ep_len += 1; s_viol += v;
            if (reward_out) reward_out[(int64_t)k * n + i] = reward;
            if (done_out) done_out[(int64_t)k * n + i] = (uint8_t)done;
            if (done) {
                s_eps += 1; s_ret += ep_ret; s_len += ep_len;
                if (c->auto_reset) { cp_reset(c, env_id, &resets, &t, x, p); ep_ret = 0; ep_len = 0; }
            }
        }
        for (int d = 0; d < 4; ++d) state[d * n + i] = x[d];
        for (int j = 0; j < 3; ++j) inertia[j * n + i] = p[j];
        ctrl_step[i] = t; episode[i] = resets; flags[i] = fl;
    }
    if (stats) { stats[0] += s_eps; stats[1] += s_ret; stats[2] += s_len; stats[3] += s_viol; stats[9] += (double)n * k_steps; }
    return 0;
}

int oracle_cartpole_reset(const cartpole_cfg* c, int64_t n, double* state, double* inertia, int32_t* ctrl_step, int32_t* episode) {
#pragma omp parallel for schedule(static)
    for (int64_t i = 0; i < n; ++i) {
        double x[4], p[3];
        int32_t t = 0, resets = episode[i];
        cp_reset(c, (uint32_t)(c->env_id_base + i), &resets, &t, x, p);
        for (int d = 0; d < 4; ++d) state[d * n + i] = x[d];
        for (int j = 0; j < 3; ++j) inertia[j * n + i] = p[j];
        ctrl_step[i] = t; episode[i] = resets;
    }
    return 0;
}

/* ======================================================================================================
 * BASELINE configs 4 and 5: the quadrotor step with the extras those configs switch on, as a scalar port.
 *   quad_type 2 (quadrotor.py:618-621 state order [x, x_dot, z, z_dot, theta, theta_dot]; the same 3-D _dynamics
 *     with roll = yaw = 0, motor map [p0, p1, p1, p0], quadrotor_utils.py:36-44),
 *   impulse / step disturbances on the action channel with per-episode random offsets (disturbances.py:89-184),
 *   per-episode mass / inertia that reach the integrator (dyn_uses_episode_inertia extension),
 *   cost rl_reward (quadrotor.py:894-913) or competition, out-of-bound termination (:979-993),
 *   3-D: the analytic gate-pass / contact scan of oracle/gate_geometry.py (extension, F5) and the Logger record
 *     (Logger.py:50-66) written per env and control step.
 * Restates oracle/quad_oracle.py (pinned on the reference's own code); checked against it in tests/test_oracle_c.py.
 * ====================================================================================================== */
#define EXT_MAX_DIST 4
typedef struct {
    oracle_cfg base;                       /* substeps ... env_id_base as above (act_low / act_high: first nu entries) */
    int32_t quad_type;                     /* 2 or 3 */
    int32_t cost_rl, rew_exponential, done_on_out_of_bound, done_on_collision;
    double rew_state_weight[12], rew_act_weight[4], u_goal[4], mse_weight[12];
    double state_low[12], state_high[12];  /* state_space bounds (out-of-bound test) */
    /* action-channel disturbances, YAML order */
    int32_t n_dist, dist_kind[EXT_MAX_DIST] /* 0 impulse, 1 step */, dist_offset[EXT_MAX_DIST] /* >= 0 fixed, -1 random */, dist_slot[EXT_MAX_DIST], max_step;
    double dist_magnitude[EXT_MAX_DIST], dist_duration[EXT_MAX_DIST], dist_decay[EXT_MAX_DIST];
    /* per-episode inertial properties: M, Ixx, Iyy, Izz = nominal + U(lo, hi) (lo == hi == 0: not randomised) */
    int32_t dyn_uses_episode_inertia;
    double prop[4], prop_lo[4], prop_hi[4], gravity_acc;
    /* gates / obstacles (3-D, gate_geometry extension) */
    int32_t gate_geometry, n_obstacles;
    double gates[8][7], obstacles[8][6], pose_lo[2], pose_hi[2];   /* pose_*[0] gates, [1] obstacles: uniform offsets of x, y, yaw */
    int32_t pose_rand[2];
} ext_cfg;

typedef struct { double m, g_force, J[3], J_inv[3]; int32_t off[EXT_MAX_DIST]; double gpose[8][3], opose[8][3]; } episode_t;

static int seg_hits_cyl(const double a[3], const double b[3], double r, double hl) {
    const double d0 = b[0] - a[0], d1 = b[1] - a[1], d2 = b[2] - a[2];
    double t0 = 0.0, t1 = 1.0;
    if (d2 == 0.0) { if (fabs(a[2]) > hl) return 0; }
    else { const double ta = (-hl - a[2]) / d2, tb = (hl - a[2]) / d2; t0 = fmax(t0, fmin(ta, tb)); t1 = fmin(t1, fmax(ta, tb)); }
    const double qa = d0 * d0 + d1 * d1, qb = a[0] * d0 + a[1] * d1, qc = a[0] * a[0] + a[1] * a[1] - r * r;
    if (qa == 0.0) { if (qc > 0.0) return 0; }
    else {
        const double disc = qb * qb - qa * qc;
        if (disc < 0.0) return 0;
        const double s = sqrt(disc);
        t0 = fmax(t0, (-qb - s) / qa); t1 = fmin(t1, (-qb + s) / qa);
    }
    return t0 <= t1 && t0 < 0.9999;
}
static int sph_box(const double p[3], double cx, double cy, double cz, double hx, double hy, double hz) {
    const double q0 = fmax(fabs(p[0] - cx) - hx, 0.0), q1 = fmax(fabs(p[1] - cy) - hy, 0.0), q2 = fmax(fabs(p[2] - cz) - hz, 0.0);
    return q0 * q0 + q1 * q1 + q2 * q2 < 0.06 * 0.06;
}
static int sph_vcyl(const double p[3], double cz, double r, double hl) {
    const double dxy = fmax(hypot(p[0], p[1]) - r, 0.0), dz = fmax(fabs(p[2] - cz) - hl, 0.0);
    return dxy * dxy + dz * dz < 0.06 * 0.06;
}
static void to_local(const double pos[3], double x, double y, double z, double yaw, double p[3]) {
    const double c = cos(yaw), s = sin(yaw), dx = pos[0] - x, dy = pos[1] - y;
    p[0] = c * dx + s * dy; p[1] = -s * dx + c * dy; p[2] = pos[2] - z;
}
/* bit 0: a ray of the current gate hit the drone; bit 1: contact (oracle/gate_geometry.py) */
static unsigned gate_scan_c(const ext_cfg* c, const episode_t* ep, const double pos[3], const double R[9], int cur_gate, int check_gate) {
    int passed = 0, collided = 0;
    const int ng = c->base.n_gates;
    if (check_gate && cur_gate < ng) {
        const double* G = c->gates[cur_gate];
        const double gx = G[0] + ep->gpose[cur_gate][0], gy = G[1] + ep->gpose[cur_gate][1], rot = G[5] + ep->gpose[cur_gate][2];
        const double height = G[6] == 0.0 ? 1.0 : 0.525, dx = 0.05 * cos(rot), dy = 0.05 * sin(rot);
        for (int i = -3; i <= 3 && !passed; ++i) {
            const double fr[3] = {gx + i * dx - pos[0], gy + i * dy - pos[1], height - 0.1875 - pos[2]};
            const double to[3] = {fr[0], fr[1], height + 0.1875 - pos[2]};
            double a[3], b[3];
            for (int r = 0; r < 3; ++r) { a[r] = R[r] * fr[0] + R[3 + r] * fr[1] + R[6 + r] * fr[2]; b[r] = R[r] * to[0] + R[3 + r] * to[1] + R[6 + r] * to[2]; }
            passed |= seg_hits_cyl(a, b, 0.06, 0.0125);
        }
    }
    for (int g = 0; g < ng && !collided; ++g) {
        const double* G = c->gates[g];
        const double gx = G[0] + ep->gpose[g][0], gy = G[1] + ep->gpose[g][1], height = G[6] == 0.0 ? 1.0 : 0.525;
        if ((pos[0] - gx) * (pos[0] - gx) + (pos[1] - gy) * (pos[1] - gy) > 0.35 * 0.35) continue;
        double p[3];
        to_local(pos, gx, gy, G[2] + height, G[5] + ep->gpose[g][2], p);
        int hit = sph_box(p, 0, 0, 0.225, 0.25, 0.025, 0.025) || sph_box(p, 0, 0, -0.225, 0.25, 0.025, 0.025) ||
                  sph_box(p, 0.225, 0, 0, 0.025, 0.025, 0.25) || sph_box(p, -0.225, 0, 0, 0.025, 0.025, 0.25);
        hit = hit || (G[6] == 0.0 ? sph_vcyl(p, -0.6, 0.05, 0.4) : sph_box(p, 0, 0, -0.4, 0.075, 0.075, 0.125));
        collided |= hit;
    }
    for (int o = 0; o < c->n_obstacles && !collided; ++o) {
        const double* O = c->obstacles[o];
        const double ox = O[0] + ep->opose[o][0], oy = O[1] + ep->opose[o][1];
        if ((pos[0] - ox) * (pos[0] - ox) + (pos[1] - oy) * (pos[1] - oy) > 0.35 * 0.35) continue;
        double p[3];
        to_local(pos, ox, oy, O[2] + 0.525, O[5] + ep->opose[o][2], p);
        collided |= sph_vcyl(p, 0.0, 0.05, 0.4) || sph_box(p, 0, 0, -0.4, 0.075, 0.075, 0.125);
    }
    if (!collided) {
        const double support = 0.0125 * fabs(R[8]) + 0.06 * sqrt(fmax(0.0, 1.0 - R[8] * R[8]));
        collided = pos[2] - support < 0.0;
    }
    return (passed ? 1u : 0u) | (collided ? 2u : 0u);
}

/* everything reset() draws for episode `epi` besides the initial state (counter-based: can be re-derived at any time) */
static void ext_episode_draws(const ext_cfg* c, episode_t* ep, uint32_t env_id, uint32_t epi) {
    const oracle_cfg* b = &c->base;
    uint32_t w[4];
    for (int k = 0; k < c->n_dist; ++k) {   /* DisturbanceList.reset: randint(max_step) (disturbances.py:122-126, 169-172) */
        ep->off[k] = c->dist_offset[k];
        if (c->dist_offset[k] < 0) {
            philox(0u, epi, 2u, (uint32_t)(c->dist_slot[k] >> 2), b->seed, env_id, w);
            ep->off[k] = (int32_t)(((uint64_t)w[c->dist_slot[k] & 3] * (uint64_t)c->max_step) >> 32);
        }
    }
    for (int o = 0; o < c->n_obstacles; ++o) {   /* quadrotor.py:371-389 */
        ep->opose[o][0] = ep->opose[o][1] = ep->opose[o][2] = 0.0;
        if (c->pose_rand[1]) { philox(0u, epi, 4u, (uint32_t)o, b->seed, env_id, w); for (int j = 0; j < 3; ++j) ep->opose[o][j] = c->pose_lo[1] + (c->pose_hi[1] - c->pose_lo[1]) * u01(w[j]); }
    }
    for (int g = 0; g < b->n_gates; ++g) {       /* quadrotor.py:398-421 */
        ep->gpose[g][0] = ep->gpose[g][1] = ep->gpose[g][2] = 0.0;
        if (c->pose_rand[0]) { philox(0u, epi, 3u, (uint32_t)g, b->seed, env_id, w); for (int j = 0; j < 3; ++j) ep->gpose[g][j] = c->pose_lo[0] + (c->pose_hi[0] - c->pose_lo[0]) * u01(w[j]); }
    }
    double prop[4];                               /* quadrotor.py:447-467 */
    philox(0u, epi, 1u, 0u, b->seed, env_id, w);
    for (int j = 0; j < 4; ++j) {
        prop[j] = c->prop[j];
        if (c->prop_lo[j] != 0.0 || c->prop_hi[j] != 0.0) prop[j] += c->prop_lo[j] + (c->prop_hi[j] - c->prop_lo[j]) * u01(w[j]);
    }
    if (c->dyn_uses_episode_inertia) {
        ep->m = prop[0]; ep->g_force = c->gravity_acc * prop[0];
        for (int j = 0; j < 3; ++j) { ep->J[j] = prop[1 + j]; ep->J_inv[j] = 1.0 / prop[1 + j]; }
    } else {
        ep->m = b->mass; ep->g_force = b->gravity_force;
        for (int j = 0; j < 3; ++j) { ep->J[j] = b->J[j]; ep->J_inv[j] = b->J_inv[j]; }
    }
}

static void ext_reset(const ext_cfg* c, env_t* e, episode_t* ep, uint32_t env_id, int32_t* resets, int32_t* t, uint32_t* flags,
                      double ang0[3]) {
    const oracle_cfg* b = &c->base;
    const uint32_t epi = (uint32_t)(*resets);
    *resets += 1; *t = 0; *flags &= 1u;   /* cnstr_violation survives reset (benchmark_env.py:218) */
    ext_episode_draws(c, ep, env_id, epi);
    uint32_t w[4];
    const int nx = c->quad_type == 2 ? 6 : 12;
    double init[12];
    for (int blk = 0; blk < 3; ++blk) {
        philox(0u, epi, 0u, (uint32_t)blk, b->seed, env_id, w);
        for (int j = 0; j < 4; ++j) {
            const int d = 4 * blk + j;
            if (d >= nx) continue;
            init[d] = b->init_state[d];
            if (b->init_lo[d] != 0.0 || b->init_hi[d] != 0.0) init[d] += b->init_lo[d] + (b->init_hi[d] - b->init_lo[d]) * u01(w[j]);
        }
    }
    if (c->quad_type == 2) {
        e->pos[0] = init[0]; e->pos[1] = 0.0; e->pos[2] = init[2]; e->vel[0] = init[1]; e->vel[1] = 0.0; e->vel[2] = init[3];
        e->pre[0] = 0.0; e->pre[1] = init[4]; e->pre[2] = 0.0;
        ang0[0] = 0.0; ang0[1] = init[5]; ang0[2] = 0.0;
    } else {
        e->pos[0] = init[0]; e->vel[0] = init[1]; e->pos[1] = init[2]; e->vel[1] = init[3]; e->pos[2] = init[4]; e->vel[2] = init[5];
        e->pre[0] = init[6]; e->pre[1] = init[7]; e->pre[2] = init[8];
        ang0[0] = init[9]; ang0[1] = init[10]; ang0[2] = init[11];
    }
    e->rates[0] = e->rates[1] = e->rates[2] = 0.0;   /* rpy_rates restart at 0 (base_aviary.py:232-233) */
    derive(e);
}

static void ext_state(const ext_cfg* c, const env_t* e, const double ang[3], double st[12]) {
    if (c->quad_type == 2) { st[0] = e->pos[0]; st[1] = e->vel[0]; st[2] = e->pos[2]; st[3] = e->vel[2]; st[4] = e->rpy[1]; st[5] = ang[1]; }
    else make_state(e, ang, st);
}

/* One Quadrotor.step(); returns done.  info bits: 1 violation, 4 collision, 8 gate passed. */
static int ext_step(const ext_cfg* c, env_t* e, const episode_t* ep, const double* raw, int32_t* t, uint32_t* flags, double st[12],
                    double* reward, unsigned* info, double rpm_out[4]) {
    const oracle_cfg* b = &c->base;
    const int nu = c->quad_type == 2 ? 2 : 4, nx = c->quad_type == 2 ? 6 : 12, n_motor = 4 / nu;
    double pre[4], thr[4], rpm[4], f[4], zt[4];
    for (int j = 0; j < nu; ++j) { pre[j] = fmin(fmax(raw[j], b->act_low[j]), b->act_high[j]); thr[j] = pre[j]; }
    for (int k = 0; k < c->n_dist; ++k) {   /* disturbances.py:128-143, 174-184; t = ctrl_step_counter (pre-increment) */
        double noise = 0.0;
        if (*t >= ep->off[k]) {
            if (c->dist_kind[k] == 1) noise = c->dist_magnitude[k];
            else {
                const int peak = (int)(ep->off[k] + c->dist_duration[k] / 2.0), kk = abs(*t - peak);
                noise = kk < c->dist_duration[k] / 2.0 ? c->dist_magnitude[k] * pow(c->dist_decay[k], kk) : 0.0;
            }
        }
        for (int j = 0; j < nu; ++j) thr[j] += noise;
    }
    double pwm[4];
    for (int j = 0; j < nu; ++j) pwm[j] = (sqrt(fmax(thr[j], 0.0) / n_motor / b->kf) - b->pwm_const) / b->pwm_scale;
    if (nu == 2) { pwm[2] = pwm[1]; pwm[3] = pwm[0]; }   /* [p0, p1, p1, p0] */
    for (int j = 0; j < 4; ++j) {
        rpm[j] = b->pwm_scale * fmin(fmax(pwm[j], b->pwm_min), b->pwm_max) + b->pwm_const;
        f[j] = rpm[j] * rpm[j] * b->kf; zt[j] = rpm[j] * rpm[j] * b->km; rpm_out[j] = rpm[j];
    }
    const double T = ((f[0] + f[1]) + f[2]) + f[3], tz = ((zt[0] - zt[1]) + zt[2]) - zt[3];
    const double tx = (((f[0] + f[1]) - f[2]) - f[3]) * (b->arm / sqrt(2.0)), ty = (((-f[0] + f[1]) + f[2]) - f[3]) * (b->arm / sqrt(2.0));
    double R[9];
    for (int s = 0; s < b->substeps; ++s) {
        matrix_from_quat(e->q, R);
        const double fw[3] = {R[2] * T, R[5] * T, R[8] * T - ep->g_force};
        const double* w = e->rates;
        const double jw[3] = {ep->J[0] * w[0], ep->J[1] * w[1], ep->J[2] * w[2]};
        const double tq[3] = {tx - (w[1] * jw[2] - w[2] * jw[1]), ty - (w[2] * jw[0] - w[0] * jw[2]), tz - (w[0] * jw[1] - w[1] * jw[0])};
        for (int a = 0; a < 3; ++a) e->vel[a] += b->dt * (fw[a] / ep->m);
        for (int a = 0; a < 3; ++a) e->rates[a] += b->dt * (ep->J_inv[a] * tq[a]);
        for (int a = 0; a < 3; ++a) e->pos[a] += b->dt * e->vel[a];
        for (int a = 0; a < 3; ++a) e->pre[a] = e->rpy[a] + b->dt * e->rates[a];
        derive(e);
    }
    ext_state(c, e, e->rates, st);
    /* _get_info: contact scan, gate progress, goal hold (quadrotor.py:1046-1130) */
    unsigned gate = (*flags >> 8) & 0xFFu, hold = *flags >> 16;
    int collided = 0, passed = 0, at_goal = 0, completed = (*flags & 8u) != 0;
    if (c->gate_geometry && c->quad_type == 3) {
        matrix_from_quat(e->q, R);
        const unsigned bits = gate_scan_c(c, ep, e->pos, R, (int)gate, 2 * (*t) > b->ctrl_freq && (int)gate < b->n_gates);
        if (bits & 1u) { gate += 1; passed = 1; }
        collided = (bits & 2u) != 0;
    }
    if (c->quad_type == 3 && (int)gate == b->n_gates) {
        const double dx = st[0] - b->goal[0], dy = st[2] - b->goal[2], dz = st[4] - b->goal[4];
        if (sqrt(dx * dx + dy * dy + dz * dz) < b->goal_tol) { at_goal = 1; if (hold < 0xFFFFu) hold += 1; } else hold = 0;
        if (hold > (unsigned)(2 * b->ctrl_freq)) completed = 1;
    }
    int done = 0;
    if (c->done_on_out_of_bound) {
        for (int d = 0; d < nx; ++d) {
            const int masked = c->quad_type == 3 ? (d == 0 || d == 2 || d == 4 || d == 6 || d == 7 || d == 8) : ((d & 1) == 0);
            if (masked && (st[d] < c->state_low[d] || st[d] > c->state_high[d])) done = 1;
        }
    }
    if (c->done_on_collision && collided) done = 1;
    if (b->done_on_completion && completed) done = 1;
    if (c->cost_rl) {
        double dist = 0.0;
        for (int d = 0; d < nx; ++d) dist += c->rew_state_weight[d] * (st[d] - b->goal[d]) * (st[d] - b->goal[d]);
        for (int j = 0; j < nu; ++j) dist += c->rew_act_weight[j] * (pre[j] - c->u_goal[j]) * (pre[j] - c->u_goal[j]);
        *reward = c->rew_exponential ? exp(-dist) : -dist;
    } else {
        *reward = 0.0;
        if (passed) *reward += 100.0;
        if (at_goal) *reward += 100.0;
        if (collided) *reward -= 1000.0;
        if (*flags & 1u) *reward -= 100.0;   /* previous step's violation (F8) */
    }
    *t += 1;
    int v = 0;
    for (int k = 0; k < b->n_bounded; ++k) {
        const int nd = b->var[k] == 0 ? nx : nu;
        for (int d = 0; d < nd; ++d) {
            const double x = b->var[k] == 0 ? st[d] : raw[d];
            const double clo = round8(b->lo[k][d] - x), chi = round8(x - b->hi[k][d]);
            v |= b->strict[k] ? (clo >= 0.0 || chi >= 0.0) : (clo > 0.0 || chi > 0.0);
        }
    }
    if (v && b->done_on_violation) done = 1;
    *flags = (v ? 1u : 0u) | (at_goal ? 2u : 0u) | (completed ? 8u : 0u) | (gate << 8) | (hold << 16);
    if (*t >= b->ctrl_steps) done = 1;
    *info = (v ? 1u : 0u) | (collided ? 4u : 0u) | (passed ? 8u : 0u);
    return done;
}

/* state [nx][n] (internal layout of the CUDA library: angles pre-round-trip, rate slots = rpy_rates), actions [k][n][nu],
 * log_out (optional) [ceil(k / log_stride)][16][n].  stats: [0] episodes [1] return [2] length [3] violation steps [5] gates
 * [6] collisions [9] env steps. */
int oracle_quad_ext_rollout(const ext_cfg* c, int64_t n, int32_t k_steps, const double* actions, double* state, int32_t* ctrl_step,
                            int32_t* episode, uint32_t* flags, double* obs_out, double* reward_out, uint8_t* done_out, double* log_out,
                            int32_t log_stride, double* stats, int32_t threads) {
    const int nu = c->quad_type == 2 ? 2 : 4, nx = c->quad_type == 2 ? 6 : 12;
    double s_eps = 0, s_ret = 0, s_len = 0, s_viol = 0, s_gates = 0, s_coll = 0;
#ifdef _OPENMP
    if (threads > 0) omp_set_num_threads(threads);
#endif
#pragma omp parallel for schedule(static) reduction(+ : s_eps, s_ret, s_len, s_viol, s_gates, s_coll)
    for (int64_t i = 0; i < n; ++i) {
        const uint32_t env_id = (uint32_t)(c->base.env_id_base + i);
        env_t e;
        episode_t ep;
        memset(&e, 0, sizeof(e));
        if (c->quad_type == 2) {
            e.pos[0] = state[0 * n + i]; e.vel[0] = state[1 * n + i]; e.pos[2] = state[2 * n + i]; e.vel[2] = state[3 * n + i];
            e.pre[1] = state[4 * n + i]; e.rates[1] = state[5 * n + i];
        } else {
            e.pos[0] = state[0 * n + i]; e.vel[0] = state[1 * n + i]; e.pos[1] = state[2 * n + i]; e.vel[1] = state[3 * n + i];
            e.pos[2] = state[4 * n + i]; e.vel[2] = state[5 * n + i];
            for (int a = 0; a < 3; ++a) { e.pre[a] = state[(6 + a) * n + i]; e.rates[a] = state[(9 + a) * n + i]; }
        }
        derive(&e);
        int32_t t = ctrl_step[i], resets = episode[i], ep_len = 0;
        uint32_t fl = flags[i];
        ext_episode_draws(c, &ep, env_id, (uint32_t)(resets - 1));
        double st[12], ep_ret = 0.0, reward = 0.0, rpm[4];
        for (int k = 0; k < k_steps; ++k) {
            unsigned info;
            const int done = ext_step(c, &e, &ep, actions + ((int64_t)k * n + i) * nu, &t, &fl, st, &reward, &info, rpm);
            ep_ret += reward; ep_len += 1; s_viol += info & 1u; s_coll += (info >> 2) & 1u; s_gates += (info >> 3) & 1u;
            if (reward_out) reward_out[(int64_t)k * n + i] = reward;
            if (done_out) done_out[(int64_t)k * n + i] = (uint8_t)done;
            if (log_out && k % log_stride == 0) {   /* Logger.py:50-66: pos3 vel3 rpy3 ang_vel3 rpm4 */
                double* L = log_out + (int64_t)(k / log_stride) * 16 * n + i;
                for (int a = 0; a < 3; ++a) { L[a * n] = e.pos[a]; L[(3 + a) * n] = e.vel[a]; L[(6 + a) * n] = e.rpy[a]; L[(9 + a) * n] = e.rates[a]; }
                for (int j = 0; j < 4; ++j) L[(12 + j) * n] = rpm[j];
            }
            if (done) {
                s_eps += 1; s_ret += ep_ret; s_len += ep_len;
                if (c->base.auto_reset) {
                    double ang0[3];
                    ext_reset(c, &e, &ep, env_id, &resets, &t, &fl, ang0);
                    ext_state(c, &e, ang0, st);
                    ep_ret = 0; ep_len = 0;
                }
            }
        }
        if (c->quad_type == 2) {
            state[0 * n + i] = e.pos[0]; state[1 * n + i] = e.vel[0]; state[2 * n + i] = e.pos[2]; state[3 * n + i] = e.vel[2];
            state[4 * n + i] = e.pre[1]; state[5 * n + i] = e.rates[1];
        } else {
            state[0 * n + i] = e.pos[0]; state[1 * n + i] = e.vel[0]; state[2 * n + i] = e.pos[1]; state[3 * n + i] = e.vel[1];
            state[4 * n + i] = e.pos[2]; state[5 * n + i] = e.vel[2];
            for (int a = 0; a < 3; ++a) { state[(6 + a) * n + i] = e.pre[a]; state[(9 + a) * n + i] = e.rates[a]; }
        }
        ctrl_step[i] = t; episode[i] = resets; flags[i] = fl;
        if (obs_out) for (int d = 0; d < nx; ++d) obs_out[d * n + i] = st[d];
    }
    if (stats) { stats[0] += s_eps; stats[1] += s_ret; stats[2] += s_len; stats[3] += s_viol; stats[5] += s_gates; stats[6] += s_coll; stats[9] += (double)n * k_steps; }
    return 0;
}

int oracle_quad_ext_reset(const ext_cfg* c, int64_t n, double* state, int32_t* ctrl_step, int32_t* episode, uint32_t* flags, double* obs_out) {
    const int nx = c->quad_type == 2 ? 6 : 12;
#pragma omp parallel for schedule(static)
    for (int64_t i = 0; i < n; ++i) {
        env_t e;
        episode_t ep;
        double st[12], ang0[3];
        int32_t t = 0, resets = episode[i];
        uint32_t fl = flags[i];
        memset(&e, 0, sizeof(e));
        ext_reset(c, &e, &ep, (uint32_t)(c->base.env_id_base + i), &resets, &t, &fl, ang0);
        ext_state(c, &e, ang0, st);
        if (c->quad_type == 2) {
            state[0 * n + i] = e.pos[0]; state[1 * n + i] = e.vel[0]; state[2 * n + i] = e.pos[2]; state[3 * n + i] = e.vel[2];
            state[4 * n + i] = e.pre[1]; state[5 * n + i] = e.rates[1];
        } else {
            state[0 * n + i] = e.pos[0]; state[1 * n + i] = e.vel[0]; state[2 * n + i] = e.pos[1]; state[3 * n + i] = e.vel[1];
            state[4 * n + i] = e.pos[2]; state[5 * n + i] = e.vel[2];
            for (int a = 0; a < 3; ++a) { state[(6 + a) * n + i] = e.pre[a]; state[(9 + a) * n + i] = e.rates[a]; }
        }
        ctrl_step[i] = t; episode[i] = resets; flags[i] = fl;
        if (obs_out) for (int d = 0; d < nx; ++d) obs_out[d * n + i] = st[d];
    }
    return 0;
}
