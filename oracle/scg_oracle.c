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
