// scg_symbolic.cu -- batched evaluation of the prior model that SymbolicModel exposes
// (math_and_models/symbolic_systems.py:63-123; Quadrotor._setup_symbolic, quadrotor.py:581-697):
//   fc_func(x, u)          continuous dynamics x_dot = f(x, u)                      SCG_SYM_FC
//   fd (RK4 / Euler)       one step of mpc_utils.py:79-101 / :54-55 over dt          SCG_SYM_FD_RK4 / _EULER
//   df_func(x, u)          Jacobians df/dx [nx*nx][n] (row-major), df/du [nx*nu][n]  SCG_SYM_DF
//   loss(x,u,Xr,Ur,Q,R)    l [n], l_x [nx][n], l_u [nu][n]                           SCG_SYM_LOSS
// One thread = one sample; x [nx][n], u [nu][n] struct-of-arrays.  The CasADi expression graph of the
// reference is plain arithmetic; the Jacobians here are its closed-form derivatives.  CasADi's CVODES
// fd_func (adaptive BDF) is not reproduced: consumers discretise with RK4.
#include "scg_host.cuh"
#include "scg_symbolic.cuh"

namespace scg {

template <typename real> struct SymP {
    real m, g, arm_s2, Ixx, Iyy, Izz, gamma;      // quadrotor: MASS, GRAVITY_ACC, L / sqrt(2), J diag, KM / KF
    real cp_l, cp_mp, cp_mc;                      // cart-pole: effective pole length, pole mass, cart mass
};

template <int KIND, typename real>
SCG_DEV void sym_f(const SymP<real>& p, const real (&x)[Dims<KIND>::NX], const real (&u)[Dims<KIND>::NU],
                   real (&f)[Dims<KIND>::NX]) {
    if constexpr (KIND == SCG_QUAD_1D) {                 // quadrotor.py:594-604
        f[0] = x[1]; f[1] = u[0] / p.m - p.g;
    } else if constexpr (KIND == SCG_QUAD_2D) {          // :605-625
        real s, c;
        m_sincos(x[4], &s, &c);
        const real T = u[0] + u[1];
        f[0] = x[1]; f[1] = s * T / p.m; f[2] = x[3]; f[3] = c * T / p.m - p.g; f[4] = x[5];
        f[5] = p.arm_s2 * (u[1] - u[0]) / p.Iyy;
    } else if constexpr (KIND == SCG_QUAD_3D) {          // :626-676
        real sph, cph, sth, cth, sps, cps;
        m_sincos(x[6], &sph, &cph); m_sincos(x[7], &sth, &cth); m_sincos(x[8], &sps, &cps);
        const real T = ((u[0] + u[1]) + u[2]) + u[3];
        const real Tm = T / p.m;
        const real P = x[9], Q = x[10], R = x[11];
        f[0] = x[1]; f[1] = (cps * sth * cph + sps * sph) * Tm;      // Rz(psi) Ry(theta) Rx(phi) third column
        f[2] = x[3]; f[3] = (sps * sth * cph - cps * sph) * Tm;
        f[4] = x[5]; f[5] = cth * cph * Tm - p.g;
        const real tth = sth / cth;
        f[6] = P + sph * tth * Q + cph * tth * R;
        f[7] = cph * Q - sph * R;
        f[8] = (sph * Q + cph * R) / cth;
        const real Mx = p.arm_s2 * (u[0] + u[1] - u[2] - u[3]);
        const real My = p.arm_s2 * (-u[0] + u[1] + u[2] - u[3]);
        const real Mz = p.gamma * (u[0] - u[1] + u[2] - u[3]);
        f[9] = (Mx - Q * R * (p.Izz - p.Iyy)) / p.Ixx;
        f[10] = (My - P * R * (p.Ixx - p.Izz)) / p.Iyy;
        f[11] = (Mz - P * Q * (p.Iyy - p.Ixx)) / p.Izz;
    } else {                                             // cart-pole (A13)
        real s, c;
        m_sincos(x[2], &s, &c);
        const real Mm = p.cp_mc + p.cp_mp, ml = p.cp_mp * p.cp_l;
        const real tmp = (u[0] + ml * x[3] * x[3] * s) / Mm;
        const real thdd = (p.g * s - c * tmp) / (p.cp_l * (real(4.0 / 3.0) - p.cp_mp * c * c / Mm));
        f[0] = x[1]; f[1] = tmp - ml * thdd * c / Mm; f[2] = x[3]; f[3] = thdd;
    }
}

template <int KIND, typename real>
SCG_DEV void load_xu(const real* __restrict__ x, const real* __restrict__ u, int64_t n, int64_t i,
                     real (&xs)[Dims<KIND>::NX], real (&us)[Dims<KIND>::NU]) {
#pragma unroll
    for (int d = 0; d < Dims<KIND>::NX; ++d) xs[d] = x[int64_t(d) * n + i];
#pragma unroll
    for (int j = 0; j < Dims<KIND>::NU; ++j) us[j] = u[int64_t(j) * n + i];
}

// op 0: f, 1: RK4 step, 4: Euler step
template <int KIND, typename real>
__global__ void __launch_bounds__(128)
k_sym_f(const SymP<real> p, int op, const real* __restrict__ x, const real* __restrict__ u, int64_t n, real dt,
        real* __restrict__ out) {
    constexpr int NX = Dims<KIND>::NX, NU = Dims<KIND>::NU;
    const int64_t i = int64_t(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i >= n) return;
    real xs[NX], us[NU], k1[NX];
    load_xu<KIND, real>(x, u, n, i, xs, us);
    sym_f<KIND, real>(p, xs, us, k1);
    if (op == SCG_SYM_FC) {
#pragma unroll
        for (int d = 0; d < NX; ++d) out[int64_t(d) * n + i] = k1[d];
        return;
    }
    if (op == SCG_SYM_FD_EULER) {
#pragma unroll
        for (int d = 0; d < NX; ++d) out[int64_t(d) * n + i] = xs[d] + dt * k1[d];
        return;
    }
    real k2[NX], k3[NX], k4[NX], y[NX];   // mpc_utils.py:93-98
#pragma unroll
    for (int d = 0; d < NX; ++d) y[d] = xs[d] + dt / real(2) * k1[d];
    sym_f<KIND, real>(p, y, us, k2);
#pragma unroll
    for (int d = 0; d < NX; ++d) y[d] = xs[d] + dt / real(2) * k2[d];
    sym_f<KIND, real>(p, y, us, k3);
#pragma unroll
    for (int d = 0; d < NX; ++d) y[d] = xs[d] + dt * k3[d];
    sym_f<KIND, real>(p, y, us, k4);
#pragma unroll
    for (int d = 0; d < NX; ++d)
        out[int64_t(d) * n + i] = xs[d] + dt / real(6) * (k1[d] + real(2) * k2[d] + real(2) * k3[d] + k4[d]);
}

// closed-form Jacobians; only non-zero entries are written, the caller zeroes the outputs
template <int KIND, typename real>
__global__ void __launch_bounds__(128)
k_sym_df(const SymP<real> p, const real* __restrict__ x, const real* __restrict__ u, int64_t n,
         real* __restrict__ A, real* __restrict__ B) {
    constexpr int NX = Dims<KIND>::NX, NU = Dims<KIND>::NU;
    const int64_t i = int64_t(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i >= n) return;
    real xs[NX], us[NU];
    load_xu<KIND, real>(x, u, n, i, xs, us);
    auto a = [&](int r, int c, real v) { A[int64_t(r * NX + c) * n + i] = v; };
    auto b = [&](int r, int c, real v) { B[int64_t(r * NU + c) * n + i] = v; };
    if constexpr (KIND == SCG_QUAD_1D) {
        a(0, 1, real(1)); b(1, 0, real(1) / p.m);
    } else if constexpr (KIND == SCG_QUAD_2D) {
        real s, c;
        m_sincos(xs[4], &s, &c);
        const real T = us[0] + us[1];
        a(0, 1, real(1)); a(2, 3, real(1)); a(4, 5, real(1));
        a(1, 4, c * T / p.m); a(3, 4, -s * T / p.m);
        b(1, 0, s / p.m); b(1, 1, s / p.m); b(3, 0, c / p.m); b(3, 1, c / p.m);
        b(5, 0, -p.arm_s2 / p.Iyy); b(5, 1, p.arm_s2 / p.Iyy);
    } else if constexpr (KIND == SCG_QUAD_3D) {
        real sph, cph, sth, cth, sps, cps;
        m_sincos(xs[6], &sph, &cph); m_sincos(xs[7], &sth, &cth); m_sincos(xs[8], &sps, &cps);
        const real Tm = (((us[0] + us[1]) + us[2]) + us[3]) / p.m;
        const real P = xs[9], Q = xs[10], R = xs[11], tth = sth / cth, sec2 = real(1) / (cth * cth);
        a(0, 1, real(1)); a(2, 3, real(1)); a(4, 5, real(1));
        a(1, 6, (-cps * sth * sph + sps * cph) * Tm); a(1, 7, cps * cth * cph * Tm); a(1, 8, (-sps * sth * cph + cps * sph) * Tm);
        a(3, 6, (-sps * sth * sph - cps * cph) * Tm); a(3, 7, sps * cth * cph * Tm); a(3, 8, (cps * sth * cph + sps * sph) * Tm);
        a(5, 6, -cth * sph * Tm); a(5, 7, -sth * cph * Tm);
        a(6, 6, (cph * Q - sph * R) * tth); a(6, 7, (sph * Q + cph * R) * sec2);
        a(6, 9, real(1)); a(6, 10, sph * tth); a(6, 11, cph * tth);
        a(7, 6, -sph * Q - cph * R); a(7, 10, cph); a(7, 11, -sph);
        a(8, 6, (cph * Q - sph * R) / cth); a(8, 7, (sph * Q + cph * R) * sth * sec2);
        a(8, 10, sph / cth); a(8, 11, cph / cth);
        a(9, 10, -R * (p.Izz - p.Iyy) / p.Ixx); a(9, 11, -Q * (p.Izz - p.Iyy) / p.Ixx);
        a(10, 9, -R * (p.Ixx - p.Izz) / p.Iyy); a(10, 11, -P * (p.Ixx - p.Izz) / p.Iyy);
        a(11, 9, -Q * (p.Iyy - p.Ixx) / p.Izz); a(11, 10, -P * (p.Iyy - p.Ixx) / p.Izz);
        const real r2[3] = {(cps * sth * cph + sps * sph) / p.m, (sps * sth * cph - cps * sph) / p.m, cth * cph / p.m};
        const real sx[4] = {1, 1, -1, -1}, sy[4] = {-1, 1, 1, -1}, sz[4] = {1, -1, 1, -1};
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            b(1, j, r2[0]); b(3, j, r2[1]); b(5, j, r2[2]);
            b(9, j, sx[j] * p.arm_s2 / p.Ixx); b(10, j, sy[j] * p.arm_s2 / p.Iyy); b(11, j, sz[j] * p.gamma / p.Izz);
        }
    } else {   // cart-pole: differentiate f = (x1, xdd, x3, thdd) in closed form
        real s, c;
        m_sincos(xs[2], &s, &c);
        const real Mm = p.cp_mc + p.cp_mp, ml = p.cp_mp * p.cp_l, w = xs[3], F = us[0];
        const real tmp = (F + ml * w * w * s) / Mm;
        const real den = p.cp_l * (real(4.0 / 3.0) - p.cp_mp * c * c / Mm);
        const real num = p.g * s - c * tmp;
        const real thdd = num / den;
        // partials of tmp, num, den
        const real dtmp_dth = ml * w * w * c / Mm, dtmp_dw = real(2) * ml * w * s / Mm, dtmp_dF = real(1) / Mm;
        const real dnum_dth = p.g * c + s * tmp - c * dtmp_dth, dnum_dw = -c * dtmp_dw, dnum_dF = -c * dtmp_dF;
        const real dden_dth = p.cp_l * (real(2) * p.cp_mp * c * s / Mm);
        const real dthdd_dth = (dnum_dth * den - num * dden_dth) / (den * den);
        const real dthdd_dw = dnum_dw / den, dthdd_dF = dnum_dF / den;
        a(0, 1, real(1)); a(2, 3, real(1));
        a(3, 2, dthdd_dth); a(3, 3, dthdd_dw); b(3, 0, dthdd_dF);
        // xdd = tmp - ml * thdd * c / Mm
        a(1, 2, dtmp_dth - ml * (dthdd_dth * c - thdd * s) / Mm);
        a(1, 3, dtmp_dw - ml * dthdd_dw * c / Mm);
        b(1, 0, dtmp_dF - ml * dthdd_dF * c / Mm);
    }
}

// quadratic cost and its gradients: l = 0.5 ex' Q ex + 0.5 eu' R eu; l_x = 0.5 (Q + Q') ex; l_u = 0.5 (R + R') eu
template <int KIND, typename real> struct LossP { real Q[Dims<KIND>::NX * Dims<KIND>::NX], R[Dims<KIND>::NU * Dims<KIND>::NU],
                                                   xr[Dims<KIND>::NX], ur[Dims<KIND>::NU]; };

template <int KIND, typename real>
__global__ void __launch_bounds__(128)
k_sym_loss(const __grid_constant__ LossP<KIND, real> p, const real* __restrict__ x, const real* __restrict__ u,
           const real* __restrict__ xr, const real* __restrict__ ur, int64_t n, real* __restrict__ l,
           real* __restrict__ lx, real* __restrict__ lu) {
    constexpr int NX = Dims<KIND>::NX, NU = Dims<KIND>::NU;
    const int64_t i = int64_t(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i >= n) return;
    real ex[NX], eu[NU];
#pragma unroll
    for (int d = 0; d < NX; ++d) ex[d] = x[int64_t(d) * n + i] - (xr ? xr[int64_t(d) * n + i] : p.xr[d]);
#pragma unroll
    for (int j = 0; j < NU; ++j) eu[j] = u[int64_t(j) * n + i] - (ur ? ur[int64_t(j) * n + i] : p.ur[j]);
    real acc = real(0);
#pragma unroll
    for (int a = 0; a < NX; ++a) {
        real row = real(0), col = real(0);
#pragma unroll
        for (int b = 0; b < NX; ++b) { row += p.Q[a * NX + b] * ex[b]; col += p.Q[b * NX + a] * ex[b]; }
        acc += real(0.5) * ex[a] * row;
        if (lx) lx[int64_t(a) * n + i] = real(0.5) * (row + col);
    }
#pragma unroll
    for (int a = 0; a < NU; ++a) {
        real row = real(0), col = real(0);
#pragma unroll
        for (int b = 0; b < NU; ++b) { row += p.R[a * NU + b] * eu[b]; col += p.R[b * NU + a] * eu[b]; }
        acc += real(0.5) * eu[a] * row;
        if (lu) lu[int64_t(a) * n + i] = real(0.5) * (row + col);
    }
    l[i] = acc;
}

template <int KIND, typename real>
static int launch_kind(const scg_config& cfg, int op, const real* x, const real* u, const real* xr, const real* ur,
                       const double* Qh, const double* Rh, int64_t n, real dt, real* out0, real* out1, real* out2,
                       cudaStream_t st) {
    constexpr int NX = Dims<KIND>::NX, NU = Dims<KIND>::NU;
    SymP<real> p;
    p.m = real(cfg.mass); p.g = real(cfg.gravity_acc); p.arm_s2 = real(cfg.arm / std::sqrt(2.0));
    p.Ixx = real(cfg.J[0]); p.Iyy = real(cfg.J[1]); p.Izz = real(cfg.J[2]); p.gamma = real(cfg.km / cfg.kf);
    p.cp_l = real(cfg.cp_length); p.cp_mp = real(cfg.cp_pole_mass); p.cp_mc = real(cfg.cp_cart_mass);
    const unsigned blocks = unsigned((n + 127) / 128);
    if (op == SCG_SYM_FC || op == SCG_SYM_FD_RK4 || op == SCG_SYM_FD_EULER) {
        k_sym_f<KIND, real><<<blocks, 128, 0, st>>>(p, op, x, u, n, dt, out0);
    } else if (op == SCG_SYM_DF) {
        if (!out1) return -2;
        cudaMemsetAsync(out0, 0, sizeof(real) * NX * NX * n, st);
        cudaMemsetAsync(out1, 0, sizeof(real) * NX * NU * n, st);
        k_sym_df<KIND, real><<<blocks, 128, 0, st>>>(p, x, u, n, out0, out1);
    } else if (op == SCG_SYM_LOSS) {
        LossP<KIND, real> lp;
        for (int a = 0; a < NX; ++a) for (int b = 0; b < NX; ++b)
            lp.Q[a * NX + b] = real(Qh ? Qh[a * NX + b] : cfg.Q[a * SCG_MAX_NX + b]);
        for (int a = 0; a < NU; ++a) for (int b = 0; b < NU; ++b)
            lp.R[a * NU + b] = real(Rh ? Rh[a * NU + b] : cfg.R[a * SCG_MAX_NU + b]);
        for (int a = 0; a < NX; ++a) lp.xr[a] = real(0);
        for (int a = 0; a < NU; ++a) lp.ur[a] = real(cfg.u_goal[a]);
        k_sym_loss<KIND, real><<<blocks, 128, 0, st>>>(lp, x, u, xr, ur, n, out0, out1, out2);
    } else return -1;
    return 0;
}

template <typename real>
int symbolic_launch(const scg_config& cfg, int op, const real* x, const real* u, const real* xr, const real* ur,
                    const double* Qh, const double* Rh, int64_t n, real dt, real* out0, real* out1, real* out2,
                    cudaStream_t st) {
    switch (cfg.env_kind) {
        case SCG_QUAD_1D: return launch_kind<SCG_QUAD_1D, real>(cfg, op, x, u, xr, ur, Qh, Rh, n, dt, out0, out1, out2, st);
        case SCG_QUAD_2D: return launch_kind<SCG_QUAD_2D, real>(cfg, op, x, u, xr, ur, Qh, Rh, n, dt, out0, out1, out2, st);
        case SCG_QUAD_3D: return launch_kind<SCG_QUAD_3D, real>(cfg, op, x, u, xr, ur, Qh, Rh, n, dt, out0, out1, out2, st);
        case SCG_CARTPOLE: return launch_kind<SCG_CARTPOLE, real>(cfg, op, x, u, xr, ur, Qh, Rh, n, dt, out0, out1, out2, st);
    }
    return -1;
}

template int symbolic_launch<float>(const scg_config&, int, const float*, const float*, const float*, const float*, const double*,
                                    const double*, int64_t, float, float*, float*, float*, cudaStream_t);
template int symbolic_launch<double>(const scg_config&, int, const double*, const double*, const double*, const double*,
                                     const double*, const double*, int64_t, double, double*, double*, double*, cudaStream_t);

}  // namespace scg
