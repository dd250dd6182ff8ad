// scg_env.cuh -- per-env device functions of the fused step: one thread owns one env, the
// state and motor commands live in registers for all PYB_FREQ / CTRL_FREQ substeps.
// Reference line numbers are under safe_control_gym/ (see SURVEY.md section 8a rows A1-A11).
#pragma once
#include "scg_params.cuh"

namespace scg {

// ------------------------------------------------------------------ attitude (row A3)
// The reference stores the attitude in PyBullet as quatFromEuler(rpy + dt * rates) and reads it
// back as eulerFromQuat(quat) before every use (base_aviary.py:565-574, 281-285, 322, 362-374).
// That round trip returns the canonical Euler angles of the same rotation, so SCG_ATT_FAST computes
// them by wrap / fold arithmetic and builds the rotation from sin/cos of the angles; only inside the
// gimbal band (|sin(pitch)| >= 0.99999, where bullet snaps the pitch) -- and always with
// SCG_ATT_ROUNDTRIP -- the literal bullet3 arithmetic is executed.  SCG_ATT_FAST_INTRINSIC (fp32
// only) additionally evaluates sin/cos on the SFU (MUFU.SIN / MUFU.COS).
template <typename real> struct Att3 {
    real rpy[3];   // wrapped angles == self.rpy after _update_and_store_kinematic_information
    real R[9];     // getMatrixFromQuaternion(self.quat), row-major; the substep loop keeps only column 2 current
    real sc[6];    // sin/cos of roll, pitch, yaw that column 2 was built from (fast path)
    int form;      // 0: all of R is current; 1: columns 0 and 1 still to be built from sc (attitude3_complete)
};

// x - 2 pi * rint(x / 2 pi): the identity (bit-exact) for x already in [-pi, pi]
SCG_DEV double wrap_pi(double x) {
    const double k = rint(x * 0.15915494309189535);
    return fma(-k, 2.4492935982947064e-16, fma(-k, 6.283185307179586, x));
}
SCG_DEV float wrap_pi(float x) {
    const float k = fast_rint(x * 0.15915494309189535f);
    return fmaf(-k, -1.7484555e-07f, fmaf(-k, 6.2831855f, x));   // 2 pi = 6.2831855f - 1.7484555e-07f
}
// wrap for SCG_ATT_FAST_INTRINSIC: one-constant reduction (an actual wrap is off by k * 1.7e-7 rad, below the MUFU error;
// still the identity inside [-pi, pi])
template <int ATT> SCG_DEV float att_wrap(float x) {
    if (ATT == SCG_ATT_FAST_INTRINSIC) return fmaf(-fast_rint(x * 0.15915494309189535f), 6.2831855f, x);
    return wrap_pi(x);
}
template <int ATT> SCG_DEV double att_wrap(double x) { return wrap_pi(x); }
template <int ATT> SCG_DEV void att_sincos(float x, float* s, float* c) {
    if (ATT == SCG_ATT_FAST_INTRINSIC) __sincosf(x, s, c);   // MUFU.SIN/COS, |err| <= 2^-21.4 on [-pi, pi]
    else sincos_pi(x, s, c);                                  // |x| <= pi after wrap_pi
}
template <int ATT> SCG_DEV void att_sincos(double x, double* s, double* c) { sincos(x, s, c); }

// The literal bullet3 sequence: getQuaternionFromEuler -> getEulerFromQuaternion / getMatrixFromQuaternion.
// Out of line: SCG_ATT_FAST reaches it only inside the gimbal band, and the hot loop must stay small.
template <typename real>
__device__ __noinline__ void attitude3_roundtrip(real p0, real p1, real p2, Att3<real>& a) {
    real sph, cph, sth, cth, sps, cps;
    m_sincos(p0 * real(0.5), &sph, &cph);
    m_sincos(p1 * real(0.5), &sth, &cth);
    m_sincos(p2 * real(0.5), &sps, &cps);
    real x = sph * cth * cps - cph * sth * sps;
    real y = cph * sth * cps + sph * cth * sps;
    real z = cph * cth * sps - sph * sth * cps;
    real w = cph * cth * cps + sph * sth * sps;
    const real len = m_sqrt(x * x + y * y + z * z + w * w);
    x /= len; y /= len; z /= len; w /= len;
    // pybullet getEulerFromQuaternion
    const real sqx = x * x, sqy = y * y, sqz = z * z, squ = w * w;
    const real sarg = real(-2) * (x * z - w * y);
    if (sarg <= real(-0.99999)) {
        a.rpy[0] = real(0); a.rpy[1] = -Consts<real>::half_pi; a.rpy[2] = real(2) * m_atan2(x, -y);
    } else if (sarg >= real(0.99999)) {
        a.rpy[0] = real(0); a.rpy[1] = Consts<real>::half_pi; a.rpy[2] = real(2) * m_atan2(-x, y);
    } else {
        a.rpy[0] = m_atan2(real(2) * (y * z + w * x), squ - sqx - sqy + sqz);
        a.rpy[1] = m_asin(sarg);
        a.rpy[2] = m_atan2(real(2) * (x * y + w * z), squ + sqx - sqy - sqz);
    }
    // pybullet getMatrixFromQuaternion
    const real s = real(2) / (sqx + sqy + sqz + squ);
    const real xs = x * s, ys = y * s, zs = z * s;
    const real wx = w * xs, wy = w * ys, wz = w * zs;
    const real xx = x * xs, xy = x * ys, xz = x * zs;
    const real yy = y * ys, yz = y * zs, zz = z * zs;
    a.R[0] = real(1) - (yy + zz); a.R[1] = xy - wz; a.R[2] = xz + wy;
    a.R[3] = xy + wz;             a.R[4] = real(1) - (xx + zz); a.R[5] = yz - wx;
    a.R[6] = xz - wy;             a.R[7] = yz + wx; a.R[8] = real(1) - (xx + yy);
    a.form = 0;
}

// rpy and column 2 of R (the thrust direction) from the pre-round-trip angles.
//
// SCG_ATT_FAST*: the round trip eulerFromQuat(quatFromEuler(pre)) returns the canonical Euler angles of
// the same rotation, so they are computed by wrap / fold arithmetic and R from their sin / cos:
//   outside the gimbal band: wrap into [-pi, pi], fold a pitch beyond +-pi/2
//       (Rz(psi+pi) Ry(pi-theta) Rx(phi+pi) == Rz(psi) Ry(theta) Rx(phi));
//   inside the band |sin(pitch)| >= 0.99999 bullet snaps roll = 0, pitch = +-pi/2 and
//       yaw = 2 atan2(-+x, +-y) of the quaternion built from the RAW angles.  Expanding the half-angle
//       products around pitch = +-pi/2 with r = (pitch -+ pi/2) / 2, |r| <= 2.3e-3, gives
//         yaw = psi -+ phi + 2 r sin(phi) / (1 +- r cos(phi))        taken modulo 4 pi into (-2 pi, 2 pi]
//       (tan r ~ r and atan u ~ u drop terms below 3e-9; checked against the literal sequence in
//       tests/test_oracle_attitude.py).  The reference never leaves the band again while
//       |pitch rate| dt stays below its half-width: envs that pitch over fall "stuck" at +-90 deg
//       until a position bound ends the episode -- about 2 % of all substeps of the BASELINE workload,
//       i.e. half of all warp-substeps see at least one such lane, so the band is handled by a few
//       predicated instructions here and not by the literal sequence.
//   fp64 keeps the literal sequence inside the band (validation build, 1e-12).
template <typename real, int ATT>
SCG_DEV void attitude3(const real (&pre)[3], Att3<real>& a) {
    if (ATT != SCG_ATT_ROUNDTRIP) {
        const real thw = att_wrap<ATT>(pre[1]);
        real th = thw, ph = pre[0], ps = pre[2];
        const bool folded = m_abs(th) > Consts<real>::half_pi;
        if (folded) {
            th = (th > real(0) ? Consts<real>::pi : -Consts<real>::pi) - th;
            ph += Consts<real>::pi; ps += Consts<real>::pi;
        }
        ph = att_wrap<ATT>(ph); ps = att_wrap<ATT>(ps);
        real sth, cth;
        att_sincos<ATT>(th, &sth, &cth);
        const bool band = !(m_abs(sth) < real(sizeof(real) == 4 ? 0.99999 : 0.99998));
        if (!band || sizeof(real) == 4) {
            real sph, cph, sps, cps;
            att_sincos<ATT>(ph, &sph, &cph);
            att_sincos<ATT>(ps, &sps, &cps);
            a.rpy[0] = ph; a.rpy[1] = th; a.rpy[2] = ps;
            a.sc[0] = sph; a.sc[1] = cph; a.sc[2] = sth; a.sc[3] = cth; a.sc[4] = sps; a.sc[5] = cps;
            a.R[2] = cps * sth * cph + sps * sph;
            a.R[5] = sps * sth * cph - cps * sph;
            a.R[8] = cth * cph;
            a.form = 1;
            if (band) {   // fp32 only (see above)
                const bool up = sth > real(0);
                const real r = real(0.5) * (thw - (up ? Consts<real>::half_pi : -Consts<real>::half_pi));
                const real sr = folded ? -sph : sph, cr = folded ? -cph : cph;   // sin / cos of the RAW roll
                const real u = up ? r * cr : -(r * cr);
                const real corr = real(2) * r * sr * (real(1) - u + u * u);       // 1 / (1 + u), |u| <= 2.3e-3
                real yaw = (up ? pre[2] - pre[0] : pre[2] + pre[0]) + corr;
                const real k = fast_rint(yaw * real(0.07957747154594767));        // modulo 4 pi
                yaw = m_fma(-k, real(-3.4969110629390343e-07), m_fma(-k, real(12.566370964050293), yaw));
                a.rpy[0] = real(0);
                a.rpy[1] = up ? Consts<real>::half_pi : -Consts<real>::half_pi;
                a.rpy[2] = yaw;
            }
            return;
        }
    }
    Att3<real> lit;   // a private copy: the out-of-line call must not pin the caller's registers to memory
    attitude3_roundtrip<real>(pre[0], pre[1], pre[2], lit);
    a = lit;
}

// columns 0 and 1 of R, needed only where the body rates p, q, r are read (quadrotor.py:862-864)
template <typename real> SCG_DEV void attitude3_complete(Att3<real>& a) {
    if (a.form == 1) {
        const real sph = a.sc[0], cph = a.sc[1], sth = a.sc[2], cth = a.sc[3], sps = a.sc[4], cps = a.sc[5];
        a.R[0] = cps * cth; a.R[1] = cps * sth * sph - sps * cph;
        a.R[3] = sps * cth; a.R[4] = sps * sth * sph + cps * cph;
        a.R[6] = -sth;      a.R[7] = cth * sph;
    }
    a.form = 0;
}

// Planar (2D) attitude: pitch alpha as the reference reports it plus the "flipped" bit that
// records roll = yaw = pi after the asin fold.  r3 = third column of R = (sin, 0, cos) of the
// physical pitch.
template <typename real> struct Att2 { real theta; real r3x, r3z; uint32_t flipped; };

template <typename real, int ATT = SCG_ATT_FAST>
SCG_DEV void attitude2(real alpha_pre, uint32_t flipped_in, Att2<real>& a) {
    if (!flipped_in && m_abs(alpha_pre) < real(1.5663)) {
        // |alpha| < pi / 2: the compact ~1-ulp pair (fp64: sincos); SCG_ATT_FAST_INTRINSIC (fp32): MUFU.SIN / MUFU.COS
        if (ATT == SCG_ATT_FAST_INTRINSIC && sizeof(real) == 4) att_sincos<ATT>(alpha_pre, &a.r3x, &a.r3z);
        else sincos_pi(alpha_pre, &a.r3x, &a.r3z);
        a.theta = alpha_pre; a.flipped = 0u;
        return;
    }
    const real phys = flipped_in ? Consts<real>::pi - alpha_pre : alpha_pre;
    real s, c;
    sincos_any(phys, &s, &c);
    a.r3x = s; a.r3z = c;
    if (s <= real(-0.99999)) { a.theta = -Consts<real>::half_pi; a.flipped = 0u; }
    else if (s >= real(0.99999)) { a.theta = Consts<real>::half_pi; a.flipped = 0u; }
    else { a.theta = m_asin(s); a.flipped = (c < real(0)) ? 1u : 0u; }
}

// ------------------------------------------------------------------ motors (rows A1-A2)
template <typename real> struct Motors {
    real T;            // sum of the four rotor forces
    real tx, ty, tz;   // body torques before the gyroscopic term
    real rpm[4];       // literal map: the commanded RPMs; fast map: the rotor FORCES (see motor_rpm)
};

// quadrotor.py:815-842, quadrotor_utils.py:21-67, base_aviary.py:538-551.
// thrust[] is the (possibly disturbed) commanded thrust per input.
template <int KIND, typename real, bool FASTMATH>
SCG_DEV void motors_from_thrust(const Hot<real>& h, const real (&thrust)[Dims<KIND>::NU], Motors<real>& m, bool clip = true) {
    constexpr int NU = Dims<KIND>::NU;
    if constexpr (FASTMATH) {
        // thrust -> PWM -> clip -> RPM -> KF rpm^2 is the identity on the rotor force apart from the clip, which maps
        // to force bounds (sqrt is monotone): f = clip(thrust / n, KF rpm_min^2, KF rpm_max^2).  Same value up to the
        // rounding of the chain it skips; the RPMs are formed only where they are logged (motor_rpm).
        real fi[NU];
#pragma unroll
        for (int j = 0; j < NU; ++j) {
            const real t = NU == 4 ? thrust[j] : thrust[j] * h.inv_n_motor;
            fi[j] = clip ? m_clip(t, h.f_min, h.f_max) : t;   // clip == false: the caller's action clip already implies it (CFG_CLIP_ONCE)
        }
        real f[4];
        if (KIND == SCG_QUAD_1D) { f[0] = f[1] = f[2] = f[3] = fi[0]; }
        else if (KIND == SCG_QUAD_2D) { f[0] = fi[0]; f[1] = fi[NU - 1]; f[2] = fi[NU - 1]; f[3] = fi[0]; }
        else { f[0] = fi[0]; f[1] = fi[1 % NU]; f[2] = fi[2 % NU]; f[3] = fi[3 % NU]; }
#pragma unroll
        for (int j = 0; j < 4; ++j) m.rpm[j] = f[j];
        const Row4<real> sr = ldc4(h.step_row);   // [dt, arm_s2, km_over_kf, inv_mass]: the same constants in one 16-byte row
        m.T = ((f[0] + f[1]) + f[2]) + f[3];
        m.tz = (((f[0] - f[1]) + f[2]) - f[3]) * sr.z;
        m.tx = (((f[0] + f[1]) - f[2]) - f[3]) * sr.y;
        m.ty = (((-f[0] + f[1]) + f[2]) - f[3]) * sr.y;
        return;
    }
    real rp[NU];
#pragma unroll
    for (int j = 0; j < NU; ++j) {
        const real t = m_max(thrust[j], real(0));
        real pwm = ((FASTMATH ? sqrt_fast(t * h.inv_nkf) : m_sqrt(t * h.inv_nkf)) - h.pwm_const) * h.inv_pwm_scale;
        pwm = m_clip(pwm, h.pwm_min, h.pwm_max);
        rp[j] = h.pwm_scale * pwm + h.pwm_const;
    }
    if (KIND == SCG_QUAD_1D) { m.rpm[0] = m.rpm[1] = m.rpm[2] = m.rpm[3] = rp[0]; }
    else if (KIND == SCG_QUAD_2D) { m.rpm[0] = rp[0]; m.rpm[1] = rp[NU - 1]; m.rpm[2] = rp[NU - 1]; m.rpm[3] = rp[0]; }
    else { m.rpm[0] = rp[0]; m.rpm[1] = rp[1 % NU]; m.rpm[2] = rp[2 % NU]; m.rpm[3] = rp[3 % NU]; }
    real f[4], zt[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) { const real r2 = m.rpm[j] * m.rpm[j]; f[j] = r2 * h.kf; zt[j] = r2 * h.km; }
    m.T = ((f[0] + f[1]) + f[2]) + f[3];
    m.tz = ((zt[0] - zt[1]) + zt[2]) - zt[3];
    m.tx = (((f[0] + f[1]) - f[2]) - f[3]) * h.arm_s2;
    m.ty = (((-f[0] + f[1]) + f[2]) - f[3]) * h.arm_s2;
}

// the RPM of rotor j as the Logger records it
template <typename real, bool FASTMATH> SCG_DEV real motor_rpm(const Hot<real>& h, const Motors<real>& m, int j) {
    if constexpr (FASTMATH) return sqrt_fast(m.rpm[j] * h.inv_kf);
    else return m.rpm[j];
}

// per-episode inertial constants (nominal in parity mode, F9)
template <typename real> struct Inertia { real inv_mass, g_over_m, J[3], J_inv[3]; real dJ[3], kJ[3]; real cp_l, cp_mp, cp_mc; };

// ------------------------------------------------------------------ cart-pole ODE (row A13)
// Frictionless cart-pole (Barto et al.), x = [x, x_dot, theta, theta_dot], theta = 0 upright,
// l = effective (half) pole length.  The env is absent from the reference fork:
// contract inferred from tests/scripts/verbose_api.yaml:1-91 -- parity unpinned.
template <typename real, int ATT = SCG_ATT_FAST>
SCG_DEV void cartpole_f(const Inertia<real>& in, real g, const real (&x)[4], real F, real (&dx)[4]) {
    real s, c;
    const real Mm = in.cp_mc + in.cp_mp, ml = in.cp_mp * in.cp_l;
    if constexpr (sizeof(real) == 4) {
        // fp32: the three divisions by the total mass become one reciprocal (hoisted out of the four RK4 stages by the
        // compiler), sin / cos by the wrap + ~1-ulp polynomial pair instead of libdevice's slow-path sincosf: ~45 instead
        // of ~90 instructions per stage, a few ulp from the literal order below (fp32 bar of the CartPole tests: 2e-4)
        const float inv_Mm = 1.f / Mm;
        if (ATT == SCG_ATT_FAST_INTRINSIC) __sincosf(att_wrap<ATT>(x[2]), &s, &c);   // MUFU.SIN / MUFU.COS: |err| <= 2^-21.4 on [-pi, pi]
        else sincos_pi(wrap_pi(x[2]), &s, &c);
        const float tmp = (F + ml * x[3] * x[3] * s) * inv_Mm;
        const float den = in.cp_l * (float(4.0 / 3.0) - in.cp_mp * c * c * inv_Mm);
        // fast_intrinsic: MUFU.RCP (2 ulp) instead of the IEEE division with its slow-path call, four times per control step
        const float thdd = ATT == SCG_ATT_FAST_INTRINSIC ? __fdividef(g * s - c * tmp, den) : (g * s - c * tmp) / den;
        dx[0] = x[1]; dx[1] = tmp - ml * thdd * c * inv_Mm; dx[2] = x[3]; dx[3] = thdd;
        return;
    }
    m_sincos(x[2], &s, &c);
    const real tmp = (F + ml * x[3] * x[3] * s) / Mm;
    const real thdd = (g * s - c * tmp) / (in.cp_l * (real(4.0 / 3.0) - in.cp_mp * c * c / Mm));
    dx[0] = x[1]; dx[1] = tmp - ml * thdd * c / Mm; dx[2] = x[3]; dx[3] = thdd;
}
template <typename real, int ATT = SCG_ATT_FAST>
SCG_DEV void cartpole_rk4(const Inertia<real>& in, real g, real (&x)[4], real F, real dt) {   // mpc_utils.py:79-101
    real k1[4], k2[4], k3[4], k4[4], y[4];
    cartpole_f<real, ATT>(in, g, x, F, k1);
#pragma unroll
    for (int d = 0; d < 4; ++d) y[d] = x[d] + dt * real(0.5) * k1[d];
    cartpole_f<real, ATT>(in, g, y, F, k2);
#pragma unroll
    for (int d = 0; d < 4; ++d) y[d] = x[d] + dt * real(0.5) * k2[d];
    cartpole_f<real, ATT>(in, g, y, F, k3);
#pragma unroll
    for (int d = 0; d < 4; ++d) y[d] = x[d] + dt * k3[d];
    cartpole_f<real, ATT>(in, g, y, F, k4);
#pragma unroll
    for (int d = 0; d < 4; ++d) x[d] += dt / real(6) * (k1[d] + real(2) * k2[d] + real(2) * k3[d] + k4[d]);
}

// ------------------------------------------------------------------ constraints (row A6)
// c = g(x) rounded to 8 decimals (constraints.py:119); violated = any(c > 0) (or >= 0 if strict).
// round(c, 8) > 0  <=>  rint(c * 1e8) > 0  <=>  c * 1e8 > 0.5 (ties go to even = 0), and
// round(c, 8) >= 0 <=>  c * 1e8 >= -0.5, so the flags need no rounding instruction.
template <typename real> SCG_DEV real round8(real c) { return m_rint(c * real(1e8)) / real(1e8); }
template <> SCG_DEV float round8<float>(float c) {
    // fp32 carries fewer than 8 decimals above |c| ~ 0.08: rounding is then the identity
    // k / 1e8 as a product and one Newton step on its residual (k is an integer below 2^23: the residual fma is exact) --
    // the division subroutine was 250 instructions per host-driven step for the 14 rows of the level0 family
    const float k = m_rint(c * 1e8f);
    const float q = k * 1e-8f;
    return m_abs(c) < 0.08f ? fmaf(fmaf(-q, 1e8f, k), 1e-8f, q) : c;
}
template <typename real> SCG_DEV bool viol_test(real c, int strict) {
    const real y = c * real(1e8);
    return strict ? (y >= real(-0.5)) : (y > real(0.5));
}

// linear / quadratic rows (uniform loads from the cold block); out of line, rare
template <int KIND, typename real>
__device__ __noinline__ void eval_dense_rows(int n_dense_rows, const Cold<real>* __restrict__ cold,
                                             const real* __restrict__ st, const real* __restrict__ raw,
                                             bool state_only, real* __restrict__ c_out, int64_t n, int64_t i,
                                             bool* violated, bool* almost) {
    constexpr int NX = Dims<KIND>::NX, NU = Dims<KIND>::NU;
    bool v = false, al = false;
    for (int r = 0; r < n_dense_rows; ++r) {
        const DenseRow<real>& dr = cold->dense[r];
        if (dr.var == SCG_VAR_INPUT && state_only) continue;
        real c = real(0);
        if (dr.quadratic) {
            const real* P = cold->P[dr.p_index];
            if (dr.var == SCG_VAR_STATE) {
                for (int a = 0; a < NX; ++a) {
                    real row = real(0);
                    for (int b = 0; b < NX; ++b) row += P[a * SCG_MAX_NX + b] * st[b];
                    c += st[a] * row;
                }
            } else {
                for (int a = 0; a < NU; ++a) {
                    real row = real(0);
                    for (int b = 0; b < NU; ++b) row += P[a * SCG_MAX_NX + b] * raw[b];
                    c += raw[a] * row;
                }
            }
        } else if (dr.var == SCG_VAR_STATE) {
            for (int d = 0; d < NX; ++d) c += dr.a[d] * st[d];
        } else {
            for (int d = 0; d < NU; ++d) c += dr.a[d] * raw[d];
        }
        c -= dr.b;
        const real rc = round8(c);
        v |= dr.strict ? (rc >= real(0)) : (rc > real(0));
        if (dr.has_tol) al |= (rc + dr.tol > real(0));
        if (c_out) c_out[int64_t(dr.out_row) * n + i] = rc;
    }
    *violated = v; *almost = al;
}

// Everything beyond the violation flag of bounded constraints -- the rounded values
// (info["constraint_values"]), tolerances (is_almost_active), symmetric and dense rows -- out of
// line: per-step outputs only exist in the host-driven scg_step mode.
template <int KIND, typename real>
__device__ __noinline__ uint32_t constraint_values_cold(const Hot<real>& h, const Cold<real>* __restrict__ cold,
                                                        const real* __restrict__ st, const real* __restrict__ raw,
                                                        bool state_only, real* __restrict__ c_out, int64_t n, int64_t i) {
    constexpr int NX = Dims<KIND>::NX, NU = Dims<KIND>::NU;
    bool violated = false, almost = false;
    for (int k = 0; k < h.n_bounded; ++k) {
        const BoundedCon<real>& bc = h.bounded[k];
        const bool is_state = bc.var == SCG_VAR_STATE;
        if (!is_state && state_only) continue;
        const int nd = is_state ? NX : NU;
        for (int d = 0; d < nd; ++d) {
            if (!((bc.mask >> d) & 1u)) continue;
            const real x = is_state ? st[d] : raw[d];
            real clo, chi, rlo, rhi;
            if (bc.symmetric) {   // constraints.py:404-406 (unrounded)
                clo = m_abs(x) - bc.hi[d]; chi = -INFINITY; rlo = clo; rhi = chi;
                violated |= bc.strict ? (clo >= real(0)) : (clo > real(0));
            } else {
                clo = bc.lo[d] - x; chi = x - bc.hi[d]; rlo = round8(clo); rhi = round8(chi);
                violated |= viol_test(clo, bc.strict) | viol_test(chi, bc.strict);
            }
            if (c_out && bc.row_lo[d] >= 0) c_out[int64_t(bc.row_lo[d]) * n + i] = rlo;
            if (c_out && bc.row_hi[d] >= 0) c_out[int64_t(bc.row_hi[d]) * n + i] = rhi;
            if (bc.has_tol) almost |= (rlo + cold->bounded_tol[k][d] > real(0)) | (rhi + cold->bounded_tol[k][SCG_MAX_NX + d] > real(0));
        }
    }
    if (h.n_dense_rows > 0) {
        bool v, al;
        eval_dense_rows<KIND, real>(h.n_dense_rows, cold, st, raw, state_only, c_out, n, i, &v, &al);
        violated |= v; almost |= al;
    }
    return (violated ? 1u : 0u) | (almost ? 2u : 0u);   // by value: the caller's flags stay in registers
}

// The same values for configs whose rows all come from plain bounded constraints (Hot::crow_*): one pass over the output
// rows instead of the object / dim / kind loops above (the host-driven step of the level0 family spent 40 % of its
// instructions there).  Same arithmetic per row: c = lo - x or x - hi, round8, viol_test.
// The row table sits in the cold block (through a reference to the kernel's parameter block handed to an out-of-line function it
// was read with generic loads: 38 instructions per row, 25 % of the host-driven step's time).  Inline: the SPEC instantiations,
// which every fused BASELINE config takes, compile the whole per-step-output path away (c_out == nullptr at compile time), and the
// kernels that keep it -- the host-driven step above all -- save the ABI call (the caller's live state spilled around it).
template <int KIND, typename real>
SCG_DEV uint32_t constraint_values_flat(const Cold<real>* __restrict__ cold, int n_crows, int strict, const real* __restrict__ v,
                                                        bool state_only, real* __restrict__ c_out, int64_t n, int64_t i) {
    constexpr int NX = Dims<KIND>::NX;
    bool violated = false;
    real* p = c_out + i;
    // one 16-byte load per row, four rows in flight (the launch is latency-bound: a rolled loop serialised load -> index -> store)
#pragma unroll 4
    for (int r = 0; r < n_crows; ++r) {
        const Row4<real> q = ldg4(cold->crow[r]);
        const int src = int(q.x);
        if (src >= NX && state_only) continue;
        const real c = q.y * (v[src] - q.z);   // sign * (x - bound): bound - x == -(x - bound) exactly
        p[int64_t(r) * n] = round8(c);
        violated |= viol_test(c, strict);
    }
    return violated ? 1u : 0u;
}

// The in-loop part: the violation flag of bounded / default constraints that merge into one box per dimension
// (no tolerances, one strictness -- CFG_COLD_CONSTRAINTS clear); anything else, and every call that wants the
// rounded values, is delegated to constraint_values_cold.
template <int KIND, typename real, bool PLAIN = false>
SCG_DEV void eval_constraints(const Hot<real>& h, const Cold<real>* __restrict__ cold,
                              const real (&st)[Dims<KIND>::NX], const real (&raw)[Dims<KIND>::NU],
                              bool state_only, real* __restrict__ c_out, int64_t n, int64_t i,
                              bool& violated, bool& almost, uint32_t cfg) {
    constexpr int NX = Dims<KIND>::NX, NU = Dims<KIND>::NU;
    violated = false; almost = false;
    if (c_out != nullptr && h.n_crows > 0) {
        real v[NX + NU];   // indexed by the row's source: a local-memory array, private to the call
#pragma unroll
        for (int d = 0; d < NX; ++d) v[d] = st[d];
#pragma unroll
        for (int d = 0; d < NU; ++d) v[NX + d] = raw[d];
        violated = constraint_values_flat<KIND, real>(cold, h.n_crows, h.crow_strict, v, state_only, c_out, n, i) != 0u;
        return;
    }
    if (c_out != nullptr || (!PLAIN && (cfg & CFG_COLD_CONSTRAINTS))) {
        real sl[NX], rl[NU];   // the out-of-line call takes copies, the caller's registers stay registers
#pragma unroll
        for (int d = 0; d < NX; ++d) sl[d] = st[d];
#pragma unroll
        for (int d = 0; d < NU; ++d) rl[d] = raw[d];
        const uint32_t bits = constraint_values_cold<KIND, real>(h, cold, sl, rl, state_only, c_out, n, i);
        violated = (bits & 1u) != 0u; almost = (bits & 2u) != 0u;
        return;
    }
    // one merged box: violated <=> the largest c = max(lo - x, x - hi) over all bounded dims rounds above 0
    // (unbounded dims carry lo = -inf, hi = +inf: every dim is evaluated, branch-free, at two subtractions and
    // one three-way max each; p, q, r only where something reads them)
    constexpr int NB = KIND == SCG_QUAD_3D ? 9 : NX;
    real w0 = -INFINITY, w1 = -INFINITY;   // two chains
    if (KIND == SCG_QUAD_3D && (cfg & CFG_BOX_POS)) {   // nothing but x, y, z is bounded (the other dims would contribute -inf)
        const Row4<real> lo = ldc4(h.pos_pack), hi = ldc4(h.pos_pack + 4);   // box_lo / box_hi[0, 2, 4] (scg_params.cuh)
        w0 = m_max3(w0, lo.x - st[0], st[0] - hi.x);
        w1 = m_max3(w1, lo.y - st[2], st[2] - hi.y);
        w0 = m_max3(w0, lo.z - st[4], st[4] - hi.z);
    } else {
#pragma unroll
        for (int d = 0; d < NB; ++d) {
            real& w = (d & 1) ? w1 : w0;
            w = m_max3(w, h.box_lo[d] - st[d], st[d] - h.box_hi[d]);
        }
    }
    if (NB < NX && !(KIND == SCG_QUAD_3D && (cfg & CFG_BOX_POS)) && (cfg & CFG_FULL_STATE)) {
#pragma unroll
        for (int d = NB; d < NX; ++d) w0 = m_max3(w0, h.box_lo[d] - st[d], st[d] - h.box_hi[d]);
    }
    if (!state_only && (cfg & CFG_UBOX)) {
#pragma unroll
        for (int d = 0; d < NU; ++d) {
            real& w = (d & 1) ? w1 : w0;
            w = m_max3(w, h.ubox_lo[d] - raw[d], raw[d] - h.ubox_hi[d]);
        }
    }
    // (positions-only box: the threshold rides in the fourth slot of the lower-bound row, same value)
    violated = m_max(w0, w1) * real(1e8) > ((KIND == SCG_QUAD_3D && (cfg & CFG_BOX_POS)) ? ldc4(h.pos_pack).w : h.box_thr);
}

// ------------------------------------------------------------------ gate progress + contact scan (row A10)
// Closed-form stand-ins for the two PyBullet queries of Quadrotor._get_info (quadrotor.py:1046-1092), evaluated on
// the integrated pose when the gate_geometry extension is on (in DYN the reference's own queries see a frozen
// drone).  Same shapes and constants as oracle/gate_geometry.py.  Out of line: GENERAL kernel only.
template <typename real> struct Geo {
    static constexpr real drone_r = real(0.06), drone_hl = real(0.0125);
};

template <typename real>
SCG_DEV bool segment_hits_cylinder(const real (&a)[3], const real (&b)[3], real r, real hl) {
    const real d0 = b[0] - a[0], d1 = b[1] - a[1], d2 = b[2] - a[2];
    real t0 = real(0), t1 = real(1);
    if (d2 == real(0)) {
        if (m_abs(a[2]) > hl) return false;
    } else {
        const real ta = (-hl - a[2]) / d2, tb = (hl - a[2]) / d2;
        t0 = m_max(t0, m_min(ta, tb)); t1 = m_min(t1, m_max(ta, tb));
    }
    const real qa = d0 * d0 + d1 * d1, qb = a[0] * d0 + a[1] * d1, qc = a[0] * a[0] + a[1] * a[1] - r * r;
    if (qa == real(0)) {
        if (qc > real(0)) return false;
    } else {
        const real disc = qb * qb - qa * qc;
        if (disc < real(0)) return false;
        const real s = m_sqrt(disc);
        t0 = m_max(t0, (-qb - s) / qa); t1 = m_min(t1, (-qb + s) / qa);
    }
    return t0 <= t1 && t0 < real(0.9999);
}

template <typename real> SCG_DEV bool sphere_box(const real (&p)[3], real cx, real cy, real cz, real hx, real hy, real hz) {
    const real q0 = m_max(m_abs(p[0] - cx) - hx, real(0)), q1 = m_max(m_abs(p[1] - cy) - hy, real(0)),
               q2 = m_max(m_abs(p[2] - cz) - hz, real(0));
    return q0 * q0 + q1 * q1 + q2 * q2 < Geo<real>::drone_r * Geo<real>::drone_r;
}
template <typename real> SCG_DEV bool sphere_vcyl(const real (&p)[3], real cz, real r, real hl) {
    const real dxy = m_max(m_sqrt(p[0] * p[0] + p[1] * p[1]) - r, real(0)), dz = m_max(m_abs(p[2] - cz) - hl, real(0));
    return dxy * dxy + dz * dz < Geo<real>::drone_r * Geo<real>::drone_r;
}

// (dx, dy, dyaw) of gate / obstacle `idx` for this episode: block idx of its reset stream (quadrotor.py:379-384, 415-421)
template <typename real>
SCG_DEV void pose_offsets(const RandP<real>& rp, uint32_t stream, uint32_t idx, const RngKey& key, uint32_t ep, real (&off)[3]) {
    off[0] = off[1] = off[2] = real(0);
    if (rp.distrib == SCG_RAND_NONE) return;
    const U4 w = philox4x32_10(0u, ep, stream, idx, key.seed, key.env);
    if (rp.distrib == SCG_RAND_UNIFORM) {
#pragma unroll
        for (int j = 0; j < 3; ++j) off[j] = rp.a + (rp.b - rp.a) * u01<real>(word_of(w, j));
    } else {
        real n[4];
        normal_pair<real>(w.x, w.y, &n[0], &n[1]);
        normal_pair<real>(w.z, w.w, &n[2], &n[3]);
#pragma unroll
        for (int j = 0; j < 3; ++j) off[j] = rp.a + rp.b * n[j];
    }
}

// ground plane against the drone's collision cylinder: its support along -z
template <typename real, bool FAST = false> SCG_DEV bool ground_contact(real pz, real r8) {
    const real s2 = m_max(real(0), real(1) - r8 * r8);
    const real support = Geo<real>::drone_hl * m_abs(r8) + Geo<real>::drone_r * (FAST ? sqrt_fast(s2) : m_sqrt(s2));   // (FAST: MUFU.SQRT, no slow-path call in the loop)
    return pz - support < real(0);
}

// ------------------------------------------------------------------ aerodynamic terms (row N4, dyn_aero extension)
// What the reference's PYB_* modes add through applyExternalForce (base_aviary.py:437-497), for the explicit integrator: world
// force and body torque (x, y) of one substep.  _ground_effect (:437-470): per propeller, height of its link (URDF offset rotated
// into the world) clipped at GND_EFF_H_CLIP, rpm^2 KF GND_EFF_COEFF (PROP_RADIUS / (4 h))^2 along the link's z axis, applied at
// the link, only while |roll| and |pitch| are below pi / 2.  _drag (:472-497): R (k * v), k = -DRAG_COEFF sum(2 pi rpm / 60) of
// the PREVIOUS substep's RPMs (`omega_prev`: last_clipped_action, zeros after a reset), handed over as a LINK_FRAME force, i.e.
// rotated into the world once more.  _downwash is zero for one drone per env.  Out of line, on copies: a rare extension.
template <typename real, bool FASTMATH>
__device__ __noinline__ void aero_terms(const Hot<real>& h, real pz, real vx, real vy, real vz, Att3<real> a, real m0, real m1, real m2,
                                        real m3, real omega_prev, real* __restrict__ out) {
    attitude3_complete<real>(a);
    const real* R = a.R;
    real fx = real(0), fy = real(0), fz = real(0), tx = real(0), ty = real(0);
    if (h.aero & 2) {
        const real mr[4] = {m0, m1, m2, m3};   // the rotor forces (fast motor map) or the RPMs (literal map), see Motors
        const bool upright = m_abs(a.rpy[0]) < Consts<real>::half_pi && m_abs(a.rpy[1]) < Consts<real>::half_pi;
        real fb = real(0);
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const real x = h.aero_prop_xy[i][0], y = h.aero_prop_xy[i][1];
            const real hgt = m_max(pz + (R[6] * x + R[7] * y), h.aero_h_clip);
            const real thrust = FASTMATH ? mr[i] : mr[i] * mr[i] * h.kf;
            const real q = h.aero_prop_radius / (real(4) * hgt);
            const real g = thrust * h.aero_gnd_coeff * (q * q);
            if (upright) { fb += g; tx += y * g; ty -= x * g; }
        }
        fx += R[2] * fb; fy += R[5] * fb; fz += R[8] * fb;
    }
    if (h.aero & 1) {
        const real d0 = -h.aero_drag[0] * omega_prev * vx, d1 = -h.aero_drag[1] * omega_prev * vy, d2 = -h.aero_drag[2] * omega_prev * vz;
        const real b0 = (R[0] * d0 + R[1] * d1) + R[2] * d2, b1 = (R[3] * d0 + R[4] * d1) + R[5] * d2, b2 = (R[6] * d0 + R[7] * d1) + R[8] * d2;
        fx += (R[0] * b0 + R[1] * b1) + R[2] * b2; fy += (R[3] * b0 + R[4] * b1) + R[5] * b2; fz += (R[6] * b0 + R[7] * b1) + R[8] * b2;
    }
    out[0] = fx; out[1] = fy; out[2] = fz; out[3] = tx; out[4] = ty;
}

// returns bit 0: a ray of the current gate hit the drone; bit 1: the drone touches a gate or an obstacle.
// Broad phase first: an object whose NOMINAL axis is farther than `broad2` (0.35 m + the largest pose offset) from the
// drone is skipped before its pose offsets are even drawn -- almost every (env, object) pair of a step.
template <typename real>
SCG_DEV uint32_t gate_scan_body(const Hot<real>& h, const Cold<real>* __restrict__ cold, RngKey key, uint32_t ep,
                                real px, real py, real pz, real r0, real r1, real r2, real r3, real r4, real r5,
                                real r6, real r7, real r8, uint32_t cur_gate, bool check_gate) {
    bool collided = false, passed = false;
    for (int g = 0; g < h.n_gates; ++g) {
        const real* G = cold->gates[g];
        const real nx = px - G[0], ny = py - G[1];
        if (nx * nx + ny * ny > cold->gate_broad2) continue;
        real gx = G[0], gy = G[1], cy = G[7], sy = G[8];
        if (cold->gate_rand.distrib != SCG_RAND_NONE) {
            real off[3];
            pose_offsets<real>(cold->gate_rand, STREAM_RESET_GATES, uint32_t(g), key, ep, off);
            gx += off[0]; gy += off[1];
            m_sincos(G[5] + off[2], &sy, &cy);
        }
        const real height = G[6] == real(0) ? real(1.0) : real(0.525);
        // (narrow phase of the rays first: they span +-0.15 m along the gate and +-0.1875 m in height around its centre, the drone's
        //  cylinder lies inside a sphere of radius sqrt(0.06^2 + 0.0125^2) < 0.0615 -- farther than that nothing can hit)
        const real cgx = px - gx, cgy = py - gy, cgz = pz - height;
        if (check_gate && uint32_t(g) == cur_gate && cgx * cgx + cgy * cgy <= real(0.2116 * 0.2116) && m_abs(cgz) <= real(0.2491)) {   // quadrotor.py:1060-1092
            const real dx = real(0.05) * cy, dy = real(0.05) * sy;
            const real zl = height - real(0.1875) - pz, zh = height + real(0.1875) - pz;
#pragma unroll 1
            for (int i = -3; i <= 3; ++i) {
                const real fx = gx + real(i) * dx - px, fy = gy + real(i) * dy - py;
                const real bx = r0 * fx + r3 * fy, by = r1 * fx + r4 * fy, bz = r2 * fx + r5 * fy;   // R^T (horizontal part)
                const real a[3] = {bx + r6 * zl, by + r7 * zl, bz + r8 * zl};
                const real b[3] = {bx + r6 * zh, by + r7 * zh, bz + r8 * zh};
                passed |= segment_hits_cylinder<real>(a, b, Geo<real>::drone_r, Geo<real>::drone_hl);
            }
        }
        const real ex = px - gx, ey = py - gy;
        if (ex * ex + ey * ey > real(0.35 * 0.35)) continue;
        const real p[3] = {cy * ex + sy * ey, -sy * ex + cy * ey, pz - (G[2] + height)};
        bool hit = sphere_box<real>(p, 0, 0, real(0.225), real(0.25), real(0.025), real(0.025)) |
                   sphere_box<real>(p, 0, 0, real(-0.225), real(0.25), real(0.025), real(0.025)) |
                   sphere_box<real>(p, real(0.225), 0, 0, real(0.025), real(0.025), real(0.25)) |
                   sphere_box<real>(p, real(-0.225), 0, 0, real(0.025), real(0.025), real(0.25));
        hit |= G[6] == real(0) ? sphere_vcyl<real>(p, real(-0.6), real(0.05), real(0.4))
                               : sphere_box<real>(p, 0, 0, real(-0.4), real(0.075), real(0.075), real(0.125));
        collided |= hit;
    }
    for (int o = 0; o < cold->n_obstacles; ++o) {
        const real* O = cold->obstacles[o];
        const real nx = px - O[0], ny = py - O[1];
        if (nx * nx + ny * ny > cold->obstacle_broad2) continue;
        real ox = O[0], oy = O[1], cy = O[6], sy = O[7];
        if (cold->obstacle_rand.distrib != SCG_RAND_NONE) {
            real off[3];
            pose_offsets<real>(cold->obstacle_rand, STREAM_RESET_OBSTACLES, uint32_t(o), key, ep, off);
            ox += off[0]; oy += off[1];
            m_sincos(O[5] + off[2], &sy, &cy);
        }
        const real ex = px - ox, ey = py - oy;
        if (ex * ex + ey * ey > real(0.35 * 0.35)) continue;
        const real p[3] = {cy * ex + sy * ey, -sy * ex + cy * ey, pz - (O[2] + real(0.525))};
        collided |= sphere_vcyl<real>(p, real(0), real(0.05), real(0.4)) |
                    sphere_box<real>(p, 0, 0, real(-0.4), real(0.075), real(0.075), real(0.125));
    }
    return (passed ? 1u : 0u) | (collided ? 2u : 0u);   // (the ground plane is the caller's: ground_contact)
}
// out of line for the LEAN / GENERAL kernels; the PLAIN + DS loop inlines the body (no call inside the packed loop)
template <typename real>
__device__ __noinline__ uint32_t gate_scan(const Hot<real>& h, const Cold<real>* __restrict__ cold, RngKey key, uint32_t ep,
                                           real px, real py, real pz, real r0, real r1, real r2, real r3, real r4, real r5,
                                           real r6, real r7, real r8, uint32_t cur_gate, bool check_gate) {
    return gate_scan_body<real>(h, cold, key, ep, px, py, pz, r0, r1, r2, r3, r4, r5, r6, r7, r8, cur_gate, check_gate);
}

// ------------------------------------------------------------------ disturbances (row A5)
// Additive, applied in YAML order (DisturbanceList.apply, disturbances.py:69-79).
// t = ctrl_step_counter (pre-increment); index = Philox counter word 0 for this call.
template <typename real, int DIM>
__device__ __noinline__ void apply_channel(const Cold<real>* __restrict__ cold, int ch, int n_dist, real (&v)[DIM],
                           int32_t t, real t_sim, const RngKey& key, uint32_t index, uint32_t episode,
                           const int32_t* __restrict__ offsets, int64_t n, int64_t i) {
    for (int k = 0; k < n_dist; ++k) {
        const DistP<real>& D = cold->dist[ch][k];
        real noise[DIM];
        if (D.kind == SCG_DIST_IMPULSE || D.kind == SCG_DIST_STEP) {
            const int32_t off = D.step_offset >= 0 ? D.step_offset : offsets[int64_t(D.slot) * n + i];
            real s = real(0);
            if (t >= off) {
                if (D.kind == SCG_DIST_STEP) s = D.magnitude;
                else {
                    const int32_t peak = static_cast<int32_t>(real(off) + D.half_duration);  // int(offset + duration / 2)
                    const int32_t kk = t >= peak ? t - peak : peak - t;
                    s = (real(kk) < D.half_duration) ? D.magnitude * m_pow(D.decay_rate, real(kk)) : real(0);
                }
            }
#pragma unroll
            for (int d = 0; d < DIM; ++d) noise[d] = s;
        } else if (D.inject) {
#pragma unroll
            for (int d = 0; d < DIM; ++d) noise[d] = D.inject[int64_t(d) * n + i];
            if (D.kind == SCG_DIST_PERIODIC) {
#pragma unroll
                for (int d = 0; d < DIM; ++d) noise[d] = D.scale * m_sin(D.omega * t_sim + noise[d]);
            }
        } else {
            const uint32_t stream = STREAM_DISTURB_BASE + 8u * ch + k;
            if (D.kind == SCG_DIST_WHITE) {
                draw_normals<real, DIM>(key, index, episode, stream, noise);
#pragma unroll
                for (int d = 0; d < DIM; ++d) noise[d] *= D.std[d];
            } else {
                draw_uniforms<real, DIM>(key, index, episode, stream, noise);
                if (D.kind == SCG_DIST_UNIFORM) {
#pragma unroll
                    for (int d = 0; d < DIM; ++d) noise[d] = D.low[d] + D.span[d] * noise[d];
                } else {
#pragma unroll
                    for (int d = 0; d < DIM; ++d)
                        noise[d] = D.scale * m_sin(D.omega * t_sim + (-Consts<real>::pi + Consts<real>::two_pi * noise[d]));
                }
            }
        }
        if (D.has_mask) {
#pragma unroll
            for (int d = 0; d < DIM; ++d) noise[d] *= D.mask[d];
        }
#pragma unroll
        for (int d = 0; d < DIM; ++d) v[d] += noise[d];
    }
}

}  // namespace scg
