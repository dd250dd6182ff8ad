// scg_pid.cuh -- the DSL CF2.x PID (quadrotor_utils.py:70-278, `PIDController`) as an on-device action source for
// closed-loop fused rollouts: position PID -> target thrust vector and attitude ->
// attitude PID -> mixer -> PWM clip -> RPM -> per-rotor thrust KF rpm^2 (competition_utils.py:338-356 for 3D,
// controllers/pid/pid.py:97-141 for 2D).  One call per env and control step; the controller's memory (integral of the
// position error, last Euler angles, integral of the rotation error: 9 values) lives in a caller-owned [9][N] buffer.
#pragma once
#include "scg_params.cuh"
#include "scg_math.cuh"
#include "scg_env.cuh"

namespace scg {

// pos / vel / rpy: the env's current state (world frame; rpy = the canonical angles the reference reads back from
// PyBullet); tpos / tvel: the reference point.  act: 4 rotor thrusts (3D) or the two pair sums (2D).
// mem: this env's nine values at stride `ld` -- a per-thread column of shared memory inside a fused rollout (rollout_range loads
// it from the caller's [9][N] buffer when a range starts and stores it back when it ends).  Inline: the gains are then
// constant-bank operands (through a reference handed to an out-of-line function they are generic loads).
// ATT: the env's attitude mode -- the fast modes take their sin / cos (|rpy| <= pi) like the integrator does.
template <int KIND, typename real, int ATT>
SCG_DEV void pid_action(const PidP<real>& P, const real (&pos)[3], const real (&vel)[3], const real (&rpy)[3], const real (&tpos)[3],
                        const real (&tvel)[3], real* __restrict__ mem, int ld, real (&act)[Dims<KIND>::NU]) {
    real ipos[3], last[3], irot[3];
#pragma unroll
    for (int j = 0; j < 3; ++j) { ipos[j] = mem[j * ld]; last[j] = mem[(3 + j) * ld]; irot[j] = mem[(6 + j) * ld]; }
    // body-to-world rotation of the current attitude: getMatrixFromQuaternion(getQuaternionFromEuler(rpy)) = Rz Ry Rx
    real sph, cph, sth, cth, sps, cps;
    if constexpr (ATT == SCG_ATT_ROUNDTRIP) { m_sincos(rpy[0], &sph, &cph); m_sincos(rpy[1], &sth, &cth); m_sincos(rpy[2], &sps, &cps); }
    else { att_sincos<ATT>(rpy[0], &sph, &cph); att_sincos<ATT>(rpy[1], &sth, &cth); att_sincos<ATT>(rpy[2], &sps, &cps); }
    const real R[9] = {cps * cth, cps * sth * sph - sps * cph, cps * sth * cph + sps * sph,
                       sps * cth, sps * sth * sph + cps * cph, sps * sth * cph - cps * sph,
                       -sth, cth * sph, cth * cph};
    // ---- _compute_force_and_euler (quadrotor_utils.py:188-237)
    real tt[3];
#pragma unroll
    for (int j = 0; j < 3; ++j) {
        const real pe = tpos[j] - pos[j], ve = tvel[j] - vel[j];
        ipos[j] = m_clip(ipos[j] + pe * P.dt, real(-2), real(2));
        if (j == 2) ipos[j] = m_clip(ipos[j], real(-0.15), real(0.15));
        tt[j] = (P.p_for[j] * pe + P.i_for[j] * ipos[j]) + P.d_for[j] * ve;
    }
    tt[2] += P.gravity;
    const real scalar = m_max(real(0), (tt[0] * R[2] + tt[1] * R[5]) + tt[2] * R[8]);
    // fp32 fast attitude modes: MUFU square root / reciprocal square root (a few ulp, like the integrator's own fast map)
    constexpr bool FASTM = sizeof(real) == 4 && ATT != SCG_ATT_ROUNDTRIP;
    const real thrust = ((FASTM ? sqrt_fast(scalar * P.inv_4kf) : m_sqrt(scalar * P.inv_4kf)) - P.pwm_const) * P.inv_pwm_scale;
    const real n2 = (tt[0] * tt[0] + tt[1] * tt[1]) + tt[2] * tt[2];
    const real inv_n = FASTM ? m_rsqrt(n2) : real(1) / m_sqrt(n2);
    const real z[3] = {tt[0] * inv_n, tt[1] * inv_n, tt[2] * inv_n};
    // target yaw 0: x_c = (1, 0, 0); y = z x x_c / |.|, x = y x z; M = [x y z] (columns)
    real y[3] = {real(0), z[2], -z[1]};
    const real y2 = y[1] * y[1] + y[2] * y[2];
    const real inv_y = FASTM ? m_rsqrt(y2) : real(1) / m_sqrt(y2);
    y[1] *= inv_y; y[2] *= inv_y;
    const real x[3] = {y[1] * z[2] - y[2] * z[1], y[2] * z[0] - y[0] * z[2], y[0] * z[1] - y[1] * z[0]};
    const real M[9] = {x[0], y[0], z[0], x[1], y[1], z[1], x[2], y[2], z[2]};
    // ---- _compute_rpms (quadrotor_utils.py:239-278): E = M^T R - R^T M, rot_e = (E21, E02, E10)
    auto mtr = [&](int a, int b) { return (M[0 + a] * R[0 + b] + M[3 + a] * R[3 + b]) + M[6 + a] * R[6 + b]; };   // (M^T R)[a][b]
    const real rot_e[3] = {mtr(2, 1) - mtr(1, 2), mtr(0, 2) - mtr(2, 0), mtr(1, 0) - mtr(0, 1)};
    real tq[3];
#pragma unroll
    for (int j = 0; j < 3; ++j) {
        const real rate_e = -(rpy[j] - last[j]) * P.inv_dt;   // target rates are 0
        last[j] = rpy[j];
        irot[j] = m_clip(irot[j] - rot_e[j] * P.dt, real(-1500), real(1500));
        if (j < 2) irot[j] = m_clip(irot[j], real(-1), real(1));
        tq[j] = m_clip((-P.p_tor[j] * rot_e[j] + P.d_tor[j] * rate_e) + P.i_tor[j] * irot[j], real(-3200), real(3200));
    }
    // MIXER_MATRIX (quadrotor_utils.py:121)
    const real mix[4] = {real(0.5) * tq[0] - real(0.5) * tq[1] + tq[2], real(0.5) * tq[0] + real(0.5) * tq[1] - tq[2],
                         -real(0.5) * tq[0] + real(0.5) * tq[1] + tq[2], -real(0.5) * tq[0] - real(0.5) * tq[1] - tq[2]};
    real f[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        const real rpm = P.pwm_scale * m_clip(thrust + mix[j], P.min_pwm, P.max_pwm) + P.pwm_const;
        f[j] = P.kf * rpm * rpm;
    }
    if constexpr (KIND == SCG_QUAD_3D) { act[0] = f[0]; act[1] = f[1]; act[2] = f[2]; act[3] = f[3]; }
    else { act[0] = f[0] + f[3]; act[1] = f[1] + f[2]; }
#pragma unroll
    for (int j = 0; j < 3; ++j) { mem[j * ld] = ipos[j]; mem[(3 + j) * ld] = last[j]; mem[(6 + j) * ld] = irot[j]; }
}

}  // namespace scg
