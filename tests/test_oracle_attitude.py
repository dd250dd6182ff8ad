"""The wrap / fold / gimbal-band arithmetic of the SCG_ATT_FAST attitude path (csrc/scg_env.cuh,
attitude3) restated in numpy and checked against the literal bullet3 round trip
eulerFromQuat(quatFromEuler(.)) of oracle/bullet_math.py -- the algebra the CUDA kernel relies on.
(The kernel itself is checked against the golden vectors in test_gpu_golden.py.)"""
import numpy as np

from oracle.bullet_math import euler_from_quat, matrix_from_quat, quat_from_euler

PI = np.pi


def wrap(x, period=2 * PI):
    return x - period * np.rint(x / period)


def fast_attitude(pre):
    """-> (rpy, third column of R) the way attitude3<FAST> computes them (fp32 band formula)."""
    thw = wrap(pre[1])
    th, ph, ps = thw, pre[0], pre[2]
    folded = abs(th) > PI / 2
    if folded:
        th = (PI if th > 0 else -PI) - th
        ph, ps = ph + PI, ps + PI
    ph, ps = wrap(ph), wrap(ps)
    sth, cth, sph, cph, sps, cps = np.sin(th), np.cos(th), np.sin(ph), np.cos(ph), np.sin(ps), np.cos(ps)
    col = np.array([cps * sth * cph + sps * sph, sps * sth * cph - cps * sph, cth * cph])
    rpy = np.array([ph, th, ps])
    if not abs(sth) < 0.99999:
        up = sth > 0
        r = 0.5 * (thw - (PI / 2 if up else -PI / 2))
        sr, cr = (-sph, -cph) if folded else (sph, cph)
        u = r * cr if up else -(r * cr)
        yaw = (pre[2] - pre[0] if up else pre[2] + pre[0]) + 2 * r * sr * (1 - u + u * u)
        rpy = np.array([0.0, PI / 2 if up else -PI / 2, wrap(yaw, 4 * PI)])
    return rpy, col


def literal(pre):
    q = quat_from_euler(*pre)
    return np.array(euler_from_quat(q)), np.array(matrix_from_quat(q)).reshape(3, 3)[:, 2]


def ang_err(a, b, period):
    d = np.abs(a - b) % period
    return np.minimum(d, period - d)


def _check(pre):
    rpy, col = fast_attitude(pre)
    rpy_l, col_l = literal(pre)
    in_band = rpy_l[0] == 0.0 and abs(abs(rpy_l[1]) - PI / 2) < 1e-15
    assert np.max(np.abs(col - col_l)) < 1e-12
    if in_band:
        assert rpy[0] == 0.0 and rpy[1] == rpy_l[1]
        assert ang_err(rpy[2], rpy_l[2], 4 * PI) < 1e-8      # linearised terms: < 3e-9
    else:
        # close to the band roll and yaw are individually ill-conditioned (1 / cos(pitch))
        tol = 1e-12 / max(abs(np.cos(rpy_l[1])), 1e-3)
        assert np.max(ang_err(rpy, rpy_l, 2 * PI)) < tol
    return in_band


def test_fast_attitude_matches_round_trip_everywhere():
    rng = np.random.default_rng(0)
    for _ in range(20000):
        pre = rng.uniform(-PI - 1.2, PI + 1.2, size=3)
        pre[1] = rng.uniform(-2.8, 2.8)
        _check(pre)


def test_gimbal_band_formula():
    rng = np.random.default_rng(1)
    hits = 0
    for _ in range(40000):
        sgn = rng.choice([-1.0, 1.0])
        pre = np.array([rng.uniform(-PI - 1.2, PI + 1.2), sgn * PI / 2 + rng.uniform(-4.4e-3, 4.4e-3),
                        rng.uniform(-PI - 1.2, PI + 1.2)])
        hits += _check(pre)
    assert hits > 30000


def test_band_is_absorbing_for_slow_pitch_rates():
    """Once snapped to +-pi/2 the next substep stays inside the band while |q| dt < acos(0.99999)."""
    dt, q = 1 / 240, 0.9   # rad/s
    rpy = np.array([0.3, PI / 2 - 1e-3, 1.0])
    for _ in range(50):
        rpy, _ = literal(rpy + dt * np.array([0.2, q, -0.1]))
        assert rpy[0] == 0.0 and rpy[1] == PI / 2
