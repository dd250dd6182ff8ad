"""PyBullet's quaternion helpers, restated in plain Python floats (fp64).

TEST INFRASTRUCTURE (see oracle/__init__.py).

pybullet ``^3.2.0`` (reference pyproject.toml:26) is not vendored in
/root/reference and not installed here, so the three C helpers the DYN path
calls are restated from the published bullet3 source
(``examples/pybullet/pybullet.c``; quaternion order is [x, y, z, w]):

* ``getQuaternionFromEuler``  -- call sites base_aviary.py:567, quadrotor.py:490
* ``getEulerFromQuaternion``  -- call site  base_aviary.py:372
* ``getMatrixFromQuaternion`` -- call sites base_aviary.py:538, quadrotor.py:862

Parity status: restated, not executed ("unpinned" for these three helpers).
"""
import math

GIMBAL_SARG = 0.99999


def quat_from_euler(roll, pitch, yaw):
    """pybullet_getQuaternionFromEuler: half-angle products, then normalise."""
    phi, the, psi = roll / 2.0, pitch / 2.0, yaw / 2.0
    sph, cph = math.sin(phi), math.cos(phi)
    sth, cth = math.sin(the), math.cos(the)
    sps, cps = math.sin(psi), math.cos(psi)
    x = sph * cth * cps - cph * sth * sps
    y = cph * sth * cps + sph * cth * sps
    z = cph * cth * sps - sph * sth * cps
    w = cph * cth * cps + sph * sth * sps
    n = math.sqrt(x * x + y * y + z * z + w * w)
    return (x / n, y / n, z / n, w / n)


def euler_from_quat(q):
    """pybullet_getEulerFromQuaternion incl. the |sarg| >= 0.99999 gimbal branch."""
    x, y, z, w = q
    sqx, sqy, sqz, squ = x * x, y * y, z * z, w * w
    sarg = -2.0 * (x * z - w * y)
    if sarg <= -GIMBAL_SARG:
        return (0.0, -0.5 * math.pi, 2.0 * math.atan2(x, -y))
    if sarg >= GIMBAL_SARG:
        return (0.0, 0.5 * math.pi, 2.0 * math.atan2(-x, y))
    return (math.atan2(2.0 * (y * z + w * x), squ - sqx - sqy + sqz),
            math.asin(sarg),
            math.atan2(2.0 * (x * y + w * z), squ + sqx - sqy - sqz))


def matrix_from_quat(q):
    """pybullet_getMatrixFromQuaternion (b3Matrix3x3::setRotation), row-major 3x3."""
    x, y, z, w = q
    d = x * x + y * y + z * z + w * w
    s = 2.0 / d
    xs, ys, zs = x * s, y * s, z * s
    wx, wy, wz = w * xs, w * ys, w * zs
    xx, xy, xz = x * xs, x * ys, x * zs
    yy, yz, zz = y * ys, y * zs, z * zs
    return ((1.0 - (yy + zz), xy - wz, xz + wy),
            (xy + wz, 1.0 - (xx + zz), yz - wx),
            (xz - wy, yz + wx, 1.0 - (xx + yy)))
