"""TEST INFRASTRUCTURE (oracle) -- analytic stand-in for the PyBullet queries of Quadrotor._get_info.

The reference asks PyBullet two geometric questions per control step (quadrotor.py:1046-1092):
  * rayTestBatch over 7 vertical segments through the current gate's opening: "did any of them hit a body
    (fraction < 0.9999)" -- in practice the drone's collision cylinder (cf2x.urdf: radius 0.06, length 0.025);
  * getContactPoints(drone, gate / obstacle / plane): "is the drone touching anything".
In DYN physics PyBullet never moves the drone, so the reference's answers are frozen at the initial pose (SURVEY.md
F5); the `gate_geometry` extension answers the same questions from the integrated pose with closed-form geometry.
**Parity unpinned**: Bullet's GJK margins are not reproduced; the shapes and dimensions are the URDFs'
(assets/portal.urdf, low_portal.urdf, obstacle.urdf, cf2x.urdf).

Shapes.  Drone: for the rays and the ground plane its collision cylinder (axis = body z); against gate frames and
obstacles a sphere of radius 0.06.  Gate (yaw only), centred at (x, y, z + height): four bars .5 x .05 x .05 forming a
0.45 m square opening, plus the support (tall: cylinder r 0.05, length 0.8 centred 0.6 below; low: box .15 x .15 x .25
centred 0.4 below).  Obstacle at (x, y, z + 0.525): cylinder r 0.05, length 0.8, plus the same box 0.4 below.
"""
import numpy as np

DRONE_R, DRONE_HL = 0.06, 0.0125
HALF_LENGTH = 0.1875          # quadrotor.py:1069
RAY_STEP = 0.05               # quadrotor.py:1070-1071
BROAD_XY2 = 0.35 ** 2         # beyond this horizontal distance from an object's axis nothing can touch


def segment_hits_cylinder(a, b, r, hl):
    """Solid cylinder x^2 + y^2 <= r^2, |z| <= hl against the segment a -> b (cylinder frame); the entry
    parameter must be below 0.9999 like the reference's hit fraction."""
    d = b - a
    t0, t1 = 0.0, 1.0
    if d[2] == 0.0:
        if abs(a[2]) > hl:
            return False
    else:
        ta, tb = (-hl - a[2]) / d[2], (hl - a[2]) / d[2]
        t0, t1 = max(t0, min(ta, tb)), min(t1, max(ta, tb))
    qa = d[0] * d[0] + d[1] * d[1]
    qb = a[0] * d[0] + a[1] * d[1]
    qc = a[0] * a[0] + a[1] * a[1] - r * r
    if qa == 0.0:
        if qc > 0.0:
            return False
    else:
        disc = qb * qb - qa * qc
        if disc < 0.0:
            return False
        s = np.sqrt(disc)
        t0, t1 = max(t0, (-qb - s) / qa), min(t1, (-qb + s) / qa)
    return bool(t0 <= t1 and t0 < 0.9999)


def gate_rays_hit_drone(gx, gy, rot, height, pos, R):
    """The 7 segments of quadrotor.py:1072-1079 against the drone's collision cylinder.  R: body-to-world."""
    pos = np.asarray(pos, dtype=float)
    R = np.asarray(R, dtype=float).reshape(3, 3)
    dx, dy = RAY_STEP * np.cos(rot), RAY_STEP * np.sin(rot)
    for i in (0, 1, -1, 2, -2, 3, -3):
        fr = np.array([gx + i * dx, gy + i * dy, height - HALF_LENGTH])
        to = np.array([gx + i * dx, gy + i * dy, height + HALF_LENGTH])
        if segment_hits_cylinder(R.T @ (fr - pos), R.T @ (to - pos), DRONE_R, DRONE_HL):
            return True
    return False


def _sphere_box(p, centre, half):
    q = np.maximum(np.abs(p - np.asarray(centre)) - np.asarray(half), 0.0)
    return float(q @ q) < DRONE_R ** 2


def _sphere_vcyl(p, centre, r, hl):
    q = p - np.asarray(centre)
    dxy = max(np.hypot(q[0], q[1]) - r, 0.0)
    dz = max(abs(q[2]) - hl, 0.0)
    return dxy * dxy + dz * dz < DRONE_R ** 2


def _local(pos, x, y, z, yaw):
    c, s = np.cos(yaw), np.sin(yaw)
    dx, dy = pos[0] - x, pos[1] - y
    return np.array([c * dx + s * dy, -s * dx + c * dy, pos[2] - z])


def drone_collides(pos, R, gates_eff, gate_types, obstacles_eff):
    """First-hit flag in the reference's scan order: gates, obstacles, ground plane (quadrotor.py:1046-1057).
    gates_eff[i] = [x, y, z + height, r, p, yaw] (EFFECTIVE_GATES_POSITIONS), obstacles_eff likewise (z + 0.525)."""
    pos = np.asarray(pos, dtype=float)
    R = np.asarray(R, dtype=float).reshape(3, 3)
    for g, typ in zip(gates_eff, gate_types):
        if (pos[0] - g[0]) ** 2 + (pos[1] - g[1]) ** 2 > BROAD_XY2:
            continue
        p = _local(pos, g[0], g[1], g[2], g[5])
        hit = (_sphere_box(p, (0, 0, 0.225), (0.25, 0.025, 0.025)) or _sphere_box(p, (0, 0, -0.225), (0.25, 0.025, 0.025)) or
               _sphere_box(p, (0.225, 0, 0), (0.025, 0.025, 0.25)) or _sphere_box(p, (-0.225, 0, 0), (0.025, 0.025, 0.25)))
        if typ == 0:
            hit = hit or _sphere_vcyl(p, (0, 0, -0.6), 0.05, 0.4)
        else:
            hit = hit or _sphere_box(p, (0, 0, -0.4), (0.075, 0.075, 0.125))
        if hit:
            return True
    for o in obstacles_eff:
        if (pos[0] - o[0]) ** 2 + (pos[1] - o[1]) ** 2 > BROAD_XY2:
            continue
        p = _local(pos, o[0], o[1], o[2], o[5])
        if _sphere_vcyl(p, (0, 0, 0), 0.05, 0.4) or _sphere_box(p, (0, 0, -0.4), (0.075, 0.075, 0.125)):
            return True
    # ground plane z = 0 against the collision cylinder: its support along -z
    support = DRONE_HL * abs(R[2, 2]) + DRONE_R * np.sqrt(max(0.0, 1.0 - R[2, 2] ** 2))
    return bool(pos[2] - support < 0.0)


VISIBILITY_RANGE = 0.45       # quadrotor.py:1094


def _dist_box(p, centre, half):
    q = np.maximum(np.abs(p - np.asarray(centre)) - np.asarray(half), 0.0)
    return float(np.sqrt(q @ q))


def _dist_vcyl(p, centre, r, hl):
    q = p - np.asarray(centre)
    dxy = max(np.hypot(q[0], q[1]) - r, 0.0)
    dz = max(abs(q[2]) - hl, 0.0)
    return float(np.sqrt(dxy * dxy + dz * dz))


def gate_in_range(pos, gate_eff, gate_type):
    """getClosestPoints(gate, drone, distance=0.45) non-empty (quadrotor.py:1095-1100): distance from the drone sphere to
    the gate's bars and support."""
    p = _local(np.asarray(pos, dtype=float), gate_eff[0], gate_eff[1], gate_eff[2], gate_eff[5])
    d = min(_dist_box(p, (0, 0, 0.225), (0.25, 0.025, 0.025)), _dist_box(p, (0, 0, -0.225), (0.25, 0.025, 0.025)),
            _dist_box(p, (0.225, 0, 0), (0.025, 0.025, 0.25)), _dist_box(p, (-0.225, 0, 0), (0.025, 0.025, 0.25)))
    d = min(d, _dist_vcyl(p, (0, 0, -0.6), 0.05, 0.4) if gate_type == 0 else _dist_box(p, (0, 0, -0.4), (0.075, 0.075, 0.125)))
    return bool(d - DRONE_R <= VISIBILITY_RANGE)
