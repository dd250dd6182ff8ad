"""Reference-trajectory tables (host side, once per env config).

Vectorised restatement of ``BenchmarkEnv._generate_trajectory`` and its figure8 / circle /
square samplers (benchmark_env.py:465-674) and of ``transform_trajectory`` /
``projection_matrix`` (math_and_models/transformations.py:52-123).  The resulting X_GOAL
table is uploaded once and shared read-only by all envs.
"""
import numpy as np

_AXES = ("x", "y", "z")


def planar_reference(traj_type="figure8", traj_length=10.0, num_cycles=1, traj_plane="xy",
                     position_offset=(0.0, 0.0), scaling=1.0, sample_time=0.01):
    """Returns (pos[T,3], vel[T,3], speed[T,1]) sampled at ``np.arange(0, traj_length, sample_time)``."""
    if traj_type not in ("circle", "square", "figure8"):
        raise ValueError("Trajectory type should be one of [circle, square, figure8].")
    ok = len(traj_plane) >= 2 and traj_plane[0] in _AXES and traj_plane[1] in _AXES and traj_plane[0] != traj_plane[1]
    if not ok:
        raise ValueError("Trajectory plane should be in form of ab, where a and b can be {x, y, z}.")
    ia, ib = _AXES.index(traj_plane[0]), _AXES.index(traj_plane[1])
    t = np.arange(0, traj_length, sample_time)
    period = traj_length / num_cycles
    w = 2.0 * np.pi / period
    if traj_type == "figure8":
        s, c = np.sin(w * t), np.cos(w * t)
        a, b = scaling * s, scaling * s * c
        da, db = scaling * w * c, scaling * w * (c ** 2 - s ** 2)
    elif traj_type == "circle":
        s, c = np.sin(w * t), np.cos(w * t)
        a, b = scaling * c, scaling * s
        da, db = -scaling * w * s, scaling * w * c
    else:
        seg_period = period / 4.0
        speed = scaling / seg_period
        cyc = t % period
        along = speed * (cyc % seg_period)
        seg = np.floor(cyc / seg_period).astype(int)
        zero, full, v = np.zeros_like(t), np.full_like(t, scaling), np.full_like(t, speed)
        a = np.select([seg == 0, seg == 1, seg == 2, seg == 3], [zero, -along, -full, -scaling + along])
        b = np.select([seg == 0, seg == 1, seg == 2, seg == 3], [along, full, scaling - along, zero])
        da = np.select([seg == 0, seg == 1, seg == 2, seg == 3], [zero, -v, zero, v])
        db = np.select([seg == 0, seg == 1, seg == 2, seg == 3], [v, zero, -v, zero])
    pos, vel = np.zeros((len(t), 3)), np.zeros((len(t), 3))
    pos[:, ia], vel[:, ia] = a + position_offset[0], da
    pos[:, ib], vel[:, ib] = b + position_offset[1], db
    return pos, vel, np.linalg.norm(vel, axis=1, keepdims=True)


def project_onto_plane(pos, vel, point, normal):
    """Orthogonal projection of a planar reference onto the plane (point, normal); the velocity
    goes through the same affine map, homogeneous 1 included, as the reference does."""
    n = np.asarray(normal[:3], dtype=np.float64)
    n = n / np.sqrt(np.sum(n * n))
    M = np.identity(4)
    M[:3, :3] -= np.outer(n, n)
    M[:3, 3] = np.dot(np.asarray(point[:3], dtype=np.float64), n) * n
    ones = np.ones((pos.shape[0], 1))
    return (np.concatenate([pos, ones], -1) @ M.T)[:, :3], (np.concatenate([vel, ones], -1) @ M.T)[:, :3]
