"""TEST INFRASTRUCTURE (oracle) -- the DSL CF2.x PID of the reference restated in numpy fp64
(envs/gym_pybullet_drones/quadrotor_utils.py:70-278, `PIDController`), the action source of the reference's
closed loops (competition/competition_utils.py:338-356 `thrusts`, controllers/pid/pid.py:66-150).

Pinned by tests/golden/pid_controller.npz: the reference's own class executed under tests/golden/ref_stubs.py
(pybullet's three quaternion helpers restated; scipy is the real one) on a recorded state sequence.

scipy's Rotation calls are written out: `from_matrix(M).as_euler('XYZ')` is the intrinsic X-Y-Z factorisation
M = Rx(a) Ry(b) Rz(c); the reference then rebuilds the rotation from those angles through a quaternion whose
components it mislabels consistently (`w, x, y, z = as_quat()` followed by `from_quat([w, x, y, z])`), i.e. it gets
M back."""
import math

import numpy as np

from . import bullet_math as bm

MIXER_MATRIX = np.array([[.5, -.5, 1], [.5, .5, -1], [-.5, .5, 1], [-.5, -.5, -1]])


def euler_xyz_intrinsic(M):
    """Rotation.from_matrix(M).as_euler('XYZ') away from gimbal lock."""
    b = math.asin(max(-1.0, min(1.0, M[0, 2])))
    a = math.atan2(-M[1, 2], M[2, 2])
    c = math.atan2(-M[0, 1], M[0, 0])
    return np.array([a, b, c])


def matrix_from_euler_xyz_intrinsic(e):
    ca, sa, cb, sb, cc, sc = math.cos(e[0]), math.sin(e[0]), math.cos(e[1]), math.sin(e[1]), math.cos(e[2]), math.sin(e[2])
    Rx = np.array([[1, 0, 0], [0, ca, -sa], [0, sa, ca]])
    Ry = np.array([[cb, 0, sb], [0, 1, 0], [-sb, 0, cb]])
    Rz = np.array([[cc, -sc, 0], [sc, cc, 0], [0, 0, 1]])
    return Rx @ Ry @ Rz


class PIDOracle:
    def __init__(self, g=9.8, m=0.027, kf=3.16e-10, km=7.94e-12, pwm2rpm_scale=0.2685, pwm2rpm_const=4070.3,
                 min_pwm=20000, max_pwm=65535, p_coeff_for=(.4, .4, 1.25), i_coeff_for=(.05, .05, .05),
                 d_coeff_for=(.2, .2, .5), p_coeff_tor=(70000., 70000., 60000.), i_coeff_tor=(.0, .0, 500.),
                 d_coeff_tor=(20000., 20000., 12000.)):
        self.GRAVITY, self.KF, self.KM = g * m, kf, km
        self.P_FOR, self.I_FOR, self.D_FOR = (np.array(v, dtype=float) for v in (p_coeff_for, i_coeff_for, d_coeff_for))
        self.P_TOR, self.I_TOR, self.D_TOR = (np.array(v, dtype=float) for v in (p_coeff_tor, i_coeff_tor, d_coeff_tor))
        self.SCALE, self.CONST, self.MIN_PWM, self.MAX_PWM = pwm2rpm_scale, pwm2rpm_const, min_pwm, max_pwm
        self.reset()

    def reset(self):   # quadrotor_utils.py:124-135
        self.control_counter = 0
        self.last_rpy = np.zeros(3)
        self.integral_pos_e = np.zeros(3)
        self.integral_rpy_e = np.zeros(3)

    def compute_control(self, dt, cur_pos, cur_quat, cur_vel, target_pos, target_rpy=np.zeros(3), target_vel=np.zeros(3),
                        target_rpy_rates=np.zeros(3)):
        """quadrotor_utils.py:137-186 -> rpm[4]."""
        self.control_counter += 1
        R = np.array(bm.matrix_from_quat(cur_quat)).reshape(3, 3)
        # _compute_force_and_euler (:188-237)
        pos_e = np.asarray(target_pos, dtype=float) - np.asarray(cur_pos, dtype=float)
        vel_e = np.asarray(target_vel, dtype=float) - np.asarray(cur_vel, dtype=float)
        self.integral_pos_e = np.clip(self.integral_pos_e + pos_e * dt, -2., 2.)
        self.integral_pos_e[2] = np.clip(self.integral_pos_e[2], -0.15, .15)
        target_thrust = self.P_FOR * pos_e + self.I_FOR * self.integral_pos_e + self.D_FOR * vel_e + np.array([0, 0, self.GRAVITY])
        scalar_thrust = max(0., float(np.dot(target_thrust, R[:, 2])))
        thrust = (math.sqrt(scalar_thrust / (4 * self.KF)) - self.CONST) / self.SCALE
        z_ax = target_thrust / np.linalg.norm(target_thrust)
        x_c = np.array([math.cos(target_rpy[2]), math.sin(target_rpy[2]), 0])
        y_ax = np.cross(z_ax, x_c) / np.linalg.norm(np.cross(z_ax, x_c))
        x_ax = np.cross(y_ax, z_ax)
        target_euler = euler_xyz_intrinsic(np.vstack([x_ax, y_ax, z_ax]).transpose())
        # _compute_rpms (:239-278)
        cur_rpy = np.array(bm.euler_from_quat(cur_quat))
        M = matrix_from_euler_xyz_intrinsic(target_euler)
        E = M.transpose() @ R - R.transpose() @ M
        rot_e = np.array([E[2, 1], E[0, 2], E[1, 0]])
        rpy_rates_e = np.asarray(target_rpy_rates, dtype=float) - (cur_rpy - self.last_rpy) / dt
        self.last_rpy = cur_rpy
        self.integral_rpy_e = np.clip(self.integral_rpy_e - rot_e * dt, -1500., 1500.)
        self.integral_rpy_e[0:2] = np.clip(self.integral_rpy_e[0:2], -1., 1.)
        torques = np.clip(-self.P_TOR * rot_e + self.D_TOR * rpy_rates_e + self.I_TOR * self.integral_rpy_e, -3200, 3200)
        pwm = np.clip(thrust + MIXER_MATRIX @ torques, self.MIN_PWM, self.MAX_PWM)
        return self.SCALE * pwm + self.CONST

    def thrusts_3d(self, dt, obs, target_pos, target_vel):
        """competition_utils.py:338-356: obs = [x, xd, y, yd, z, zd, phi, theta, psi, p, q, r] -> 4 thrusts."""
        rpm = self.compute_control(dt, [obs[0], obs[2], obs[4]], bm.quat_from_euler(obs[6], obs[7], obs[8]),
                                   [obs[1], obs[3], obs[5]], target_pos, target_vel=target_vel)
        return self.KF * rpm ** 2

    def thrusts_2d(self, dt, obs, target_pos, target_vel):
        """controllers/pid/pid.py:97-141: obs = [x, xd, z, zd, theta, thetad] -> [f0 + f3, f1 + f2]."""
        rpm = self.compute_control(dt, [obs[0], 0, obs[2]], bm.quat_from_euler(0.0, obs[4], 0.0), [obs[1], 0, obs[3]],
                                   target_pos, target_vel=target_vel)
        f = self.KF * rpm ** 2
        return np.array([f[0] + f[3], f[1] + f[2]])
