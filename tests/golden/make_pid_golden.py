"""Generates tests/golden/pid_controller.npz: the reference's own PIDController
(/root/reference/safe_control_gym/envs/gym_pybullet_drones/quadrotor_utils.py:70-278) executed in THIS container under
tests/golden/ref_stubs.py (pybullet's quaternion helpers restated, real scipy) on a random but reproducible sequence
of states and targets.  Not read on the GPU box; the fixture is committed.

    python tests/golden/make_pid_golden.py
"""
import importlib.util
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
import ref_stubs  # noqa: E402,F401  (installs the stand-in modules)

REF = "/root/reference/safe_control_gym/envs/gym_pybullet_drones/quadrotor_utils.py"


def main():
    ref_stubs.install()
    spec = importlib.util.spec_from_file_location("ref_quadrotor_utils", REF)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    import pybullet as p
    rng = np.random.default_rng(2024)
    T, dt = 60, 1 / 60
    ctrl = mod.PIDController()
    pos = rng.uniform(-1, 1, size=(T, 3)) * [1, 1, 0.5] + [0, 0, 1]
    vel = rng.normal(0, 0.3, size=(T, 3))
    rpy = rng.uniform(-0.5, 0.5, size=(T, 3))
    tpos = rng.uniform(-1, 1, size=(T, 3)) * [1, 1, 0.5] + [0, 0, 1]
    tvel = rng.normal(0, 0.2, size=(T, 3))
    rpm = np.zeros((T, 4))
    for t in range(T):
        quat = np.array(p.getQuaternionFromEuler(rpy[t]))
        out, _, _ = ctrl.compute_control(control_timestep=dt, cur_pos=pos[t], cur_quat=quat, cur_vel=vel[t],
                                         cur_ang_vel=np.zeros(3), target_pos=tpos[t], target_vel=tvel[t])
        rpm[t] = out
    np.savez(os.path.join(HERE, "pid_controller.npz"), dt=dt, pos=pos, vel=vel, rpy=rpy, target_pos=tpos, target_vel=tvel, rpm=rpm)
    print("wrote pid_controller.npz; rpm range", rpm.min(), rpm.max())


if __name__ == "__main__":
    main()
