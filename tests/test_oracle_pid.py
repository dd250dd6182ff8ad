"""oracle/pid_oracle.py against the reference's own PIDController (quadrotor_utils.py:70-278) recorded in
tests/golden/pid_controller.npz by tests/golden/make_pid_golden.py."""
import os

import numpy as np

from oracle import bullet_math as bm
from oracle.pid_oracle import PIDOracle

G = np.load(os.path.join(os.path.dirname(__file__), "golden", "pid_controller.npz"))


def test_pid_oracle_matches_reference_class():
    ctrl = PIDOracle()
    dt = float(G["dt"])
    unsat = 0
    for t in range(G["rpm"].shape[0]):
        q = bm.quat_from_euler(*G["rpy"][t])
        rpm = ctrl.compute_control(dt, G["pos"][t], q, G["vel"][t], G["target_pos"][t], target_vel=G["target_vel"][t])
        assert np.max(np.abs(rpm - G["rpm"][t])) <= 1e-9 * 2e4, (t, rpm, G["rpm"][t])   # 1e-9 relative to full scale
        unsat += int(np.sum((G["rpm"][t] > 9440.4) & (G["rpm"][t] < 21666.4)))
    assert unsat >= 40   # the sequence exercises the unsaturated mixer, not only the PWM clip
