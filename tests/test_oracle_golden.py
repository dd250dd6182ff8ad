"""The numpy oracle against the golden vectors produced by the reference's own code
(tests/golden/make_golden.py).  Bar: bit-exact flags, <= 1e-13 relative on floats."""
import numpy as np
import pytest

from golden_util import SCENARIOS, load, replay_oracle


@pytest.mark.parametrize("name", SCENARIOS)
def test_oracle_reproduces_reference(name):
    g = load(name)
    ev, env = replay_oracle(g)
    assert len(ev) == len(g["ev_type"])
    for k in ["type", "done", "viol", "trunc", "goal"]:
        assert (np.array([e[k] for e in ev]) == g["ev_" + k]).all(), k
    for k in ["obs", "rew", "mse", "cvals", "state", "kin"]:
        a = np.array([e[k] for e in ev], dtype=float)
        b = g["ev_" + k]
        m = np.isfinite(b)
        assert (np.isfinite(a) == m).all(), k
        if m.any():
            rel = np.abs(a[m] - b[m]) / np.maximum(1.0, np.abs(b[m]))
            assert rel.max() <= 1e-13, (k, rel.max())
    assert env.draws.pos == len(g["draws"])  # every recorded draw consumed, in order
    # what the reset draws turned into: the poses the reference handed to loadURDF / EFFECTIVE_GATES_POSITIONS and the
    # overridden inertial properties (pins the draw ORDER obstacles -> gates -> inertia -> init, quadrotor.py:371-488)
    for k in ["gates", "obstacles", "inertia"]:
        a, b = np.array(env.reset_log[k], dtype=float), g["reset_" + k]
        assert a.shape == b.shape, k
        if a.size:
            assert np.max(np.abs(a - b)) <= 1e-15, (k, np.max(np.abs(a - b)))
    np.testing.assert_array_equal(env.action_low, g["action_low"])
    np.testing.assert_array_equal(env.state_high, g["state_high"])
    np.testing.assert_allclose(env.X_GOAL, g["X_GOAL"], rtol=0, atol=1e-15)


def test_tumble_hits_fold_and_wrap_branches():
    """The tumble scenario must actually drive pitch through the asin fold so the
    Euler<->quaternion round trip is exercised beyond the identity regime."""
    g = load("quad3d_tumble")
    rpy = g["ev_kin"][:, 6:9]
    jumps = np.abs(np.diff(rpy, axis=0))
    assert (jumps[:, 0] > 3.0).any() or (jumps[:, 2] > 3.0).any()  # roll / yaw flipped by pi
    assert np.abs(rpy[:, 1]).max() > 1.2
