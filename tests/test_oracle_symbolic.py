"""Self-consistency of the symbolic-model oracle (oracle/symbolic_oracle.py): Jacobians against finite
differences, equilibria, RK4 against an adaptive integrator, loss gradients.  CPU only."""
import numpy as np
import pytest
from scipy.integrate import solve_ivp

from oracle import symbolic_oracle as so

PRM = dict(m=0.03454, g=9.8, L=0.0397, Ixx=1.4e-5, Iyy=1.4e-5, Izz=2.17e-5, KF=3.16e-10, KM=7.94e-12,
           cp_l=0.5, cp_mp=0.1, cp_mc=1.0)


def _sample(kind, rng):
    nx, nu = so.DIMS[kind]
    return rng.normal(0, 0.3, nx), (rng.uniform(0.05, 0.1, nu) if kind != 10 else rng.normal(0, 2.0, nu))


@pytest.mark.parametrize("kind", [1, 2, 3, 10])
def test_jacobians_match_finite_differences(kind):
    rng = np.random.default_rng(kind)
    for _ in range(5):
        x, u = _sample(kind, rng)
        A, B = so.df(kind, PRM, x, u)
        eps = 1e-6
        for j in range(len(x)):
            d = np.zeros(len(x)); d[j] = eps
            fd = (so.fc(kind, PRM, x + d, u) - so.fc(kind, PRM, x - d, u)) / (2 * eps)
            assert np.max(np.abs(A[:, j] - fd)) <= 1e-6 * max(1.0, np.abs(A).max())
        for j in range(len(u)):
            d = np.zeros(len(u)); d[j] = eps
            fd = (so.fc(kind, PRM, x, u + d) - so.fc(kind, PRM, x, u - d)) / (2 * eps)
            assert np.max(np.abs(B[:, j] - fd)) <= 1e-6 * max(1.0, np.abs(B).max())


def test_hover_is_an_equilibrium_of_every_quadrotor_model():
    for kind in (1, 2, 3):
        nx, nu = so.DIMS[kind]
        u = np.ones(nu) * PRM["m"] * PRM["g"] / nu       # U_GOAL (quadrotor.py:262)
        assert np.max(np.abs(so.fc(kind, PRM, np.zeros(nx), u))) < 1e-14


def test_3d_model_is_not_the_dyn_integrator():
    """The symbolic 3D model carries the Euler kinematic matrix, DYN does not (SURVEY.md A12 note):
    with roll = 0.3 a pure body pitch rate also turns yaw."""
    x = np.zeros(12); x[6] = 0.3; x[10] = 1.0
    f = so.fc(3, PRM, x, np.ones(4) * PRM["m"] * PRM["g"] / 4)
    assert abs(f[8] - np.sin(0.3)) < 1e-15 and abs(f[7] - np.cos(0.3)) < 1e-15


@pytest.mark.parametrize("kind", [2, 3, 10])
def test_rk4_close_to_adaptive_integration(kind):
    rng = np.random.default_rng(10 + kind)
    x, u = _sample(kind, rng)
    dt = 1 / 60
    ref = solve_ivp(lambda t, y: so.fc(kind, PRM, y, u), (0, dt), x, rtol=1e-10, atol=1e-12).y[:, -1]
    assert np.max(np.abs(so.fd_rk4(kind, PRM, x, u, dt) - ref)) < 1e-6     # the loose cross-check for CVODES' role
    assert np.max(np.abs(so.fd_euler(kind, PRM, x, u, dt) - ref)) < 5e-2


def test_loss_gradients():
    rng = np.random.default_rng(3)
    x, u, xr, ur = rng.normal(size=6), rng.normal(size=2), rng.normal(size=6), rng.normal(size=2)
    Q, R = rng.normal(size=(6, 6)), rng.normal(size=(2, 2))
    out = so.loss(x, u, xr, ur, Q, R)
    eps = 1e-6
    for j in range(6):
        d = np.zeros(6); d[j] = eps
        fd = (so.loss(x + d, u, xr, ur, Q, R)["l"] - so.loss(x - d, u, xr, ur, Q, R)["l"]) / (2 * eps)
        assert abs(fd - out["l_x"][j]) < 1e-6
    assert np.allclose(out["l_xx"], 0.5 * (Q + Q.T)) and np.allclose(out["l_uu"], 0.5 * (R + R.T))
