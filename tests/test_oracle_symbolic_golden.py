"""Rows A12 / N4: the numpy restatement of the prior models (oracle/symbolic_oracle.py) against golden vectors produced by
the reference's OWN ``Quadrotor._setup_symbolic`` + ``SymbolicModel`` (tests/golden/make_symbolic_golden.py: CasADi replaced
by the expression-graph stand-in tests/golden/mini_casadi.py) -- f, both Jacobians, the linearised field and the loss with
all its derivatives to rounding; the discrete models (CasADi: CVODES; fixture: LSODA at rtol 1e-13) against fixed-step RK4,
whose truncation error over one control period is what the bound states."""
import os

import numpy as np
import pytest

from oracle import symbolic_oracle as so

G = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "symbolic_model.npz"))


def prm_of(qt):
    m, g, L, ixx, iyy, izz, kf, km = G[f"q{qt}_prm"]
    return dict(m=m, g=g, L=L, Ixx=ixx, Iyy=iyy, Izz=izz, KF=kf, KM=km)


def rel(a, b):
    a, b = np.asarray(a, float), np.asarray(b, float)
    return float(np.max(np.abs(a - b) / np.maximum(1.0, np.abs(b))))


@pytest.mark.parametrize("qt", [1, 2, 3])
def test_fc_df_and_linearised_field_match_the_reference_expressions(qt):
    prm, p = prm_of(qt), f"q{qt}_"
    for i in range(len(G[p + "X"])):
        x, u, xe, ue = G[p + "X"][i], G[p + "U"][i], G[p + "Xe"][i], G[p + "Ue"][i]
        f = so.fc(qt, prm, x, u)
        A, B = so.df(qt, prm, x, u)
        assert rel(f, G[p + "fc"][i]) <= 1e-13
        assert rel(A, G[p + "dfdx"][i]) <= 1e-12 and rel(B, G[p + "dfdu"][i]) <= 1e-12
        assert rel(f + A @ (xe - x) + B @ (ue - u), G[p + "fc_lin"][i]) <= 1e-12   # symbolic_systems.py:93-99


@pytest.mark.parametrize("qt", [1, 2, 3])
def test_loss_and_derivatives_match_the_reference_expressions(qt):
    p = f"q{qt}_"
    for i in range(len(G[p + "X"])):
        for Qn, pre in (("Q", "l"), ("Qa", "la")):   # symmetric and non-symmetric Q
            out = so.loss(G[p + "X"][i], G[p + "U"][i], G[p + "Xr"][i], G[p + "Ur"][i], G[p + Qn], G[p + "R"])
            assert rel(out["l"], np.ravel(G[p + pre][i])[0]) <= 1e-13
            assert rel(out["l_x"], np.ravel(G[p + pre + "_x"][i])) <= 1e-12
            assert rel(out["l_xx"], G[p + pre + "_xx"][i]) <= 1e-12
        assert rel(out["l_u"], np.ravel(G[p + "l_u"][i])) <= 1e-12 and rel(out["l_uu"], G[p + "l_uu"][i]) <= 1e-12
        assert rel(out["l_xu"], G[p + "l_xu"][i]) <= 1e-15


@pytest.mark.parametrize("qt,tol", [(1, 1e-12), (2, 1e-5), (3, 2e-5)])
def test_rk4_discrete_model_against_the_integrated_flow(qt, tol):
    """fd_func: one RK4 step over dt = 1 / 60 s against the ODE's flow (what CasADi's integrator converges to; CVODES' own default
    tolerances are reltol 1e-6 / abstol 1e-8).  Measured at the fixture's points (thrust imbalances of 30 %, i.e. angular
    accelerations up to 180 rad/s^2): 3e-16 (1D: the field is affine), 2.5e-6 (2D), 6.3e-6 (3D)."""
    prm, p = prm_of(qt), f"q{qt}_"
    dt = float(G[p + "dt"])
    worst = max(rel(so.fd_rk4(qt, prm, G[p + "X"][i], G[p + "U"][i], dt), G[p + "fd"][i]) for i in range(len(G[p + "X"])))
    print(f"quad_type {qt}: RK4 vs flow over dt = {dt:.5f}: {worst:.2e}")
    assert worst <= tol
