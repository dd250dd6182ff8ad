"""SymbolicModel facade -> scg_symbolic_eval / scg_symbolic_loss (csrc/scg_symbolic.cu) against
oracle/symbolic_oracle.py: fc, RK4 / Euler discretisation, closed-form Jacobians (vs complex-step), loss.
fp64 <= 1e-12, fp32 <= 1e-5 relative (unit floor)."""
import numpy as np
import pytest
import torch

from oracle import symbolic_oracle as so

pytestmark = pytest.mark.gpu
N = 257


def _env(kind, dtype):
    if kind == 10:
        from safe_control_gym_b200.envs.cartpole import BatchedCartPole
        env = BatchedCartPole(num_envs=4, dtype=dtype, prior_prop=dict(pole_length=0.4, pole_mass=0.12, cart_mass=0.9))
        prm = dict(g=9.8, m=1.0, cp_l=0.4, cp_mp=0.12, cp_mc=0.9)
    else:
        from safe_control_gym_b200.envs.quadrotor import BatchedQuadrotor
        ip = {1: [0.03], 2: [0.031, 1.5e-5], 3: [0.032, 1.3e-5, 1.5e-5, 2.2e-5]}[kind]
        ti = dict(stabilization_goal=[0, 0, 1][:max(kind, 2)], stabilization_goal_tolerance=0.05)
        env = BatchedQuadrotor(num_envs=4, dtype=dtype, quad_type=kind, inertial_prop=ip, task_info=ti)
        prm = dict(m=env.MASS, g=env.GRAVITY_ACC, L=env.L, Ixx=env.J[0, 0], Iyy=env.J[1, 1], Izz=env.J[2, 2], KF=env.KF, KM=env.KM)
    return env, prm


def _rel(got, want):
    got, want = np.asarray(got, dtype=np.float64), np.asarray(want, dtype=np.float64)
    return float(np.max(np.abs(got - want) / np.maximum(1.0, np.abs(want))))


@pytest.mark.parametrize("dtype,tol", [(torch.float64, 1e-12), (torch.float32, 1e-5)])
@pytest.mark.parametrize("kind", [1, 2, 3, 10])
def test_fc_fd_df(kind, dtype, tol):
    env, prm = _env(kind, dtype)
    nx, nu = so.DIMS[kind]
    rng = np.random.default_rng(kind)
    x = rng.normal(0, 0.4, size=(N, nx))
    u = rng.uniform(0.03, 0.12, size=(N, nu)) if kind != 10 else rng.normal(0, 3.0, size=(N, nu))
    if dtype == torch.float32:   # give both sides the same representable inputs
        x, u = x.astype(np.float32).astype(np.float64), u.astype(np.float32).astype(np.float64)
    sym = env.symbolic
    f = sym.fc_func(x, u)
    xr = sym.fd_func(x0=x, p=u)["xf"]
    xe = sym.fd_euler_func(x0=x, p=u)["xf"]
    A, B = sym.df_func(x, u)
    assert f.shape == (N, nx) and A.shape == (N, nx, nx) and B.shape == (N, nx, nu)
    dt = env.CTRL_TIMESTEP
    for i in range(0, N, 16):
        assert _rel(f[i], so.fc(kind, prm, x[i], u[i])) <= tol
        assert _rel(xr[i], so.fd_rk4(kind, prm, x[i], u[i], dt)) <= tol
        assert _rel(xe[i], so.fd_euler(kind, prm, x[i], u[i], dt)) <= tol
        Ao, Bo = so.df(kind, prm, x[i], u[i])
        # the Jacobian entries scale with 1 / inertia (1e5): relative to the matrix scale
        assert np.max(np.abs(A[i] - Ao)) <= tol * max(1.0, np.abs(Ao).max())
        assert np.max(np.abs(B[i] - Bo)) <= tol * max(1.0, np.abs(Bo).max())
    # single-sample call style of the reference: fc_func(x, u) with 1-D arrays
    f1 = sym.fc_func(x[0], u[0])
    assert f1.shape == (nx,) and _rel(f1, so.fc(kind, prm, x[0], u[0])) <= tol
    env.close()


@pytest.mark.parametrize("dtype,tol", [(torch.float64, 1e-12), (torch.float32, 1e-5)])
def test_loss(dtype, tol):
    env, _ = _env(2, dtype)
    rng = np.random.default_rng(0)
    x, u = rng.normal(size=(N, 6)), rng.normal(size=(N, 2))
    xr, ur = rng.normal(size=6), rng.normal(size=2)
    Q, R = rng.normal(size=(6, 6)), rng.normal(size=(2, 2))
    if dtype == torch.float32:
        x, u, xr, ur = (a.astype(np.float32).astype(np.float64) for a in (x, u, xr, ur))
        Q, R = Q.astype(np.float32).astype(np.float64), R.astype(np.float32).astype(np.float64)
    out = env.symbolic.loss(x, u, xr, ur, Q, R)
    for i in range(0, N, 32):
        o = so.loss(x[i], u[i], xr, ur, Q, R)
        scale = max(1.0, abs(o["l"]))
        assert abs(out["l"][i] - o["l"]) <= 10 * tol * scale
        assert _rel(out["l_x"][i], o["l_x"]) <= 10 * tol and _rel(out["l_u"][i], o["l_u"]) <= 10 * tol
    assert np.allclose(out["l_xx"], 0.5 * (Q + Q.T)) and out["l_xu"].shape == (6, 2)
    env.close()


def test_device_tensors_stay_on_device():
    env, _ = _env(3, torch.float32)
    x = torch.zeros(8, 12, device="cuda")
    u = torch.full((8, 4), float(env.MASS * 9.8 / 4), device="cuda")
    f = env.symbolic.fc_func(x, u)
    assert f.is_cuda and f.shape == (8, 12) and float(f.abs().max()) < 1e-5   # hover equilibrium
    env.close()


@pytest.mark.parametrize("kind", [2, 3])
def test_fd_linear_func(kind):
    """SymbolicModel.fd_linear_func (symbolic_systems.py:100-106): at the linearisation point with u_eval = u it is the exact
    flow of the affine field x' = f + A (x_eval - x): one RK4 step of it against the matrix-exponential solution, and it
    agrees with the non-linear fd_func to second order in dt."""
    from scipy.linalg import expm
    env, prm = _env(kind, torch.float64)
    nx, nu = so.DIMS[kind]
    rng = np.random.default_rng(40 + kind)
    x = rng.normal(0, 0.2, size=(5, nx)); u = rng.uniform(0.05, 0.1, size=(5, nu))
    xe = x + rng.normal(0, 0.01, size=x.shape); ue = u + rng.normal(0, 0.002, size=u.shape)
    dt = 1.0 / 240
    got = env.symbolic.fd_linear_func(x0=xe, p=np.concatenate([ue, x, u], axis=-1), dt=dt)["xf"]
    f = env.symbolic.fc_func(x, u)
    A, B = env.symbolic.df_func(x, u)
    for i in range(5):
        c = f[i] + B[i] @ (ue[i] - u[i]) - A[i] @ x[i]
        M = np.zeros((nx + 1, nx + 1)); M[:nx, :nx] = A[i]; M[:nx, nx] = c   # d/dt [y; 1] = M [y; 1]
        want = (expm(M * dt) @ np.append(xe[i], 1.0))[:nx]
        assert _rel(got[i], want) <= 1e-10
    nonlin = env.symbolic.fd_func(x0=xe, p=ue, dt=dt)["xf"]
    assert _rel(got, nonlin) <= 1e-4
    single = env.symbolic.fd_linear_func(x0=xe[0], p=np.concatenate([ue[0], x[0], u[0]]), dt=dt)["xf"]
    assert single.shape == (nx,) and _rel(single, got[0]) <= 1e-14
    env.close()


@pytest.mark.parametrize("dtype,tol", [(torch.float64, 1e-12), (torch.float32, 2e-5)])
@pytest.mark.parametrize("kind", [1, 2, 3])
def test_device_model_matches_the_reference_expressions(kind, dtype, tol):
    """The device functions against golden vectors produced by the reference's OWN ``_setup_symbolic`` + ``SymbolicModel``
    (tests/golden/symbolic_model.npz, make_symbolic_golden.py: CasADi replaced by an expression-graph stand-in): fc_func, both
    Jacobians of df_func, fc_linear_func, loss with its gradients / Hessians; fd_func / fd_linear_func (RK4 over dt) against
    the integrated flow within RK4's truncation error (see tests/test_oracle_symbolic_golden.py for the measured values)."""
    import os
    from safe_control_gym_b200.envs.quadrotor import BatchedQuadrotor
    G = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "symbolic_model.npz"))
    p = f"q{kind}_"
    ti = dict(stabilization_goal=[0, 0, 1][:max(kind, 2)], stabilization_goal_tolerance=0.05)
    env = BatchedQuadrotor(num_envs=4, dtype=dtype, quad_type=kind, ctrl_freq=60, pyb_freq=240, task_info=ti)   # nominal URDF constants
    sm = env.symbolic
    assert abs(sm.dt - float(G[p + "dt"])) < 1e-15
    X, U, Xe, Ue = G[p + "X"], G[p + "U"], G[p + "Xe"], G[p + "Ue"]
    scale = 1.0 if dtype == torch.float64 else 300.0     # fp32: |f| reaches 200 rad/s^2 at these points (unit floor in _rel)
    assert _rel(sm.fc_func(X, U), G[p + "fc"]) <= tol * scale
    A, B = sm.df_func(X, U)
    assert _rel(A, G[p + "dfdx"]) <= tol * scale and _rel(B, G[p + "dfdu"]) <= tol * scale
    assert _rel(sm.fc_linear_func(Xe, Ue, X, U), G[p + "fc_lin"]) <= tol * scale
    for Qn, pre in (("Q", "l"), ("Qa", "la")):
        out = sm.loss(X, U, G[p + "Xr"], G[p + "Ur"], G[p + Qn], G[p + "R"])
        assert _rel(out["l"], G[p + pre].reshape(-1)) <= tol * 10
        assert _rel(out["l_x"], G[p + pre + "_x"].reshape(len(X), -1)) <= tol * 10
        assert _rel(out["l_xx"], G[p + pre + "_xx"][0]) <= 1e-12
    assert _rel(out["l_u"], G[p + "l_u"].reshape(len(X), -1)) <= tol * 10 and _rel(out["l_uu"], G[p + "l_uu"][0]) <= 1e-12
    rk4 = {1: 1e-12, 2: 1e-5, 3: 2e-5}[kind]
    assert _rel(sm.fd_func(x0=X, p=U)["xf"], G[p + "fd"]) <= max(rk4, tol * scale)
    assert _rel(sm.fd_linear_func(x0=Xe, p=np.concatenate([Ue, X, U], axis=-1))["xf"], G[p + "fd_lin"]) <= max(rk4, tol * scale)
    env.close()
