"""CPU restatement of the prior models SymbolicModel exposes -- TEST INFRASTRUCTURE ONLY.

Follows safe_control_gym/envs/gym_pybullet_drones/quadrotor.py:581-697 (_setup_symbolic: the 1D / 2D / 3D
ODEs, written with the explicit Rz @ Ry @ Rx product of math_and_models/transformations.py:125-182) and
safe_control_gym/math_and_models/symbolic_systems.py:63-123 (fc / df / loss).  The CasADi expression graph is
plain arithmetic; its Jacobians are reproduced here by complex-step differentiation of the restated f
(accurate to rounding, and independent of the closed-form derivatives in csrc/scg_symbolic.cu).  PINNED: fc, both
Jacobians, the linearised field and the loss with all derivatives equal, to 1e-12, what the reference's OWN
``_setup_symbolic`` + ``SymbolicModel`` code evaluates to when CasADi is replaced by an expression-graph stand-in
(tests/golden/symbolic_model.npz, make_symbolic_golden.py, mini_casadi.py; tests/test_oracle_symbolic_golden.py).  The
discrete model is fixed-step RK4 (controllers/mpc/mpc_utils.py:79-101); CasADi's CVODES ``fd_func`` (adaptive BDF at
reltol 1e-6, symbolic_systems.py:70-73) is not reproducible bit for bit by anything else -- RK4 is held to the ODE's
integrated flow within its truncation error (2.5e-6 / 6.3e-6 relative for 2D / 3D over 1 / 60 s at the fixture's points).  The cart-pole model is absent from the
reference fork: standard frictionless cart-pole, parity unpinned.
"""
import numpy as np

DIMS = {1: (2, 1), 2: (6, 2), 3: (12, 4), 10: (4, 1)}


def rot_xyz(phi, theta, psi):
    """csRotXYZ = Rz(psi) @ Ry(theta) @ Rx(phi) (transformations.py:125-182)."""
    c, s = np.cos, np.sin
    rz = np.array([[c(psi), -s(psi), 0], [s(psi), c(psi), 0], [0, 0, 1]], dtype=complex if np.iscomplexobj(psi) else float)
    ry = np.array([[c(theta), 0, s(theta)], [0, 1, 0], [-s(theta), 0, c(theta)]], dtype=rz.dtype)
    rx = np.array([[1, 0, 0], [0, c(phi), -s(phi)], [0, s(phi), c(phi)]], dtype=rz.dtype)
    return rz @ ry @ rx


def fc(kind, prm, x, u):
    """x_dot = f(x, u) for one sample; works on complex inputs (complex-step differentiation)."""
    x = np.asarray(x)
    u = np.asarray(u)
    dt = complex if (np.iscomplexobj(x) or np.iscomplexobj(u)) else float
    m, g = prm["m"], prm["g"]
    if kind == 1:
        return np.array([x[1], u[0] / m - g], dtype=dt)
    if kind == 2:
        T = u[0] + u[1]
        return np.array([x[1], np.sin(x[4]) * T / m, x[3], np.cos(x[4]) * T / m - g, x[5],
                         prm["L"] * (u[1] - u[0]) / prm["Iyy"] / np.sqrt(2)], dtype=dt)
    if kind == 3:
        phi, theta, psi, p, q, r = x[6], x[7], x[8], x[9], x[10], x[11]
        J = np.diag([prm["Ixx"], prm["Iyy"], prm["Izz"]])
        Jinv = np.diag([1.0 / prm["Ixx"], 1.0 / prm["Iyy"], 1.0 / prm["Izz"]])
        gamma = prm["KM"] / prm["KF"]
        Rob = rot_xyz(phi, theta, psi)
        acc = Rob @ np.array([0, 0, u[0] + u[1] + u[2] + u[3]], dtype=dt) / m - np.array([0, 0, g])
        l = prm["L"]
        Mb = np.array([l / np.sqrt(2.0) * (u[0] + u[1] - u[2] - u[3]), l / np.sqrt(2.0) * (-u[0] + u[1] + u[2] - u[3]),
                       gamma * (u[0] - u[1] + u[2] - u[3])], dtype=dt)
        w = np.array([p, q, r], dtype=dt)
        skew = np.array([[0, -r, q], [r, 0, -p], [-q, p, 0]], dtype=dt)
        rate_dot = Jinv @ (Mb - skew @ J @ w)
        W = np.array([[1, np.sin(phi) * np.tan(theta), np.cos(phi) * np.tan(theta)],
                      [0, np.cos(phi), -np.sin(phi)],
                      [0, np.sin(phi) / np.cos(theta), np.cos(phi) / np.cos(theta)]], dtype=dt)
        ang_dot = W @ w
        return np.array([x[1], acc[0], x[3], acc[1], x[5], acc[2], ang_dot[0], ang_dot[1], ang_dot[2],
                         rate_dot[0], rate_dot[1], rate_dot[2]], dtype=dt)
    if kind == 10:
        l, mp, mc = prm["cp_l"], prm["cp_mp"], prm["cp_mc"]
        s, c = np.sin(x[2]), np.cos(x[2])
        Mm, ml = mc + mp, mp * l
        tmp = (u[0] + ml * x[3] * x[3] * s) / Mm
        thdd = (g * s - c * tmp) / (l * (4.0 / 3.0 - mp * c * c / Mm))
        return np.array([x[1], tmp - ml * thdd * c / Mm, x[3], thdd], dtype=dt)
    raise ValueError(kind)


def fd_rk4(kind, prm, x, u, dt):
    """mpc_utils.py:93-98."""
    k1 = fc(kind, prm, x, u)
    k2 = fc(kind, prm, x + dt / 2 * k1, u)
    k3 = fc(kind, prm, x + dt / 2 * k2, u)
    k4 = fc(kind, prm, x + dt * k3, u)
    return x + dt / 6 * (k1 + 2 * k2 + 2 * k3 + k4)


def fd_euler(kind, prm, x, u, dt):
    return x + dt * fc(kind, prm, x, u)


def df(kind, prm, x, u, h=1e-30):
    """(df/dx, df/du) by complex-step differentiation of fc."""
    nx, nu = DIMS[kind]
    x = np.asarray(x, dtype=float)
    u = np.asarray(u, dtype=float)
    A, B = np.zeros((nx, nx)), np.zeros((nx, nu))
    for j in range(nx):
        xc = x.astype(complex)
        xc[j] += 1j * h
        A[:, j] = fc(kind, prm, xc, u).imag / h
    for j in range(nu):
        uc = u.astype(complex)
        uc[j] += 1j * h
        B[:, j] = fc(kind, prm, x, uc).imag / h
    return A, B


def loss(x, u, Xr, Ur, Q, R):
    """symbolic_systems.py:114-123 for cost_func = 0.5 ex'Q ex + 0.5 eu'R eu (quadrotor.py:682)."""
    ex, eu = np.asarray(x, float) - np.asarray(Xr, float), np.asarray(u, float) - np.asarray(Ur, float)
    Q, R = np.asarray(Q, float), np.asarray(R, float)
    return {"l": 0.5 * ex @ Q @ ex + 0.5 * eu @ R @ eu, "l_x": 0.5 * (Q + Q.T) @ ex, "l_xx": 0.5 * (Q + Q.T),
            "l_u": 0.5 * (R + R.T) @ eu, "l_uu": 0.5 * (R + R.T), "l_xu": np.zeros((len(ex), len(eu)))}
