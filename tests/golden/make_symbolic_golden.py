"""Golden vectors for SymbolicModel (rows A12 / N4) from the reference's OWN code:

    python tests/golden/make_symbolic_golden.py      ->  tests/golden/symbolic_model.npz

``Quadrotor._setup_symbolic`` (quadrotor.py:581-697) and ``SymbolicModel`` (symbolic_systems.py:8-123) are executed
from /root/reference with CasADi replaced by tests/golden/mini_casadi.py (expression graphs, symbolic Jacobians,
``Function``; the integrator is scipy's LSODA at rtol 1e-13 in place of CVODES) and pybullet / gym by ref_stubs.py.
Recorded per quad_type at random points: fc_func, df_func (dfdx, dfdu), fc_linear_func, loss (l, l_x, l_xx, l_u, l_uu,
l_xu), fd_func and fd_linear_func (the ODE's flow over dt), plus the physical constants the model was built from."""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
import mini_casadi  # noqa: E402
import ref_stubs  # noqa: E402

REF = "/root/reference"
NX = {1: (2, 1), 2: (6, 2), 3: (12, 4)}


def main():
    Quadrotor = ref_stubs.install(REF, casadi=mini_casadi)
    rng = np.random.default_rng(2024)
    out = {}
    for qt in (1, 2, 3):
        kw = dict(task_info=dict(stabilization_goal=[0, 0, 1], stabilization_goal_tolerance=0.05)) if qt == 3 else {}
        env = Quadrotor(quad_type=qt, physics="dyn", ctrl_freq=60, pyb_freq=240, gui=False, seed=1, **kw)
        sm = env.symbolic
        nx, nu = NX[qt]
        assert (sm.nx, sm.nu) == (nx, nu)
        hover = env.MASS * env.GRAVITY_ACC / nu
        k = 8
        X = rng.normal(0, 0.4, size=(k, nx))
        if qt == 3:
            X[:, 7] = np.clip(X[:, 7], -1.2, 1.2)   # away from tan(theta) singularities
        U = hover * (1 + rng.normal(0, 0.3, size=(k, nu)))
        Xe, Ue = X + rng.normal(0, 0.05, size=X.shape), U + hover * rng.normal(0, 0.05, size=U.shape)
        Xr, Ur = rng.normal(0, 0.3, size=(k, nx)), hover * np.ones((k, nu))
        Q = rng.normal(0, 1, size=(nx, nx)); Q = Q @ Q.T + np.eye(nx)
        R = rng.normal(0, 1, size=(nu, nu)); R = R @ R.T + np.eye(nu)
        Qa = Q + 0.3 * rng.normal(0, 1, size=(nx, nx))          # a non-symmetric weight as well: l_x = 0.5 (Q + Q') e
        rec = {n: [] for n in ("fc", "dfdx", "dfdu", "fc_lin", "fd", "fd_lin", "l", "l_x", "l_xx", "l_u", "l_uu", "l_xu", "la", "la_x", "la_xx")}
        for i in range(k):
            x, u = X[i], U[i]
            rec["fc"].append(np.asarray(sm.fc_func(x, u)).reshape(-1))
            A, B = sm.df_func(x, u)
            rec["dfdx"].append(np.asarray(A)); rec["dfdu"].append(np.asarray(B))
            rec["fc_lin"].append(np.asarray(sm.fc_linear_func(Xe[i], Ue[i], x, u)).reshape(-1))
            rec["fd"].append(np.asarray(sm.fd_func(x0=x, p=u)["xf"]).reshape(-1))
            rec["fd_lin"].append(np.asarray(sm.fd_linear_func(x0=Xe[i], p=np.concatenate([Ue[i], x, u]))["xf"]).reshape(-1))
            ls = sm.loss(x=x, u=u, Xr=Xr[i], Ur=Ur[i], Q=Q, R=R)
            for n in ("l", "l_x", "l_xx", "l_u", "l_uu", "l_xu"):
                rec[n].append(np.asarray(ls[n]))
            la = sm.loss(x=x, u=u, Xr=Xr[i], Ur=Ur[i], Q=Qa, R=R)
            rec["la"].append(np.asarray(la["l"])); rec["la_x"].append(np.asarray(la["l_x"])); rec["la_xx"].append(np.asarray(la["l_xx"]))
        pre = f"q{qt}_"
        out.update({pre + "X": X, pre + "U": U, pre + "Xe": Xe, pre + "Ue": Ue, pre + "Xr": Xr, pre + "Ur": Ur, pre + "Q": Q, pre + "R": R,
                    pre + "Qa": Qa, pre + "dt": np.float64(sm.dt),
                    pre + "prm": np.array([env.MASS, env.GRAVITY_ACC, env.L, env.J[0, 0], env.J[1, 1], env.J[2, 2], env.KF, env.KM])})
        out.update({pre + n: np.array(v, dtype=np.float64) for n, v in rec.items()})
        print(f"quad_type {qt}: fc[0] = {np.round(rec['fc'][0][:6], 5)}, |dfdx| = {np.abs(rec['dfdx'][0]).sum():.6f}, l = {float(np.ravel(rec['l'][0])[0]):.6f}")
    np.savez_compressed(os.path.join(HERE, "symbolic_model.npz"), **out)


if __name__ == "__main__":
    main()
