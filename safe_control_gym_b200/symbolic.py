"""SymbolicModel facade (safe_control_gym/math_and_models/symbolic_systems.py:8-123).

The reference builds a CasADi expression graph and exposes ``cs.Function`` objects.  CasADi is a third-party
dependency (and absent here); what controllers call on this path is numeric evaluation at given points:
``fc_func(x, u)``, ``fd_func(x0=, p=)['xf']``, ``df_func(x, u)``, ``loss(x, u, Xr, Ur, Q, R)``, ``g_func``.
This facade keeps those names and argument orders and evaluates whole batches on the GPU through
``scg_symbolic_eval`` / ``scg_symbolic_loss`` (csrc/scg_symbolic.cu).  Accepted inputs: numpy arrays or torch
tensors, one sample ``[nx]`` / ``[nu]`` or a batch ``[N, nx]`` / ``[N, nu]``; the result has the input's kind
(numpy in -> numpy out, fp64; cuda tensor in -> cuda tensor out).

``fd_func`` is fixed-step RK4 over ``dt`` (controllers/mpc/mpc_utils.py:79-101), not CasADi's CVODES
integrator (symbolic_systems.py:70-73; adaptive BDF -- unreproducible bit-for-bit).
``fd_linear_func`` likewise (RK4 over the linearised field).  ``x_sym`` / ``u_sym`` / ``x_dot`` / ``cost_func`` are kept as
labels for introspection (tests/scripts/verbose_api.py prints them); there is no expression graph.
"""
import ctypes as C

import numpy as np
import torch

from . import _lib as L


class SymbolicModel:
    def __init__(self, env, dt=None, state_labels=None, input_labels=None, handle=None):
        self._env = env
        self._handle = handle   # a separate scg_env* holding prior parameters (cart-pole), else the env's own
        self.nx, self.nu, self.ny = env.state_dim, env.action_dim, env.state_dim
        self.dt = env.CTRL_TIMESTEP if dt is None else dt
        self.integration_algo = "rk4"
        self.x_sym = list(state_labels or getattr(env, "STATE_LABELS", [f"x{i}" for i in range(self.nx)]))
        self.u_sym = list(input_labels or getattr(env, "ACTION_LABELS", [f"u{i}" for i in range(self.nu)]))
        self.y_sym = self.x_sym
        self.x_dot = [s + "_dot" for s in self.x_sym]
        # cost (quadrotor.py:676-682): labels only, like x_sym -- `loss` evaluates it
        self.cost_func = "0.5 * (X - Xr)' Q (X - Xr) + 0.5 * (U - Ur)' R (U - Ur)"
        self.Q, self.R, self.Xr, self.Ur = "Q", "R", "Xr", "Ur"

    # ------------------------------------------------------------------ plumbing
    def _prep(self, x, u):
        env = self._env
        as_numpy = not isinstance(x, torch.Tensor)
        xt = torch.as_tensor(np.asarray(x, dtype=np.float64) if as_numpy else x, device=env.device).to(env.dtype)
        ut = torch.as_tensor(np.asarray(u, dtype=np.float64) if not isinstance(u, torch.Tensor) else u, device=env.device).to(env.dtype)
        single = xt.dim() == 1
        xt, ut = xt.reshape(-1, self.nx), ut.reshape(-1, self.nu)
        if ut.shape[0] == 1 and xt.shape[0] > 1:
            ut = ut.expand(xt.shape[0], -1)
        assert xt.shape[0] == ut.shape[0], "x and u must hold the same number of samples"
        return xt.t().contiguous(), ut.t().contiguous(), xt.shape[0], single, as_numpy

    def _out(self, t, single, as_numpy):
        if single:
            t = t[0]
        return t.double().cpu().numpy() if as_numpy else t

    def _eval(self, op, x, u, dt, n_out0, n_out1=0):
        env = self._env
        xs, us, n, single, as_numpy = self._prep(x, u)
        out0 = torch.empty(n_out0, n, dtype=env.dtype, device=env.device)
        out1 = torch.empty(n_out1, n, dtype=env.dtype, device=env.device) if n_out1 else None
        with torch.cuda.device(env.device):
            L.check(env.lib.scg_symbolic_eval(self._handle or env._h, op, C.c_void_p(xs.data_ptr()), C.c_void_p(us.data_ptr()), n, float(dt),
                                              C.c_void_p(out0.data_ptr()), C.c_void_p(out1.data_ptr()) if n_out1 else None,
                                              env._stream()))
        return out0, out1, n, single, as_numpy

    # ------------------------------------------------------------------ symbolic_systems.py:63-75
    def fc_func(self, x, u):
        """Continuous dynamics x_dot = f(x, u)."""
        out, _, n, single, as_numpy = self._eval(L.SYM_FC, x, u, 0.0, self.nx)
        return self._out(out.t(), single, as_numpy)

    def fd_func(self, x0=None, p=None, dt=None):
        """Discrete dynamics over ``dt`` (RK4); CasADi-integrator call style: ``fd_func(x0=x, p=u)['xf']``."""
        out, _, n, single, as_numpy = self._eval(L.SYM_FD_RK4, x0, p, self.dt if dt is None else dt, self.nx)
        return {"xf": self._out(out.t(), single, as_numpy)}

    def fd_euler_func(self, x0=None, p=None, dt=None):
        """Forward-Euler discretisation (controllers/mpc/mpc_utils.py:54-55)."""
        out, _, n, single, as_numpy = self._eval(L.SYM_FD_EULER, x0, p, self.dt if dt is None else dt, self.nx)
        return {"xf": self._out(out.t(), single, as_numpy)}

    def g_func(self, x, u):
        """Observation model y = x (quadrotor.py:604, 625, 676)."""
        return x

    # ------------------------------------------------------------------ symbolic_systems.py:77-123
    def df_func(self, x, u):
        """Jacobians (df/dx [nx, nx], df/du [nx, nu]) at (x, u)."""
        A, B, n, single, as_numpy = self._eval(L.SYM_DF, x, u, 0.0, self.nx * self.nx, self.nx * self.nu)
        A = A.t().reshape(n, self.nx, self.nx)
        B = B.t().reshape(n, self.nx, self.nu)
        return self._out(A, single, as_numpy), self._out(B, single, as_numpy)

    def dg_func(self, x, u):
        xs, us, n, single, as_numpy = self._prep(x, u)
        eye = torch.eye(self.nx, dtype=xs.dtype, device=xs.device).expand(n, -1, -1)
        zero = torch.zeros(n, self.nx, self.nu, dtype=xs.dtype, device=xs.device)
        return self._out(eye, single, as_numpy), self._out(zero, single, as_numpy)

    def fc_linear_func(self, x_eval, u_eval, x, u):
        """f(x, u) + df/dx (x_eval - x) + df/du (u_eval - u) (symbolic_systems.py:93-99); composed on the
        host side from fc_func and df_func."""
        f = self.fc_func(x, u)
        A, B = self.df_func(x, u)
        if isinstance(f, np.ndarray):
            dx, du = np.asarray(x_eval, float) - np.asarray(x, float), np.asarray(u_eval, float) - np.asarray(u, float)
            return f + np.einsum("...ij,...j->...i", A, dx) + np.einsum("...ij,...j->...i", B, du)
        return f + torch.einsum("...ij,...j->...i", A, x_eval - x) + torch.einsum("...ij,...j->...i", B, u_eval - u)

    def fd_linear_func(self, x0=None, p=None, dt=None):
        """Discrete LINEARISED dynamics (symbolic_systems.py:100-106): the ODE x_eval' = f(x, u) + A (x_eval - x) + B (u_eval - u),
        linearised at (x, u), integrated over ``dt`` from x_eval = x0; CasADi call style ``fd_linear_func(x0=x_eval,
        p=vertcat(u_eval, x, u))['xf']`` -- p is the concatenation [u_eval (nu), x (nx), u (nu)] along the last axis.
        Fixed-step RK4 like ``fd_func`` (the reference integrates with CVODES); f, A, B come from the device (fc_func, df_func),
        the four stages of the affine field are composed on the caller's side."""
        dt = self.dt if dt is None else dt
        as_numpy = not isinstance(x0, torch.Tensor)
        xe = np.asarray(x0, dtype=np.float64) if as_numpy else x0
        pp = np.asarray(p, dtype=np.float64) if not isinstance(p, torch.Tensor) else p
        ue, x, u = pp[..., :self.nu], pp[..., self.nu:self.nu + self.nx], pp[..., self.nu + self.nx:]
        f = self.fc_func(x, u)
        A, B = self.df_func(x, u)
        ein = np.einsum if as_numpy else torch.einsum
        c = f + ein("...ij,...j->...i", B, ue - u) - ein("...ij,...j->...i", A, x)   # field(y) = A y + c

        def field(y):
            return ein("...ij,...j->...i", A, y) + c
        k1 = field(xe)
        k2 = field(xe + 0.5 * dt * k1)
        k3 = field(xe + 0.5 * dt * k2)
        k4 = field(xe + dt * k3)
        return {"xf": xe + dt / 6.0 * (k1 + 2.0 * k2 + 2.0 * k3 + k4)}

    def g_linear_func(self, x_eval, u_eval, x, u):
        return x_eval

    def loss(self, x, u, Xr, Ur, Q, R):
        """Quadratic cost and derivatives (symbolic_systems.py:114-123): dict with l, l_x, l_xx, l_u, l_uu, l_xu."""
        env = self._env
        xs, us, n, single, as_numpy = self._prep(x, u)
        xr, ur, _, _, _ = self._prep(np.array(np.broadcast_to(np.asarray(Xr, float), (n, self.nx))) if not isinstance(Xr, torch.Tensor) else Xr,
                                     np.array(np.broadcast_to(np.asarray(Ur, float), (n, self.nu))) if not isinstance(Ur, torch.Tensor) else Ur)
        Qh = np.ascontiguousarray(np.asarray(Q, dtype=np.float64).reshape(self.nx, self.nx))
        Rh = np.ascontiguousarray(np.asarray(R, dtype=np.float64).reshape(self.nu, self.nu))
        l = torch.empty(n, dtype=env.dtype, device=env.device)
        lx = torch.empty(self.nx, n, dtype=env.dtype, device=env.device)
        lu = torch.empty(self.nu, n, dtype=env.dtype, device=env.device)
        dp = C.POINTER(C.c_double)
        with torch.cuda.device(env.device):
            L.check(env.lib.scg_symbolic_loss(self._handle or env._h, C.c_void_p(xs.data_ptr()), C.c_void_p(us.data_ptr()),
                                              C.c_void_p(xr.data_ptr()), C.c_void_p(ur.data_ptr()),
                                              Qh.ctypes.data_as(dp), Rh.ctypes.data_as(dp), n, C.c_void_p(l.data_ptr()),
                                              C.c_void_p(lx.data_ptr()), C.c_void_p(lu.data_ptr()), env._stream()))
        lxx, luu = 0.5 * (Qh + Qh.T), 0.5 * (Rh + Rh.T)
        out = {"l": self._out(l, single, as_numpy) if not single else (float(l[0]) if as_numpy else l[0]),
               "l_x": self._out(lx.t(), single, as_numpy), "l_u": self._out(lu.t(), single, as_numpy),
               "l_xx": lxx, "l_uu": luu, "l_xu": np.zeros((self.nx, self.nu))}
        return out


def quadrotor_symbolic_model(env):
    return SymbolicModel(env)
