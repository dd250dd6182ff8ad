"""A small stand-in for the part of CasADi the reference's SymbolicModel uses (casadi ^3.5 is a third-party
dependency that is absent here): expression graphs over scalars, ``MX.sym / vertcat / blockcat / skew / sin / cos /
tan / sqrt``, ``Function``, ``jacobian`` (symbolic differentiation of the graph) and ``integrator``.

Used ONLY by tests/golden/make_symbolic_golden.py, in the build container, so that the reference's OWN
``Quadrotor._setup_symbolic`` (quadrotor.py:581-697) and ``SymbolicModel`` (symbolic_systems.py:8-123) run and
their functions can be evaluated on numbers: what is recorded is the arithmetic of the expressions the
reference builds, evaluated in fp64.  ``integrator`` integrates the ODE with scipy's LSODA at rtol 1e-13 /
atol 1e-15 (CasADi's default is CVODES; both converge to the flow of the ODE, neither is reproducible bit
for bit by the other).  Nothing here ships, nothing here is read on the GPU box.
"""
import math

import numpy as np


class E:
    """Scalar expression node."""
    __array_ufunc__ = None          # numpy scalars on the left defer to the reflected operators below
    __array_priority__ = 1000
    __slots__ = ("op", "a", "b", "name")

    def __init__(self, op, a=None, b=None, name=None):
        self.op, self.a, self.b, self.name = op, a, b, name

    # -- construction
    def _bin(self, op, o, swap=False):
        o = o if isinstance(o, E) else float(o)
        return E(op, o, self) if swap else E(op, self, o)

    def __add__(self, o): return self._bin("+", o)
    def __radd__(self, o): return self._bin("+", o, True)
    def __sub__(self, o): return self._bin("-", o)
    def __rsub__(self, o): return self._bin("-", o, True)
    def __mul__(self, o): return self._bin("*", o)
    def __rmul__(self, o): return self._bin("*", o, True)
    def __truediv__(self, o): return self._bin("/", o)
    def __rtruediv__(self, o): return self._bin("/", o, True)
    def __neg__(self): return E("neg", self)
    def __repr__(self): return self.name if self.op == "var" else f"({self.op} ...)"


def _ev(e, env, memo):
    if not isinstance(e, E):
        return float(e)
    k = id(e)
    if k in memo:
        return memo[k]
    if e.op == "var":
        v = env[e]
    elif e.op == "neg":
        v = -_ev(e.a, env, memo)
    elif e.op in ("sin", "cos", "tan", "sqrt"):
        v = getattr(math, e.op)(_ev(e.a, env, memo))
    else:
        x, y = _ev(e.a, env, memo), _ev(e.b, env, memo)
        v = x + y if e.op == "+" else x - y if e.op == "-" else x * y if e.op == "*" else x / y
    memo[k] = v
    return v


def _d(e, var, memo):
    """d e / d var as an expression (constants folded so that the graphs stay small)."""
    if not isinstance(e, E):
        return 0.0
    k = id(e)
    if k in memo:
        return memo[k]
    if e.op == "var":
        r = 1.0 if e is var else 0.0
    elif e.op == "neg":
        r = _neg(_d(e.a, var, memo))
    elif e.op == "sin":
        r = _mul(E("cos", e.a), _d(e.a, var, memo))
    elif e.op == "cos":
        r = _neg(_mul(E("sin", e.a), _d(e.a, var, memo)))
    elif e.op == "tan":
        c = E("cos", e.a)
        r = _div(_d(e.a, var, memo), _mul(c, c))
    elif e.op == "sqrt":
        r = _div(_d(e.a, var, memo), _mul(2.0, e))
    else:
        da, db = _d(e.a, var, memo), _d(e.b, var, memo)
        if e.op == "+":
            r = _add(da, db)
        elif e.op == "-":
            r = _add(da, _neg(db))
        elif e.op == "*":
            r = _add(_mul(da, e.b), _mul(e.a, db))
        else:
            r = _add(_div(da, e.b), _neg(_div(_mul(e.a, db), _mul(e.b, e.b))))
    memo[k] = r
    return r


def _is0(x): return not isinstance(x, E) and x == 0.0
def _is1(x): return not isinstance(x, E) and x == 1.0
def _neg(x): return (-x) if not isinstance(x, E) else E("neg", x)
def _add(x, y): return y if _is0(x) else x if _is0(y) else (x + y)
def _mul(x, y): return 0.0 if _is0(x) or _is0(y) else y if _is1(x) else x if _is1(y) else (x * y)
def _div(x, y): return 0.0 if _is0(x) else (x / y)


class MX:
    """Matrix of scalar expressions (numpy object array underneath)."""
    __array_ufunc__ = None
    __array_priority__ = 1000

    def __init__(self, data):
        a = np.empty(np.shape(data), dtype=object)
        a[...] = data
        self.d = a.reshape(a.shape[0], -1) if a.ndim else a.reshape(1, 1)

    @staticmethod
    def sym(name, n=1, m=1):
        return MX([[E("var", name=f"{name}[{i},{j}]" if n * m > 1 else name) for j in range(m)] for i in range(n)])

    @property
    def shape(self): return self.d.shape
    @property
    def T(self): return MX(self.d.T)
    def __getitem__(self, i): return MX(np.atleast_2d(self.d.reshape(-1)[i] if self.d.shape[1] == 1 else self.d[i]))
    def __repr__(self): return f"MX{self.d.shape}"

    @staticmethod
    def _raw(o):
        if isinstance(o, MX):
            return o.d
        if isinstance(o, np.ndarray):
            return o.astype(object).reshape(o.shape[0], -1) if o.ndim else float(o)
        return o if isinstance(o, E) else float(o)

    def _ew(self, o, f):
        r = MX._raw(o)
        a, b = self.d, r
        if isinstance(b, np.ndarray) and b.shape != a.shape:    # 1 x 1 against a matrix: broadcast like casadi
            if a.shape == (1, 1):
                a = np.broadcast_to(a, b.shape)
            elif b.shape == (1, 1):
                b = np.broadcast_to(b, a.shape)
        return MX(f(a, b))

    def __add__(self, o): return self._ew(o, lambda a, b: a + b)
    def __radd__(self, o): return self._ew(o, lambda a, b: b + a)
    def __sub__(self, o): return self._ew(o, lambda a, b: a - b)
    def __rsub__(self, o): return self._ew(o, lambda a, b: b - a)
    def __mul__(self, o): return self._ew(o, lambda a, b: a * b)
    def __rmul__(self, o): return self._ew(o, lambda a, b: b * a)
    def __truediv__(self, o): return self._ew(o, lambda a, b: a / b)
    def __rtruediv__(self, o): return self._ew(o, lambda a, b: b / a)
    def __neg__(self): return MX(-self.d)

    def __matmul__(self, o):
        b = MX._raw(o)
        return MX(_matmul(self.d, b))

    def __rmatmul__(self, o):
        return MX(_matmul(MX._raw(o), self.d))


def _matmul(a, b):
    out = np.empty((a.shape[0], b.shape[1]), dtype=object)
    for i in range(a.shape[0]):
        for j in range(b.shape[1]):
            acc = 0.0
            for k in range(a.shape[1]):
                acc = _add(acc, _mul(a[i, k], b[k, j]))
            out[i, j] = acc
    return out


def _as_mx(x):
    return x if isinstance(x, MX) else MX(np.atleast_2d(np.asarray(x, dtype=object)).reshape(-1, 1) if np.ndim(x) <= 1 else x)


def vertcat(*args):
    return MX(np.concatenate([_as_mx(a).d for a in args], axis=0))


def blockcat(rows):
    return MX(np.concatenate([np.concatenate([_as_mx(c).d if isinstance(c, MX) else np.array([[c]], dtype=object) for c in r], axis=1)
                              for r in rows], axis=0))


def skew(v):
    a = _as_mx(v).d.reshape(-1)
    return MX([[0.0, -a[2], a[1]], [a[2], 0.0, -a[0]], [-a[1], a[0], 0.0]])


def _unary(op):
    def f(x):
        if isinstance(x, MX):
            return MX(np.vectorize(lambda e: E(op, e) if isinstance(e, E) else getattr(math, op)(e), otypes=[object])(x.d))
        if isinstance(x, E):
            return E(op, x)
        return getattr(math, op)(float(x))
    return f


sin, cos, tan, sqrt = _unary("sin"), _unary("cos"), _unary("tan"), _unary("sqrt")


def jacobian(expr, var):
    e, v = _as_mx(expr).d.reshape(-1), _as_mx(var).d.reshape(-1)
    out = np.empty((len(e), len(v)), dtype=object)
    for j, x in enumerate(v):
        memo = {}
        for i, f in enumerate(e):
            out[i, j] = _d(f, x, memo)
    return MX(out)


class Function:
    """cs.Function(name, inputs, outputs, input names, output names): positional call -> array (one output) or tuple;
    keyword call -> dict of arrays, like CasADi."""

    def __init__(self, name, ins, outs, in_names=None, out_names=None):
        self.name, self.ins, self.outs = name, [_as_mx(i) for i in ins], [_as_mx(o) for o in outs]
        self.in_names, self.out_names = in_names, out_names

    def _eval(self, vals):
        env = {}
        for sym, val in zip(self.ins, vals):
            val = np.asarray(val, dtype=np.float64)
            val = val.reshape(sym.d.shape) if val.size == sym.d.size else np.broadcast_to(val, sym.d.shape)
            for s, x in zip(sym.d.reshape(-1), val.reshape(-1)):
                env[s] = float(x)
        memo = {}
        return [np.array([[_ev(e, env, memo) for e in row] for row in o.d], dtype=np.float64) for o in self.outs]

    def __call__(self, *args, **kw):
        if kw:
            res = self._eval([kw[n] for n in self.in_names])
            return dict(zip(self.out_names, res))
        res = self._eval(args)
        return res[0] if len(res) == 1 else tuple(res)


def integrator(name, algo, dae, opts):
    """fd(x0=, p=)['xf']: the ODE's flow over opts['tf'] (scipy LSODA at tight tolerances stands for CVODES)."""
    from scipy.integrate import solve_ivp
    f = Function(name + "_ode", [dae["x"], dae["p"]], [dae["ode"]])
    tf = float(opts["tf"])

    def fd(x0=None, p=None):
        x0 = np.asarray(x0, dtype=np.float64).reshape(-1)
        pv = np.asarray(p, dtype=np.float64).reshape(-1)
        sol = solve_ivp(lambda t, x: f(x, pv).reshape(-1), (0.0, tf), x0, method="LSODA", rtol=1e-13, atol=1e-15)
        return {"xf": sol.y[:, -1].reshape(-1, 1)}
    return fd
