"""Host-side constraint objects: the ``ConstraintList`` API of the reference
(safe_control_gym/envs/constraints.py:21-612) over plain matrices.

The per-step evaluation for the batch happens inside the CUDA step kernel
(csrc/scg_env.cuh: eval_constraints); these objects (a) parse the YAML entries into the
``scg_constraint`` records of the C ABI and (b) keep the attribute surface controllers read:
``constraints / state_constraints / input_constraints / num_constraints``, per-constraint
``A / b / lower_bounds / upper_bounds / P / sym_func / constraint_filter / tolerance`` and
``get_*_symbolic_models()``.  ``sym_func`` is a Python callable over numpy arrays or torch
tensors (no CasADi here), so MPC-style consumers can still evaluate g(x).
"""
import numpy as np

from . import _lib as L

STATE, INPUT, INPUT_AND_STATE = "state", "input", "input_and_state"


def _round(x, decimals):
    if hasattr(x, "detach"):  # torch tensor
        import torch
        return torch.round(x, decimals=decimals)
    return np.round(x, decimals=decimals)


class Constraint:
    """One g(x) <= 0 block.  ``form`` selects the expression; rows are ordered as the reference
    orders them (bounded: all lower rows, then all upper rows)."""

    form = None

    def __init__(self, env, constrained_variable, strict=False, active_dims=None, tolerance=None, rounding=8, **_):
        if constrained_variable not in (STATE, INPUT, INPUT_AND_STATE):
            raise NotImplementedError("[ERROR] invalid constrained_variable (use STATE, INPUT or INPUT_AND_STATE).")
        if constrained_variable == INPUT_AND_STATE:
            raise NotImplementedError("input_and_state constraints cannot be evaluated by the reference either "
                                      "(ragged np.array, constraints.py:169-170); not supported")
        self.constrained_variable = constrained_variable
        self.rounding = rounding
        self.strict = bool(strict)
        full = env.state_dim if constrained_variable == STATE else env.action_dim
        if active_dims is not None:
            if isinstance(active_dims, int):
                active_dims = [active_dims]
            assert isinstance(active_dims, (list, np.ndarray)), "[ERROR] active_dims is not a list/array."
            assert len(active_dims) <= full, "[ERROR] more active_dim than constrainable self.dim"
            assert all(isinstance(n, (int, np.integer)) for n in active_dims), "[ERROR] non-integer active_dim."
            assert all(n < full for n in active_dims), "[ERROR] active_dim not stricly smaller than self.dim."
            assert len(active_dims) == len(set(active_dims)), "[ERROR] duplicates in active_dim"
            self.active_dims = [int(n) for n in active_dims]
        else:
            self.active_dims = list(range(full))
        self.constraint_filter = np.eye(full)[self.active_dims]
        self.dim = len(self.active_dims)
        self.tolerance = None if tolerance is None else np.array(tolerance, ndmin=1, dtype=float)

    # -- evaluation on host arrays (x: [..., dim_full]) ------------------------------
    def _select(self, x):
        return x[..., self.active_dims]

    def sym_func(self, x):
        raise NotImplementedError

    def get_symbolic_model(self):
        return self.sym_func

    def get_env_constraint_var(self, env):
        return env.state if self.constrained_variable == STATE else env.current_raw_input_action

    def get_value(self, env):
        return _round(self.sym_func(self.get_env_constraint_var(env)), self.rounding)

    def is_violated(self, env, c_value=None):
        c = self.get_value(env) if c_value is None else c_value
        flag = (c >= 0.0) if self.strict else (c > 0.0)
        return flag.any(-1) if getattr(flag, "ndim", 1) > 1 else bool(flag.any())

    def is_almost_active(self, env, c_value=None):
        if self.tolerance is None:
            return False
        c = self.get_value(env) if c_value is None else c_value
        tol = self.tolerance
        if hasattr(c, "detach"):
            import torch
            tol = torch.as_tensor(tol, dtype=c.dtype, device=c.device)
        flag = (c + tol) > 0.0
        return flag.any(-1) if getattr(flag, "ndim", 1) > 1 else bool(flag.any())

    def check_tolerance_shape(self):
        if self.tolerance is not None and len(self.tolerance) != self.num_constraints:
            raise ValueError("[ERROR] the tolerance dimension does not match the number of constraints.")

    def _base_record(self, form):
        r = L.Constraint()
        r.form, r.var = form, (L.VAR_STATE if self.constrained_variable == STATE else L.VAR_INPUT)
        r.strict, r.n_rows, r.n_active = int(self.strict), int(self.num_constraints), int(self.dim)
        for j, d in enumerate(self.active_dims):
            r.active_dims[j] = d
        if self.tolerance is not None:
            r.has_tolerance = 1
            for j, t in enumerate(self.tolerance):
                r.tolerance[j] = float(t)
        return r


class LinearConstraint(Constraint):
    """A @ F @ x - b <= 0 (constraints.py:226-270)."""

    def __init__(self, env, A, b, constrained_variable, strict=False, active_dims=None, tolerance=None, **kw):
        super().__init__(env, constrained_variable, strict=strict, active_dims=active_dims, tolerance=tolerance)
        self.A = np.array(A, ndmin=1, dtype=float)
        self.b = np.array(b, ndmin=1, dtype=float)
        assert self.A.ndim == 2 and self.A.shape[1] == self.dim, "[ERROR] A has the wrong dimension!"
        assert self.b.shape[0] == self.A.shape[0], "[ERROR] Dimension 0 of b does not match A!"
        self.num_constraints = self.A.shape[0]
        self.check_tolerance_shape()

    def sym_func(self, x):
        A, b = self.A, self.b
        if hasattr(x, "detach"):
            import torch
            A = torch.as_tensor(A, dtype=x.dtype, device=x.device)
            b = torch.as_tensor(b, dtype=x.dtype, device=x.device)
        else:
            x = np.asarray(x, dtype=float)
        return self._select(x) @ A.T - b

    def to_record(self):
        if self.num_constraints > L.MAX_CON_ROWS:
            raise ValueError("linear constraint has too many rows for the C ABI")
        r = self._base_record(L.CON_LINEAR)
        for i in range(self.num_constraints):
            for j in range(self.dim):
                r.A[i * self.dim + j] = float(self.A[i, j])
            r.b[i] = float(self.b[i])
        return r


class BoundedConstraint(LinearConstraint):
    """lb <= F x <= ub as [-I; I] F x - [-lb; ub] <= 0 (constraints.py:273-304)."""

    def __init__(self, env, lower_bounds, upper_bounds, constrained_variable, strict=False, active_dims=None,
                 tolerance=None, **kw):
        self.lower_bounds = np.array(lower_bounds, ndmin=1, dtype=float)
        self.upper_bounds = np.array(upper_bounds, ndmin=1, dtype=float)
        n = self.lower_bounds.shape[0]
        super().__init__(env, np.vstack((-np.eye(n), np.eye(n))), np.hstack((-self.lower_bounds, self.upper_bounds)),
                         constrained_variable, strict=strict, active_dims=active_dims, tolerance=tolerance)

    def to_record(self):
        r = self._base_record(L.CON_BOUNDED)
        for j in range(self.dim):
            r.lower[j], r.upper[j] = float(self.lower_bounds[j]), float(self.upper_bounds[j])
        return r


class DefaultConstraint(BoundedConstraint):
    """Bounds taken from the env's float32 state / action space, promoted to float64
    (constraints.py:307-368)."""

    def __init__(self, env, constrained_variable, lower_bounds=None, upper_bounds=None, strict=False, tolerance=None, **kw):
        if constrained_variable == STATE:
            space = env.state_space if hasattr(env, "state_space") else env.observation_space
        elif constrained_variable == INPUT:
            space = env.action_space
        else:
            raise NotImplementedError("[ERROR] DefaultConstraint can only be of type STATE or INPUT")
        ub = space.high if upper_bounds is None else np.array(upper_bounds, ndmin=1)
        lb = space.low if lower_bounds is None else np.array(lower_bounds, ndmin=1)
        assert len(ub) == space.shape[0], "[ERROR]: Upper bound must have length equal to space dimension."
        assert len(lb) == space.shape[0], "[ERROR]: Lower bound must have length equal to space dimension."
        super().__init__(env, lb.astype(np.float64), ub.astype(np.float64), constrained_variable, strict=strict,
                         active_dims=None, tolerance=tolerance)


class QuadraticContstraint(Constraint):
    """x' F' P F x - b <= 0 (constraints.py:179-223; the class name keeps the reference's spelling)."""

    def __init__(self, env, P, b, constrained_variable, strict=False, active_dims=None, tolerance=None, **kw):
        super().__init__(env, constrained_variable, strict=strict, active_dims=active_dims, tolerance=tolerance)
        self.P = np.array(P, ndmin=1, dtype=float)
        assert self.P.shape == (self.dim, self.dim), "[ERROR] P has the wrong dimension!"
        assert isinstance(b, float), "[ERROR] b is not a scalar!"
        self.b = b
        self.num_constraints = 1
        self.check_tolerance_shape()

    def sym_func(self, x):
        P = self.P
        if hasattr(x, "detach"):
            import torch
            P = torch.as_tensor(P, dtype=x.dtype, device=x.device)
        else:
            x = np.asarray(x, dtype=float)
        xs = self._select(x)
        return ((xs @ P) * xs).sum(-1, keepdims=True) - self.b

    def to_record(self):
        r = self._base_record(L.CON_QUADRATIC)
        for i in range(self.dim):
            for j in range(self.dim):
                r.A[i * self.dim + j] = float(self.P[i, j])
        r.b[0] = float(self.b)
        return r


class SymmetricStateConstraint(BoundedConstraint):
    """|F x| - bound <= 0, value NOT rounded (constraints.py:371-415); cart-pole RL envs only."""

    def __init__(self, env, constrained_variable, bound, strict=False, active_dims=None, tolerance=None, **kw):
        assert bound is not None
        self.bound = np.array(bound, ndmin=1, dtype=float)
        assert env.NAME == "cartpole", "[ERROR] SymmetricStateConstraint is meant for CartPole environments"
        assert env.COST == "rl_reward", "[ERROR] SymmetricStateConstraint is meant for RL environments"
        tol, self._tol = tolerance, tolerance
        super().__init__(env, -self.bound, self.bound, constrained_variable, strict=strict, active_dims=active_dims,
                         tolerance=None)
        self.tolerance = None if tol is None else np.array(tol, ndmin=1, dtype=float)
        self.num_constraints = self.bound.shape[0]
        if self.tolerance is not None and len(self.tolerance) != len(self.bound):
            raise ValueError("[ERROR] the tolerance dimension does not match the number of constraints.")

    def sym_func(self, x):
        xs = self._select(x if hasattr(x, "detach") else np.asarray(x, dtype=float))
        bound = self.bound
        if hasattr(x, "detach"):
            import torch
            bound = torch.as_tensor(bound, dtype=x.dtype, device=x.device)
        return abs(xs) - bound

    def get_value(self, env):
        return self.sym_func(env.state)

    def to_record(self):
        r = self._base_record(L.CON_SYMMETRIC)
        for j in range(self.dim):
            r.lower[j], r.upper[j] = -float(self.bound[j]), float(self.bound[j])
        return r


class ConstraintList:
    """Ordered collection (constraints.py:426-583)."""

    def __init__(self, constraints):
        self.constraints = constraints
        self.constraint_lengths = [c.num_constraints for c in constraints]
        self.constraint_indices = np.cumsum(self.constraint_lengths[:-1])
        self.num_constraints = sum(self.constraint_lengths)
        self.state_constraints = [c for c in constraints if c.constrained_variable == STATE]
        self.num_state_constraints = sum(c.num_constraints for c in self.state_constraints)
        self.input_constraints = [c for c in constraints if c.constrained_variable == INPUT]
        self.num_input_constraints = sum(c.num_constraints for c in self.input_constraints)
        self.input_state_constraints = []
        self.num_input_state_constraints = 0

    def __len__(self):
        return len(self.constraints)

    def get_all_symbolic_models(self):
        return [c.get_symbolic_model() for c in self.constraints]

    def get_state_constraint_symbolic_models(self):
        return [c.get_symbolic_model() for c in self.state_constraints]

    def get_input_constraint_symbolic_models(self):
        return [c.get_symbolic_model() for c in self.input_constraints]

    def get_input_and_state_constraint_symbolic_models(self):
        return []

    def get_stacked_symbolic_model(self, env=None):
        """Callable (x, u) -> stacked g; replaces the cs.Function of constraints.py:486-501."""
        def stacked(x, u):
            parts = [c.sym_func(x if c.constrained_variable == STATE else u) for c in self.constraints]
            if hasattr(parts[0], "detach"):
                import torch
                return torch.cat(parts, -1)
            return np.concatenate(parts, -1)
        return stacked

    def _cat(self, parts):
        if parts and hasattr(parts[0], "detach"):
            import torch
            return torch.cat(parts, -1)
        return np.concatenate(parts, -1) if parts else np.zeros(0)

    def get_values(self, env, only_state=False):
        cons = self.state_constraints if only_state else self.constraints
        return self._cat([c.get_value(env) for c in cons])

    def get_violations(self, env, only_state=False):
        cons = self.state_constraints if only_state else self.constraints
        return [c.is_violated(env) for c in cons]

    def _split(self, c_value):
        idx = [int(i) for i in self.constraint_indices]
        if hasattr(c_value, "detach"):
            import torch
            sizes = self.constraint_lengths
            return torch.split(c_value, sizes, dim=-1)
        return np.split(c_value, idx, axis=-1)

    def is_violated(self, env, c_value=None):
        if c_value is not None:
            flags = [c.is_violated(env, c_value=v) for c, v in zip(self.constraints, self._split(c_value))]
        else:
            flags = [c.is_violated(env) for c in self.constraints]
        out = flags[0]
        for f in flags[1:]:
            out = out | f
        return out

    def is_almost_active(self, env, c_value=None):
        if c_value is not None:
            flags = [c.is_almost_active(env, c_value=v) for c, v in zip(self.constraints, self._split(c_value))]
        else:
            flags = [c.is_almost_active(env) for c in self.constraints]
        out = flags[0]
        for f in flags[1:]:
            out = out | f
        return out

    def to_records(self):
        return [c.to_record() for c in self.constraints]


GENERAL_CONSTRAINTS = {"linear_constraint": LinearConstraint, "quadratic_constraint": QuadraticContstraint,
                       "bounded_constraint": BoundedConstraint, "default_constraint": DefaultConstraint}


def create_constraint_list(constraint_specs, available_constraints, env):
    """constraints.py:594-612."""
    out = []
    for spec in constraint_specs:
        assert isinstance(spec, dict), "[ERROR]: Each constraint must be specified as a dict."
        assert "constraint_form" in spec, "[ERROR]: Each constraint must have a key 'constraint_form'"
        form = spec["constraint_form"]
        assert form in available_constraints, "[ERROR]. constraint not in list of available constraints"
        out.append(available_constraints[form](env, **{k: v for k, v in spec.items() if k != "constraint_form"}))
    return ConstraintList(out)
