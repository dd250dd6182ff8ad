"""Stand-ins for the third-party modules the reference hard-imports.

Used ONLY by tests/golden/make_golden.py, in the build container, to execute the
reference's own Python (``/root/reference/safe_control_gym``) without pybullet / gym /
casadi / matplotlib installed.  Nothing here ships, nothing here is read on the GPU box.

* ``pybullet``: a state store (pose + velocity per body id) plus the three quaternion
  helpers restated in oracle/bullet_math.py.  Collision queries return "nothing hit",
  which is what DYN sees for ``getContactPoints`` (no stepSimulation).
* ``gym``: ``Env``, ``spaces.Box`` (float32 casting as gym 0.21) and
  ``utils.seeding.np_random`` (gym 0.21: sha512-hashed seed -> RandomState), with every
  draw recorded so the oracle / kernel can replay the identical numbers.
* ``casadi``, ``matplotlib``: MagicMock (symbolic objects are built but never evaluated).  The one
  casadi Function the env step EVALUATES -- ``SymbolicModel.loss`` inside ``_get_reward`` for
  ``cost: quadratic`` (quadrotor.py:915-930) -- gets a numeric stand-in that computes the expression the
  reference builds it from (quadrotor.py:682, symbolic_systems.py:114-123), so the reference's own
  ``_get_reward`` / ``_get_done`` run for that cost too.
"""
import hashlib
import os
import struct
import sys
import types
from unittest import mock

import numpy as np

REPO = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, REPO)
from oracle import bullet_math as bm  # noqa: E402

DRAW_LOG = []  # (method, value) in call order, across all RandomState proxies


class RecordingRandomState:
    def __init__(self, rng):
        self._rng = rng

    def __getattr__(self, name):
        fn = getattr(self._rng, name)
        if not callable(fn):
            return fn

        def wrapped(*a, **k):
            out = fn(*a, **k)
            DRAW_LOG.append((name, np.array(out, dtype=np.float64).copy()))
            return out
        return wrapped


def _gym_np_random(seed=None):
    """gym 0.21 utils/seeding.py: create_seed / hash_seed / _int_list_from_bigint."""
    if seed is None:
        seed = int.from_bytes(os.urandom(8), "little")
    digest = hashlib.sha512(str(seed).encode("utf8")).digest()[:8]
    a, b = struct.unpack("2I", digest)
    big = a + (b << 32)
    ints = []
    while big > 0:
        big, mod = divmod(big, 2 ** 32)
        ints.append(mod)
    rng = np.random.RandomState()
    rng.seed(ints if ints else [0])
    return RecordingRandomState(rng), seed


class Box:
    def __init__(self, low, high, shape=None, dtype=np.float32):
        if shape is None:
            shape = np.asarray(low).shape
        self.dtype = np.dtype(dtype)
        self.shape = tuple(shape)
        self.low = np.full(self.shape, low, dtype=np.float64).astype(self.dtype) if np.isscalar(low) \
            else np.asarray(low).astype(self.dtype)
        self.high = np.full(self.shape, high, dtype=np.float64).astype(self.dtype) if np.isscalar(high) \
            else np.asarray(high).astype(self.dtype)
        self._rng = np.random.RandomState(0)

    def seed(self, seed=None):
        self._rng = np.random.RandomState(None if seed is None else int(seed) % (2 ** 32))
        return [seed]

    def sample(self):
        return self._rng.uniform(self.low, self.high).astype(self.dtype)


class FakeBullet(types.ModuleType):
    DIRECT, GUI, WORLD_FRAME, URDF_USE_INERTIA_FROM_FILE = 2, 1, 1, 2

    def __init__(self):
        super().__init__("pybullet")
        self.bodies = {}
        self.next_id = 0

    def __getattr__(self, name):  # everything not modelled is a no-op
        if name.startswith("__"):
            raise AttributeError(name)
        return lambda *a, **k: None

    def connect(self, *a, **k):
        return 0

    def resetSimulation(self, **k):
        self.bodies = {}
        self.next_id = 0

    def loadURDF(self, path, basePosition=(0, 0, 0), baseOrientation=(0, 0, 0, 1), **k):
        i = self.next_id
        self.next_id += 1
        self.bodies[i] = dict(pos=tuple(float(v) for v in basePosition), quat=tuple(float(v) for v in baseOrientation),
                              vel=(0.0, 0.0, 0.0), angv=(0.0, 0.0, 0.0), urdf=os.path.basename(str(path)))
        return i

    def resetBasePositionAndOrientation(self, i, pos, quat, **k):
        self.bodies[i]["pos"] = tuple(float(v) for v in pos)
        self.bodies[i]["quat"] = tuple(float(v) for v in quat)

    def resetBaseVelocity(self, i, vel, angv, **k):
        self.bodies[i]["vel"] = tuple(float(v) for v in vel)
        self.bodies[i]["angv"] = tuple(float(v) for v in angv)

    def getBasePositionAndOrientation(self, i, **k):
        return self.bodies[i]["pos"], self.bodies[i]["quat"]

    def getBaseVelocity(self, i, **k):
        return self.bodies[i]["vel"], self.bodies[i]["angv"]

    def getQuaternionFromEuler(self, rpy, **k):
        return bm.quat_from_euler(float(rpy[0]), float(rpy[1]), float(rpy[2]))

    def getEulerFromQuaternion(self, q, **k):
        return bm.euler_from_quat(tuple(float(v) for v in q))

    def getMatrixFromQuaternion(self, q, **k):
        m = bm.matrix_from_quat(tuple(float(v) for v in q))
        return tuple(v for row in m for v in row)

    def getContactPoints(self, **k):
        return ()

    def getClosestPoints(self, **k):
        return ()

    def rayTestBatch(self, rayFromPositions, rayToPositions, **k):
        return [(-1, -1, 1.0, (0, 0, 0), (0, 0, 0)) for _ in rayFromPositions]


class _NumpyCompat:
    """numpy proxy for the reference's numpy-1.x-only calls (np.array(..., copy=False))."""

    def __getattr__(self, name):
        return getattr(np, name)

    @staticmethod
    def array(*a, **k):
        k.pop("copy", None)
        return np.array(*a, **k)


def install(reference_root="/root/reference", casadi=None):
    """``casadi``: a module to install in place of the MagicMock (tests/golden/mini_casadi.py, make_symbolic_golden.py)."""
    if not hasattr(np, "round_"):
        np.round_ = np.round  # removed in numpy 2.x
    gym = types.ModuleType("gym")
    gym.Env = object
    spaces = types.ModuleType("gym.spaces")
    spaces.Box = Box
    utils = types.ModuleType("gym.utils")
    seeding = types.ModuleType("gym.utils.seeding")
    seeding.np_random = _gym_np_random
    utils.seeding = seeding
    gym.spaces, gym.utils = spaces, utils
    sys.modules.update({"gym": gym, "gym.spaces": spaces, "gym.utils": utils, "gym.utils.seeding": seeding})
    sys.modules["pybullet"] = FakeBullet()
    pd = types.ModuleType("pybullet_data")
    pd.getDataPath = lambda: "/nonexistent"
    sys.modules["pybullet_data"] = pd
    for name in ("casadi", "matplotlib", "matplotlib.pyplot"):
        sys.modules[name] = mock.MagicMock()
    if casadi is not None:
        sys.modules["casadi"] = casadi
    # Import the reference's sub-packages without running safe_control_gym/__init__.py
    # (which pulls in every controller: torch, cvxpy, gpytorch, ...).
    pkg = types.ModuleType("safe_control_gym")
    pkg.__path__ = [os.path.join(reference_root, "safe_control_gym")]
    sys.modules["safe_control_gym"] = pkg
    envs = types.ModuleType("safe_control_gym.envs")
    envs.__path__ = [os.path.join(reference_root, "safe_control_gym", "envs")]
    sys.modules["safe_control_gym.envs"] = envs
    import safe_control_gym.math_and_models.transformations as tr
    tr.np = _NumpyCompat()
    from safe_control_gym.envs.gym_pybullet_drones.quadrotor import Quadrotor

    class DynQuadrotor(Quadrotor):
        """F3: ``_dynamics`` reads the never-assigned ``self.TIMESTEP``; the only consistent
        reading is one call per PyBullet substep, i.e. TIMESTEP = PYB_TIMESTEP."""

        @property
        def TIMESTEP(self):
            return self.PYB_TIMESTEP

        # F10: step() calls _get_info() BEFORE _get_done() (quadrotor.py:553-555), and _get_info reads
        # self.goal_reached (:1013), which only _get_done assigns (:965): stabilisation + cost "quadratic" raises
        # AttributeError on the first step as shipped.  The minimal reading that lets the reference's own code run is
        # "not reached yet" before the first _get_done; info["goal_reached"] then reports the PREVIOUS step's flag
        # (and survives reset(), which never clears it).
        goal_reached = False

        def _setup_symbolic(self):
            super()._setup_symbolic()   # builds MagicMock expressions
            if casadi is not None:      # a working casadi stand-in: the reference's own SymbolicModel.loss evaluates
                return

            def loss(x, u, Xr, Ur, Q, R):
                """cost_func of quadrotor.py:682 evaluated on numbers (casadi would do the same arithmetic):
                0.5 (X - Xr)' Q (X - Xr) + 0.5 (U - Ur)' R (U - Ur)."""
                ex = np.asarray(x, dtype=np.float64).reshape(-1) - np.asarray(Xr, dtype=np.float64).reshape(-1)
                eu = np.asarray(u, dtype=np.float64).reshape(-1) - np.asarray(Ur, dtype=np.float64).reshape(-1)
                Qm, Rm = np.asarray(Q, dtype=np.float64), np.asarray(R, dtype=np.float64)
                return {"l": 0.5 * ex @ Qm @ ex + 0.5 * eu @ Rm @ eu}
            self.symbolic.loss = loss

    return DynQuadrotor
