"""Duck-typed stand-in for ``gym.spaces.Box`` (gym is not a dependency of this package).

Keeps what callers of the reference read: ``low / high / shape / dtype / sample() / seed() /
contains()``; bounds are stored in float32 exactly like ``spaces.Box(dtype=np.float32)``
(quadrotor.py:716-733, 790-806), which is what the clip / constraint constants derive from.
"""
import numpy as np


class Box:
    def __init__(self, low, high, shape=None, dtype=np.float32):
        self.dtype = np.dtype(dtype)
        if shape is None:
            shape = np.asarray(low).shape
        self.shape = tuple(shape)
        self.low = np.broadcast_to(np.asarray(low, dtype=np.float64), self.shape).astype(self.dtype)
        self.high = np.broadcast_to(np.asarray(high, dtype=np.float64), self.shape).astype(self.dtype)
        self._rng = np.random.default_rng()

    def seed(self, seed=None):
        self._rng = np.random.default_rng(seed)
        return [seed]

    def sample(self):
        lo = np.where(np.isfinite(self.low), self.low, -1e3)
        hi = np.where(np.isfinite(self.high), self.high, 1e3)
        big = np.finfo(np.float32).max / 4
        lo, hi = np.maximum(lo, -big), np.minimum(hi, big)
        return self._rng.uniform(lo, hi).astype(self.dtype)

    def contains(self, x):
        x = np.asarray(x)
        return x.shape == self.shape and bool(np.all(x >= self.low) and np.all(x <= self.high))

    def __repr__(self):
        return f"Box({self.low.min()}, {self.high.max()}, {self.shape}, {self.dtype})"
