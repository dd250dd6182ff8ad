"""Host-side disturbance descriptors (reference: safe_control_gym/envs/disturbances.py).

The additive noise itself is generated and applied inside the CUDA step kernel
(csrc/scg_env.cuh: apply_channel) from counter-based Philox draws, or from tensors the
caller injects for parity runs.  These classes validate the YAML entries exactly like
the reference constructors do and turn them into ``scg_disturbance`` records.
"""
import numpy as np

from . import _lib as L


class Disturbance:
    kind = None

    def __init__(self, env, dim, mask=None, **kwargs):
        self.dim = dim
        self.mask = None if mask is None else np.asarray(mask, dtype=float)
        if self.mask is not None:
            assert self.dim == len(self.mask)
        self.slot = 0
        self.max_step = int(env.EPISODE_LEN_SEC / env.CTRL_TIMESTEP)

    def to_record(self):
        r = L.Disturbance()
        r.kind, r.slot, r.max_step, r.step_offset = self.kind, self.slot, self.max_step, 0
        if self.mask is not None:
            r.has_mask = 1
            for d, m in enumerate(self.mask):
                r.mask[d] = float(m)
        return r


class ImpulseDisturbance(Disturbance):
    """disturbances.py:89-143."""
    kind = L.DIST_IMPULSE

    def __init__(self, env, dim, mask=None, magnitude=1, step_offset=None, duration=1, decay_rate=1, **kwargs):
        super().__init__(env, dim, mask)
        assert duration >= 1
        assert decay_rate > 0 and decay_rate <= 1
        self.magnitude, self.step_offset, self.duration, self.decay_rate = magnitude, step_offset, duration, decay_rate

    def to_record(self):
        r = super().to_record()
        r.magnitude, r.duration, r.decay_rate = float(self.magnitude), float(self.duration), float(self.decay_rate)
        r.step_offset = -1 if self.step_offset is None else int(self.step_offset)
        return r


class StepDisturbance(Disturbance):
    """disturbances.py:146-184."""
    kind = L.DIST_STEP

    def __init__(self, env, dim, mask=None, magnitude=1, step_offset=None, **kwargs):
        super().__init__(env, dim, mask)
        self.magnitude, self.step_offset = magnitude, step_offset

    def to_record(self):
        r = super().to_record()
        r.magnitude = float(self.magnitude)
        r.step_offset = -1 if self.step_offset is None else int(self.step_offset)
        return r


def _vec(value, dim, what):
    if isinstance(value, float):
        return np.asarray([value] * dim)
    if isinstance(value, list):
        return np.asarray(value, dtype=float)
    raise ValueError(f"[ERROR] {what} must be specified as a float or list.")


class UniformNoise(Disturbance):
    """disturbances.py:186-212."""
    kind = L.DIST_UNIFORM

    def __init__(self, env, dim, mask=None, low=0.0, high=1.0, **kwargs):
        super().__init__(env, dim, mask)
        self.low = _vec(low, dim, "UniformNoise.__init__(): low")
        self.high = _vec(high, dim, "UniformNoise.__init__(): high")
        assert len(self.low) == dim and len(self.high) == dim

    def to_record(self):
        r = super().to_record()
        for d in range(self.dim):
            r.low[d], r.high[d] = float(self.low[d]), float(self.high[d])
        return r


class WhiteNoise(Disturbance):
    """disturbances.py:215-248."""
    kind = L.DIST_WHITE

    def __init__(self, env, dim, mask=None, std=1.0, **kwargs):
        super().__init__(env, dim, mask)
        self.std = _vec(std, dim, "WhiteNoise.__init__(): std")
        assert self.dim == len(self.std), "std shape should be the same as dim."

    def to_record(self):
        r = super().to_record()
        for d in range(self.dim):
            r.std[d] = float(self.std[d])
        return r


class PeriodicNoise(Disturbance):
    """disturbances.py:260-288 (the reference constructor drops ``mask``)."""
    kind = L.DIST_PERIODIC

    def __init__(self, env, dim, mask=None, scale=1.0, frequency=1.0, **kwargs):
        super().__init__(env, dim)
        self.scale, self.frequency = scale, frequency

    def to_record(self):
        r = super().to_record()
        r.scale, r.frequency = float(self.scale), float(self.frequency)
        return r


class DisturbanceList:
    def __init__(self, disturbances):
        self.disturbances = disturbances

    def __len__(self):
        return len(self.disturbances)

    def __iter__(self):
        return iter(self.disturbances)


DISTURBANCE_TYPES = {"impulse": ImpulseDisturbance, "step": StepDisturbance, "uniform": UniformNoise,
                     "white_noise": WhiteNoise, "periodic": PeriodicNoise}


def create_disturbance_list(disturbance_specs, shared_args, env):
    """disturbances.py:315-333."""
    out = []
    if len(disturbance_specs) > L.MAX_DIST:
        raise ValueError(f"at most {L.MAX_DIST} disturbances per channel are supported")
    for spec in disturbance_specs:
        assert "disturbance_func" in spec, "[ERROR]: Every distrubance must specify a disturbance_func."
        func = spec["disturbance_func"]
        assert func in DISTURBANCE_TYPES, "[ERROR] in BenchmarkEnv._setup_disturbances(), disturbance type not available."
        out.append(DISTURBANCE_TYPES[func](env, **shared_args, **{k: v for k, v in spec.items() if k != "disturbance_func"}))
    return DisturbanceList(out)
