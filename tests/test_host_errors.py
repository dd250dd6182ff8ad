"""Error conventions of the reference that the facade must keep (last row): they are raised from
the constructor BEFORE anything touches the GPU, so they are checked here without one."""
import pytest

import kat_cases
from safe_control_gym_b200.envs.quadrotor import BatchedQuadrotor


def _make(**over):
    return BatchedQuadrotor(num_envs=4, **kat_cases.base(**over))


@pytest.mark.parametrize("over,exc,text", [
    (dict(pyb_freq=70), ValueError, "pyb_freq is not divisible"),                          # benchmark_env.py:154-155
    (dict(init_state=[1.0, 2.0]), ValueError, "init_state incorrect format"),              # quadrotor.py:224
    (dict(inertial_prop=[1.0, 2.0]), ValueError, "inertial_prop incorrect format"),        # quadrotor.py:256
    (dict(info_mse_metric_state_weight=[1, 2]), ValueError, "info_mse_metric_state_weight"),  # quadrotor.py:193,200
    (dict(gates=[[0, 0, 0, 0, 0, 0, 5]]), ValueError, "Unknown gate type"),                 # quadrotor.py:414
    (dict(randomized_gates_and_obstacles=True), ValueError, "gates_and_obstacles_randomization_info"),
    (dict(constraints=[dict(constraint_form="nope")]), AssertionError, "not in list of available"),   # constraints.py:604-608
    (dict(disturbances={"action": [dict(disturbance_func="nope")]}), AssertionError, "disturbance type not available"),  # disturbances.py:326-328
    (dict(disturbances={"bogus": [dict(disturbance_func="impulse")]}), AssertionError, "disturbance mode not available"),  # benchmark_env.py:320
    (dict(task="hovering"), ValueError, "not a valid Task"),
    (dict(cost="nope"), ValueError, "not a valid Cost"),
    (dict(physics="pyb"), NotImplementedError, "dyn"),
    (dict(gui=True), NotImplementedError, "unavailable on this box"),
])
def test_constructor_errors(over, exc, text):
    with pytest.raises(exc, match=text):
        _make(**over)


def test_no_cpu_fallback():
    """Without a CUDA device the env refuses to exist (there is no CPU product path)."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("needs a host without a GPU")
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        _make()
