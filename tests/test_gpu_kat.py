"""SURVEY.md section 8(c) known-answer cases through the C ABI: one env, reference scalar signatures
(``num_envs=1, squeeze=True``), fp64 with the literal attitude sequence and the fp64 / fp32 fast paths."""
import pytest
import torch

import kat_cases

pytestmark = pytest.mark.gpu


def _maker(dtype, attitude):
    from safe_control_gym_b200.envs.quadrotor import BatchedQuadrotor

    def make(**kw):
        return BatchedQuadrotor(num_envs=1, squeeze=True, dtype=dtype, attitude=attitude, **kw)
    return make


@pytest.mark.parametrize("dtype,attitude,tol", [(torch.float64, "roundtrip", 1e-13), (torch.float64, "fast", 1e-13),
                                                (torch.float32, "fast", 2e-6), (torch.float32, "fast_intrinsic", 2e-6)])
@pytest.mark.parametrize("case", kat_cases.ALL, ids=lambda f: f.__name__)
def test_cuda_known_answers(case, dtype, attitude, tol):
    case(_maker(dtype, attitude), tol)
