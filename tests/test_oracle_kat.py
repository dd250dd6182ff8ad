"""SURVEY.md section 8(c) known-answer cases against the numpy oracle (CPU)."""
import pytest

import kat_cases
from oracle.quad_oracle import QuadOracle


def _make(**kw):
    return QuadOracle(**kw)


@pytest.mark.parametrize("case", kat_cases.ALL, ids=lambda f: f.__name__)
def test_oracle_known_answers(case):
    case(_make, 1e-13)
