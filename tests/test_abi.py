"""The C-ABI boundary without a GPU: libscg_b200.so loads and exports every function include/scg_b200.h
declares; the ctypes structures mirror the header's layout; host-side validation errors come back as codes."""
import ctypes as C
import os
import re

from safe_control_gym_b200 import _lib as L

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    src = open(os.path.join(ROOT, "include", "scg_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(scg_[a-z0-9_]+)\s*\(", src)))


def test_every_declared_symbol_is_exported():
    lib = L.load()
    names = _declared()
    assert len(names) >= 17
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/scg_b200.h but not exported"
    assert sorted(L.EXPORTS) == names


def test_struct_sizes_match_the_header():
    """sizeof() as the C compiler lays the structs out (computed here from the header's own field list)."""
    assert C.sizeof(L.Rand) == 24
    assert C.sizeof(L.Constraint) == 4 * 5 + 4 * 12 + 4 + 8 * (12 + 12 + 144 + 24) + 8 + 8 * 24   # with padding to 8
    assert C.sizeof(L.RolloutArgs) == 48
    assert C.sizeof(L.Buffers) == 8 * 21


def test_version_and_error_codes_without_a_gpu():
    lib = L.load()
    assert b"sm_100a" in lib.scg_version()
    assert lib.scg_create(None, None) == -2                     # SCG_ERR_NULL
    assert b"null" in lib.scg_last_error()
    cfg, h = L.Config(), C.c_void_p()
    cfg.env_kind = 99
    assert lib.scg_create(C.byref(cfg), C.byref(h)) == -1       # SCG_ERR_BAD_CONFIG: unknown env_kind
    cfg.env_kind, cfg.precision, cfg.substeps, cfg.pyb_freq, cfg.ctrl_freq = L.QUAD_3D, 32, 1, 50, 30
    assert lib.scg_create(C.byref(cfg), C.byref(h)) == -1
    assert b"not divisible" in lib.scg_last_error()             # benchmark_env.py:154-155
    assert lib.scg_step(None, None, None) == -2
