"""Philox4x32-10 known-answer vectors (Random123 kat_vectors) + transform sanity."""
import numpy as np

from oracle import philox as px


def test_philox_kat():
    kat = [((0, 0, 0, 0), (0, 0), (0x6627e8d5, 0xe169c58d, 0xbc57ac4c, 0x9b00dbd8)),
           ((0xffffffff,) * 4, (0xffffffff,) * 2, (0x408f276d, 0x41c83b0e, 0xa20bc7c6, 0x6d5451fd)),
           ((0x243f6a88, 0x85a308d3, 0x13198a2e, 0x03707344), (0xa4093822, 0x299f31d0),
            (0xd16cfe09, 0x94fdcceb, 0x5001e420, 0x24126ea1))]
    for c, k, want in kat:
        got = tuple(int(v) for v in px.philox4x32_10(c, k))
        assert got == want


def test_transforms():
    w = px.philox4x32_10((np.arange(20000), 0, 16, 0), (1337, 5))
    u = px.u01(w[0])
    assert 0.0 < u.min() and u.max() < 1.0 and abs(u.mean() - 0.5) < 0.01
    n0, n1 = px.normal_pair(w[0], w[1])
    assert abs(n0.mean()) < 0.03 and abs(n0.std() - 1) < 0.03 and abs(np.mean(n0 * n1)) < 0.03
    r = px.randint(w[2], 1980)
    assert r.min() >= 0 and r.max() < 1980
