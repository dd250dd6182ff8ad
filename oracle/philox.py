"""Philox4x32-10 and the draw transforms shared with the CUDA kernels.

TEST INFRASTRUCTURE (see oracle/__init__.py).

The reference draws from gym 0.21's ``np.random.RandomState`` (MT19937;
benchmark_env.py:231, disturbances.py:123,170,208,241).  A sequential stream
cannot be reproduced by millions of lock-step envs, so the build defines a
counter-based contract instead:

    key     = (seed, global_env_id)
    counter = (index, episode, stream, block)

``index`` is the ctrl step (or observation-call index), ``stream`` names the
consumer (see STREAM_*), ``block`` extends a draw beyond 4 values.  One Philox
call yields four 32-bit words:

    u01(word)     = ((word >> 8) + 0.5) * 2**-24          (exact in fp32 and fp64)
    uniform(a, b) = a + (b - a) * u01
    normal pair   = sqrt(-2 ln u01(w0)) * (cos, sin)(2 pi u01(w1))   (Box-Muller)
    randint(n)    = (word * n) >> 32

Known-answer vectors (Random123 kat_vectors, philox4x32 10 rounds) are checked
in tests/test_oracle_philox.py.
"""
import numpy as np

M0 = np.uint64(0xD2511F53)
M1 = np.uint64(0xCD9E8D57)
W0 = 0x9E3779B9
W1 = 0xBB67AE85
MASK32 = np.uint64(0xFFFFFFFF)

# Stream ids (counter word 2).  Keep in sync with csrc/scg_rng.cuh.
STREAM_RESET_INIT = 0      # init-state randomisation, value i of block b = label 4*b+i
STREAM_RESET_INERTIA = 1   # M, Ixx, Iyy, Izz
STREAM_RESET_OFFSETS = 2   # impulse / step disturbance offsets, slot i
STREAM_RESET_GATES = 3     # block g: (dx, dy, dyaw, -) of gate g
STREAM_RESET_OBSTACLES = 4 # block o: (dx, dy, dyaw, -) of obstacle o
STREAM_DISTURB_BASE = 16   # + channel * 8 + index of the disturbance in its channel
CHANNEL_OBSERVATION = 0
CHANNEL_ACTION = 1
CHANNEL_DYNAMICS = 2


def philox4x32_10(counter, key):
    """counter: 4 arrays (or ints) of uint32; key: 2 arrays of uint32. Returns 4 uint32 arrays."""
    c = [np.asarray(v, dtype=np.uint64) & MASK32 for v in counter]
    c = list(np.broadcast_arrays(*c, *[np.asarray(k, dtype=np.uint64) for k in key]))
    k0 = c[4].copy() & MASK32
    k1 = c[5].copy() & MASK32
    c = [v.copy() for v in c[:4]]
    for r in range(10):
        p0 = M0 * c[0]
        p1 = M1 * c[2]
        hi0, lo0 = p0 >> np.uint64(32), p0 & MASK32
        hi1, lo1 = p1 >> np.uint64(32), p1 & MASK32
        c = [(hi1 ^ c[1] ^ k0) & MASK32, lo1, (hi0 ^ c[3] ^ k1) & MASK32, lo0]
        k0 = (k0 + np.uint64(W0)) & MASK32
        k1 = (k1 + np.uint64(W1)) & MASK32
    return [v.astype(np.uint32) for v in c]


def u01(words):
    return ((np.asarray(words, dtype=np.uint32) >> np.uint32(8)).astype(np.float64) + 0.5) * (2.0 ** -24)


def normal_pair(w0, w1):
    r = np.sqrt(-2.0 * np.log(u01(w0)))
    a = 2.0 * np.pi * u01(w1)
    return r * np.cos(a), r * np.sin(a)


def randint(word, n):
    return ((np.asarray(word, dtype=np.uint64) * np.uint64(n)) >> np.uint64(32)).astype(np.int64)


class PhiloxDraws:
    """Per-env draw provider used by the oracle envs (same contract as the kernels)."""

    def __init__(self, seed, env_id):
        self.seed = int(seed) & 0xFFFFFFFF
        self.env_id = int(env_id) & 0xFFFFFFFF

    def words(self, index, episode, stream, block=0):
        w = philox4x32_10((index, episode, stream, block), (self.seed, self.env_id))
        return [int(np.asarray(v).reshape(-1)[0]) for v in w]

    def uniforms(self, n, index, episode, stream):
        """n values of u01, value i = word i%4 of block i//4."""
        out = []
        for b in range((n + 3) // 4):
            out.extend(float(u01(w)) for w in self.words(index, episode, stream, b))
        return np.array(out[:n])

    def normals(self, n, index, episode, stream):
        out = []
        for b in range((n + 3) // 4):
            w = self.words(index, episode, stream, b)
            n0, n1 = normal_pair(w[0], w[1])
            n2, n3 = normal_pair(w[2], w[3])
            out.extend([float(n0), float(n1), float(n2), float(n3)])
        return np.array(out[:n])

    def randint(self, n, slot, episode, stream):
        w = self.words(0, episode, stream, slot // 4)
        return int(randint(w[slot % 4], n))
