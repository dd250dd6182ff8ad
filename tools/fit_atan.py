"""Coefficients of atan2_fast (csrc/scg_math.cuh): atan(a) ~= a * P(a^2) on [0, 1], P of degree 7 with
P(0) = 1, iteratively re-weighted least squares on Chebyshev nodes (near-minimax), error evaluated
in float32 Horner arithmetic.   python tools/fit_atan.py"""
import numpy as np


def fit(deg, iters=60):
    k = np.arange(4000)
    a = 0.5 * (1 - np.cos(np.pi * (k + 0.5) / 4000))
    z = a * a
    y = np.arctan(a) / a
    V = np.vander(z, deg + 1, increasing=True)[:, 1:]
    w = np.ones_like(a)
    for _ in range(iters):
        c = np.linalg.lstsq(V * w[:, None], (y - 1) * w, rcond=None)[0]
        err = np.abs((V @ c + 1) * a - np.arctan(a))
        w = w * (1 + 2 * err / err.max())
        w /= w.mean()
    return np.concatenate([[1.0], c])


if __name__ == "__main__":
    c = fit(7)
    a = np.linspace(0, 1, 200001).astype(np.float32)
    z = a * a
    p = np.float32(c[-1])
    for cc in c[-2::-1]:
        p = p * z + np.float32(cc)
    print("max abs err (float32 Horner):", np.abs((p * a).astype(np.float64) - np.arctan(a.astype(np.float64))).max())
    print(", ".join(f"{np.float32(x):.10e}f" for x in c))
