"""oracle/mp_truth.py — TEST INFRASTRUCTURE ONLY: a high-precision evaluation of the discretised line solution.

The reference (bin/EMtool_analytical@0xc771-0xc8ae) forms the modal factor as (exp(x) - 1.0) / (lambda a)
with x = lambda_k a t, which loses every digit once |x| approaches 1e-16; the GPU path reproduces that arithmetic bit for
bit on purpose.  To say which side is further from the mathematically exact answer where they differ — and how wrong
both are — this module evaluates

    sigma_n(t) = R_0 t / N + sum_{k>=1} 2 R_k expm1(lambda_k a t) / (N lambda_k a) cos(pi k (2n+1) / (2N)),
    lambda_k = -4 sin^2(k pi / (2N))   (= 2 cos(k pi / N) - 2 without the cancellation),
    R_k = sum_v r_v cos(pi (2v+1) k / (2N)),   r = B u,

with the modal factor in mpmath (40 digits, mpmath.expm1) and the trigonometric sums in 80-bit long double on exactly
reduced integer angles.  Only tests import it.  Cost O(N (S + nodes)), so a handful of nodes of a 65 536-node line takes
seconds.
"""
from __future__ import annotations

import numpy as np

LD = np.longdouble
PI_LD = LD("3.14159265358979323846264338327950288419716939937510")


def _cos_frac(num: np.ndarray, den: int) -> np.ndarray:
    """cos(pi * num / den) for integer arrays, angle reduced exactly to [0, pi/2] before the one long-double division."""
    num = np.asarray(num, dtype=np.int64) % (2 * den)
    num = np.where(num > den, 2 * den - num, num)            # cos is even around pi
    neg = num > den // 2 if den % 2 == 0 else 2 * num > den
    num = np.where(neg, den - num, num)                      # cos(pi - x) = -cos x
    c = np.cos(PI_LD * num.astype(LD) / LD(den))
    return np.where(neg, -c, c)


def modal_factors(N: int, a: float, ts, digits: int = 40) -> np.ndarray:
    """[T, N] long double: f_0 = t, f_k = expm1(lambda_k a t) / (lambda_k a), exact to long-double rounding."""
    import mpmath as mp
    mp.mp.dps = digits
    ts = [mp.mpf(float(t)) for t in np.atleast_1d(ts)]
    am = mp.mpf(float(a))
    out = np.empty((len(ts), N), dtype=LD)
    half = mp.pi / (2 * N)
    for k in range(N):
        if k == 0:
            for j, t in enumerate(ts):
                out[j, 0] = LD(mp.nstr(t, 25))
            continue
        s = mp.sin(k * half)
        la = -4 * s * s * am
        for j, t in enumerate(ts):
            out[j, k] = LD(mp.nstr(mp.expm1(la * t) / la, 25))
    return out


def truth_nodes(N: int, a: float, r_idx, r_val, ts, nodes) -> np.ndarray:
    """sigma at the given nodes and timestamps -> float64 [T, len(nodes)] (rounded once from long double).
    r_idx / r_val: the non-zeros of r = B u (as the reference's SpMV produces them)."""
    r_idx = np.asarray(r_idx, dtype=np.int64)
    r_val = np.asarray(r_val, dtype=np.float64).astype(LD)
    nodes = np.asarray(nodes, dtype=np.int64)
    k = np.arange(N, dtype=np.int64)
    R = np.zeros(N, dtype=LD)
    for v, rv in zip(r_idx, r_val):
        R += rv * _cos_frac((2 * v + 1) * k, 2 * N)
    f = modal_factors(N, a, ts)
    w = np.full(N, LD(2) / LD(N), dtype=LD)
    w[0] = LD(1) / LD(N)
    Y = f * (R * w)[None, :]                                   # [T, N]
    out = np.empty((Y.shape[0], len(nodes)), dtype=np.float64)
    for i, n in enumerate(nodes):
        c = _cos_frac(k * (2 * int(n) + 1), 2 * N)
        out[:, i] = (Y * c[None, :]).sum(axis=1).astype(np.float64)
    return out
