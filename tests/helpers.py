"""Seeded input generators shared by the oracle tests and the GPU parity tests."""

from __future__ import annotations

import numpy as np


def finf(vals):
    return np.array([np.inf if v == "inf" else (-np.inf if v == "-inf" else v) for v in vals], dtype=np.float32)


def uniform(n, m, seed=0):
    return np.random.default_rng(seed).random((n, m), dtype=np.float32)


def integer_grid(n, m, seed=0, levels=16):
    """Heavy ties and duplicates."""
    return np.random.default_rng(seed).integers(0, levels, size=(n, m)).astype(np.float32)


def dtlz2(n, m=3, d=12, seed=0, converged=False):
    """DTLZ2 objectives of uniform decision vectors (reference formula: problems/numerical/dtlz.py:84-103)."""
    rng = np.random.default_rng(seed)
    x = rng.random((n, d))
    if converged:
        x[:, m - 1 :] = 0.5 + (rng.random((n, d - m + 1)) - 0.5) * 2e-3
    g = np.sum((x[:, m - 1 :] - 0.5) ** 2, axis=1, keepdims=True)
    ones = np.ones((n, 1))
    cosx = np.cos(x[:, : m - 1] * np.pi / 2)
    sinx = np.sin(x[:, : m - 1] * np.pi / 2)
    f = (1 + g) * np.fliplr(np.cumprod(np.hstack([ones, cosx]), axis=1)) * np.hstack([ones, sinx[:, ::-1]])
    return f.astype(np.float32)


def nasty(n, m, seed=0):
    """Uniform data salted with NaN, +-inf, -0.0, duplicated rows and tied coordinates."""
    rng = np.random.default_rng(seed)
    x = rng.random((n, m), dtype=np.float32)
    if n >= 16:
        k = max(1, n // 16)
        idx = rng.permutation(n)
        x[idx[:k]] = np.round(x[idx[:k]] * 4) / 4  # ties
        x[idx[k : 2 * k]] = x[idx[2 * k : 3 * k]]  # exact duplicates
        x[idx[3 * k], rng.integers(m)] = np.nan
        x[idx[3 * k + 1]] = np.nan
        x[idx[3 * k + 2]] = np.inf
        x[idx[3 * k + 3]] = -np.inf
        x[idx[3 * k + 4], 0] = np.inf
        x[idx[3 * k + 5], m - 1] = -np.inf
        x[idx[3 * k + 6], 0] = -0.0
        x[idx[3 * k + 7], 0] = 0.0
    return x


def chain(n, m):
    """A single dominance chain: n fronts of one member each (deepest possible peel)."""
    v = np.arange(n, dtype=np.float32)[:, None]
    return np.repeat(v, m, axis=1)


def lsmop1(n, m=10, d=None, seed=0):
    """LSMOP1 objectives (reference: problems/numerical/lsmop.py:127-141) of decision vectors uniform in
    [0, 1]^(m-1) x [0, 10]^(d-m+1), the box of tests/test_lsmop.py:10-14; d = 100 m as the reference's
    base class defaults to (lsmop.py:40-43)."""
    from oracle import lsmop_oracle

    d = d or 100 * m
    rng = np.random.default_rng(seed)
    x = rng.random((n, d), dtype=np.float32)
    x[:, m - 1 :] *= np.float32(10)
    return lsmop_oracle.lsmop1(x, m)
