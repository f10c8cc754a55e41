"""NumPy restatement of the reference's SBX, polynomial mutation, clip and DTLZ2 (test infrastructure only).

The formulas follow the reference line by line; the uniforms come from a NumPy Philox-4x32-10 laid out
exactly like csrc/variation.cuh (JAX's threefry streams are not reproducible here, so the random
numbers are an input of the comparison, not part of it).
  src/evox/operators/crossover/sbx.py:8-41, src/evox/operators/mutation/pm_mutation.py:6-54,
  src/evox/algorithms/mo/nsga2.py:76-84, src/evox/problems/numerical/dtlz.py:84-103
"""

from __future__ import annotations

import numpy as np

STREAM_SBX, STREAM_SBX_PAIR, STREAM_PM = 1, 2, 3


def philox4x32_10(c0, c1, key0, key1):
    """Vectorised Philox-4x32-10 with counter (c0, c1, 0, 0); returns four uint32 arrays."""
    c0 = np.asarray(c0, dtype=np.uint64) & 0xFFFFFFFF
    c1 = np.full_like(c0, c1, dtype=np.uint64)
    c2 = np.zeros_like(c0)
    c3 = np.zeros_like(c0)
    k0, k1 = np.uint64(key0 & 0xFFFFFFFF), np.uint64(key1 & 0xFFFFFFFF)
    M0, M1 = np.uint64(0xD2511F53), np.uint64(0xCD9E8D57)
    mask = np.uint64(0xFFFFFFFF)
    for _ in range(10):
        p0, p1 = M0 * c0, M1 * c2
        n0 = ((p1 >> np.uint64(32)) ^ c1 ^ k0) & mask
        n1 = p1 & mask
        n2 = ((p0 >> np.uint64(32)) ^ c3 ^ k1) & mask
        n3 = p0 & mask
        c0, c1, c2, c3 = n0, n1, n2, n3
        k0 = (k0 + np.uint64(0x9E3779B9)) & mask
        k1 = (k1 + np.uint64(0xBB67AE85)) & mask
    return c0.astype(np.uint32), c1.astype(np.uint32), c2.astype(np.uint32), c3.astype(np.uint32)


def u01(bits):
    return (bits >> np.uint32(8)).astype(np.float32) * np.float32(1.0 / 16777216.0)


def simulated_binary(x, pro_c, dis_c, seed, gen):
    """sbx.py:8-41, type 1."""
    f = np.float32
    n, d = x.shape
    half = n // 2
    p1, p2 = x[:half], x[half : half * 2]
    el = np.arange(half * d).reshape(half, d)
    r0, r1, r2, _ = philox4x32_10(el, STREAM_SBX, seed, gen)
    q0, _, _, _ = philox4x32_10(np.arange(half), STREAM_SBX_PAIR, seed, gen)
    mu = u01(r0)
    with np.errstate(all="ignore"):
        beta = np.where(mu <= 0.5, np.power(f(2) * mu, f(1) / f(dis_c + 1)), np.power(f(2) - f(2) * mu, f(-1) / f(dis_c + 1)))
    beta = beta * np.where(r1 & 1, f(-1), f(1))                        # :19-21
    beta = np.where(u01(r2) < 0.5, f(1), beta)                         # :22
    beta = np.where((u01(q0) > f(pro_c))[:, None], f(1), beta)         # :23-25
    beta = beta.astype(np.float32)
    o1 = (p1 + p2) / f(2) + beta * (p1 - p2) / f(2)                    # :27-33
    o2 = (p1 + p2) / f(2) - beta * (p1 - p2) / f(2)
    out = np.vstack([o1, o2]).astype(np.float32)
    if n % 2:
        out = np.vstack([out, x[-1:]])                                 # :35-36
    return out


def polynomial(x, lb, ub, pro_m, dis_m, seed, gen):
    """pm_mutation.py:6-54."""
    f = np.float32
    n_all, d = x.shape
    pop = x[: (n_all // 2) * 2].copy() if n_all > 1 else x.copy()
    n = pop.shape[0]
    el = np.arange(n * d).reshape(n, d)
    r0, r1, _, _ = philox4x32_10(el, STREAM_PM, seed, gen)
    site = u01(r0) < f(pro_m) / f(d)
    mu = u01(r1)
    lower, upper = np.tile(lb, (n, 1)), np.tile(ub, (n, 1))
    pop = np.maximum(np.minimum(pop, upper), lower)
    with np.errstate(all="ignore"):
        temp = site & (mu <= 0.5)
        norm = np.where(temp, (pop - lower) / (upper - lower), f(0))
        pop = np.where(temp, pop + (upper - lower) * (np.power(f(2) * mu + (f(1) - f(2) * mu) * np.power(f(1) - norm, f(dis_m + 1)),
                                                               f(1) / f(dis_m + 1)) - f(1)), pop).astype(np.float32)
        temp = site & (mu > 0.5)
        norm = np.where(temp, (upper - pop) / (upper - lower), f(0))
        pop = np.where(temp, pop + (upper - lower) * (f(1) - np.power(f(2) * (f(1) - mu) + f(2) * (mu - f(0.5)) * np.power(f(1) - norm, f(dis_m + 1)),
                                                                    f(1) / f(dis_m + 1))), pop).astype(np.float32)
    if n_all % 2 and n_all > 1:
        pop = np.vstack([pop, x[-1:]])
    return pop


def nsga2_ask(pop, lb, ub, seed, gen, pro_c=1.0, dis_c=20.0, pro_m=1.0, dis_m=20.0):
    """nsga2.py:76-84: crossover -> mutation -> clip."""
    pop = np.asarray(pop, dtype=np.float32)
    lb, ub = np.asarray(lb, np.float32), np.asarray(ub, np.float32)
    x = simulated_binary(pop, pro_c, dis_c, seed, gen)
    x = polynomial(x, lb, ub, pro_m, dis_m, seed, gen)
    return np.clip(x, lb, ub).astype(np.float32)


def dtlz2(X, m):
    """dtlz.py:84-103."""
    X = np.asarray(X, dtype=np.float32)
    n = X.shape[0]
    g = np.sum((X[:, m - 1 :] - np.float32(0.5)) ** 2, axis=1, keepdims=True, dtype=np.float32)
    c = np.maximum(np.cos(X[:, : m - 1] * np.float32(np.pi / 2)), 0)
    f = (np.tile(1 + g, (1, m)) * np.fliplr(np.cumprod(np.c_[np.ones((n, 1), np.float32), c], axis=1))
         * np.c_[np.ones((n, 1), np.float32), np.sin(X[:, m - 2 :: -1] * np.float32(np.pi / 2))])
    return f.astype(np.float32)


def igd(front, pf):
    """metrics/igd.py:8-10: mean over the true front of the distance to the nearest obtained point."""
    d = np.sqrt(((pf[:, None, :] - front[None, :, :]) ** 2).sum(-1))
    return float(d.min(axis=1).mean())
