"""NumPy restatement of NSGA-III's environmental selection after the ranking (test infrastructure only).

Follows /root/reference/src/evox/algorithms/mo/nsga3.py:93-211 line by line, with the reference points of
/root/reference/src/evox/operators/sampling/uniform.py:10-60 and ``cos_dist`` of src/evox/utils/common.py:82-87.
Everything is evaluated in float32 like the reference's jnp arrays.  The niche loop (:156-201) draws with
``jax.random.randint(keys[selected_number], (), 0, rho_last[group])``; oracle/jax_prng.py restates that
sampler, so the loop is reproduced exactly rather than statistically.

Parity status: the reference holds NO test that pins any value of this function
(tests/test_multi_objective_algorithms.py only runs it), and JAX cannot be run here: "parity unpinned" --
semantics come from reading the code.  Threefry itself is pinned (Random123 vector, the 128 x 8 crowding
golden input); ``split`` / ``randint`` are restated from the published JAX source.
"""

from __future__ import annotations

import warnings
from itertools import combinations
from math import comb

import numpy as np

from . import jax_prng
from .nds_oracle import non_dominated_sort

F32 = np.float32


def uniform_sampling(n: int, m: int) -> np.ndarray:
    """``UniformSampling(n, m)()[0]`` -- uniform.py:19-60 (Das and Dennis, two layers when h1 < m)."""
    h1 = 1
    while comb(h1 + m, m - 1) <= n:
        h1 += 1
    w = (np.array(list(combinations(range(1, h1 + m), m - 1))) - np.tile(np.arange(m - 1), (comb(h1 + m - 1, m - 1), 1)) - 1)
    w = ((np.c_[w, np.zeros((w.shape[0], 1)) + h1] - np.c_[np.zeros((w.shape[0], 1)), w]) / h1).astype(F32)
    if h1 < m:
        h2 = 0
        while comb(h1 + m - 1, m - 1) + comb(h2 + m, m - 1) <= n:
            h2 += 1
        if h2 > 0:
            w2 = (np.array(list(combinations(range(1, h2 + m), m - 1)))
                  - np.tile(np.arange(m - 1), (comb(h2 + m - 1, m - 1), 1)) - 1)
            w2 = ((np.c_[w2, np.zeros((w2.shape[0], 1)) + h2] - np.c_[np.zeros((w2.shape[0], 1)), w2]) / h2).astype(F32)
            w = np.r_[w, (w2 / F32(2.0) + F32(1.0 / (2.0 * m))).astype(F32)]
    return np.maximum(w, F32(1e-6)).astype(F32)


def normalise(fit: np.ndarray, rank: np.ndarray, last_rank: int):
    """nsga3.py:99-133: NaN-mask, ideal point, extreme points, hyperplane intercepts (or the worst point)."""
    n, m = fit.shape
    ranked = np.where((rank <= last_rank)[:, None], fit.astype(F32), F32(np.nan))
    with np.errstate(all="ignore"), warnings.catch_warnings():
        warnings.simplefilter("ignore", RuntimeWarning)
        ideal = np.nanmin(ranked, axis=0)
        ranked = (ranked - ideal).astype(F32)
        weight = (np.eye(m) + 1e-6).astype(F32)
        extreme_ind = np.zeros(m, dtype=np.int64)
        for i in range(m):
            scal = np.nanmax((ranked / weight[i]).astype(F32), axis=1)  # all-NaN rows stay NaN
            extreme_ind[i] = np.nanargmin(scal)
        extreme = ranked[extreme_ind]
        # jnp.linalg.matrix_rank of a matrix holding inf / NaN yields NaN singular values, `S > tol` is False
        # everywhere and the rank is 0 (NumPy raises instead): the worst-point branch is taken
        full_rank = bool(np.all(np.isfinite(extreme))) and np.linalg.matrix_rank(extreme) == m
        if full_rank:
            plane = np.linalg.solve(extreme, np.ones(m, dtype=F32)).astype(F32)
            nadir = (F32(1) / plane).astype(F32)
        else:
            nadir = np.nanmax(ranked, axis=0)
    return ranked, ideal.astype(F32), extreme_ind, nadir.astype(F32), bool(full_rank)


def associate(ranked: np.ndarray, nadir: np.ndarray, ref: np.ndarray, chunk: int = 2048):
    """nsga3.py:135-143: perpendicular distance to every reference direction, nearest one (NaN skipped).
    Returns (group_id with len(ref) for rows without a direction, group_dist with NaN there, and the gap to
    the runner-up direction -- used by the tests to tell a rounding flip from an error)."""
    n, m = ranked.shape
    r = ref.shape[0]
    with np.errstate(all="ignore"):
        norm_fit = (ranked / nadir).astype(F32)
        xn = np.linalg.norm(norm_fit, axis=-1, keepdims=True).astype(F32)
        xu = (norm_fit / xn).astype(F32)
        yu = (ref / np.linalg.norm(ref, axis=-1, keepdims=True).astype(F32)).astype(F32)
        gid = np.full(n, r, dtype=np.int64)
        gdist = np.full(n, np.nan, dtype=F32)
        gap = np.full(n, np.inf, dtype=F32)
        for a in range(0, n, chunk):
            cosd = (xu[a:a + chunk] @ yu.T).astype(F32)
            d = (xn[a:a + chunk] * np.sqrt((F32(1) - cosd * cosd).astype(F32))).astype(F32)
            allnan = np.all(np.isnan(d), axis=1)
            dd = np.where(np.isnan(d), np.inf, d)
            best = np.argmin(dd, axis=1)
            bd = dd[np.arange(dd.shape[0]), best]
            dd[np.arange(dd.shape[0]), best] = np.inf
            second = dd.min(axis=1)
            gid[a:a + chunk] = np.where(allnan, r, best)
            gdist[a:a + chunk] = np.where(allnan, np.nan, bd)
            gap[a:a + chunk] = np.where(allnan, np.inf, second - bd)
    return gid, gdist, gap


def niche_select(rank: np.ndarray, last_rank: int, gid: np.ndarray, gdist: np.ndarray, nref: int, pop_size: int, key):
    """nsga3.py:144-211: niche counts, the select loop (one survivor per iteration), survivors in index order."""
    n = rank.shape[0]
    rank = rank.astype(np.int64).copy()
    rho = np.bincount(np.where(rank < last_rank, gid, nref), minlength=nref + 1)[:nref].astype(np.float64)
    rho_last = np.bincount(np.where(rank == last_rank, gid, nref), minlength=nref + 1)[:nref].astype(np.int64)
    group_id = np.where(rank == last_rank, gid.astype(np.float64), np.inf)
    group_dist = np.where(rank == last_rank, gdist.astype(np.float64), np.inf)
    selected = int(rho.sum())
    rho = np.where(rho_last == 0, np.inf, rho)
    keys = jax_prng.split(key, pop_size + 1)
    while selected < pop_size:
        group = int(np.argmin(rho))
        cand = np.where(group_id == group, group_dist, np.inf)
        if rho[group] == 0 or rho_last[group] == 1:
            c = int(np.argmin(cand))
        else:
            order = np.sort(np.where(np.isinf(cand), np.inf, np.arange(n)))
            c = int(order[jax_prng.randint_scalar(keys[selected], 0, int(rho_last[group]))])
        rank[c] = last_rank - 1
        group_id[c] = np.nan
        rho_last[group] -= 1
        rho[group] = np.inf if rho_last[group] == 0 else rho[group] + 1
        selected += 1
    idx = np.nonzero(rank < last_rank)[0][:pop_size].astype(np.int32)
    return idx, keys[0]


def nsga3_select(fit: np.ndarray, pop_size: int, ref: np.ndarray, key, rank: np.ndarray | None = None):
    """The whole of nsga3.py:96-211 on merged fitness: returns (selected_idx, next_key, details)."""
    fit = np.asarray(fit, dtype=F32)
    if rank is None:
        rank = non_dominated_sort(fit)
    last_rank = int(np.sort(rank, kind="stable")[pop_size])
    ranked, ideal, extreme_ind, nadir, full_rank = normalise(fit, rank, last_rank)
    gid, gdist, gap = associate(ranked, nadir, ref)
    idx, next_key = niche_select(rank, last_rank, gid, gdist, ref.shape[0], pop_size, key)
    details = dict(rank=rank, last_rank=last_rank, ideal=ideal, extreme_ind=extreme_ind, nadir=nadir, full_rank=full_rank,
                   group_id=gid, group_dist=gdist, gap=gap)
    return idx, next_key, details
