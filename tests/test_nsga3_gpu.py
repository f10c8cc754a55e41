"""GPU parity of the NSGA-III selection (nsga3.py:96-211) against the NumPy restatement, through the C ABI.

The path mixes exact integer logic with float32 arithmetic whose association differs between NumPy and the
kernels (dot products, an M x M solve), so the comparison is layered:
  1. ranks, last_rank, ideal point, extreme rows: exact; nadir point: rtol 1e-4;
  2. association: with the CUDA ideal / nadir fed to the oracle, the nearest direction must agree except
     where the oracle's runner-up is within 1e-5 (relative to the distance scale) -- a rounding flip;
     distances rtol 1e-4;
  3. niche loop: with the CUDA ranks / directions / distances fed to the oracle's loop (which draws with the
     restated jax.random.split / randint), the survivors must be IDENTICAL, including the next key.
Populations holding +-inf are left out: there the reference itself runs on float -> int conversions of inf and
out-of-bounds scatters (nsga3.py:166-176 with every candidate distance infinite), which no restatement can pin.
"""

import numpy as np
import pytest

import helpers as h

pytestmark = pytest.mark.gpu
torch = pytest.importorskip("torch")


def _case(kind, n, m, seed):
    if kind == "uniform_nan":  # a few rows with a NaN objective: kept (rank 0) but without a direction
        x = h.uniform(n, m, seed)
        x[::97, seed % m] = np.nan
        return x
    if kind == "lsmop1":
        return h.lsmop1(n, m, seed=seed)
    if kind == "dtlz2":
        return h.dtlz2(n, m, m + 9, seed)
    return getattr(h, kind)(n, m, seed)


CASES = [("dtlz2", 200, 3, 1), ("uniform", 200, 3, 2), ("dtlz2", 4000, 3, 3), ("uniform", 6000, 5, 4), ("lsmop1", 4000, 10, 5),
         ("uniform", 3000, 10, 6), ("integer_grid", 2000, 3, 7), ("dtlz2", 1000, 2, 8), ("uniform_nan", 2001, 4, 9)]


@pytest.mark.parametrize("kind,n,m,seed", CASES, ids=[f"{c[0]}_{c[1]}x{c[2]}" for c in CASES])
def test_nsga3_select_layered(kind, n, m, seed):
    from evox_myc_b200 import nsga3
    from oracle import c_oracle, nsga3_oracle as o3

    x = _case(kind, n, m, seed)
    pop = n // 2
    ref = o3.uniform_sampling(pop, m)
    assert np.array_equal(ref, nsga3.UniformSampling(pop, m)()[0])
    key = (seed * 7919, 12345 + seed)
    idx, next_key, d = nsga3.nsga3_select(x, pop, ref, key, return_details=True)

    rank = c_oracle.non_dominated_sort(x)
    assert np.array_equal(d["rank"], rank)
    last = int(np.sort(rank, kind="stable")[pop])
    ranked, ideal, ext, nadir, full = o3.normalise(x, rank, last)
    ideal_g, nadir_g, ext_g = d["norm"][:m], d["norm"][m:2 * m], d["norm"][2 * m:].astype(np.int64)
    assert np.array_equal(ideal_g, ideal, equal_nan=True)
    assert np.array_equal(ext_g, ext)
    np.testing.assert_allclose(nadir_g, nadir, rtol=1e-4, equal_nan=True)

    # 2. association with the kernel's own normalisation.  d = |x| sqrt(1 - cos^2) cancels when a row lies close to a
    #    direction: an error of a few float32 ulps in cos (dot products are associated differently) becomes
    #    |x|^2 * 1e-6 / d in d, which is the bound used here for distances and for "the runner-up was a tie".
    gid_o, gd_o, gap = o3.associate(ranked, nadir_g, ref)
    gid_g, gd_g = d["group_id"].astype(np.int64), d["group_dist"]
    with np.errstate(all="ignore"):
        xn = np.linalg.norm((ranked / nadir_g).astype(np.float32), axis=1)
    bound = 1e-4 * np.abs(np.nan_to_num(gd_o)) + 1e-6 * np.nan_to_num(xn) ** 2 / np.maximum(np.nan_to_num(gd_o), 1e-6) + 1e-7
    differ = gid_g != gid_o
    # a row (numerically) ON a direction: 1 - cos^2 is rounding noise, sqrt of a negative one is NaN and that
    # direction is skipped (the reference's formula, nsga3.py:136-138) -- which implementation skips is decided by
    # the last bit of cos.  Such rows are told apart in float64 and accepted either way.
    on_a_direction = np.zeros(len(differ), dtype=bool)
    for i in np.nonzero(differ)[0]:
        v = (ranked[i].astype(np.float64) / nadir_g.astype(np.float64))
        for g in (gid_g[i], gid_o[i]):
            if g < ref.shape[0]:
                w = ref[g].astype(np.float64)
                c = v @ w / np.linalg.norm(v) / np.linalg.norm(w)
                on_a_direction[i] |= (1.0 - c * c) < 4e-6
    strict = differ & ~on_a_direction
    assert np.all(gap[strict] <= 2 * bound[strict]), f"{int(strict.sum())} rows go to another direction, not all near-ties"
    same = ~differ & np.isfinite(gd_o)
    assert np.array_equal(np.isnan(gd_g), np.isnan(gd_o))
    assert np.all(np.abs(gd_g[same] - gd_o[same]) <= bound[same])

    # 3. the select loop on the kernel's own ranks / directions / distances
    want_idx, want_key = o3.niche_select(rank, last, gid_g, gd_g, ref.shape[0], pop, key)
    assert np.array_equal(idx, want_idx)
    assert tuple(next_key) == tuple(want_key)
    assert len(np.unique(idx)) == pop and np.all(np.diff(idx) > 0)
    assert np.all(rank[idx] <= last)
    if not np.isnan(x).any():  # rows without a direction are not counted by the loop, which then over-selects and the
        assert np.all(np.isin(np.nonzero(rank < last)[0], idx))  # [:pop_size] cut of nsga3.py:205 drops the last rows


def test_nsga3_config5_full_size():
    """BASELINE config 5: NSGA-III on LSMOP1, M = 10, pop = 32768 (N = 65536, 29315 reference points)."""
    from evox_myc_b200 import nsga3
    from oracle import c_oracle, nsga3_oracle as o3

    n, m, pop = 65536, 10, 32768
    x = h.lsmop1(n, m, seed=61)
    ref = nsga3.UniformSampling(pop, m)()[0]
    assert ref.shape == (29315, 10)
    key = (0, 42)
    idx, next_key, d = nsga3.nsga3_select(x, pop, ref, key, return_details=True)
    rank = c_oracle.non_dominated_sort(x)
    assert np.array_equal(d["rank"], rank)
    last = int(np.sort(rank, kind="stable")[pop])
    want_idx, want_key = o3.niche_select(rank, last, d["group_id"].astype(np.int64), d["group_dist"], ref.shape[0], pop, key)
    assert np.array_equal(idx, want_idx) and tuple(next_key) == tuple(want_key)
    # association spot check on a slice (the full N x R distance matrix is 7.7 GB in float32)
    ranked, ideal, ext, nadir, full = o3.normalise(x, rank, last)
    rows = np.nonzero(rank <= last)[0][:: 37][:600]
    gid_o, gd_o, gap = o3.associate(ranked[rows], d["norm"][m:2 * m], ref)
    differ = d["group_id"][rows] != gid_o
    with np.errstate(all="ignore"):
        xn = np.linalg.norm((ranked[rows] / d["norm"][m:2 * m]).astype(np.float32), axis=1)
    bound = 1e-4 * np.abs(np.nan_to_num(gd_o)) + 1e-6 * np.nan_to_num(xn) ** 2 / np.maximum(np.nan_to_num(gd_o), 1e-6) + 1e-7
    assert np.all(gap[differ] <= 2 * bound[differ])


def test_nsga3_generation_loop_runs_and_improves():
    """A few generations of the NSGA3 mirror on DTLZ2 (the reference's own algorithm test asserts no values)."""
    from evox_myc_b200 import nsga2, nsga3

    d, m, pop = 12, 3, 200
    algo = nsga3.NSGA3(np.zeros(d), np.ones(d), m, pop)
    prob = nsga2.DTLZ2(d=d, m=m)
    state = algo.setup(3)
    x, state = algo.init_ask(state)
    f, _ = prob.evaluate(None, x)
    state = algo.init_tell(state, f)
    g0 = float((state.fitness ** 2).sum(1).sqrt().mean())
    for _ in range(30):
        x, state = algo.ask(state)
        f, _ = prob.evaluate(None, x)
        state = algo.tell(state, f)
    g1 = float((state.fitness ** 2).sum(1).sqrt().mean())
    assert state.population.shape == (pop, d) and g1 < g0  # moves towards the unit sphere (the DTLZ2 front)
