"""Randomised differential test: many small populations of random shape and tie structure, GPU vs oracle.

Small sizes hit the paths large benchmarks never do: populations smaller than a tile, a single chunk,
runs of identical rows spanning slabs, NaN / inf rows, all-equal objectives, pop_size at both ends."""

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
torch = pytest.importorskip("torch")


def _case(rng):
    n = int(rng.choice([1, 2, 3, 31, 32, 33, 127, 129, 500, 1023, 1025, 2047, 2500, 4095, 4097, 6000]))
    m = int(rng.integers(1, 9))
    kind = rng.integers(0, 5)
    if kind == 0:
        x = rng.random((n, m), dtype=np.float32)
    elif kind == 1:
        x = rng.integers(0, int(rng.choice([2, 3, 8, 50])), size=(n, m)).astype(np.float32)  # heavy ties
    elif kind == 2:
        x = np.repeat(rng.random((max(1, n // 7), m), dtype=np.float32), 7, axis=0)[:n]       # runs of duplicates
        if x.shape[0] < n:
            x = np.concatenate([x, rng.random((n - x.shape[0], m), dtype=np.float32)])
        x = x[rng.permutation(n)]
    elif kind == 3:
        x = rng.normal(size=(n, m)).astype(np.float32)
        x[rng.random((n, m)) < 0.02] = np.nan
        x[rng.random((n, m)) < 0.02] = np.inf
        x[rng.random((n, m)) < 0.02] = -np.inf
        x[rng.random((n, m)) < 0.02] = -0.0
    else:
        x = np.full((n, m), 3.0, np.float32)                                                 # one big tie
        x[:, 0] = rng.integers(0, 3, size=n)
    return np.ascontiguousarray(x, dtype=np.float32)


@pytest.mark.parametrize("seed", range(6))
def test_many_small_random_cases(seed):
    import evox_myc_b200 as e
    from oracle import c_oracle

    rng = np.random.default_rng(1000 + seed)
    for _ in range(25):
        x = _case(rng)
        n = x.shape[0]
        want = c_oracle.non_dominated_sort(x)
        got = e.non_dominated_sort(x)
        assert np.array_equal(got, want), (x.shape, int(np.sum(got != want)))
        mask = rng.random(n) < rng.random()
        wc = c_oracle.crowding_distance(x, mask)
        gc = e.crowding_distance(x, mask)
        assert np.array_equal(np.isnan(gc), np.isnan(wc)) and np.array_equal(np.isinf(gc), np.isinf(wc))
        fin = np.isfinite(wc)
        np.testing.assert_allclose(gc[fin], wc[fin], rtol=1e-6, atol=0)
        assert np.array_equal(np.sign(gc[~fin & ~np.isnan(wc)]), np.sign(wc[~fin & ~np.isnan(wc)]))
        if n >= 2:
            pop = int(rng.integers(0, n))
            wo, _, _ = c_oracle.nsga2_select(x, pop)
            go = e.nsga2_select(x, pop)
            assert np.array_equal(go, wo), (x.shape, pop)
