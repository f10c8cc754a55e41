"""Property tests of the oracle (hypothesis): NumPy restatement vs its C twin on adversarial floats, and
invariants of Pareto ranking that hold for any input."""

import numpy as np
from hypothesis import given, settings, strategies as st
from hypothesis.extra import numpy as hnp

from oracle import c_oracle as c, nds_oracle as o

vals = st.one_of(
    st.floats(width=32, allow_nan=True, allow_infinity=True),
    st.sampled_from([0.0, -0.0, 1.0, 2.0, 3.0, float("inf"), float("-inf"), float("nan")]),
    st.integers(0, 4).map(float),
)
pops = st.tuples(st.integers(1, 40), st.integers(1, 5)).flatmap(
    lambda nm: hnp.arrays(np.float32, nm, elements=vals))


@settings(max_examples=150, deadline=None)
@given(pops)
def test_numpy_and_c_agree(x):
    r_np, r_c = o.non_dominated_sort(x), c.non_dominated_sort(x)
    assert np.array_equal(r_np, r_c)
    mask = (np.arange(x.shape[0]) % 3) != 0
    assert np.array_equal(o.crowding_distance(x, mask), c.crowding_distance(x, mask), equal_nan=True)
    if x.shape[0] > 1:
        pop = x.shape[0] // 2
        assert np.array_equal(o.nsga2_select(x, pop)[0], c.nsga2_select(x, pop)[0])


@settings(max_examples=100, deadline=None)
@given(pops)
def test_rank_invariants(x):
    rank = o.non_dominated_sort(x)
    d = o.dominate_relation(x, x)
    assert np.array_equal(rank == 0, ~d.any(axis=0))          # front 0 = the non-dominated set
    ii, jj = np.nonzero(d)
    assert np.all(rank[jj] > rank[ii])                          # dominated => strictly later front
    worst = np.full(x.shape[0], -1)
    np.maximum.at(worst, jj, rank[ii])
    assert np.array_equal(rank, worst + 1)                      # exactly one after the worst dominator
    assert not d[np.isnan(x).any(axis=1)].any() and not d[:, np.isnan(x).any(axis=1)].any()  # NaN rows are isolated
