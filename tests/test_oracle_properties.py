"""NumPy restatement vs its C twin, plus properties of Pareto ranking that need no reference."""

import numpy as np
import pytest

import helpers as h
from oracle import c_oracle as c, nds_oracle as o

CASES = [
    ("uniform", lambda: h.uniform(1500, 3, 1)),
    ("uniform_m1", lambda: h.uniform(257, 1, 2)),
    ("uniform_m2", lambda: h.uniform(1200, 2, 3)),
    ("uniform_m10", lambda: h.uniform(700, 10, 4)),
    ("grid", lambda: h.integer_grid(1000, 3, 5, levels=6)),
    ("nasty", lambda: h.nasty(1111, 4, 6)),
    ("dtlz2", lambda: h.dtlz2(900, 3, 12, 7)),
    ("chain", lambda: h.chain(70, 2)),
    ("one", lambda: h.uniform(1, 3, 8)),
    ("two_equal", lambda: np.ones((2, 3), np.float32)),
]


@pytest.mark.parametrize("name,gen", CASES)
def test_c_twin_matches_numpy(name, gen):
    x = gen()
    r_np = o.non_dominated_sort(x)
    r_c = c.non_dominated_sort(x)
    assert np.array_equal(r_np, r_c)
    rng = np.random.default_rng(11)
    for mask in (None, rng.random(x.shape[0]) < 0.5, r_np == r_np.max(), np.zeros(x.shape[0], bool)):
        d_np = o.crowding_distance(x, mask)
        d_c = c.crowding_distance(x, mask)
        assert np.array_equal(d_np, d_c, equal_nan=True), name
    if x.shape[0] > 2:
        pop = x.shape[0] // 2
        a, _, _ = o.nsga2_select(x, pop)
        b, _, _ = c.nsga2_select(x, pop)
        assert np.array_equal(a, b)
    assert np.array_equal(o.dominate_relation(x[:50], x[:70]), c.dominate_relation(x[:50], x[:70]))


@pytest.mark.parametrize("name,gen", CASES[:7])
def test_rank_properties(name, gen):
    x = gen()
    rank = o.non_dominated_sort(x)
    d = o.dominate_relation(x, x)
    # front 0 is the non-dominated set
    assert np.array_equal(rank == 0, ~d.any(axis=0))
    # a dominated point ranks strictly after each of its dominators, and exactly one after the worst of them
    ii, jj = np.nonzero(d)
    assert np.all(rank[jj] > rank[ii])
    worst = np.full(x.shape[0], -1)
    np.maximum.at(worst, jj, rank[ii])
    assert np.array_equal(rank, worst + 1)
    # permutation invariance
    p = np.random.default_rng(3).permutation(x.shape[0])
    assert np.array_equal(o.non_dominated_sort(x[p]), rank[p])


def test_method_argument_is_validated():
    x = h.uniform(10, 2)
    for mth in ("auto", "scan", "full map-reduce", "host"):
        o.non_dominated_sort(x, mth)
    with pytest.raises(AssertionError):
        o.non_dominated_sort(x, "bogus")


def test_crowding_edge_cases():
    x = h.uniform(9, 2, 5)
    # num_valid in {0, 1, 2}
    assert np.all(np.isneginf(o.crowding_distance(x, np.zeros(9, bool))))
    m1 = np.zeros(9, bool); m1[4] = True
    d = o.crowding_distance(x, m1)
    assert np.isposinf(d[4]) and np.all(np.isneginf(np.delete(d, 4)))
    m2 = m1.copy(); m2[7] = True
    d = o.crowding_distance(x, m2)
    assert np.isposinf(d[4]) and np.isposinf(d[7])
    # range == 0 -> interior is 0/0 = NaN
    y = np.ones((5, 2), np.float32)
    d = o.crowding_distance(y)
    assert np.isposinf(d[0]) and np.isposinf(d[4]) and np.all(np.isnan(d[1:4]))
