"""GPU parity: NSGA-II survivor selection (nsga2.py:90-98) and the operator form non_dominate."""

import numpy as np
import pytest

import helpers as h

pytestmark = pytest.mark.gpu
torch = pytest.importorskip("torch")

CASES = [
    ("uniform_m3", lambda: h.uniform(4000, 3, 1), 2000),
    ("uniform_m2", lambda: h.uniform(3001, 2, 2), 1500),
    ("uniform_m5", lambda: h.uniform(6000, 5, 3), 3000),
    ("grid_ties", lambda: h.integer_grid(4000, 3, 4, levels=6), 2000),
    ("nasty", lambda: h.nasty(5000, 4, 5), 2500),
    ("dtlz2", lambda: h.dtlz2(8192, 3, 12, 6), 4096),
    ("dtlz2_converged", lambda: h.dtlz2(4096, 3, 12, 7, converged=True), 2048),
    ("m18_full_matrix_mode", lambda: h.uniform(3000, 18, 12), 1500),
    ("pop_1", lambda: h.uniform(100, 3, 8), 1),
    ("pop_n_minus_1", lambda: h.uniform(100, 3, 9), 99),
    ("config1_pop100", lambda: h.dtlz2(200, 3, 12, 10), 100),
    ("config3_pop16384", lambda: h.dtlz2(32768, 3, 12, 11), 16384),
]


@pytest.mark.parametrize("name,gen,pop", CASES, ids=[c[0] for c in CASES])
def test_survivor_set_identical(name, gen, pop):
    import evox_myc_b200 as e
    from oracle import c_oracle

    x = gen()
    want_order, want_rank, want_crowd = c_oracle.nsga2_select(x, pop)
    order, rank, crowd = e.nsga2_select(x, pop, return_details=True)
    assert np.array_equal(rank, want_rank)
    fin = np.isfinite(want_crowd)
    assert np.array_equal(np.isfinite(crowd), fin)
    np.testing.assert_allclose(crowd[fin], want_crowd[fin], rtol=1e-6, atol=0)
    # identical survivors in identical order (ties broken by the stable sort)
    assert np.array_equal(order, want_order), f"{np.sum(order != want_order)} of {pop} survivor slots differ"


def test_small_against_numpy_restatement():
    import evox_myc_b200 as e
    from oracle import nds_oracle

    for seed in range(4):
        x = h.nasty(300, 3, seed)
        want, _, _ = nds_oracle.nsga2_select(x, 150)
        assert np.array_equal(e.nsga2_select(x, 150), want)


def test_pop_size_out_of_range_is_an_error():
    import evox_myc_b200 as e

    x = h.uniform(10, 2, 1)
    with pytest.raises(Exception):
        e.nsga2_select(x, 10)  # rank[order[10]] does not exist


def test_non_dominate_operator():
    import evox_myc_b200 as e
    from oracle import nds_oracle

    for seed, (n, m, k) in enumerate([(500, 3, 250), (1000, 2, 300), (64, 4, 64), (300, 3, 1)]):
        fit = h.uniform(n, m, seed)
        pop = np.random.default_rng(seed).random((n, 7), dtype=np.float32)
        wp, wf = nds_oracle.non_dominate(pop, fit, k)
        gp, gf = e.non_dominate(pop, fit, k)
        assert np.array_equal(gp, wp) and np.array_equal(gf, wf)
        gp2, gf2 = e.NonDominate(k)(torch.from_numpy(pop).cuda(), torch.from_numpy(fit).cuda())
        assert gp2.is_cuda and np.array_equal(gp2.cpu().numpy(), wp) and np.array_equal(gf2.cpu().numpy(), wf)


def test_nsga3_call_site():
    """nsga3.py:96-98: rank and last_rank = rank[argsort(rank)[pop_size]]."""
    import evox_myc_b200 as e
    from oracle import nds_oracle

    for seed, (n, m, pop) in enumerate([(2000, 10, 1000), (3000, 3, 1500), (500, 5, 499), (64, 2, 0)]):
        x = h.uniform(n, m, seed)
        want_rank, want_last = nds_oracle.nsga3_last_rank(x, pop)
        rank, last = e.nsga3_last_rank(x, pop)
        assert np.array_equal(rank, want_rank) and last == int(want_last)
        r2, l2 = e.nsga3_last_rank(torch.from_numpy(x).cuda(), pop)
        assert np.array_equal(r2.cpu().numpy(), want_rank) and int(l2) == int(want_last)
