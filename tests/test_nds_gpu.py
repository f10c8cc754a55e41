"""GPU parity: non_dominated_sort through the C ABI vs the oracle -- ranks must be bit-exact."""

import ctypes

import numpy as np
import pytest

import helpers as h

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")


def _oracle_rank(x):
    from oracle import c_oracle, nds_oracle

    return nds_oracle.non_dominated_sort(x) if x.shape[0] <= 1500 else c_oracle.non_dominated_sort(x)


def _gpu_rank(x):
    import evox_myc_b200 as e

    return e.non_dominated_sort(x)


def test_reference_kat(kats):
    k = kats["nds_11x2"]
    r = _gpu_rank(np.array(k["x"], dtype=np.float32))
    assert r.dtype == np.int32 and r.tolist() == k["rank"]


def test_reference_monitor_kat_int32(kats):
    k = kats["monitor_pf_int32"]
    f1 = np.array(k["fitness1"], dtype=np.int32)  # int32 fitness is cast to float32 by the wrapper
    pf1 = f1[_gpu_rank(f1) == 0]
    assert pf1.tolist() == k["pf1"]
    f2 = np.concatenate([pf1.astype(np.float32), np.array(k["fitness2"], dtype=np.float32)])
    assert f2[_gpu_rank(f2) == 0].tolist() == k["pf2"]


CASES = [
    ("one", lambda: h.uniform(1, 3, 1)),
    ("two_equal", lambda: np.ones((2, 3), np.float32)),
    ("n31", lambda: h.uniform(31, 2, 2)),
    ("n33", lambda: h.uniform(33, 5, 3)),
    ("n127_m1", lambda: h.uniform(127, 1, 4)),
    ("n129", lambda: h.nasty(129, 3, 5)),
    ("n1000_m2", lambda: h.uniform(1000, 2, 6)),
    ("n1025_m3", lambda: h.uniform(1025, 3, 7)),
    ("n4097_m5", lambda: h.uniform(4097, 5, 8)),
    ("n5000_m10", lambda: h.uniform(5000, 10, 9)),
    ("n3000_m13", lambda: h.uniform(3000, 13, 10)),
    ("n3000_m16", lambda: h.nasty(3000, 16, 18)),
    ("n2500_m20_full_matrix_mode", lambda: h.nasty(2500, 20, 19)),   # m > 16: unsorted full-matrix kernels
    ("n5000_m32_full_matrix_mode", lambda: h.uniform(5000, 32, 20)),
    ("grid", lambda: h.integer_grid(6000, 3, 11, levels=8)),
    ("grid_m2", lambda: h.integer_grid(5000, 2, 12, levels=64)),
    ("nasty", lambda: h.nasty(7001, 4, 13)),
    ("dtlz2", lambda: h.dtlz2(8192, 3, 12, 14)),
    ("dtlz2_converged", lambda: h.dtlz2(8192, 3, 12, 15, converged=True)),
    ("chain", lambda: h.chain(700, 2)),
    ("all_nan", lambda: np.full((100, 3), np.nan, np.float32)),
    ("all_equal", lambda: np.zeros((5000, 3), np.float32)),
    ("uniform_16k_m3", lambda: h.uniform(16384, 3, 16)),
    ("uniform_20k_m2_deep", lambda: h.uniform(20000, 2, 17)),
]


@pytest.mark.parametrize("name,gen", CASES, ids=[c[0] for c in CASES])
def test_rank_bit_exact(name, gen):
    x = gen()
    want = _oracle_rank(x)
    got = _gpu_rank(x)
    assert got.dtype == np.int32
    assert np.array_equal(got, want), f"{name}: {np.sum(got != want)} of {x.shape[0]} ranks differ"


def test_device_tensor_in_device_tensor_out():
    import evox_myc_b200 as e

    x = h.uniform(3000, 3, 21)
    r = e.non_dominated_sort(torch.from_numpy(x).cuda(), method="full map-reduce")
    assert r.is_cuda and r.dtype == torch.int32
    assert np.array_equal(r.cpu().numpy(), _oracle_rank(x))
    with pytest.raises(AssertionError):
        e.non_dominated_sort(torch.from_numpy(x).cuda(), method="bogus")


def test_empty_population():
    import evox_myc_b200 as e

    r = e.non_dominated_sort(torch.empty((0, 3), device="cuda"))
    assert r.shape == (0,)


def test_repeated_calls_share_workspace():
    """The workspace keeps the previous call's matrix; a second call must not see stale bits."""
    a = h.uniform(5000, 3, 31)
    b = h.integer_grid(3000, 3, 32, levels=4)
    for x in (a, b, a, b):
        assert np.array_equal(_gpu_rank(x), _oracle_rank(x))


def test_dominate_relation_dense():
    import evox_myc_b200 as e
    from oracle import nds_oracle

    x = h.nasty(300, 3, 41)
    y = h.nasty(500, 3, 42)
    got = e.dominate_relation(x, y)
    assert got.dtype == bool and np.array_equal(got, nds_oracle.dominate_relation(x, y))


def test_workspace_too_small_is_an_error():
    from evox_myc_b200 import _lib

    L = _lib.lib()
    x = torch.rand(1000, 3, device="cuda")
    rank = torch.empty(1000, dtype=torch.int32, device="cuda")
    ws = torch.empty(1024, dtype=torch.uint8, device="cuda")
    rc = L.evoxb_nds_rank_f32(x.data_ptr(), 1000, 3, rank.data_ptr(), ws.data_ptr(), ws.numel(), None)
    assert rc == -2 and b"workspace too small" in L.evoxb_last_error()


@pytest.mark.parametrize("n,m", [(65536, 3)])
def test_config2_full_size_properties(n, m):
    """BASELINE config 2 at full size: oracle comparison (C twin, a few seconds) + size-independent properties."""
    import evox_myc_b200 as e

    x = h.uniform(n, m, 51)
    got = _gpu_rank(x)
    assert np.array_equal(got, _oracle_rank(x))
    # permutation invariance at full size
    p = np.random.default_rng(5).permutation(n)
    assert np.array_equal(_gpu_rank(x[p]), got[p])


def test_config4_full_size():
    """BASELINE config 4 at full size (N = 262144, M = 5): ranks and survivor order against the C oracle
    (about 15 s of host time), plus properties that need no oracle."""
    import evox_myc_b200 as e
    from oracle import c_oracle

    n, m = 262144, 5
    x = h.uniform(n, m, 52)
    order, rank, crowd = e.nsga2_select(x, n // 2, return_details=True)
    w_order, w_rank, w_crowd = c_oracle.nsga2_select(x, n // 2)
    assert np.array_equal(rank, w_rank)
    assert np.array_equal(order, w_order)
    fin = np.isfinite(w_crowd)
    assert np.array_equal(np.isfinite(crowd), fin)
    np.testing.assert_allclose(crowd[fin], w_crowd[fin], rtol=1e-6, atol=0)
    # size-independent properties: every rank class is an antichain's worth of work -- check the cheap ones
    assert rank.min() == 0 and len(np.unique(rank)) == rank.max() + 1        # no empty front
    assert len(np.unique(order)) == n // 2                                    # survivors are distinct rows
    assert rank[order].max() <= np.sort(rank)[n // 2]                         # nobody beyond the cut-off rank
    p = np.random.default_rng(9).permutation(n)
    assert np.array_equal(e.non_dominated_sort(x[p]), rank[p])                # permutation invariance


def test_odd_large_size_many_chunks():
    """N not a multiple of anything, 74 column chunks (exercises the 4-entries-per-lane prefix variant)."""
    import evox_myc_b200 as e
    from oracle import c_oracle

    x = h.nasty(150001, 4, 53)
    assert np.array_equal(e.non_dominated_sort(x), c_oracle.non_dominated_sort(x))


def test_config5_full_size_lsmop1():
    """BASELINE config 5 at its own size: N = 65536 (pop 32768), M = 10, LSMOP1 objectives with d = 1000
    (tests/helpers.lsmop1, restating lsmop.py:127-141).  Ranks bit-exact and the NSGA-III cut-off rank
    (nsga3.py:96-98) against the C oracle."""
    import evox_myc_b200 as e
    from oracle import c_oracle

    n, m = 65536, 10
    x = h.lsmop1(n, m, seed=54)
    want = c_oracle.non_dominated_sort(x)
    rank, last = e.nsga3_last_rank(x, n // 2)
    assert np.array_equal(rank, want)
    assert int(last) == int(np.sort(want, kind="stable")[n // 2])
    p = np.random.default_rng(10).permutation(n)
    assert np.array_equal(e.non_dominated_sort(x[p]), want[p])


def test_config5_shape_uniform_m10():
    """Same shape on uniform-random fitness (a single huge front plus a thin tail at M = 10)."""
    from oracle import c_oracle

    x = h.uniform(65536, 10, 55)
    assert np.array_equal(_gpu_rank(x), c_oracle.non_dominated_sort(x))


def test_deep_front_stress_m2_full_size():
    """SURVEY.md 8(d) deep-front stress: uniform M = 2 at N = 65536 (about 500 fronts): ranks, survivor
    order and crowding against the C oracle."""
    import evox_myc_b200 as e
    from oracle import c_oracle

    n = 65536
    x = h.uniform(n, 2, 56)
    order, rank, crowd = e.nsga2_select(x, n // 2, return_details=True)
    w_order, w_rank, w_crowd = c_oracle.nsga2_select(x, n // 2)
    assert int(w_rank.max()) > 300
    assert np.array_equal(rank, w_rank)
    assert np.array_equal(order, w_order)
    fin = np.isfinite(w_crowd)
    assert np.array_equal(np.isfinite(crowd), fin)
    np.testing.assert_allclose(crowd[fin], w_crowd[fin], rtol=1e-6, atol=0)


def test_triangular_storage_beyond_262144_rows():
    """Above 262144 rows only the triangle of the bit matrix is stored (row i keeps the chunks from its own on):
    same ranks, checked at 270000 x 6 against the C oracle (the sharded layout is covered in test_sharded_gpu)."""
    import evox_myc_b200 as e
    from oracle import c_oracle

    x = h.uniform(270000, 6, 57)
    assert np.array_equal(e.non_dominated_sort(x), c_oracle.non_dominated_sort(x))
