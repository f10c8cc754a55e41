"""GPU parity of the level form of the dominance build (csrc/levels.cuh, dominance_build_lvl_kernel): ranks must be
bit-exact with the oracle whichever way the matrix is built -- float compares, 15-bit levels + fix-up, or levels
abandoned on the device for the float kernel -- and the path each data set takes is checked too."""


import numpy as np
import pytest

import helpers as h

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")


def _oracle_rank(x):
    from oracle import c_oracle, nds_oracle

    return nds_oracle.non_dominated_sort(x) if x.shape[0] <= 1500 else c_oracle.non_dominated_sort(x)


@pytest.fixture
def build_mode(monkeypatch):
    def set_mode(mode):
        if mode is None:
            monkeypatch.delenv("EVOXB_BUILD_MODE", raising=False)
        else:
            monkeypatch.setenv("EVOXB_BUILD_MODE", mode)

    return set_mode


def _rank_and_path(x):
    import evox_myc_b200 as e
    from evox_myc_b200 import operators

    r = e.non_dominated_sort(x)
    return r, operators.nds_build_path(x.shape[0], x.shape[1])


def _special_values(n, m, seed):
    """Every coordinate drawn from a handful of special floats: heavy ties, both zeros, denormals, infinities, NaN."""
    rng = np.random.default_rng(seed)
    vals = np.array([-np.inf, -1.0, -1e-45, -0.0, 0.0, 1e-45, 1.1754944e-38, 0.5, 1.0, 3.4028235e38, np.inf, np.nan],
                    dtype=np.float32)
    p = np.array([1, 3, 2, 4, 4, 2, 2, 4, 3, 1, 1, 0.2])
    return vals[rng.choice(len(vals), size=(n, m), p=p / p.sum())]


def _narrow(n, m, seed):
    """Values packed into [1, 1 + 2^-10): 8192 distinct floats per objective inside ONE first-level bin of 2^20 keys,
    i.e. about 37 distinct values per level.  Few rows: fix-up still affordable; many rows: the device falls back."""
    rng = np.random.default_rng(seed)
    return (1.0 + rng.random((n, m)) * 2.0**-10).astype(np.float32)


def _crowded(n, m, seed):
    """Half the rows uniform in [0, 1), the other half on 1000 adjacent floats just above 1: those share about 14
    levels, far too crowded for the fix-up, so the device falls back to the float kernel."""
    rng = np.random.default_rng(seed)
    x = rng.random((n, m)).astype(np.float32)
    x[: n // 2] = (1.0 + rng.integers(0, 1000, size=(n // 2, m)) * 2.0**-23).astype(np.float32)
    return x


def _skewed(n, m, seed):
    rng = np.random.default_rng(seed)
    return np.exp(rng.normal(0.0, 6.0, size=(n, m))).astype(np.float32) * np.where(rng.random((n, m)) < 0.3, -1, 1).astype(np.float32)


LEVEL_CASES = [
    ("n1_m4", lambda: h.uniform(1, 4, 1)),
    ("n33_m2", lambda: h.uniform(33, 2, 2)),
    ("n129_m3", lambda: h.nasty(129, 3, 3)),
    ("n1000_m4", lambda: h.nasty(1000, 4, 4)),
    ("n4097_m5", lambda: h.nasty(4097, 5, 5)),
    ("n9000_m6", lambda: h.nasty(9000, 6, 6)),
    ("n5000_m7", lambda: h.uniform(5000, 7, 7)),
    ("n5000_m10", lambda: h.nasty(5000, 10, 8)),
    ("n3000_m16", lambda: h.nasty(3000, 16, 9)),
    ("n20000_m5", lambda: h.nasty(20000, 5, 10)),
    ("special_m4", lambda: _special_values(6000, 4, 11)),
    ("special_m5", lambda: _special_values(9000, 5, 12)),
    ("grid_m5", lambda: h.integer_grid(8000, 5, 13, levels=8)),
    ("skewed_m5", lambda: _skewed(12000, 5, 14)),
    ("narrow_m4", lambda: _narrow(12000, 4, 15)),
    ("narrow_few_m4", lambda: _narrow(3000, 4, 18)),
    ("crowded_m5", lambda: _crowded(10000, 5, 19)),
    ("chain_m4", lambda: h.chain(3000, 4)),
    ("dtlz2_m5", lambda: h.dtlz2(8192, 5, 14, 16)),
    ("lsmop1_m10", lambda: h.lsmop1(6000, 10, seed=17)),
]


@pytest.mark.parametrize("name,make", LEVEL_CASES, ids=[c[0] for c in LEVEL_CASES])
def test_level_build_matches_oracle(name, make, build_mode):
    x = make()
    build_mode("levels")
    r, path = _rank_and_path(x)
    assert path in ("levels", "levels->float")
    want = _oracle_rank(x)
    assert np.array_equal(r, want), f"{name}: {int((r != want).sum())} ranks differ on path {path}"
    build_mode("exact")
    r2, path2 = _rank_and_path(x)
    assert path2 == "float" and np.array_equal(r2, want)


def test_paths_taken(build_mode):
    """Uniform data stays on the levels; integer-valued (pure levels) too; values crowded far below the resolution of
    32768 levels make the device fall back -- and the ranks are right in every case."""
    build_mode("levels")
    for x, want_path in [
        (h.uniform(30000, 5, 21), "levels"),
        (h.integer_grid(30000, 4, 22, levels=50), "levels"),
        (_narrow(30000, 4, 23), "levels->float"),
        (_crowded(40000, 4, 24), "levels->float"),
    ]:
        r, path = _rank_and_path(x)
        assert path == want_path
        assert np.array_equal(r, _oracle_rank(x))


def test_default_choice_by_size(build_mode):
    build_mode(None)
    x = h.uniform(65536, 4, 31)
    r, path = _rank_and_path(x)
    assert path == "levels" and np.array_equal(r, _oracle_rank(x))
    x = h.uniform(65536, 3, 32)
    assert _rank_and_path(x)[1] == "float"
    x = h.uniform(32768, 5, 33)
    assert _rank_and_path(x)[1] == "float"


def test_full_size_both_paths_agree(build_mode):
    """BASELINE configs[3] (262144 x 5): the level build and the float build give the same ranks (the float build is
    checked against the oracle at this size in test_nds_gpu.py)."""
    x = h.uniform(262144, 5, 41)
    build_mode("levels")
    r1, p1 = _rank_and_path(x)
    build_mode("exact")
    r2, p2 = _rank_and_path(x)
    assert (p1, p2) == ("levels", "float")
    assert np.array_equal(r1, r2)


def test_triangular_storage_with_levels(build_mode):
    """Above 262144 rows the matrix is stored as a triangle (TriRows): the fix-up addresses it through vstart()."""
    x = h.uniform(270000, 4, 51)
    build_mode("levels")
    r1, p1 = _rank_and_path(x)
    build_mode("exact")
    r2, _ = _rank_and_path(x)
    assert p1 == "levels" and np.array_equal(r1, r2)
