"""CPU checks of the level form of the dominance build (csrc/levels.cuh, dominance_build_lvl_kernel), as executable
specification: a NumPy restatement of the level map, of the guard-bit compare and of the fix-up, run against the
oracle's dominance relation.  The CUDA kernels are checked bit for bit on the GPU (tests/test_levels_gpu.py); what is
pinned here is the arithmetic they rely on:

  * level(x) is monotone in the float order (-0 == +0), so "all levels <=" is a superset of "all floats <=";
  * (column word | guard bits) - row word keeps both guard bits iff both 15-bit halves compare <=, whatever the
    borrows do;
  * clearing the pairs that share a level in some objective but compare the other way as floats leaves exactly the
    reference's relation (utils/common.py:95-104 restricted to the presorted predicate).
"""

import ctypes

import numpy as np
import pytest

import helpers as h

BINS, LEVELS = 4096, 32768
SPARE = LEVELS - BINS
GUARD = np.uint32(0x80008000)


def float_to_key(x):
    """Order-preserving u32 image of a float32 (csrc/sort.cuh: float_to_key): NaN last, -0 == +0."""
    u = np.ascontiguousarray(x, dtype=np.float32).view(np.uint32).copy()
    u[u == 0x80000000] = 0
    neg = (u & 0x80000000) != 0
    key = np.where(neg, ~u, u | np.uint32(0x80000000)).astype(np.uint32)
    key[np.isnan(x)] = 0xFFFFFFFF
    return key


def levels_of(x):
    """levels.cuh: level_hist_kernel + level_plan_kernel + level_of, one objective.  NaN -> level 0 (never used)."""
    key = float_to_key(x)
    ok = key != 0xFFFFFFFF
    hist = np.bincount((key[ok] >> 20).astype(np.int64), minlength=BINS)
    total = int(hist.sum())
    nsub = np.where(hist > 0, 1 + (hist.astype(np.uint64) * SPARE) // max(total, 1), 0).astype(np.uint64)
    base = np.concatenate([[0], np.cumsum(nsub)[:-1]]).astype(np.uint64)
    assert int(nsub.sum()) <= LEVELS
    b = (key >> 20).astype(np.int64)
    lv = base[b] + (((key & 0xFFFFF).astype(np.uint64) << 12) * nsub[b] >> 32)
    return np.where(ok, lv, 0).astype(np.uint32), key


DATA = [
    ("uniform", lambda: h.uniform(3000, 1, 1)[:, 0]),
    ("nasty", lambda: h.nasty(3000, 3, 2)[:, 1]),
    ("grid", lambda: h.integer_grid(3000, 1, 3, levels=50)[:, 0]),
    ("lognormal_signed", lambda: (np.exp(np.random.default_rng(4).normal(0, 8, 3000)) *
                                  np.random.default_rng(5).choice([-1.0, 1.0], 3000)).astype(np.float32)),
    ("narrow", lambda: (1.0 + np.random.default_rng(6).random(3000) * 2.0 ** -12).astype(np.float32)),
    ("specials", lambda: np.random.default_rng(7).choice(
        np.array([-np.inf, -1.0, -1e-45, -0.0, 0.0, 1e-45, 1.1754944e-38, 1.0, 3.4028235e38, np.inf], dtype=np.float32), 3000)),
]


@pytest.mark.parametrize("name,make", DATA, ids=[d[0] for d in DATA])
def test_level_map_is_monotone(name, make):
    x = make()
    lv, key = levels_of(x)
    ok = ~np.isnan(x)
    assert lv.max() < LEVELS
    order = np.argsort(key[ok], kind="stable")
    assert np.all(np.diff(lv[ok][order].astype(np.int64)) >= 0)          # monotone in the key order ...
    i, j = np.random.default_rng(0).integers(0, ok.sum(), size=(2, 20000))
    xs, ls = x[ok], lv[ok]
    le = xs[i] <= xs[j]
    assert np.all(ls[i][le] <= ls[j][le])                                 # ... hence x_i <= x_j => level_i <= level_j
    eq = xs[i] == xs[j]                                                   # (-0 == +0 included)
    assert np.all(ls[i][eq] == ls[j][eq])


def test_guard_bit_compare_equals_the_two_field_compares():
    rng = np.random.default_rng(1)
    a = rng.integers(0, LEVELS, size=(200000, 2), dtype=np.uint32)
    b = rng.integers(0, LEVELS, size=(200000, 2), dtype=np.uint32)
    b[:50000] = a[:50000]                                                 # plenty of equal halves
    b[50000:100000, 0] = a[50000:100000, 0]
    row = a[:, 0] | (a[:, 1] << 16)
    col = (b[:, 0] | (b[:, 1] << 16)) | GUARD
    d = (col - row).astype(np.uint32)                                     # wraps like the IMAD
    got = (d & GUARD) == GUARD
    want = (a[:, 0] <= b[:, 0]) & (a[:, 1] <= b[:, 1])
    assert np.array_equal(got, want)
    # the AND over several words, as the kernel folds them before the test
    d2 = np.roll(d, 1)
    assert np.array_equal(((d & d2) & GUARD) == GUARD, want & np.roll(want, 1))


@pytest.mark.parametrize("kind,n,m,seed", [("uniform", 700, 4, 1), ("nasty", 900, 5, 2), ("integer_grid", 600, 4, 3),
                                           ("uniform", 500, 10, 4)])
def test_levels_plus_fixup_give_the_reference_relation(kind, n, m, seed):
    """superset from the levels, minus the same-level pairs that compare the other way as floats, == the relation
    the float kernel builds: all_{k>=1}(x_ik <= x_jk) (the uid test supplies objective 0 and strictness, and is the
    same in both kernels)."""
    x = getattr(h, kind)(n, m, seed)
    if kind == "uniform":                                                 # crowd the levels so that the fix-up has work
        x = (1.0 + x * 2.0 ** -9).astype(np.float32)
    clean = ~np.isnan(x).any(axis=1)                                      # NaN rows are rejected by the uid test
    x = x[clean]
    want = np.ones((len(x), len(x)), dtype=bool)
    sup = np.ones_like(want)
    wrong = np.zeros_like(want)
    for k in range(1, m):
        lv, _ = levels_of(x[:, k])
        want &= x[:, k][:, None] <= x[:, k][None, :]
        sup &= lv[:, None] <= lv[None, :]
        wrong |= (lv[:, None] == lv[None, :]) & (x[:, k][:, None] > x[:, k][None, :])   # what level_fixup_find records
    assert np.all(sup[want])                                              # superset
    assert np.array_equal(sup & ~wrong, want)                             # exact after the fix-up
    if kind == "uniform":
        assert (sup & wrong).sum() > 0                                    # the case really exercises the fix-up


def test_build_path_host_side_answer(monkeypatch):
    """evoxb_nds_build_path with a null workspace is pure host logic: which shapes the level build is planned for."""
    from evox_myc_b200 import _lib

    L = _lib.lib()
    path = ctypes.c_int(-1)

    def planned(n, m):
        _lib.check(L.evoxb_nds_build_path(None, n, m, None, ctypes.byref(path)))
        return path.value

    monkeypatch.delenv("EVOXB_BUILD_MODE", raising=False)
    assert planned(262144, 5) == 1 and planned(65536, 10) == 1 and planned(65536, 4) == 1
    assert planned(65536, 3) == 0 and planned(65535, 5) == 0 and planned(262144, 20) == 0 and planned(0, 5) == 0
    monkeypatch.setenv("EVOXB_BUILD_MODE", "levels")
    assert planned(100, 2) == 1 and planned(100, 1) == 0 and planned(100, 17) == 0
    monkeypatch.setenv("EVOXB_BUILD_MODE", "exact")
    assert planned(262144, 5) == 0
    # the workspace grows by the level buffers only where they are used
    need_a, need_b = ctypes.c_size_t(), ctypes.c_size_t()
    _lib.check(L.evoxb_nds_workspace_bytes(65536, 5, ctypes.byref(need_a)))
    monkeypatch.setenv("EVOXB_BUILD_MODE", "levels")
    _lib.check(L.evoxb_nds_workspace_bytes(65536, 5, ctypes.byref(need_b)))
    assert 0 < need_b.value - need_a.value < (32 << 20)


def test_level_map_monotone_on_arbitrary_floats():
    """Property form of the monotonicity above on arbitrary float32 bit patterns (hypothesis): denormals, both
    zeros, infinities, NaN (excluded from the order, as in the kernels), huge dynamic range."""
    hyp = pytest.importorskip("hypothesis")
    from hypothesis import given, settings
    from hypothesis import strategies as st
    from hypothesis.extra import numpy as hnp

    @settings(max_examples=60, deadline=None)
    @given(hnp.arrays(np.uint32, st.integers(2, 400), elements=st.integers(0, 2**32 - 1)))
    def check(bits):
        x = bits.view(np.float32)
        lv, key = levels_of(x)
        ok = ~np.isnan(x)
        if ok.sum() < 2:
            return
        xs, ls = x[ok], lv[ok].astype(np.int64)
        le = xs[:, None] <= xs[None, :]
        assert np.all((ls[:, None] <= ls[None, :])[le])
        assert ls.max() < LEVELS

    check()
