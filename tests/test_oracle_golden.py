"""The oracle against every golden vector the reference's own tests hold for this path."""

import numpy as np
import pytest

from helpers import finf
from oracle import c_oracle, jax_prng, nds_oracle as o

IMPLS = [("numpy", o), ("c", c_oracle)]


@pytest.mark.parametrize("name,impl", IMPLS)
def test_nds_kat(kats, name, impl):
    k = kats["nds_11x2"]
    rank = impl.non_dominated_sort(np.array(k["x"], dtype=np.float32))
    assert rank.dtype == np.int32
    assert rank.tolist() == k["rank"]


@pytest.mark.parametrize("name,impl", IMPLS)
def test_crowding_kat(kats, name, impl):
    k = kats["crowd_5x2"]
    d = impl.crowding_distance(np.array(k["x"], dtype=np.float32))
    np.testing.assert_allclose(d, finf(k["dist"]), rtol=1e-6)


@pytest.mark.parametrize("name,impl", IMPLS)
def test_crowding_masked_kat(kats, name, impl):
    k = kats["crowd_masked_9x2"]
    d = impl.crowding_distance(np.array(k["x"], dtype=np.float32), np.array(k["mask"]))
    np.testing.assert_allclose(d, finf(k["dist"]), rtol=1e-6)


@pytest.mark.parametrize("name,impl", IMPLS)
def test_monitor_front_extraction_int32(kats, name, impl):
    # tests/test_monitors.py:30-48 -- archive update: concat, rank == 0 on int32 / mixed fitness
    k = kats["monitor_pf_int32"]
    f1 = np.array(k["fitness1"], dtype=np.int32)
    pf1 = f1[impl.non_dominated_sort(f1.astype(np.float32)) == 0]
    assert pf1.tolist() == k["pf1"]
    f2 = np.concatenate([pf1.astype(np.float32), np.array(k["fitness2"], dtype=np.float32)])
    pf2 = f2[impl.non_dominated_sort(f2) == 0]
    assert pf2.tolist() == k["pf2"]


def test_threefry_known_answer():
    # Random123 known-answer vector for Threefry-2x32-20: zero key, zero counter
    x0, x1 = jax_prng.threefry2x32((0, 0), np.zeros(1, np.uint32), np.zeros(1, np.uint32))
    assert (int(x0[0]), int(x1[0])) == (0x6B200159, 0x99BA4EFE)


@pytest.mark.parametrize("name,impl", IMPLS)
def test_crowding_golden_128x8(kats, name, impl):
    # tests/test_crowding_distance.py:24-124: the input is regenerated with oracle/jax_prng.py
    k = kats["crowd_front0_128x8_prng314"]
    x = jax_prng.normal_f32(314, (128, 8))
    rank = impl.non_dominated_sort(x)
    front = x[rank == 0]
    gold = finf(k["dist"])
    assert front.shape[0] == gold.shape[0] == 90
    d = impl.crowding_distance(front)
    assert np.array_equal(np.isinf(d), np.isinf(gold))
    fin = np.isfinite(gold)
    np.testing.assert_allclose(d[fin], gold[fin], rtol=k["rtol"])


def test_lsmop1_spot_values(kats):
    # tests/test_lsmop.py:10-29: the input is regenerated with oracle/jax_prng.py
    from oracle import lsmop_oracle

    k = kats["lsmop1_100x300_prng0"]
    upper = np.hstack([np.ones(2), 10 * np.ones(298)]).astype(np.float32)
    data = jax_prng.uniform_array_f32(0, (100, 300), 0.0, upper)
    r = lsmop_oracle.lsmop1(data, 3)
    assert list(r.shape) == k["shape"]
    for i, j, v in k["spot"]:
        assert abs(float(r[i, j]) - v) < k["atol"]
