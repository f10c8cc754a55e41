"""GPU tests of the NSGA-II generation loop: variation and DTLZ2 kernels against the NumPy restatement
(same Philox uniforms => same offspring up to powf rounding), statistical properties of the operators,
and convergence of the whole loop on DTLZ2 (the reference's own algorithm tests assert no more)."""

import numpy as np
import pytest

import helpers as h

pytestmark = pytest.mark.gpu
torch = pytest.importorskip("torch")


def test_philox_known_answer():
    # Random123 kat_vectors: philox4x32-10, counter = key = 0
    from oracle import variation_oracle as vo

    r = vo.philox4x32_10(np.zeros(1, np.uint64), 0, 0, 0)
    assert [int(v[0]) for v in r] == [0x6627E8D5, 0xE169C58D, 0xBC57AC4C, 0x9B00DBD8]


@pytest.mark.parametrize("n,d", [(200, 12), (201, 12), (4096, 7), (2, 3)])
def test_variation_matches_restatement(n, d):
    from evox_myc_b200 import nsga2
    from oracle import variation_oracle as vo

    rng = np.random.default_rng(n + d)
    lb = (rng.random(d) - 0.5).astype(np.float32)
    ub = (lb + 0.5 + rng.random(d)).astype(np.float32)
    pop = (lb + (ub - lb) * rng.random((n, d))).astype(np.float32)
    got = nsga2.variation(torch.from_numpy(pop).cuda(), torch.from_numpy(lb).cuda(), torch.from_numpy(ub).cuda(), 1234, 7)
    want = vo.nsga2_ask(pop, lb, ub, 1234, 7)
    np.testing.assert_allclose(got.cpu().numpy(), want, rtol=2e-5, atol=2e-6)


def test_variation_statistics():
    from evox_myc_b200 import nsga2

    n, d = 20000, 12
    lb, ub = torch.zeros(d, device="cuda"), torch.ones(d, device="cuda")
    pop = torch.rand((n, d), device="cuda")
    out = nsga2.variation(pop, lb, ub, 99, 3, nsga2.SimulatedBinary(), nsga2.Polynomial((lb, ub), pro_m=0.0))
    assert float(out.min()) >= 0.0 and float(out.max()) <= 1.0
    # without mutation and away from the bounds SBX preserves the parents' mean (sbx.py:27-33)
    half = n // 2
    inner = ((out > 0) & (out < 1)).all(dim=1)
    both = inner[:half] & inner[half:]
    mean_p = (pop[:half] + pop[half:]) / 2
    mean_c = (out[:half] + out[half:]) / 2
    assert torch.allclose(mean_p[both], mean_c[both], atol=1e-6)
    # about half of the genes are copied unchanged (beta = 1 with probability 1/2, sbx.py:22)
    same = ((out[:half] - pop[:half]).abs() < 1e-6).float().mean().item()  # (a+b)/2 + (a-b)/2 rounds, so not ==
    assert 0.45 < same < 0.56
    # mutation alone touches about pro_m / d of the genes
    out2 = nsga2.variation(pop, lb, ub, 99, 4, nsga2.SimulatedBinary(pro_c=0.0), nsga2.Polynomial((lb, ub), pro_m=1.0))
    changed = ((out2 - pop).abs() > 1e-6).float().mean().item()
    assert 0.8 / d < changed < 1.2 / d
    # different generation counter => different offspring; same arguments => identical offspring
    a = nsga2.variation(pop, lb, ub, 99, 5)
    assert torch.equal(a, nsga2.variation(pop, lb, ub, 99, 5)) and not torch.equal(a, nsga2.variation(pop, lb, ub, 99, 6))


@pytest.mark.parametrize("n,m,d", [(1000, 3, 12), (777, 2, 11), (300, 5, 14), (64, 10, 19)])
def test_dtlz2_matches_restatement(n, m, d):
    from evox_myc_b200 import nsga2
    from oracle import variation_oracle as vo

    X = np.random.default_rng(n).random((n, d), dtype=np.float32)
    f, _ = nsga2.DTLZ2(d=d, m=m).evaluate(None, torch.from_numpy(X).cuda())
    np.testing.assert_allclose(f.cpu().numpy(), vo.dtlz2(X, m), rtol=1e-5, atol=1e-6)


def test_tell_matches_oracle_selection():
    """One tell() on fixed inputs gives the survivors the oracle's nsga2_tell gives."""
    from evox_myc_b200 import nsga2
    from oracle import nds_oracle, variation_oracle as vo

    pop_size, d, m = 300, 12, 3
    rng = np.random.default_rng(5)
    P, Q = rng.random((pop_size, d), dtype=np.float32), rng.random((pop_size, d), dtype=np.float32)
    fP, fQ = vo.dtlz2(P, m), vo.dtlz2(Q, m)
    algo = nsga2.NSGA2(np.zeros(d), np.ones(d), m, pop_size)
    st = nsga2.State(population=torch.from_numpy(P).cuda(), fitness=torch.from_numpy(fP).cuda(),
                     next_generation=torch.from_numpy(Q).cuda(), key=0, generation=1)
    st = algo.tell(st, torch.from_numpy(fQ).cuda())
    wp, wf = nds_oracle.nsga2_tell(P, Q, fP, fQ, pop_size)
    assert np.array_equal(st.population.cpu().numpy(), wp) and np.array_equal(st.fitness.cpu().numpy(), wf)


def test_nsga2_converges_on_dtlz2_config1():
    """BASELINE config 1 shape (M=3, D=12, pop=100, 100 generations): IGD against the unit-sphere front."""
    from evox_myc_b200 import nsga2
    from oracle import variation_oracle as vo

    algo = nsga2.NSGA2(np.zeros(12), np.ones(12), 3, 100)
    state = nsga2.run(algo, nsga2.DTLZ2(m=3), 100, key=42)
    fit = state.fitness.cpu().numpy()
    rng = np.random.default_rng(0)
    pf = np.abs(rng.normal(size=(2000, 3)))
    pf /= np.linalg.norm(pf, axis=1, keepdims=True)
    start = vo.dtlz2(np.random.default_rng(1).random((100, 12), dtype=np.float32), 3)
    assert vo.igd(fit, pf) < 0.12 < vo.igd(start, pf)
    assert np.abs(np.linalg.norm(fit, axis=1) - 1).mean() < 0.05  # on the sphere: g -> 0


def test_lsmop1_reference_spot_values(kats):
    """tests/test_lsmop.py:10-29 through the CUDA kernel (input regenerated with oracle/jax_prng.py)."""
    from evox_myc_b200 import nsga2
    from oracle import jax_prng

    k = kats["lsmop1_100x300_prng0"]
    upper = np.hstack([np.ones(2), 10 * np.ones(298)]).astype(np.float32)
    data = jax_prng.uniform_array_f32(0, (100, 300), 0.0, upper)
    prob = nsga2.LSMOP1(d=300, m=3)
    r, _ = prob.evaluate(None, torch.from_numpy(data).cuda())
    r = r.cpu().numpy()
    assert list(r.shape) == k["shape"]
    for i, j, v in k["spot"]:
        assert abs(float(r[i, j]) - v) < k["atol"]


@pytest.mark.parametrize("n,m,d", [(1000, 3, 300), (257, 10, 1000), (64, 2, 40), (33, 16, 1600)])
def test_lsmop1_matches_restatement(n, m, d):
    from evox_myc_b200 import nsga2
    from oracle import lsmop_oracle

    rng = np.random.default_rng(n + m)
    x = rng.random((n, d), dtype=np.float32)
    x[:, m - 1 :] *= np.float32(10)
    prob = nsga2.LSMOP1(d=d, m=m)
    assert (prob.sublen, prob.len) == lsmop_oracle.lsmop_groups(d, m)
    got, _ = prob.evaluate(None, torch.from_numpy(x).cuda())
    want = lsmop_oracle.lsmop1(x, m)
    # float32 sums of up to ~700 squares in a different association: 1e-5 relative
    np.testing.assert_allclose(got.cpu().numpy(), want, rtol=1e-5, atol=1e-30)
