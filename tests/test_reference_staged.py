"""The reference's own numpy samplers, staged into oracle/_ref by build() (oracle/ref_python.py), against the oracle
restatement: the same check the golden vectors make, run live wherever the staged copy exists (the build container and,
because oracle/_ref travels with the snapshot, the GPU box).  Also covers the checker-side helpers of bench.py."""
import os

import numpy as np
import pytest

from oracle import oracle, ref_python


@pytest.fixture(scope="module")
def staged():
    ref_python.stage()
    if not ref_python.available():
        pytest.skip("oracle/_ref not staged: /root/reference is not mounted here")
    return ref_python._import_reference()


@pytest.mark.parametrize("n,seed", [(12, 5), (20, 11)])
def test_oracle_equals_the_live_reference_on_a_fresh_case(staged, n, seed):
    """a case that is NOT among the committed goldens: klein.py:181-220 after np.random.seed(seed) vs the oracle on
    RandomState(seed).random_sample -- integer coefficients bit for bit"""
    Klein, _ = staged
    rng = np.random.RandomState(100 + n)
    B = (rng.randint(-3, 4, size=(n, n)) + 2 * n * np.eye(n)).astype(np.float64)
    sigma = 1.5 * 2 * n
    lat = ref_python.ShimLattice(B)
    smp = Klein(lat, sigma)
    smp._sample_cache = type("NoCache", (dict,), {"__contains__": lambda s, k: False,
                                                  "__setitem__": lambda s, k, v: None, "__len__": lambda s: 0})()
    S = 25
    np.random.seed(seed)
    V = smp.sample(S)
    prob = oracle.Problem(B, sigma)
    U = oracle.uniform_stream(seed, S * prob.ups)
    Xo, Vo, _ = oracle.klein_replay(prob, U, S)
    assert np.array_equal(V, Vo)
    assert np.array_equal(np.rint(np.linalg.solve(B, V.T).T).astype(np.int64), Xo)


def test_reference_timing_legs_run(staged):
    rng = np.random.RandomState(3)
    B = (rng.randint(-3, 4, size=(16, 16)) + 32 * np.eye(16)).astype(np.float64)
    k = ref_python.time_klein(B, 48.0, procs=2, samples_per_proc=5)
    assert k["unit"] == "samples/s" and k["value"] > 0 and k["procs"] == 2 and k["units_per_proc"] == 5
    m = ref_python.time_imhk(B, 48.0, procs=2, steps_per_proc=3, thin=2)
    assert m["unit"] == "chain-steps/s" and m["units_per_proc"] == 6 and m["value"] > 0


def test_conditional_counts_accepts_the_law_and_rejects_a_biased_sampler():
    from scipy import stats
    rng = np.random.RandomState(8)
    n = 24
    B = (rng.randint(-3, 4, size=(n, n)) + 2 * n * np.eye(n)).astype(np.float64)
    sigma = 1.5 * 2 * n
    prob = oracle.Problem(B, sigma)
    S = 20000
    X, _ = oracle.klein_seeded(prob, S, seed=3, chains=4, want_v=False, threads=4)
    obs, exp = oracle.conditional_counts(prob, X, K=5, FB=4, threads=4)
    assert obs.shape == exp.shape == (n, 4, 11)
    assert obs.sum() == S * n and abs(exp.sum() - S * n) < 1e-6 * S * n
    # inside a frac bin observed and expected totals agree (every table sums to one)
    assert np.allclose(obs.sum(axis=2), exp.sum(axis=2), rtol=0, atol=1e-6 * S)

    def agg(o, e):
        keep = e >= 8
        return stats.chi2.sf(np.sum((o[keep] - e[keep]) ** 2 / e[keep]), keep.sum() - o.shape[0] * o.shape[1])

    assert agg(obs, exp) > 1e-3
    wide = oracle.Problem(B, sigma * 1.03)
    Xw, _ = oracle.klein_seeded(wide, S, seed=3, chains=4, want_v=False, threads=4)
    ow, ew = oracle.conditional_counts(prob, Xw, K=5, FB=4, threads=4)
    assert agg(ow, ew) < 1e-6
    # coarse bins: width 3 pools offsets {-1, 0, 1} into the centre cell
    o3, e3 = oracle.conditional_counts(prob, X, K=2, FB=1, threads=2, width=np.full(n, 3, dtype=np.int32))
    assert o3.sum() == S * n
    assert np.array_equal(o3[:, 0, 2], (obs[:, :, 4:7].sum(axis=(1, 2))))


def test_bench_helpers():
    import bench
    traffic, src = bench.ncu_traffic()
    assert traffic and traffic > 1e9 and src.startswith("profiles/") and src.endswith("ncu_full_summary.json")
    info = bench.host_info()
    assert info["nproc"] >= 1 and "affinity" in info
