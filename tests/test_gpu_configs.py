"""The BASELINE.json configurations C1-C5 as GPU parity cases.

C1  Z^10, sigma = 5, 10 000 samples                          replay bit-exact (golden klein_z10) + native statistics
C2  LLL-reduced random lattice n = 64 (and the recorded n = 16, 32 runs), Klein + IMHK with many chains:
    per-coordinate spread against the statistics the reference RECORDED for exactly these bases
    (results/Klein_LLL_n=*.json -> tests/golden/klein_lll_recorded.json)
C3  q-ary LWE lattice n = 256, m = 512, q = 3329              V == B x exactly; replay vs oracle; marginals vs oracle
C4  NTRU n = 512, q = 12289 trapdoor basis (1024-dim)         replay vs oracle at full size; native size-independent
                                                              properties (V == B x, E||v||^2, chi-square of marginals)
C5  dimension sweep n = 16 .. 1024                            marginal TVD GPU vs oracle within sampling noise
"""
import json
import os

import numpy as np
import pytest
from scipy import stats

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))


@pytest.fixture(scope="module")
def lgs():
    import __graft_entry__ as g
    g.build()
    import lattice_gaussian_mcmc_b200 as m
    return m


@pytest.fixture(scope="module")
def oracle():
    from oracle import oracle as o
    return o


@pytest.fixture(scope="module")
def ntru512():
    from lattice_gaussian_mcmc_b200.ntru import ntru_lattice
    return ntru_lattice(512, 12289, seed=0)


def _recorded():
    return json.load(open(os.path.join(HERE, "golden", "klein_lll_recorded.json")))["cases"]


def _marginal_tvd_ratio(Xa, Xb):
    """max over coordinates of TVD(marginal a, marginal b) / sqrt(bins / S)."""
    S = min(len(Xa), len(Xb))
    worst = 0.0
    for d in range(Xa.shape[1]):
        lo = min(Xa[:, d].min(), Xb[:, d].min())
        hi = max(Xa[:, d].max(), Xb[:, d].max())
        a = np.bincount(Xa[:, d] - lo, minlength=hi - lo + 1) / len(Xa)
        b = np.bincount(Xb[:, d] - lo, minlength=hi - lo + 1) / len(Xb)
        bins = int(((a + b) > 0).sum())
        worst = max(worst, 0.5 * np.abs(a - b).sum() / np.sqrt(bins / S))
    return worst


# ------------------------------------------------------------------------------------------------- C1
def test_c1_identity_z10_quickstart(lgs):
    lat = lgs.lattices.identity_lattice(10)
    S, sigma = 10000, 5.0
    smp = lgs.KleinSampler(lat, sigma, seed=0)
    V = smp.sample(S)
    assert V.shape == (S, 10) and V.dtype == np.float64 and V.flags["C_CONTIGUOUS"]
    assert np.array_equal(V, np.rint(V))
    assert np.all(np.abs(V.mean(axis=0)) < 3 * sigma / np.sqrt(S))            # tests/unit/test_samplers.py:56-70
    assert np.all(np.abs(V.std(axis=0) / sigma - 1) < 0.03)
    # chi-square, alpha = 0.01, 1-D histogram against the discrete Gaussian (tests/unit/test_samplers.py:90-131)
    k = np.arange(-25, 26)
    p = np.exp(-k ** 2 / (2 * sigma ** 2))
    p /= p.sum()
    obs = np.bincount((np.clip(V[:, 3], -25, 25) + 25).astype(int), minlength=51)
    keep = p * S >= 5
    chi2 = np.sum((obs[keep] - p[keep] * S) ** 2 / (p[keep] * S))
    assert stats.chi2.sf(chi2, keep.sum() - 1) > 0.001
    # seed twice => identical arrays (tests/unit/test_samplers.py:217-233)
    assert np.array_equal(V, lgs.KleinSampler(lat, sigma, seed=0).sample(S))


# ------------------------------------------------------------------------------------------------- C2
@pytest.mark.parametrize("n", [16, 32, 64])
def test_c2_lll_reduced_lattice_matches_reference_recorded_statistics(lgs, n):
    """Same basis (seed 0+n, LLL), same sigma = 1.5 max||b*||, same N = 50 000 as the reference's recorded run:
    per-coordinate std and mean of the integer coefficients must agree within the sampling noise of two
    independent 50k-sample estimates (std: relative 1/sqrt(2N) each; mean: std/sqrt(N) each).
    n = 32: the recorded Gram-Schmidt profile is reproduced to 1e-14 (tests/test_host_setup.py) but the recorded
    per-coordinate stds sit up to 15 % off sigma sqrt(diag((B^T B)^-1)) of our reduced basis while the oracle is
    within 0.6 % of it -- Sage's LLL broke a size-reduction tie differently there (same GS norms, different column
    Gram matrix), so only the theoretical spread is checked for that case."""
    rec = _recorded()[str(n)]
    lat = lgs.lattices.lll_lattice(n)
    sigma = 1.5 * lat.max_gram_schmidt_norm
    assert abs(sigma - rec["sigma"]) < 1e-9 * sigma
    N = 50000
    X = lgs.KleinSampler(lat, sigma, seed=n).sample_coefficients(N)
    std, mean = X.std(axis=0), X.mean(axis=0)
    B = lat.basis
    theory = sigma * np.sqrt(np.diag(np.linalg.inv(B.T @ B)))      # cov(x) = sigma^2 (B^T B)^-1 above smoothing
    assert np.max(np.abs(std / theory - 1)) < 0.025, np.max(np.abs(std / theory - 1))
    if n == 32:
        return
    ref_std, ref_mean = np.array(rec["all_stds"]), np.array(rec["all_means"])
    assert np.max(np.abs(std / ref_std - 1)) < 5 * np.sqrt(2) / np.sqrt(2 * N), np.max(np.abs(std / ref_std - 1))
    assert np.all(np.abs(mean - ref_mean) < 5 * np.sqrt(2) * ref_std / np.sqrt(N))


def test_c2_imhk_many_chains_on_the_lll_lattice(lgs):
    """n = 64, 4096 chains, burn_in = 1000, thin = 10 (keep shortened to 4 per chain): reference weights accept
    everything (the reference records acceptance 1.000), and the kept states have the recorded spread."""
    rec = _recorded()["64"]
    lat = lgs.lattices.lll_lattice(64)
    sigma = rec["sigma"]
    im = lgs.IMHKSampler(lat, sigma, burn_in=0, n_chains=4096, seed=64)
    X = im.sample_coefficients(4, thin=10, burn_in=1000)
    assert X.shape == (4096, 4, 64)
    assert im.total_proposals == 4096 * (1000 + 40)
    assert im.acceptance_rate > 0.9999
    flat = X.reshape(-1, 64)
    ref_std = np.array(rec["all_stds"])
    assert np.max(np.abs(flat.std(axis=0) / ref_std - 1)) < 0.04
    # chains are independent and keyed by global chain id: two half-size chain sets reproduce the full set
    a = lgs.IMHKSampler(lat, sigma, burn_in=0, n_chains=2048, seed=64, first_chain_id=0)
    b = lgs.IMHKSampler(lat, sigma, burn_in=0, n_chains=2048, seed=64, first_chain_id=2048)
    Xa = a.sample_coefficients(1, thin=3, burn_in=5)
    Xb = b.sample_coefficients(1, thin=3, burn_in=5)
    whole = lgs.IMHKSampler(lat, sigma, burn_in=0, n_chains=4096, seed=64).sample_coefficients(1, thin=3, burn_in=5)
    assert np.array_equal(np.concatenate([Xa, Xb]), whole)


# ------------------------------------------------------------------------------------------------- C3
def test_c3_qary_lwe_lattice(lgs, oracle):
    lat = lgs.lattices.qary_lattice(256, 512, 3329, seed=0)
    q = 3329
    sigma = 1.5 * q
    B = lat.basis
    # replay: the ordered path on the recorded stream against the oracle (window truncated to 1001 entries on the
    # sigma_i = sigma coordinates, 3 sigma_i = 1.5 on the q-coordinates)
    prob = oracle.Problem(B, sigma)
    S = 48
    U = oracle.uniform_stream(3, S * prob.ups)
    Xo, Vo, lqo = oracle.klein_replay(prob, U, S)
    rp = lgs.KleinSampler(lat, sigma, mode="replay", uniforms=U)
    V, X, lq = rp.sample_with_coefficients(S)
    assert np.array_equal(X, Xo) and np.array_equal(V, Vo)
    assert np.max(np.abs(lq - lqo) / np.abs(lqo)) < 1e-12
    # native, batched: exact lattice membership and the two marginal regimes against oracle samples
    N = 1 << 15
    smp = lgs.KleinSampler(lat, sigma, seed=1)
    Vn, Xn, _ = smp.sample_with_coefficients(N)
    assert np.array_equal(Vn, Xn.astype(np.float64) @ B.T)
    top = np.rint(Vn[:, :256]).astype(np.int64)
    low = np.rint(Vn[:, 256:]).astype(np.int64)
    A = np.rint(B[:256, 256:]).astype(np.int64)
    assert np.all((top - low @ A.T) % q == 0)                      # x1 = A x2 (mod q)
    Xs, _ = oracle.klein_seeded(prob, 4000, seed=5, chains=os.cpu_count(), threads=os.cpu_count())
    assert _marginal_tvd_ratio(Xs[:, ::37], Xn[:4000, ::37]) < 1.2
    # truncated window: |x - round(mean)| <= 500 always, and both tails are reached
    Rd = np.abs(np.diag(smp.R))
    wide = np.where(sigma / Rd > 50)[0]
    assert wide.size == 256
    spread = Xn[:, wide].std(axis=0)
    assert np.all(spread < 500 / np.sqrt(3) * 1.05) and np.all(spread > 250)


# ------------------------------------------------------------------------------------------------- C4
def test_c4_ntru512_replay_matches_oracle_at_full_size(lgs, oracle, ntru512):
    sigma = 165.74
    prob = oracle.Problem(ntru512.basis, sigma)
    S = 6
    U = oracle.uniform_stream(2026, S * prob.ups)
    Xo, Vo, lqo = oracle.klein_replay(prob, U, S)
    smp = lgs.KleinSampler(ntru512, sigma, mode="replay", uniforms=U)
    V, X, lq = smp.sample_with_coefficients(S)
    assert np.array_equal(X, Xo), int(np.sum(X != Xo))
    assert np.array_equal(V, Vo)
    assert np.max(np.abs(lq - lqo) / np.abs(lqo)) < 1e-12
    # IMHK on the recorded stream: accept bits and log-weights
    steps, chains = 3, 2
    per = oracle.imhk_stream_len(prob, steps)
    Ui = oracle.uniform_stream(7, chains * per)
    ref = oracle.imhk_replay(prob, Ui, chains, steps, thin=1, exact_logz=False)
    im = lgs.IMHKSampler(ntru512, sigma, burn_in=0, n_chains=chains, mode="replay", uniforms=Ui)
    Vi = im.sample(steps, thin=1)
    assert np.array_equal(Vi, ref["V"])
    assert np.array_equal(im._last_accept, ref["accept"])
    assert np.max(np.abs(im._last_logw - ref["logw"]) / np.abs(ref["logw"])) < 1e-12


def test_c4_ntru512_native_properties(lgs, oracle, ntru512):
    sigma, n = 165.74, 1024
    B = ntru512.basis
    N = 1 << 15
    smp = lgs.KleinSampler(ntru512, sigma, seed=4, path="blocked")
    V, X, lq = smp.sample_with_coefficients(N)
    assert np.array_equal(V, X.astype(np.float64) @ B.T)                      # exact: integer basis
    # lattice membership through the NTRU relation is implied by V = B x; norm statistics:
    r2 = np.einsum("ij,ij->i", V, V)
    assert abs(r2.mean() / (n * sigma ** 2) - 1) < 0.02
    assert np.all(np.abs(V.mean(axis=0)) < 5 * sigma / np.sqrt(N))
    # log q by-product equals the density recomputed from the coefficients by the ordered kernel (klein.py:222-271)
    lc = smp.compute_log_density_batch(X[:256])
    assert np.max(np.abs(lc - lq[:256]) / np.abs(lc)) < 1e-10
    # marginals of the first and last drawn coordinates against the exact conditional law of the LAST coordinate
    # (x_{n-1} ~ D_{Z, sigma_{n-1}, 0}) and against oracle samples for a spread of coordinates
    Rd = np.abs(np.diag(smp.R))
    s_last = sigma / Rd[-1]
    k = np.arange(-12, 13)
    p = np.exp(-k ** 2 / (2 * s_last ** 2))
    p /= p.sum()
    obs = np.bincount(np.clip(X[:, -1], -12, 12) + 12, minlength=25)
    keep = p * N >= 5
    chi2 = np.sum((obs[keep] - p[keep] * N) ** 2 / (p[keep] * N))
    assert stats.chi2.sf(chi2, keep.sum() - 1) > 0.001
    prob = oracle.Problem(B, sigma)
    Xs, _ = oracle.klein_seeded(prob, 1500, seed=9, chains=os.cpu_count(), threads=os.cpu_count())
    assert _marginal_tvd_ratio(Xs[:, ::64], X[:1500, ::64]) < 1.2


# ------------------------------------------------------------------------------------------------- C5
@pytest.mark.parametrize("n,S", [(16, 20000), (64, 8000), (128, 4000), (256, 2000), (512, 1000), (1024, 600)])
def test_c5_dimension_sweep_marginals_vs_oracle(lgs, oracle, n, S):
    lat = lgs.lattices.lll_lattice(n) if n <= 64 else lgs.lattices.dd_lattice(n)
    sigma = 1.5 * lat.max_gram_schmidt_norm
    prob = oracle.Problem(lat.basis, sigma)
    Xs, _ = oracle.klein_seeded(prob, S, seed=n, chains=os.cpu_count(), threads=os.cpu_count())
    smp = lgs.KleinSampler(lat, sigma, seed=n + 1)
    V, X, _ = smp.sample_with_coefficients(S)
    ref = X.astype(np.float64) @ lat.basis.T
    assert np.max(np.abs(V - ref)) <= 1e-12 * np.max(np.abs(ref))
    step = max(1, n // 16)
    assert _marginal_tvd_ratio(Xs[:, ::step], X[:, ::step]) < 1.2
    assert abs(np.mean(np.einsum("ij,ij->i", V, V)) / (n * sigma ** 2) - 1) < 0.1   # tests/test_samplers.py:336-360
