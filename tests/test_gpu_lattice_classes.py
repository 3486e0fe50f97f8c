"""the reference-named lattice classes drive the samplers (README.md:37-77 with the imports swapped),
and the NTRU PUBLIC basis [[q I, 0], [H, I]] (ntru_clean.py:116-134) -- sigma_i = sigma / q (three-entry windows) and
sigma_i = sigma (1001-entry truncated windows) in one lattice -- replays bit for bit against the oracle."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


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


def test_readme_quick_start_with_swapped_imports(lgs):
    from lattice_gaussian_mcmc_b200.lattices import IdentityLattice, QaryLattice
    from lattice_gaussian_mcmc_b200.ntru import NTRULattice
    from lattice_gaussian_mcmc_b200.samplers import IMHKSampler, KleinSampler
    np.random.seed(0)
    lattice = IdentityLattice(dimension=10)                      # README.md:41-48
    sampler = KleinSampler(lattice, sigma=5.0)
    sample = sampler.sample()
    assert sample.shape == (1, 10) and np.array_equal(sample, np.rint(sample))
    imhk = IMHKSampler(lattice, sigma=5.0)                       # README.md:53-62
    samples = imhk.sample(n_samples=500, burn_in=100, thin=10)
    assert len(samples) == 500 and samples.shape == (500, 10)
    assert 0.9 < imhk.acceptance_rate <= 1.0
    assert np.all(np.abs(samples.mean(axis=0)) < 5 * 5.0 / np.sqrt(500))
    n, m, q = 64, 128, 3329                                      # README.md:69-72 at a test size
    lwe_lattice = QaryLattice.from_lwe_instance(n, m, q, alpha=3.2 / np.sqrt(q))
    V = KleinSampler(lwe_lattice, sigma=1.5 * q).sample(64)
    assert V.shape == (64, m) and np.all((V[:, :m - n] - V[:, m - n:] @ lwe_lattice.A.T) % q == 0)
    ntru = NTRULattice(n=64, q=12289, sigma=4.05 * np.sqrt(8))   # README.md:75-76
    ntru.generate_basis()
    smp = KleinSampler(ntru, sigma=1.3 * ntru.max_gram_schmidt_norm)
    V, X, _ = smp.sample_with_coefficients(256)
    assert np.array_equal(V, X.astype(np.float64) @ ntru.basis.T)
    H = ntru.get_public_basis()[64:, :64]
    assert np.all((V[:, :64] - V[:, 64:] @ H) % 12289 == 0)      # members of the lattice the public key defines


@pytest.mark.parametrize("n,S", [(64, 256), (512, 4)])
def test_ntru_public_basis_replay_matches_the_oracle(lgs, oracle, n, S):
    from lattice_gaussian_mcmc_b200.ntru import NTRULattice
    q, sigma = 12289, 165.74
    L = NTRULattice(n=n, q=q, sigma=4.05 * np.sqrt(512 / n), seed=1).generate_basis()
    pub = L.public_lattice()
    prob = oracle.Problem(pub.basis, sigma)
    U = oracle.uniform_stream(33 + n, S * prob.ups)
    Xo, Vo, lqo = oracle.klein_replay(prob, U, S)
    smp = lgs.KleinSampler(pub, sigma, mode="replay", uniforms=U)
    V, X, lq = smp.sample_with_coefficients(S)
    assert np.array_equal(X, Xo), int(np.sum(X != Xo))
    assert np.array_equal(V, Vo)
    assert np.max(np.abs(lq - lqo) / np.maximum(np.abs(lqo), 1e-300)) < 1e-12
    # both regimes are present: sigma_i = sigma (wide) and sigma / q (the draw is the rounded centre)
    si = sigma / np.abs(np.diag(smp.R))
    assert np.sum(si > 100) == n and np.sum(si < 0.05) == n


def test_ntru_public_basis_native_samples(lgs, oracle):
    """native Philox draws on the public basis: exact lattice points, the wide coordinates span their truncated window,
    the narrow ones sit on the rounded centre, marginals agree with oracle samples"""
    from lattice_gaussian_mcmc_b200.ntru import NTRULattice
    n, q, sigma = 64, 12289, 165.74
    L = NTRULattice(n=n, q=q, sigma=4.05 * np.sqrt(8), seed=1).generate_basis()
    pub = L.public_lattice()
    smp = lgs.KleinSampler(pub, sigma, seed=6)
    N = 20000
    V, X, lq = smp.sample_with_coefficients(N)
    assert np.array_equal(V, X.astype(np.float64) @ pub.basis.T)
    H = L.get_public_basis()[n:, :n]
    assert np.all((V[:, :n] - V[:, n:] @ H) % q == 0)
    wide = X[:, n:]                                   # coefficients of the (h x^i, x^i) vectors: sigma_i = sigma
    assert np.all(np.abs(wide.std(axis=0) / sigma - 1) < 0.05) and np.abs(wide).max() <= 501
    assert np.all(np.abs(V[:, :n]) <= q)              # the q-vectors reduce the first half to (about) one period
    prob = oracle.Problem(pub.basis, sigma)
    Xs, _ = oracle.klein_seeded(prob, 3000, seed=9, chains=os.cpu_count(), threads=os.cpu_count())
    for d in (n, n + 17, 2 * n - 1):                  # two-sample check of wide marginals: means and spreads
        assert abs(X[:, d].mean() - Xs[:, d].mean()) < 5 * sigma * np.sqrt(1 / N + 1 / 3000)
        assert abs(X[:, d].std() / Xs[:, d].std() - 1) < 0.08
    lc = smp.compute_log_density_batch(X[:128])
    assert np.max(np.abs(lc - lq[:128]) / np.abs(lc)) < 1e-10
