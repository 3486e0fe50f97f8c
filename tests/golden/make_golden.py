#!/usr/bin/env python
"""Mint golden vectors for the Klein / IMHK path by running the REFERENCE ITSELF.

Run in the build container, where the reference is mounted read-only at /root/reference:

    python tests/golden/make_golden.py            # writes tests/golden/*.npz

The reference has no checked-in bit-level vectors, so these are produced by importing
`src.samplers` from the reference tree (nothing is copied from it), driving RefinedKleinSampler / IMHKSampler
through a duck-typed lattice (the four attributes the samplers read: basis, dimension,
min_gram_schmidt_norm, smoothing_parameter(), name) after np.random.seed(seed), and recording

  * the inputs (basis -- columns are basis vectors, as klein.py:59,218 uses it --, sigma, center, seed),
  * the uniform stream np.random.RandomState(seed).random_sample(N) the run consumed,
  * every 1-D draw in call order: (mean, sigma_i, x_i)   [wrapping _sample_1d_discrete_gaussian],
  * the lattice points returned, the Klein log-density of each, and for IMHK every importance log-weight
    and accept bit,
  * numpy / scipy versions.

The reference's 6-decimal table cache (klein.py:148-162) makes draws weakly history dependent; each case is
also run with the cache neutralised and the number of differing draws is stored as `cache_diff_draws`
(0 for every case below).
"""
import os
import sys

sys.dont_write_bytecode = True
REF = os.environ.get("LGS_REFERENCE", "/root/reference")
sys.path.insert(0, REF)

import numpy as np  # noqa: E402
import scipy  # noqa: E402

from src.samplers.klein import RefinedKleinSampler  # noqa: E402
from src.samplers.imhk import IMHKSampler  # noqa: E402

OUT = os.path.dirname(os.path.abspath(__file__))


class ShimLattice:
    """The duck type the reference samplers consume."""

    def __init__(self, basis, name="shim"):
        self.basis = np.array(basis, dtype=np.float64)
        self.dimension = self.basis.shape[0]
        self.name = name
        # row Gram-Schmidt minimum, as lattices/simple.py:39-58,79-82 computes it
        gs = np.zeros_like(self.basis)
        for i in range(self.dimension):
            v = self.basis[i].copy()
            for j in range(i):
                v -= (self.basis[i] @ gs[j]) / (gs[j] @ gs[j]) * gs[j]
            gs[i] = v
        self.min_gram_schmidt_norm = float(np.min(np.linalg.norm(gs, axis=1)))

    def smoothing_parameter(self):
        return 0.0


class NoCache(dict):
    def __contains__(self, k):
        return False

    def __setitem__(self, k, v):
        pass

    def __len__(self):
        return 0


def instrument(sampler, log):
    orig = sampler._sample_1d_discrete_gaussian

    def wrapped(mean, sigma):
        x = orig(mean, sigma)
        log.append((float(mean), float(sigma), int(x)))
        return x

    sampler._sample_1d_discrete_gaussian = wrapped


def run_klein(basis, sigma, center, seed, S, no_cache=False):
    lat = ShimLattice(basis)
    smp = RefinedKleinSampler(lat, sigma, center)
    if no_cache:
        smp._sample_cache = NoCache()
        # the lru_cache on _compute_1d_probabilities keys on exact floats: transparent, leave it
    draws = []
    instrument(smp, draws)
    np.random.seed(seed)
    V = smp.sample(S) if S > 1 else smp.sample(1)
    logq = np.array([smp.compute_log_density(v) for v in V])
    return smp, V, logq, draws


def klein_case(name, basis, sigma, seed, S, center=None):
    basis = np.array(basis, dtype=np.float64)
    n = basis.shape[0]
    smp, V, logq, draws = run_klein(basis, sigma, center, seed, S)
    _, V2, _, draws2 = run_klein(basis, sigma, center, seed, S, no_cache=True)
    cache_diff = sum(1 for a, b in zip(draws, draws2) if a[2] != b[2]) + abs(len(draws) - len(draws2))
    per = len(draws) // S
    assert per * S == len(draws)
    d = np.array(draws, dtype=np.float64).reshape(S, per, 3)
    # coefficient vectors in coordinate order 0..n-1; coordinates that took the rounding path
    # (sigma_i < 1e-10, klein.py:201-204) never reach _sample_1d_discrete_gaussian, recover all by solve
    X = np.rint(np.linalg.solve(basis, V.T).T).astype(np.int64)
    assert np.array_equal(basis @ X.T, V.T) or np.allclose(basis @ X.T, V.T, rtol=1e-12, atol=1e-9)
    U = np.random.RandomState(seed).random_sample(len(draws))
    np.savez_compressed(
        os.path.join(OUT, name + ".npz"),
        kind="klein", basis=basis, sigma=float(sigma), center=smp.center, seed=int(seed), S=int(S),
        uniforms=U, draw_mean=d[:, :, 0], draw_sigma=d[:, :, 1], draw_x=d[:, :, 2].astype(np.int64),
        X=X, V=V, logq=logq, R=smp.R, Q=smp.Q, t=smp.center_transformed,
        cache_diff_draws=int(cache_diff), numpy=np.__version__, scipy=scipy.__version__)
    print(f"{name}: n={n} S={S} draws/sample={per} cache_diff_draws={cache_diff} "
          f"logq[0]={logq[0]:.6f}")


def imhk_case(name, basis, sigma, seed, steps, thin, center=None):
    basis = np.array(basis, dtype=np.float64)
    n = basis.shape[0]
    lat = ShimLattice(basis)
    smp = IMHKSampler(lat, sigma, center, burn_in=0)
    draws, logws = [], []
    instrument(smp.proposal_sampler, draws)
    orig_w = smp._compute_importance_weight

    def wrapped_w(pt):
        w, lw = orig_w(pt)
        logws.append(float(lw))
        return w, lw

    smp._compute_importance_weight = wrapped_w
    np.random.seed(seed)
    states, accepts = [], []
    for _ in range(steps):
        st, acc = smp.step()
        states.append(st)
        accepts.append(bool(acc))
    states = np.array(states)
    # the generator state after the run tells how many uniforms were consumed
    per = n
    n_uniform = per + steps * (per + 1)
    chk = np.random.RandomState(seed)
    chk.random_sample(n_uniform)
    assert np.random.get_state()[2] == chk.get_state()[2] and \
        np.array_equal(np.random.get_state()[1], chk.get_state()[1]), "stream length model wrong"
    U = np.random.RandomState(seed).random_sample(n_uniform)
    kept = states[thin - 1::thin]
    Xk = np.rint(np.linalg.solve(basis, kept.T).T).astype(np.int64)
    # the same chain through the public sample(num_samples, thin) entry must give the kept states
    smp2 = IMHKSampler(lat, sigma, center, burn_in=0)
    np.random.seed(seed)
    V_api = smp2.sample(steps // thin, thin=thin)
    assert np.array_equal(V_api, kept)
    np.savez_compressed(
        os.path.join(OUT, name + ".npz"),
        kind="imhk", basis=basis, sigma=float(sigma), center=smp.center, seed=int(seed), steps=int(steps),
        thin=int(thin), uniforms=U, V_keep=kept, X_keep=Xk, accept=np.array(accepts, dtype=np.uint8),
        logw=np.array(logws), klein_log_partition=float(smp.klein_log_partition),
        acceptance_rate=float(smp2.stats.acceptance_rate), R=smp.proposal_sampler.R,
        t=smp.proposal_sampler.center_transformed, numpy=np.__version__, scipy=scipy.__version__)
    print(f"{name}: n={n} steps={steps} thin={thin} acc={np.mean(accepts):.3f} "
          f"logw spread={np.ptp(logws):.2e} klp={smp.klein_log_partition:.12f}")


def dd_basis(n, seed):
    """Diagonally dominant integer basis, the family SURVEY.md appendix A.3/A.5 probes with."""
    rng = np.random.RandomState(seed)
    return (rng.randint(-3, 4, size=(n, n)) + 2 * n * np.eye(n)).astype(np.float64)


def main():
    # C1 shape: Z^10, sigma = 5 (README quick start), fewer samples
    klein_case("klein_z10", np.eye(10), 5.0, seed=0, S=64)
    # the 2-D bases the reference's own validation records (COMPREHENSIVE_BENCHMARK_RESULTS.md:75-134)
    klein_case("klein_2d_a", [[4, 1], [1, 3]], 2.0, seed=1, S=200)
    klein_case("klein_2d_b", [[1, 0.1], [0, 1]], 2.0, seed=2, S=200)
    klein_case("klein_2d_c", [[9, 3], [4, 7]], 2.0, seed=3, S=200, center=[0.3, -1.7])
    # diagonally dominant integer bases (appendix A.3), one with an off-lattice centre
    klein_case("klein_dd8", dd_basis(8, 8), 15.0, seed=123, S=50)
    klein_case("klein_dd24", dd_basis(24, 24), 15.0 * 3, seed=124, S=40,
               center=np.random.RandomState(5).uniform(-20, 20, 24))
    # a generic real (non-integer) basis
    rb = np.random.RandomState(16).normal(size=(16, 16)) * 3 + 12 * np.eye(16)
    klein_case("klein_real16", rb, 20.0, seed=7, S=40)
    # C2/C5 shape: n = 64, sigma = 1.5 max||b*||
    b64 = dd_basis(64, 64)
    r64 = np.abs(np.diag(np.linalg.qr(b64)[1]))
    klein_case("klein_dd64", b64, 1.5 * r64.max(), seed=64, S=12)
    # truncated window: sigma_i > 50 => 1001-entry table centred on round(mean) (klein.py:121-126)
    klein_case("klein_trunc", np.diag([1.0, 2.0, 0.5, 1.0]), 80.0, seed=9, S=50, center=[0.2, 0.0, 7.7, -3.4])
    # tiny conditional sigma: rounding path (sigma_i < 1e-10) and the max(3, precision) branch (sigma_i < 0.1)
    klein_case("klein_tiny", np.diag([1e12, 40.0, 1.0]), 2.0, seed=10, S=50, center=[3.3e11, 17.2, 0.4])
    # half-integer means: two equal maxima inside logsumexp (m = 2)
    klein_case("klein_half", np.eye(3), 1.3, seed=11, S=100, center=[0.5, 0.5, -1.5])
    # IMHK (appendix A.4) -- fresh sampler, every step recorded
    imhk_case("imhk_dd8", dd_basis(8, 8), 15.0, seed=200, steps=60, thin=3)
    imhk_case("imhk_real16", rb, 20.0, seed=201, steps=24, thin=2, center=np.linspace(-3, 3, 16))
    imhk_case("imhk_2d", [[4, 1], [1, 3]], 2.0, seed=202, steps=100, thin=1)


if __name__ == "__main__":
    main()
