"""Replay-mode parity on the GPU, through the C ABI (host forms of include/lgs_b200.h):

  * against the golden vectors minted by the reference itself (tests/golden/*.npz): integer coefficients bit-exact,
    lattice points exact for integer bases (<= 1e-12 relative otherwise), Klein log-density and IMHK log-weights
    <= 1e-12 relative, accept bits equal;
  * against the CPU oracle on the same recorded uniform stream at sizes the oracle finishes in seconds.

Tolerance for floating point: REL = 1e-12 (BASELINE.json north_star).  Integer work: bit-exact, mismatches counted.
"""
import os

import numpy as np
import pytest

from conftest import GOLDEN, golden_files

pytestmark = pytest.mark.gpu

REL = 1e-12


def _rel(a, b):
    a, b = np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64)
    return float(np.max(np.abs(a - b) / np.maximum(np.abs(b), 1e-300))) if a.size else 0.0


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


@pytest.mark.parametrize("fname", golden_files("klein"))
def test_klein_replay_matches_reference_golden(lgs, fname):
    g = np.load(os.path.join(GOLDEN, fname))
    lat = lgs.SimpleLattice(g["basis"])
    smp = lgs.KleinSampler(lat, float(g["sigma"]), g["center"], mode="replay", uniforms=g["uniforms"])
    assert _rel(smp.R, g["R"]) <= REL and _rel(smp.center_transformed, g["t"]) <= REL
    S = int(g["S"])
    assert smp._engine.ups == g["draw_x"].shape[1]
    V, X, lq = smp.sample_with_coefficients(S)
    mism = int(np.sum(np.any(X != g["X"], axis=1)))
    assert mism == 0, f"{mism} of {S} samples differ from the reference"
    if np.all(g["basis"] == np.rint(g["basis"])) and np.abs(g["basis"]).max() < 2**40:
        assert np.array_equal(V, g["V"])
    assert _rel(V, g["V"]) <= REL
    assert _rel(lq, g["logq"]) <= REL
    assert V.dtype == np.float64 and V.shape == (S, lat.dimension) and V.flags["C_CONTIGUOUS"]


def test_klein_replay_from_the_global_numpy_stream(lgs):
    """np.random.seed(s); sampler.sample(N) -- the reference's own reproducibility idiom
    (tests/unit/test_samplers.py:217-233) -- gives the reference's points."""
    g = np.load(os.path.join(GOLDEN, "klein_dd24.npz"))
    lat = lgs.SimpleLattice(g["basis"])
    smp = lgs.KleinSampler(lat, float(g["sigma"]), g["center"], mode="replay")
    np.random.seed(int(g["seed"]))
    V = smp.sample(int(g["S"]))
    assert np.array_equal(V, g["V"])
    np.random.seed(int(g["seed"]))
    V2 = smp.sample(int(g["S"]))
    assert np.array_equal(V, V2)
    # sample(1) has shape (1, n) (klein.py:334-335); sample_single() has shape (n,)
    np.random.seed(int(g["seed"]))
    one = smp.sample(1)
    assert one.shape == (1, lat.dimension) and np.array_equal(one[0], g["V"][0])
    assert smp.sample_single().shape == (lat.dimension,)


@pytest.mark.parametrize("fname", golden_files("imhk"))
def test_imhk_replay_matches_reference_golden(lgs, fname):
    g = np.load(os.path.join(GOLDEN, fname))
    lat = lgs.SimpleLattice(g["basis"])
    steps, thin = int(g["steps"]), int(g["thin"])
    im = lgs.IMHKSampler(lat, float(g["sigma"]), g["center"], burn_in=0, mode="replay", uniforms=g["uniforms"])
    V = im.sample(steps // thin, thin=thin)
    assert V.shape == g["V_keep"].shape
    assert _rel(V, g["V_keep"]) <= REL
    assert np.array_equal(im._last_accept[0], g["accept"])
    assert _rel(im._last_logw[0], g["logw"]) <= REL
    assert abs(im.klein_log_partition - float(g["klein_log_partition"])) <= 1e-12 * abs(float(g["klein_log_partition"]))
    assert im.acceptance_rate == float(g["acceptance_rate"]) == 1.0
    assert im.stats.acceptance_rate == 1.0 and im.stats.samples_generated == steps // thin
    assert im.total_proposals == steps and im.accepted_proposals == steps
    assert np.array_equal(im.current_state, V[-1])
    Xk = im.current_coeffs
    assert np.array_equal(Xk, g["X_keep"][-1])


def test_imhk_step_by_step_equals_sample(lgs):
    g = np.load(os.path.join(GOLDEN, "imhk_dd8.npz"))
    lat = lgs.SimpleLattice(g["basis"])
    im = lgs.IMHKSampler(lat, float(g["sigma"]), g["center"], burn_in=0, mode="replay", uniforms=g["uniforms"])
    thin = int(g["thin"])
    states = []
    for _ in range(int(g["steps"])):
        st, acc = im.step()
        assert acc is True
        states.append(st)
    assert np.array_equal(np.array(states)[thin - 1::thin], g["V_keep"])


@pytest.mark.parametrize("n,S,seed", [(32, 2000, 1), (64, 1500, 2), (128, 600, 3)])
def test_klein_replay_matches_oracle_dd(lgs, oracle, n, S, seed):
    from lattice_gaussian_mcmc_b200.lattices import dd_lattice
    lat = dd_lattice(n, seed=n + seed)
    sigma = 1.5 * lat.max_gram_schmidt_norm
    center = np.random.RandomState(seed).uniform(-30, 30, n)
    prob = oracle.Problem(lat.basis, sigma, center)
    U = oracle.uniform_stream(1000 + seed, S * prob.ups)
    Xo, Vo, lqo = oracle.klein_replay(prob, U, S, exact_logz=False)
    smp = lgs.KleinSampler(lat, sigma, center, mode="replay", uniforms=U)
    V, X, lq = smp.sample_with_coefficients(S)
    mism = int(np.sum(np.any(X != Xo, axis=1)))
    assert mism == 0, f"{mism} of {S} samples differ from the oracle"
    assert np.array_equal(V, Vo)
    assert _rel(lq, lqo) <= REL


def test_klein_replay_real_basis_and_qary_shape(lgs, oracle):
    rng = np.random.RandomState(5)
    B = rng.normal(size=(40, 40)) * 2 + 9 * np.eye(40)
    for basis, sigma in ((B, 14.0),):
        prob = oracle.Problem(basis, sigma, np.linspace(-2, 2, 40))
        S = 500
        U = oracle.uniform_stream(11, S * prob.ups)
        Xo, Vo, lqo = oracle.klein_replay(prob, U, S, exact_logz=False)
        smp = lgs.KleinSampler(lgs.SimpleLattice(basis), sigma, np.linspace(-2, 2, 40), mode="replay", uniforms=U)
        V, X, lq = smp.sample_with_coefficients(S)
        assert int(np.sum(np.any(X != Xo, axis=1))) == 0
        assert _rel(V, Vo) <= REL and _rel(lq, lqo) <= REL
    # q-ary shape (config C3 scaled down): R = B, sigma_i in {sigma/q, sigma}; short and long windows mixed
    from lattice_gaussian_mcmc_b200.lattices import qary_basis
    A = np.random.RandomState(0).randint(0, 97, size=(12, 12))
    Bq = qary_basis(A, 97)
    prob = oracle.Problem(Bq, 1.5 * 97)
    S = 300
    U = oracle.uniform_stream(12, S * prob.ups)
    Xo, Vo, _ = oracle.klein_replay(prob, U, S, want_logq=False)
    smp = lgs.KleinSampler(lgs.SimpleLattice(Bq), 1.5 * 97, mode="replay", uniforms=U)
    V, X, _ = smp.sample_with_coefficients(S)
    assert np.array_equal(X, Xo) and np.array_equal(V, Vo)


def test_imhk_replay_many_chains_matches_oracle(lgs, oracle):
    from lattice_gaussian_mcmc_b200.lattices import dd_lattice
    lat = dd_lattice(24, seed=7)
    sigma = 40.0
    center = np.random.RandomState(3).uniform(-5, 5, 24)
    prob = oracle.Problem(lat.basis, sigma, center)
    chains, burn, keep, thin = 96, 5, 4, 3
    steps = burn + keep * thin
    per = oracle.imhk_stream_len(prob, steps)
    U = oracle.uniform_stream(99, chains * per)
    ref = oracle.imhk_replay(prob, U, chains, steps, thin=1, exact_logz=False)
    im = lgs.IMHKSampler(lat, sigma, center, burn_in=0, n_chains=chains, mode="replay", uniforms=U)
    V = im.sample(n_samples=keep, burn_in=burn, thin=thin)
    assert V.shape == (chains, keep, 24)
    kept = ref["V"][:, burn + thin - 1::thin][:, :keep]
    assert np.array_equal(V, kept)
    assert np.array_equal(im._last_accept, ref["accept"])
    assert _rel(im._last_logw, ref["logw"]) <= REL
    assert im.total_proposals == chains * steps


def test_replay_edge_cases(lgs, oracle):
    lat = lgs.SimpleLattice(np.eye(3))
    smp = lgs.KleinSampler(lat, 2.0, mode="replay", uniforms=np.zeros(0))
    V = smp.sample(0)
    assert V.shape == (0, 3)
    with pytest.raises(ValueError):
        smp.sample(2)  # recorded stream exhausted
    # u = 0 and u just below 1: first / last table entry with mass
    prob = oracle.Problem(np.eye(3), 2.0)
    U = np.array([0.0, 0.0, 0.0, np.nextafter(1.0, 0), np.nextafter(1.0, 0), np.nextafter(1.0, 0), 0.5, 0.5, 0.5])
    Xo, _, _ = oracle.klein_replay(prob, U, 3)
    smp.set_uniform_stream(U)
    X = smp.sample_coefficients(3)
    assert np.array_equal(X, Xo)
    assert X[0].max() <= -15 and X[1].min() >= 15 and np.all(X[2] == 0)


def test_constructor_errors_match_reference(lgs):
    lat = lgs.SimpleLattice(np.eye(3))
    with pytest.raises(ValueError):
        lgs.KleinSampler(lat, 0.0)
    with pytest.raises(ValueError):
        lgs.KleinSampler(lat, -1.0)
    with pytest.raises(ValueError):
        lgs.KleinSampler(lat, 1.0, center=np.zeros(4))
    with pytest.raises(ValueError):
        lgs.IMHKSampler(lat, -2.0)
    smp = lgs.KleinSampler(lat, 1.0)
    assert smp.compute_log_density(np.array([0.5, 0.0, 0.0])) == -np.inf  # not a lattice point (klein.py:237-240)


@pytest.mark.parametrize("name", ["dd8", "dd24"])
def test_m6_spectral_gap_and_diagnose_convergence_match_the_reference(lgs, name):
    """M6 (imhk.py:252-313) in replay mode against values the reference itself produced after np.random.seed
    (tests/golden/m6_imhk_diag.npz, minted by make_golden_m6.py): same stream consumption order
    (k Klein samples; then chain init + num steps; then 100 Klein samples), same numbers."""
    import os
    g = np.load(os.path.join(os.path.dirname(__file__), "golden", "m6_imhk_diag.npz"))
    B, sigma = g[f"{name}_basis"], float(g[f"{name}_sigma"])
    k, num = int(g[f"{name}_k"]), int(g[f"{name}_num"])
    im = lgs.IMHKSampler(lgs.SimpleLattice(B), sigma, burn_in=5, mode="replay", uniforms=g[f"{name}_uniforms"])
    assert im.klein_log_partition == pytest.approx(float(g[f"{name}_klein_log_partition"]), rel=1e-13)
    gap = im.estimate_spectral_gap(k)
    assert gap == pytest.approx(float(g[f"{name}_gap"]), rel=1e-9)
    d = im.diagnose_convergence(num)
    assert d["acceptance_rate"] == float(g[f"{name}_acceptance_rate"])
    assert d["spectral_gap_estimate"] == pytest.approx(float(g[f"{name}_gap2"]), rel=1e-9)
    np.testing.assert_allclose(d["empirical_mean"], g[f"{name}_empirical_mean"], rtol=0, atol=1e-12)
    np.testing.assert_allclose(d["empirical_std"], g[f"{name}_empirical_std"], rtol=1e-13)
    np.testing.assert_array_equal(d["theoretical_std"], g[f"{name}_theoretical_std"])
    assert im.proposal_sampler._upos == g[f"{name}_uniforms"].size          # the whole recorded stream, no more


def test_small_api_surface_matches_the_reference(lgs):
    """B2 / B3 / K6 / K7 / M1 of SURVEY 8a -- the small methods around the hot path -- in replay mode against values the
    reference itself produced (tests/golden/surface_api.npz, minted by make_golden_surface.py)."""
    import json
    import os
    g = np.load(os.path.join(os.path.dirname(__file__), "golden", "surface_api.npz"))
    B, c, sigma = g["basis"], g["center"], float(g["sigma"])
    S, k = int(g["S"]), int(g["k"])
    lat = lgs.SimpleLattice(B, name="dd8-surface")
    smp = lgs.RefinedKleinSampler(lat, sigma, c, mode="replay", uniforms=g["uniforms"])
    V = smp.sample(S)
    assert np.array_equal(V, g["V"])
    X = smp.sample_coefficients(k)                       # the kernel's coefficients == the reference's lstsq + round
    assert X.shape == (k, 8) and np.array_equal(X, g["X"])
    for p, lw, w, ld in zip(g["points"], g["log_gaussian_weight"], g["gaussian_weight"], g["log_density"]):
        assert smp.log_gaussian_weight(p) == pytest.approx(lw, rel=1e-14)
        assert smp.gaussian_weight(p) == pytest.approx(w, rel=1e-12)
        assert smp.compute_log_density(p) == pytest.approx(ld, rel=1e-12)
    assert smp.compute_log_density(g["off_point"]) == -np.inf and float(g["log_density_off"]) == -np.inf
    np.testing.assert_allclose(smp.empirical_mean(V), g["empirical_mean"], rtol=0, atol=1e-13)
    np.testing.assert_allclose(smp.empirical_covariance(V), g["empirical_covariance"], rtol=1e-12, atol=1e-12)
    np.testing.assert_array_equal(smp.theoretical_covariance(), g["theoretical_covariance"])
    want = json.loads(str(g["diagnostic_info"]))
    got = smp.diagnostic_info()
    assert set(got) >= set(want) - {"cache_size"}
    for key, val in want.items():
        if key == "cache_size":
            continue                                     # the reference's table cache has no counterpart
        if isinstance(val, (str, bool)):
            assert got[key] == val
        else:
            assert got[key] == pytest.approx(val, rel=1e-12)
    st = smp.get_stats()
    assert json.loads(str(g["stats"])) == {"samples_generated": st.samples_generated, "acceptance_rate": st.acceptance_rate}
    assert repr(smp) == str(g["repr_klein"])
    smp.reset_stats()
    assert smp.get_stats().samples_generated == 0
    # default burn-in estimate (imhk.py:68-91) incl. its 10 000 cap, and klein_log_partition (93-100)
    for key, val in json.loads(str(g["burn_in"])).items():
        nn, ratio = key.split(":")
        im = lgs.IMHKSampler(lgs.SimpleLattice(np.eye(int(nn)) * 10.0), 10.0 * float(ratio))
        assert im.burn_in == val
    im = lgs.IMHKSampler(lat, sigma, c, burn_in=17, mode="replay", uniforms=g["imhk_uniforms"])
    assert im.klein_log_partition == pytest.approx(float(g["imhk_klein_log_partition"]), rel=1e-14)
    assert isinstance(repr(im), str)                     # the reference raises TypeError here before any sampling
    assert bool(g["repr_imhk_raises_before_sampling"])
    im.sample(3)
    assert repr(im) == str(g["repr_imhk"])
