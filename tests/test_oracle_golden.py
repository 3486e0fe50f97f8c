"""The CPU oracle (oracle/lgs_oracle.c) against golden vectors minted by the reference itself
(tests/golden/make_golden.py).  This is what pins the oracle: integer coefficients bit-exact, lattice points
exact, Klein log-densities / IMHK log-weights within 1e-12 relative, accept bits equal."""
import os

import numpy as np
import pytest

from conftest import GOLDEN, golden_files
from oracle import oracle

REL = 1e-12


def _rel(a, b):
    a, b = np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64)
    return np.max(np.abs(a - b) / np.maximum(np.abs(b), 1e-300)) if a.size else 0.0


def test_uniform_stream_is_numpy_legacy_mt19937():
    for seed in (0, 1, 123, 2**31 + 5):
        got = oracle.uniform_stream(seed, 2000)
        want = np.random.RandomState(seed).random_sample(2000)
        assert np.array_equal(got, want)


def test_pairwise_sum_matches_numpy_bitwise():
    rng = np.random.RandomState(3)
    for n in list(range(0, 40)) + [127, 128, 129, 255, 256, 257, 500, 1000, 1001]:
        a = rng.random_sample(n) * np.exp(rng.uniform(-30, 0, n))
        assert oracle.np_sum(a) == float(np.sum(a)), n


@pytest.mark.parametrize("fname", golden_files("klein"))
def test_klein_replay_matches_reference(fname):
    g = np.load(os.path.join(GOLDEN, fname))
    prob = oracle.Problem(g["basis"], float(g["sigma"]), g["center"])
    # host precompute is the same LAPACK call the reference makes
    assert _rel(prob.R, g["R"]) <= REL and _rel(prob.t, g["t"]) <= REL
    S = int(g["S"])
    assert prob.ups == g["draw_x"].shape[1]
    # the recorded stream is RandomState(seed).random_sample
    U = oracle.uniform_stream(int(g["seed"]), S * prob.ups)
    assert np.array_equal(U, g["uniforms"])
    X, V, logq = oracle.klein_replay(prob, U, S)
    mism = int(np.sum(np.any(X != g["X"], axis=1)))
    assert mism == 0, f"{mism} of {S} samples differ from the reference"
    if np.all(g["basis"] == np.rint(g["basis"])) and np.abs(g["basis"]).max() < 2**40:
        assert np.array_equal(V, g["V"])
    assert _rel(V, g["V"]) <= REL
    assert _rel(logq, g["logq"]) <= REL


@pytest.mark.parametrize("fname", golden_files("imhk"))
def test_imhk_replay_matches_reference(fname):
    g = np.load(os.path.join(GOLDEN, fname))
    prob = oracle.Problem(g["basis"], float(g["sigma"]), g["center"])
    steps, thin = int(g["steps"]), int(g["thin"])
    U = oracle.uniform_stream(int(g["seed"]), oracle.imhk_stream_len(prob, steps))
    assert np.array_equal(U, g["uniforms"])
    out = oracle.imhk_replay(prob, U, chains=1, steps=steps, thin=thin)
    assert np.array_equal(out["X"][0], g["X_keep"])
    assert _rel(out["V"][0], g["V_keep"]) <= REL
    assert np.array_equal(out["accept"][0], g["accept"])
    assert _rel(out["logw"][0], g["logw"]) <= REL
    # reference quirk: the weight is the constant klein_log_partition
    assert np.max(np.abs(out["logw"][0] - float(g["klein_log_partition"]))) < 1e-10
    assert int(out["accepted"][0]) == steps


def test_seeded_equals_replay_of_own_stream():
    g = np.load(os.path.join(GOLDEN, "klein_dd8.npz"))
    prob = oracle.Problem(g["basis"], float(g["sigma"]), g["center"])
    X1, V1 = oracle.klein_seeded(prob, int(g["S"]), int(g["seed"]))
    assert np.array_equal(X1, g["X"]) and np.array_equal(V1, g["V"])


def test_samplez_draws_match_reference_draw_log():
    """every recorded 1-D draw (mean, sigma_i, u) -> x_i, in isolation"""
    for fname in ("klein_real16.npz", "klein_trunc.npz", "klein_half.npz", "klein_tiny.npz"):
        g = np.load(os.path.join(GOLDEN, fname))
        S, per = g["draw_x"].shape
        U = g["uniforms"].reshape(S, per)
        bad = 0
        for s in range(S):
            for k in range(per):
                x, used = oracle.samplez(float(g["draw_mean"][s, k]), float(g["draw_sigma"][s, k]), float(U[s, k]))
                bad += (x != int(g["draw_x"][s, k])) or not used
        assert bad == 0, fname


def test_diagnostics_oracle_reproduces_reference_golden():
    """oracle/diag_oracle.py against outputs of the reference's own convergence_diag / mcmc_diag functions
    (tests/golden/diag_ar1.npz, minted by tests/golden/make_golden_diag.py)"""
    from oracle import diag_oracle as do
    g = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "diag_ar1.npz"))
    X = g["X"]
    for a in range(X.shape[0]):
        for c in range(X.shape[1]):
            acf = do.compute_autocorrelation(X[a, c])
            assert np.max(np.abs(acf - g["acf"][a, c])) < 1e-12
            tau, _ = do.integrated_autocorrelation_time(X[a, c])
            assert abs(tau - g["tau"][a, c]) < 1e-11 * max(1.0, abs(g["tau"][a, c]))
            assert abs(do.effective_sample_size(X[a, c]) - g["ess"][a, c]) < 1e-9 * g["ess"][a, c]
        assert abs(do.gelman_rubin_statistic(list(X[a])) - g["rhat"][a]) < 1e-13
    # multivariate ESS = minimum over coordinates (mcmc_diag.py:98-105)
    xm = np.stack([X[0, 0], X[3, 0]], axis=1)
    assert do.effective_sample_size(xm) == min(do.effective_sample_size(xm[:, 0]), do.effective_sample_size(xm[:, 1]))


def test_tvd_and_acceptance_rate_match_the_reference_golden(golden_dir):
    """D1: compute_tvd / compute_acceptance_rate -- the oracle restatement and the product's host drop-in
    against outputs of the reference's own functions (tests/golden/tvd_d1.npz, minted by make_golden_tvd.py)."""
    import os
    from oracle import diag_oracle
    from lattice_gaussian_mcmc_b200 import diagnostics as D
    g = np.load(os.path.join(golden_dir, "tvd_d1.npz"))
    for name in g["names"]:
        a, b, want = g[f"{name}_a"], g[f"{name}_b"], g[f"{name}_tvd"]
        for bins, w in zip((None, 16, 50), want):
            assert diag_oracle.compute_tvd(a, b, bins) == pytest.approx(w, rel=0, abs=1e-15)
            assert D.compute_tvd(a, b, bins) == pytest.approx(w, rel=0, abs=1e-14)
    for acc, w in zip(g["accepted"], g["acceptance_rate"]):
        assert diag_oracle.compute_acceptance_rate(acc) == w and D.compute_acceptance_rate(acc) == w
    assert D.compute_acceptance_rate(g["accepted"]) == float(g["acceptance_rate_all"])
