"""Parity evidence on the BENCHMARKED kernel (klein_chain_kernel, csrc/lgs_chain.cu), through the C ABI.

1. Replay: a handle created with path="blocked" sends lgs_klein_replay / lgs_imhk_replay through the chain kernel's
   REPLAY instantiation -- the DMMA-contracted conditional means feed the reference's table SampleZ on the reference's
   recorded uniforms (klein.py:143-220, imhk.py:141-177).  Checked against every golden minted from the reference and
   against the oracle at BASELINE sizes (C3: S = 4096; C4: S = 4096 Klein, 256 chains x 20 IMHK steps).  The
   contraction sums the conditional mean in another order than klein.py:191-193, so a CDF comparison can flip with
   probability ~1e-15 per draw: mismatching SAMPLES are counted and printed, the bound is 1e-6 per draw, and for the
   samples that agree everything else must be exact (V = B x exactly for integer bases, log q within 1e-12).
2. Native mode: the conditional law of EVERY coordinate of the chain kernel's Philox output, tested against the
   reference's own tables at the recomputed conditional means (oracle.conditional_counts) -- 2^17 samples, all 1024
   coordinates of C4, chi-square per coordinate binned by frac(m_i), Bonferroni-corrected; and the same test must
   REJECT a sampler that is off by 1 % in sigma (its power is demonstrated, not assumed).
"""
import os
import time

import numpy as np
import pytest
from scipy import stats

from conftest import GOLDEN, golden_files

pytestmark = pytest.mark.gpu

REL = 1e-12
FLIP_BOUND = 1e-6   # mismatching draws per draw


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


@pytest.fixture(scope="module")
def ntru512():
    from lattice_gaussian_mcmc_b200.ntru import ntru_lattice
    return ntru_lattice(512, 12289, seed=0)


def _compare_klein(tag, X, V, lq, Xo, Vo, lqo, basis, ups):
    """mismatch count reported; samples that agree must agree in everything"""
    S = X.shape[0]
    bad = np.any(X != Xo, axis=1)
    mism = int(bad.sum())
    print(f"[chain replay] {tag}: {mism} of {S} samples differ ({S * ups} draws)")
    assert mism <= max(0, int(FLIP_BOUND * S * ups)), f"{tag}: {mism} mismatching samples"
    ok = ~bad
    integer = np.all(basis == np.rint(basis)) and np.abs(basis).max() < 2 ** 40
    if integer:
        assert np.array_equal(V, X.astype(np.float64) @ basis.T), f"{tag}: V != B x"
        assert np.array_equal(V[ok], Vo[ok])
    assert _rel(V[ok], Vo[ok]) <= REL
    if lq is not None and lqo is not None:
        assert _rel(lq[ok], lqo[ok]) <= REL
    return mism


# --------------------------------------------------------------------------------------------- goldens
@pytest.mark.parametrize("fname", golden_files("klein"))
def test_chain_kernel_replay_matches_reference_golden(lgs, fname):
    g = np.load(os.path.join(GOLDEN, fname))
    lat = lgs.SimpleLattice(g["basis"])
    try:
        smp = lgs.KleinSampler(lat, float(g["sigma"]), g["center"], mode="replay", uniforms=g["uniforms"],
                               path="blocked")
    except Exception as e:  # pragma: no cover
        raise AssertionError(f"{fname}: {e}")
    if not smp._engine.replay_on_chain:
        pytest.skip("the chain kernel does not serve this lattice (clamped sigma_i): the ordered kernel replays it")
    S = int(g["S"])
    V, X, lq = smp.sample_with_coefficients(S)
    mism = _compare_klein(fname, X, V, lq, g["X"], g["V"], g["logq"], g["basis"], smp._engine.ups)
    assert mism == 0, f"{fname}: {mism} of {S} samples differ from the reference"


@pytest.mark.parametrize("fname", golden_files("imhk"))
def test_chain_kernel_imhk_replay_matches_reference_golden(lgs, fname):
    g = np.load(os.path.join(GOLDEN, fname))
    lat = lgs.SimpleLattice(g["basis"])
    steps, thin = int(g["steps"]), int(g["thin"])
    im = lgs.IMHKSampler(lat, float(g["sigma"]), g["center"], burn_in=0, mode="replay", uniforms=g["uniforms"],
                         path="blocked")
    assert im._engine.replay_on_chain
    V = im.sample(steps // thin, thin=thin)
    assert V.shape == g["V_keep"].shape
    assert _rel(V, g["V_keep"]) <= REL
    assert np.array_equal(im._last_accept[0], g["accept"])
    assert _rel(im._last_logw[0], g["logw"]) <= REL
    assert im.total_proposals == steps and im.accepted_proposals == steps
    assert np.array_equal(im.current_state, V[-1])
    assert np.array_equal(im.current_coeffs, g["X_keep"][-1])


def test_chain_kernel_imhk_replay_stepwise_and_split_calls(lgs, oracle):
    """the stream position of every chain persists over launches of one call (keep > 1 => one launch per kept state)
    and restarts per call; exact weights use the table normalisers"""
    from lattice_gaussian_mcmc_b200.lattices import dd_lattice
    lat = dd_lattice(24, seed=7)
    center = np.random.RandomState(3).uniform(-5, 5, 24)
    chains, burn, keep, thin = 200, 5, 4, 3
    steps = burn + keep * thin
    for exact in (False, True):
        sigma = 18.0 if exact else 40.0    # sigma_i ~ 0.4: the normalisers vary with the means => real rejections
        prob = oracle.Problem(lat.basis, sigma, center)
        per = oracle.imhk_stream_len(prob, steps)
        U = oracle.uniform_stream(99, chains * per)
        ref = oracle.imhk_replay(prob, U, chains, steps, thin=1, exact_logz=exact)
        im = lgs.IMHKSampler(lat, sigma, center, burn_in=0, n_chains=chains, mode="replay", uniforms=U,
                             path="blocked", weight_mode="exact" if exact else "reference")
        V = im.sample(n_samples=keep, burn_in=burn, thin=thin)
        kept = ref["V"][:, burn + thin - 1::thin][:, :keep]
        if not exact:   # (the oracle's exact_logz flag only changes its log q; its weights stay the reference's)
            assert np.array_equal(V, kept)
            assert np.array_equal(im._last_accept, ref["accept"])
            assert _rel(im._last_logw, ref["logw"]) <= REL
        else:
            # exact weights on the recorded stream: the ordered kernel is the comparison (same arithmetic, same stream)
            io = lgs.IMHKSampler(lat, sigma, center, burn_in=0, n_chains=chains, mode="replay", uniforms=U,
                                 weight_mode="exact")
            Vo = io.sample(n_samples=keep, burn_in=burn, thin=thin)
            assert np.array_equal(V, Vo)
            assert np.array_equal(im._last_accept, io._last_accept)
            assert 0.1 < im._last_accept.mean() < 0.999, im._last_accept.mean()
            assert _rel(im._last_logw, io._last_logw) <= 1e-11
        assert im.total_proposals == chains * steps


# --------------------------------------------------------------------------------------------- C3 / C4 at size
def test_c3_qary_replay_through_the_chain_kernel(lgs, oracle):
    lat = lgs.lattices.qary_lattice(256, 512, 3329, seed=0)
    sigma = 1.5 * 3329
    prob = oracle.Problem(lat.basis, sigma)
    S = 4096
    U = oracle.uniform_stream(33, S * prob.ups)
    Xo, Vo, lqo = oracle.klein_replay(prob, U, S, exact_logz=False, threads=os.cpu_count())
    smp = lgs.KleinSampler(lat, sigma, mode="replay", uniforms=U, path="blocked")
    assert smp._engine.replay_on_chain
    t0 = time.perf_counter()
    V, X, lq = smp.sample_with_coefficients(S)
    dt = time.perf_counter() - t0
    print(f"[chain replay] C3 q-ary n=512: {S} samples in {dt * 1e3:.1f} ms (host call, first use)")
    _compare_klein("C3 q-ary n=512", X, V, lq, Xo, Vo, lqo, lat.basis, prob.ups)


def test_c4_ntru512_replay_through_the_chain_kernel(lgs, oracle, ntru512):
    sigma = 165.74
    prob = oracle.Problem(ntru512.basis, sigma)
    S = 4096
    U = oracle.uniform_stream(2026, S * prob.ups)
    Xo, Vo, lqo = oracle.klein_replay(prob, U, S, exact_logz=False, threads=os.cpu_count())
    smp = lgs.KleinSampler(ntru512, sigma, mode="replay", uniforms=U, path="blocked")
    assert smp._engine.replay_on_chain
    V, X, lq = smp.sample_with_coefficients(S)
    _compare_klein("C4 NTRU-512 Klein", X, V, lq, Xo, Vo, lqo, ntru512.basis, prob.ups)
    # IMHK: 256 chains x 20 steps, kept every 5th, on the recorded stream
    chains, steps, thin = 256, 20, 5
    per = oracle.imhk_stream_len(prob, steps)
    Ui = oracle.uniform_stream(7, chains * per)
    ref = oracle.imhk_replay(prob, Ui, chains, steps, thin=thin, exact_logz=False, threads=os.cpu_count())
    im = lgs.IMHKSampler(ntru512, sigma, burn_in=0, n_chains=chains, mode="replay", uniforms=Ui, path="blocked")
    Vi = im.sample(steps // thin, thin=thin)
    bad = np.any(Vi.reshape(chains, -1) != ref["V"].reshape(chains, -1), axis=1)
    print(f"[chain replay] C4 NTRU-512 IMHK: {int(bad.sum())} of {chains} chains differ "
          f"({chains * (steps + 1) * prob.ups} draws)")
    assert int(bad.sum()) <= max(0, int(FLIP_BOUND * chains * (steps + 1) * prob.ups))
    assert np.array_equal(Vi[~bad], ref["V"][~bad])
    assert np.array_equal(im._last_accept, ref["accept"])
    assert _rel(im._last_logw, ref["logw"]) <= REL
    assert im.total_proposals == chains * steps


# --------------------------------------------------------------------------------------------- native: conditional law
def _conditional_pvalues(obs, exp, min_exp=8.0):
    """chi-square per coordinate: cells are (frac bin, clipped offset); inside a frac bin observed and expected counts
    have the same total, so each bin with c usable cells contributes c - 1 degrees of freedom.  Cells with a small
    expectation are pooled inside their frac bin."""
    n, FB, W = obs.shape
    pv = np.ones(n)
    chi = np.zeros(n)
    dof = np.zeros(n, dtype=np.int64)
    for i in range(n):
        c_tot, d_tot = 0.0, 0
        for f in range(FB):
            o, e = obs[i, f].astype(np.float64), exp[i, f]
            if e.sum() < 2 * min_exp:
                continue
            keep = e >= min_exp
            o2, e2 = o[keep], e[keep]
            if (~keep).any():
                ro, re = o[~keep].sum(), e[~keep].sum()
                if re >= min_exp:
                    o2, e2 = np.append(o2, ro), np.append(e2, re)
                else:   # too little even pooled: fold into the smallest kept cell
                    j = int(np.argmin(e2))
                    o2[j] += ro
                    e2[j] += re
            if len(e2) < 2:
                continue
            c_tot += float(np.sum((o2 - e2) ** 2 / e2))
            d_tot += len(e2) - 1
        chi[i], dof[i] = c_tot, d_tot
        if d_tot > 0:
            pv[i] = stats.chi2.sf(c_tot, d_tot)
    return pv, chi, dof


def test_c4_native_conditional_law_of_every_coordinate(lgs, oracle, ntru512):
    """2^17 native samples of the chain kernel on the headline lattice: for all 1024 coordinates the offsets
    x_i - round(m_i), binned by frac(m_i), follow the reference's table law at the recomputed means."""
    sigma, n = 165.74, 1024
    S = 1 << 17
    smp = lgs.KleinSampler(ntru512, sigma, seed=20261104, path="blocked")
    X = smp.sample_coefficients(S)
    prob = oracle.Problem(ntru512.basis, sigma)
    t0 = time.perf_counter()
    obs, exp = oracle.conditional_counts(prob, X, K=6, FB=4, threads=os.cpu_count())
    assert obs.sum() == S * n and abs(exp.sum() - S * n) < 1e-6 * S * n
    pv, chi, dof = _conditional_pvalues(obs, exp)
    print(f"[conditional law] C4: {S} samples x {n} coordinates in {time.perf_counter() - t0:.1f} s; "
          f"min p = {pv.min():.3g} at coordinate {int(pv.argmin())}, dof per coordinate {dof.min()}..{dof.max()}, "
          f"sum chi2 / sum dof = {chi.sum() / dof.sum():.4f}")
    assert dof.min() >= 8
    alpha = 0.01
    assert pv.min() > alpha / n, f"coordinate {int(pv.argmin())}: p = {pv.min():.3g} (Bonferroni bound {alpha / n:.3g})"
    # the p-values themselves must look uniform (a small uniform bias on all coordinates would show here first)
    assert stats.kstest(pv, "uniform").pvalue > 1e-3
    # aggregate: chi2 summed over all coordinates against its own degrees of freedom
    assert stats.chi2.sf(chi.sum(), dof.sum()) > 1e-3 and stats.chi2.cdf(chi.sum(), dof.sum()) > 1e-6
    # power: the same statistic rejects a sampler whose sigma is 1 % too wide (tested against the law at sigma)
    wide = lgs.KleinSampler(ntru512, sigma * 1.01, seed=5, path="blocked")
    Xw = wide.sample_coefficients(S)
    obs_w, exp_w = oracle.conditional_counts(prob, Xw, K=6, FB=4, threads=os.cpu_count())
    pv_w, chi_w, dof_w = _conditional_pvalues(obs_w, exp_w)
    print(f"[conditional law] power check (sigma x 1.01): min p = {pv_w.min():.3g}, "
          f"{int((pv_w < alpha / n).sum())} coordinates rejected, aggregate p = "
          f"{stats.chi2.sf(chi_w.sum(), dof_w.sum()):.3g}")
    assert stats.chi2.sf(chi_w.sum(), dof_w.sum()) < 1e-12


@pytest.mark.parametrize("name", ["c3", "lll64", "dd256"])
def test_native_conditional_law_other_configs(lgs, oracle, name):
    """the same check on the q-ary lattice (C3: sigma_i in {1.5, 4993.5}; wide offsets binned coarsely), the LLL
    lattice of C2 and a mid-size lattice of the C5 sweep"""
    if name == "c3":
        lat, sigma, S, K, width = lgs.lattices.qary_lattice(256, 512, 3329, seed=0), 1.5 * 3329, 1 << 15, 5, None
    elif name == "lll64":
        lat = lgs.lattices.lll_lattice(64)
        sigma, S, K, width = 1.5 * lat.max_gram_schmidt_norm, 1 << 18, 6, None
    else:
        lat = lgs.lattices.dd_lattice(256)
        sigma, S, K, width = 1.5 * lat.max_gram_schmidt_norm, 1 << 16, 6, None
    smp = lgs.KleinSampler(lat, sigma, seed=77, path="blocked")
    X = smp.sample_coefficients(S)
    prob = oracle.Problem(lat.basis, sigma)
    if name == "c3":   # sigma_i = sigma coordinates: offsets spread over +-500 => bins of 100
        sig_i = sigma / np.abs(np.diag(prob.R))
        width = np.where(sig_i > 50, 100, 1).astype(np.int32)
    obs, exp = oracle.conditional_counts(prob, X, K=K, FB=4, threads=os.cpu_count(), width=width)
    pv, chi, dof = _conditional_pvalues(obs, exp)
    n = lat.dimension
    print(f"[conditional law] {name}: min p = {pv.min():.3g}, sum chi2 / sum dof = {chi.sum() / dof.sum():.4f}")
    assert pv.min() > 0.01 / n
    assert stats.chi2.sf(chi.sum(), dof.sum()) > 1e-3
