"""Wide truncated windows (sigma_i >= 50: the reference cuts its table to max_range + 1 = 1001 entries, klein.py:122-126).
The native sampler draws them from closed-form window masses (Euler-Maclaurin, lgs_device.cuh window_mass_em) instead of
an O(window) walk.  It is the same inverse transform u -> x: checked draw for draw against the summed form
(LGS_TRUNC_CLOSED_FORM=0) on the same Philox keys, and against the reference's table by chi-square."""
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


def _make(lgs, closed, *args, **kw):
    old = os.environ.get("LGS_TRUNC_CLOSED_FORM")
    os.environ["LGS_TRUNC_CLOSED_FORM"] = "1" if closed else "0"
    try:
        return lgs.KleinSampler(*args, **kw)
    finally:
        if old is None:
            del os.environ["LGS_TRUNC_CLOSED_FORM"]
        else:
            os.environ["LGS_TRUNC_CLOSED_FORM"] = old


@pytest.mark.parametrize("path", ["auto", "ordered"])
@pytest.mark.parametrize("sigma,center", [(50.0, 0.0), (80.0, 0.2), (151.0, -3.5), (4993.5, 0.49), (2.0e5, 17.25)])
def test_closed_form_draws_equal_the_summed_form_draw_for_draw(lgs, sigma, center, path):
    """1-D: every u maps to the same integer (a disagreement needs u within ~1e-14 of a bin edge)"""
    lat = lgs.SimpleLattice(np.array([[1.0]]))
    c = np.array([center])
    a = _make(lgs, True, lat, sigma, c, seed=7, path=path)
    b = _make(lgs, False, lat, sigma, c, seed=7, path=path)
    S = 400000
    Xa = a.sample_coefficients(S)
    Xb = b.sample_coefficients(S)
    assert np.array_equal(Xa, Xb)
    assert Xa.min() >= round(center) - 500 and Xa.max() <= round(center) + 500
    if sigma >= 1000:                  # an almost flat law over the 1001 entries: both ends are reached
        assert Xa.min() == round(center) - 500 and Xa.max() == round(center) + 500


def test_qary_lattice_mixed_regimes_and_log_density(lgs):
    """BASELINE config C3 shape (sigma_i in {1.5, 4993}): coefficients, lattice points and log q agree between the two
    forms on the chain kernel, and the exact-weight IMHK normalisers (window masses) agree to 1e-13"""
    lat = lgs.lattices.qary_lattice(64, 128, 3329, seed=1)
    sigma = 1.5 * 3329
    a = _make(lgs, True, lat, sigma, seed=5)
    b = _make(lgs, False, lat, sigma, seed=5)
    Va, Xa, qa = a.sample_with_coefficients(6000)
    Vb, Xb, qb = b.sample_with_coefficients(6000)
    assert np.array_equal(Xa, Xb) and np.array_equal(Va, Vb)
    np.testing.assert_allclose(qa, qb, rtol=1e-12)
    span = Xa.max(axis=0) - Xa.min(axis=0)
    # the truncated half (sampled first, centre 0): its 1001-entry window, end to end; the other half follows A x / q
    assert ((span >= 990) & (span <= 1000)).sum() == 64 and (span > 1000).sum() == 64


@pytest.mark.parametrize("sigma", [30.0, 60.0])
def test_wide_windows_at_precision_20_keep_the_summed_form_draws(lgs, sigma):
    """A handle made with precision = 20 (what adaptive_precision_sample passes) truncates from sigma_i = 25 on.  The
    chain kernel's per-coordinate tables hold their FP64 accuracy only while a * dmax = 500 / (2 sigma_i^2) <= 0.1
    (sigma_i >= 50): sigma_i = 30 must take the closed-form draw instead, sigma_i = 60 the tables; both give the summed
    form's integers on the same Philox keys."""
    n = 64
    lat = lgs.SimpleLattice(np.eye(n))
    c = np.linspace(-0.5, 0.5, n)
    a = _make(lgs, True, lat, sigma, c, seed=11, precision=20, path="blocked")
    b = _make(lgs, False, lat, sigma, c, seed=11, precision=20, path="blocked")
    Xa = a.sample_coefficients(8192)
    Xb = b.sample_coefficients(8192)
    assert np.array_equal(Xa, Xb)
    assert Xa.max() - Xa.min() > 4 * sigma and np.abs(Xa).max() <= 501


def test_truncated_window_chi_square_against_the_reference_table(lgs):
    from oracle import oracle
    sigma, center = 300.0, 0.3
    lat = lgs.SimpleLattice(np.array([[1.0]]))
    smp = _make(lgs, True, lat, sigma, np.array([center]), seed=11)
    S = 2_000_000
    x = smp.sample_coefficients(S)[:, 0]
    sup, lp = oracle.table(center, sigma)
    p = np.exp(lp)
    assert len(p) == 1001 and abs(p.sum() - 1) < 1e-9
    assert x.min() >= sup[0] and x.max() <= sup[-1]
    counts = np.bincount(x - sup[0], minlength=len(p))
    assert counts.sum() == S
    chi2 = ((counts - S * p) ** 2 / (S * p)).sum()
    dof = len(p) - 1
    assert chi2 < dof + 3.1 * np.sqrt(2 * dof), (chi2, dof)          # alpha ~ 0.001
