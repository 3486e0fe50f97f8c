"""Device-side chain diagnostics (lgs_chain_diagnostics / lgs_gelman_rubin_sums) against the reference's
own numbers (tests/golden/diag_ar1.npz) and the oracle restatement, through the C-ABI.

Tolerances: the kernel sums the lagged products directly, the reference goes through an FFT: acf agrees to 1e-11
absolute, tau_int to 1e-10 relative unless the Sokal window lands on a different lag (it cannot here: the margin
k - c * tau is checked), R-hat to 1e-12.
"""
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


def test_diagnostics_match_reference_golden(lgs):
    import torch
    from lattice_gaussian_mcmc_b200.diagnostics import ChainDiagnostics
    from lattice_gaussian_mcmc_b200.lattices import identity_lattice
    g = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "diag_ar1.npz"))
    X = g["X"]                                   # [4 correlation strengths, 12 chains, 400 steps]
    A, C, K = X.shape
    n = 6
    smp = lgs.KleinSampler(identity_lattice(n), 3.0, seed=1)
    # coordinate a of "lattice point" t of chain c carries series a; the other coordinates are noise
    V = np.random.RandomState(0).normal(size=(C, K, n))
    dims = [4, 0, 2, 5]
    for a in range(A):
        V[:, :, dims[a]] = X[a]
    Vd = torch.from_numpy(V).cuda()
    out = ChainDiagnostics(smp).run(Vd, dims, c_window=5.0, acf_len=K // 4 + 1)
    acf = out["acf"].cpu().numpy()
    tau = out["tau_int"].cpu().numpy()
    for a in range(A):
        assert np.max(np.abs(acf[:, a, :] - g["acf"][a])) < 1e-11
        assert np.max(np.abs(tau[:, a] - g["tau"][a]) / np.abs(g["tau"][a])) < 1e-10
        assert np.max(np.abs(out["ess"][:, a].cpu().numpy() - g["ess"][a]) / g["ess"][a]) < 1e-10
        assert np.allclose(out["mean"][:, a].cpu().numpy(), X[a].mean(axis=1), rtol=0, atol=1e-12)
        assert np.allclose(out["var"][:, a].cpu().numpy(), X[a].var(axis=1, ddof=1), rtol=1e-13)
        assert abs(float(out["r_hat"][a]) - g["rhat"][a]) < 1e-12
    assert np.array_equal(out["ess_min"].cpu().numpy(), out["ess"].cpu().numpy().min(axis=1))
    # without the ACF output the kernel stops at the Sokal window: same tau
    out2 = ChainDiagnostics(smp).run(Vd, dims)
    assert np.array_equal(out2["tau_int"].cpu().numpy(), tau)
    assert out2["acf"] is None


def test_diagnostics_on_imhk_output_match_oracle(lgs):
    """kept states straight from lgs_imhk_run on the device -> diagnostics, against the oracle on the copied states"""
    import torch
    from oracle import diag_oracle as do
    from lattice_gaussian_mcmc_b200.diagnostics import ChainDiagnostics
    from lattice_gaussian_mcmc_b200.lattices import dd_lattice
    n, chains, keep = 24, 64, 200
    lat = dd_lattice(n, seed=5)
    sigma = 0.3 * lat.max_gram_schmidt_norm           # exact weights reject here: correlated chains
    im = lgs.IMHKSampler(lat, sigma, burn_in=0, n_chains=chains, seed=3, weight_mode="exact", path="blocked")
    eng = im._engine
    V = torch.empty((chains, keep, n), dtype=torch.float64, device="cuda")
    st = torch.cuda.current_stream().cuda_stream
    eng.imhk_run_dev(20, keep, 1, "fast", "exact", V=V.data_ptr(), stream=st)
    dims = [0, n // 2, n - 1]
    out = ChainDiagnostics(im).run(V, dims, acf_len=11)
    Vh = V.cpu().numpy()
    tau = out["tau_int"].cpu().numpy()
    win = out["window"].cpu().numpy()
    for c in range(0, chains, 7):
        for j, d in enumerate(dims):
            t_ref, w_ref = do.integrated_autocorrelation_time(Vh[c, :, d])
            assert win[c, j] == w_ref
            assert abs(tau[c, j] - t_ref) < 1e-10 * abs(t_ref)
            assert np.max(np.abs(out["acf"][c, j].cpu().numpy() - do.compute_autocorrelation(Vh[c, :, d])[:11])) < 1e-11
    for j, d in enumerate(dims):
        r_ref = do.gelman_rubin_statistic([Vh[c, :, d] for c in range(chains)])
        assert abs(float(out["r_hat"][j]) - r_ref) < 1e-11
    assert 0.05 < im.acceptance_rate < 0.999


def test_diagnostics_argument_errors(lgs):
    import torch
    from lattice_gaussian_mcmc_b200.diagnostics import ChainDiagnostics
    from lattice_gaussian_mcmc_b200.lattices import identity_lattice
    smp = lgs.KleinSampler(identity_lattice(4), 2.0, seed=1)
    d = ChainDiagnostics(smp)
    V = torch.zeros((3, 8, 4), dtype=torch.float64, device="cuda")
    with pytest.raises(Exception):
        d.run(V, [4])                              # coordinate out of range
    with pytest.raises(ValueError):
        d.run(V.cpu(), [0])
    with pytest.raises(Exception):
        d.run(torch.zeros((3, 1, 4), dtype=torch.float64, device="cuda"), [0])   # a series needs two points


def test_device_tvd_matches_the_reference_golden(lgs):
    """D1 on device-resident samples: lgs_histogram with unit bins == the reference's exact-value counting
    (convergence_diag.py:15-72; tests/golden/tvd_d1.npz holds the reference's own outputs)."""
    import os
    import torch
    from lattice_gaussian_mcmc_b200 import diagnostics as D
    from lattice_gaussian_mcmc_b200.lattices import identity_lattice
    g = np.load(os.path.join(os.path.dirname(__file__), "golden", "tvd_d1.npz"))
    for name in g["names"]:
        a, b, want = g[f"{name}_a"], g[f"{name}_b"], float(g[f"{name}_tvd"][0])
        if a.ndim == 1:
            a, b = a[:, None], b[:, None]
        smp = lgs.KleinSampler(identity_lattice(a.shape[1]), 5.0, seed=1)
        A = torch.from_numpy(np.ascontiguousarray(a)).cuda()
        B = torch.from_numpy(np.ascontiguousarray(b)).cuda()
        assert D.device_tvd(smp, A, B) == pytest.approx(want, rel=0, abs=1e-14)
    # and on the sampler's own output: two independent native batches of Z^10 are close, a wider sigma is not
    lat = identity_lattice(10)
    s5, s7 = lgs.KleinSampler(lat, 5.0, seed=2), lgs.KleinSampler(lat, 7.0, seed=3)
    S = 40000
    V = [torch.empty((S, 10), dtype=torch.float64, device="cuda") for _ in range(3)]
    st = torch.cuda.current_stream().cuda_stream
    s5._engine.klein_sample_dev(S, 2, 0, "fast", V=V[0].data_ptr(), stream=st)
    s5._engine.klein_sample_dev(S, 2, S, "fast", V=V[1].data_ptr(), stream=st)
    s7._engine.klein_sample_dev(S, 3, 0, "fast", V=V[2].data_ptr(), stream=st)
    same, wider = D.device_tvd(s5, V[0], V[1]), D.device_tvd(s5, V[0], V[2])
    assert same == pytest.approx(D.compute_tvd(V[0].cpu().numpy(), V[1].cpu().numpy()), abs=1e-14)
    assert same < 0.03 < 0.1 < wider


def test_device_tvd_default_call_at_the_headline_dimension(lgs):
    """compute_tvd's default -- every coordinate, exact-value counting (convergence_diag.py:15-72) -- on device-resident
    NTRU-512 samples: 1024 marginals over a range of ~3000 integers (the shared-memory histogram alone
    could hold 13 of them).  Must equal the host function on the same arrays."""
    import torch
    from lattice_gaussian_mcmc_b200.diagnostics import compute_tvd, device_tvd
    from lattice_gaussian_mcmc_b200.ntru import ntru_lattice
    lat = ntru_lattice(512, 12289, seed=0)
    smp = lgs.KleinSampler(lat, 165.74, seed=1, path="blocked")
    S = 3000
    Va = torch.empty((S, 1024), dtype=torch.float64, device="cuda")
    Vb = torch.empty((S, 1024), dtype=torch.float64, device="cuda")
    st = torch.cuda.current_stream().cuda_stream
    smp._engine.klein_sample_dev(S, 1, 0, "fast", V=Va.data_ptr(), stream=st)
    smp._engine.klein_sample_dev(S, 2, 0, "fast", V=Vb.data_ptr(), stream=st)
    got = device_tvd(smp, Va, Vb)
    want = compute_tvd(Va.cpu().numpy(), Vb.cpu().numpy())
    assert abs(got - want) < 1e-12, (got, want)
    assert 0.1 < got < 1.0          # 3000 samples over ~1000 occupied values per marginal: mostly sampling noise
    # a subset of coordinates through the shared-memory variant, more than 256 through the chunked one
    dims = np.arange(0, 1024, 3)
    got2 = device_tvd(smp, Va, Vb, dims=dims)
    want2 = compute_tvd(Va.cpu().numpy()[:, dims], Vb.cpu().numpy()[:, dims])
    assert abs(got2 - want2) < 1e-12
    # histogram accumulation with few bins and many coordinates: 300 tracked dims x 16 bins
    from lattice_gaussian_mcmc_b200.distributed import host_histogram
    dims3 = np.arange(300, dtype=np.int32)
    hist = torch.zeros((300, 16), dtype=torch.int64, device="cuda")
    smp._engine.histogram_dev(Va.data_ptr(), S, dims3, -800.0, 100.0, 16, hist.data_ptr(), stream=st)
    assert torch.equal(hist.cpu(), host_histogram(Va.cpu(), dims3, -800.0, 100.0, 16))


def test_histogram_fused_into_the_sampling_kernel_equals_lgs_histogram(lgs):
    """lgs_attach_histogram: the lattice-point epilogue of the chain kernel counts the tracked coordinates of what it
    writes -- Klein batches (several tiles, a ragged last one), IMHK kept states (step-by-step and step-parallel
    launches) -- into the same bins as lgs_histogram run afterwards on the returned V."""
    import torch
    from lattice_gaussian_mcmc_b200.lattices import dd_lattice
    lat = dd_lattice(96, seed=3)
    sigma = 1.5 * lat.max_gram_schmidt_norm
    smp = lgs.KleinSampler(lat, sigma, seed=9, path="blocked")
    eng = smp._engine
    n = 96
    dims = np.array([0, 5, 63, 64, 95], dtype=np.int32)
    lo, width, nbins = -3.0 * sigma, sigma / 5, 32          # narrower than the support: both clamps are exercised
    st = torch.cuda.current_stream().cuda_stream

    def separate(V, S):
        h = torch.zeros((len(dims), nbins), dtype=torch.int64, device="cuda")
        eng.histogram_dev(V.data_ptr(), S, dims, lo, width, nbins, h.data_ptr(), stream=st)
        torch.cuda.synchronize()
        return h.cpu().numpy()

    fused = torch.zeros((len(dims), nbins), dtype=torch.int64, device="cuda")
    eng.attach_histogram(dims, lo, width, nbins, fused.data_ptr())
    S = 5 * 128 + 37
    V = torch.empty((S, n), dtype=torch.float64, device="cuda")
    eng.klein_sample_dev(S, 9, 0, "fast", V=V.data_ptr(), stream=st)
    torch.cuda.synchronize()
    want = separate(V, S)
    assert fused.sum().item() == S * len(dims)
    assert np.array_equal(fused.cpu().numpy(), want)
    # accumulation + IMHK kept states (3 kept states of 300 chains: the step-parallel route for so few chains)
    C, keep = 300, 3
    eng.imhk_reset(C, seed=4)
    Vk = torch.empty((C, keep, n), dtype=torch.float64, device="cuda")
    eng.imhk_run_dev(5, keep, 2, "fast", "exact", V=Vk.data_ptr(), stream=st)
    torch.cuda.synchronize()
    want = want + separate(Vk, C * keep)
    assert np.array_equal(fused.cpu().numpy(), want)
    old = os.environ.get("LGS_IMHK_STEPPAR")
    os.environ["LGS_IMHK_STEPPAR"] = "0"                       # ... and the step-by-step launches
    try:
        eng.imhk_run_dev(0, keep, 2, "fast", "exact", V=Vk.data_ptr(), stream=st)
        torch.cuda.synchronize()
    finally:
        if old is None:
            del os.environ["LGS_IMHK_STEPPAR"]
        else:
            os.environ["LGS_IMHK_STEPPAR"] = old
    want = want + separate(Vk, C * keep)
    assert np.array_equal(fused.cpu().numpy(), want)
    # polling the chains (what IMHKSampler.step / current_state / get_state do) and the host routes of Klein batches
    # are not new device-resident samples: the histogram stays as it is
    Vs = torch.empty((C, n), dtype=torch.float64, device="cuda")
    cnt = torch.zeros(2 * C, dtype=torch.int64, device="cuda")
    L, hnd = eng._lib, eng._h
    for _ in range(2):
        assert L.lgs_imhk_get_state(hnd, cnt.data_ptr(), None, Vs.data_ptr(), None, st) == 0
    eng.imhk_state()
    Vh = np.empty((S, n))
    assert L.lgs_klein_sample_host(hnd, S, 9, 0, 1, Vh.ctypes.data, None, None) == 0
    Xh = np.empty((S, n), dtype=np.int64)
    assert L.lgs_klein_sample_host(hnd, S, 9, 0, 1, Vh.ctypes.data, Xh.ctypes.data, None) == 0   # fp64 staging route
    torch.cuda.synchronize()
    assert np.array_equal(fused.cpu().numpy(), want)
    # detached: nothing is counted any more
    eng.attach_histogram(None, 0.0, 1.0, 1, 0)
    eng.klein_sample_dev(S, 9, S, "fast", V=V.data_ptr(), stream=st)
    torch.cuda.synchronize()
    assert np.array_equal(fused.cpu().numpy(), want)
