"""lgs_chain_diagnostics_static: chain diagnostics for hosts that may not allocate device memory.

Why the entry point exists.  lgs_chain_diagnostics cudaMalloc's a work array (the tracked series, gathered per chain)
the first time it runs and again whenever a larger request arrives.  Three kinds of user cannot live with that: a
process in which a framework has already claimed the whole GPU for its own pool (one more cudaMalloc fails, or takes
memory somebody else was promised), a monitor that runs next to live sampling (cudaMalloc and cudaFree stall every
stream of the device), and code in a phase where the runtime allows no allocation at all.  They pass their own input
and output buffers to lgs_chain_diagnostics_static, whose work array is a fixed 32 MiB array inside the library.

What is checked here, through the C-ABI (ctypes on liblgs_b200.so, no Python wrapper in between):
  * two calls in a row on the static entry point -- different shapes, different data, the second smaller than the
    first so that it runs over the first call's leftovers -- each give what lgs_chain_diagnostics gives on the same
    input.  Both entry points run the same kernels, so the numbers should be identical; the asserts use the
    tolerances of tests/test_gpu_diagnostics.py (acf 1e-11 absolute, the statistics 1e-10 relative);
  * the second call leaves the device's free memory where it was (nothing was allocated or freed);
  * a request beyond LGS_DIAG_STATIC_MAX_DOUBLES is refused with LGS_E_INVALID and a message that names the limit,
    and the call after the refusal still works.

This file is named so that pytest collects it last: with `pytest -x` a failure here hides no other test.  It does not
depend on anything the other tests leave behind (nor on a fresh process): the shared array is overwritten, never
assumed to be zero.
"""
import ctypes
import re

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

ACF_ATOL = 1e-11      # tests/test_gpu_diagnostics.py
STAT_RTOL = 1e-10


@pytest.fixture(scope="module")
def lgs():
    import __graft_entry__ as g
    g.build()
    import lattice_gaussian_mcmc_b200 as m
    return m


def _header_limit():
    import os
    from conftest import ROOT
    txt = open(os.path.join(ROOT, "include", "lgs_b200.h")).read()
    return int(re.search(r"#define\s+LGS_DIAG_STATIC_MAX_DOUBLES\s+(\d+)", txt).group(1))


def _ar1(rng, chains, keep, n, phi):
    """correlated series in every coordinate (a constant series would give nan, as in the reference)"""
    e = rng.normal(size=(chains, keep, n))
    x = np.empty_like(e)
    x[:, 0] = e[:, 0]
    for t in range(1, keep):
        x[:, t] = phi * x[:, t - 1] + e[:, t]
    return x


def _err(L, h):
    msg = L.lgs_last_error(h)
    return msg.decode() if msg else ""


def _call(fn, h, V, dims, stats, acf, acf_len, stream):
    chains, keep, _ = V.shape
    d = np.ascontiguousarray(dims, dtype=np.int32)
    return fn(h, V.data_ptr(), chains, keep, d.ctypes.data_as(ctypes.POINTER(ctypes.c_int32)), d.size, 5.0, 0,
              stats.data_ptr(), acf.data_ptr() if acf is not None else None, acf_len, stream)


def test_static_entry_point_matches_the_allocating_one_on_two_calls_in_a_row(lgs):
    import torch
    from lattice_gaussian_mcmc_b200 import _capi
    from lattice_gaussian_mcmc_b200.lattices import identity_lattice
    L = _capi.lib()
    n = 6
    smp = lgs.KleinSampler(identity_lattice(n), 3.0, seed=1)
    h = smp._engine._h
    rng = np.random.RandomState(20261106)
    # call A: 24 chains x 400 kept states x 4 tracked coordinates; call B: smaller in every extent, other data
    VA = torch.from_numpy(_ar1(rng, 24, 400, n, 0.7)).cuda()
    VB = torch.from_numpy(_ar1(rng, 5, 64, n, 0.3)).cuda()
    dimsA, dimsB = [4, 0, 2, 5], [1, 3]
    lenA, lenB = 400 // 4 + 1, 64 // 4 + 1
    F = 5                                                          # LGS_DIAG_FIELDS

    def outputs(V, dims, acf_len):
        return (torch.full((V.shape[0], len(dims), F), float("nan"), dtype=torch.float64, device="cuda"),
                torch.full((V.shape[0], len(dims), acf_len), float("nan"), dtype=torch.float64, device="cuda"))
    sA, aA = outputs(VA, dimsA, lenA)
    sB, aB = outputs(VB, dimsB, lenB)
    rA, raA = outputs(VA, dimsA, lenA)
    rB, raB = outputs(VB, dimsB, lenB)
    st = torch.cuda.current_stream().cuda_stream
    torch.cuda.synchronize()

    # the two static calls back to back on one stream (ordered, so they may share the array)
    assert _call(L.lgs_chain_diagnostics_static, h, VA, dimsA, sA, aA, lenA, st) == 0, _err(L, h)
    torch.cuda.synchronize()
    free_before = torch.cuda.mem_get_info()[0]
    assert _call(L.lgs_chain_diagnostics_static, h, VB, dimsB, sB, aB, lenB, st) == 0, _err(L, h)
    torch.cuda.synchronize()
    assert torch.cuda.mem_get_info()[0] == free_before, "the static entry point changed the device's free memory"

    # the existing entry point on the same inputs
    assert _call(L.lgs_chain_diagnostics, h, VA, dimsA, rA, raA, lenA, st) == 0, _err(L, h)
    assert _call(L.lgs_chain_diagnostics, h, VB, dimsB, rB, raB, lenB, st) == 0, _err(L, h)
    torch.cuda.synchronize()

    for got, want, gacf, wacf in ((sA, rA, aA, raA), (sB, rB, aB, raB)):
        got, want, gacf, wacf = (t.cpu().numpy() for t in (got, want, gacf, wacf))
        assert np.all(np.isfinite(want)) and np.all(np.isfinite(wacf))
        assert np.max(np.abs(gacf - wacf)) <= ACF_ATOL
        assert np.max(np.abs(got - want) / np.maximum(np.abs(want), 1e-300)) <= STAT_RTOL
        assert np.array_equal(got[..., 3], want[..., 3])          # the Sokal window is an integer: same lag
    # and without the ACF output (the series kernel stops at the window): same statistics again
    s2, _ = outputs(VA, dimsA, 1)
    assert _call(L.lgs_chain_diagnostics_static, h, VA, dimsA, s2, None, 0, st) == 0, _err(L, h)
    torch.cuda.synchronize()
    assert np.max(np.abs(s2.cpu().numpy() - rA.cpu().numpy()) / np.abs(rA.cpu().numpy())) <= STAT_RTOL


def test_static_entry_point_refuses_what_does_not_fit(lgs):
    import torch
    from lattice_gaussian_mcmc_b200 import _capi
    from lattice_gaussian_mcmc_b200.lattices import identity_lattice
    L = _capi.lib()
    cap = _header_limit()
    assert cap == 4 * 1024 * 1024
    n = 4
    smp = lgs.KleinSampler(identity_lattice(n), 2.0, seed=1)
    h = smp._engine._h
    V = torch.from_numpy(_ar1(np.random.RandomState(3), 3, 16, n, 0.5)).cuda()
    stats = torch.zeros((3, 2, 5), dtype=torch.float64, device="cuda")
    d = np.array([0, 3], dtype=np.int32)
    dp = d.ctypes.data_as(ctypes.POINTER(ctypes.c_int32))
    st = torch.cuda.current_stream().cuda_stream
    fn = L.lgs_chain_diagnostics_static
    # one value too many (the sizes are checked before anything is read, so V need not be that large), and sizes whose
    # product does not fit 64 bits
    for chains, keep in ((cap // (2 * 16) + 1, 16), (2 ** 62, 2 ** 40), (1, cap)):
        rc = fn(h, V.data_ptr(), chains, keep, dp, 2, 5.0, 0, stats.data_ptr(), None, 0, st)
        assert rc == _capi.LGS_E_INVALID
        msg = _err(L, h)
        assert "LGS_DIAG_STATIC_MAX_DOUBLES" in msg and str(cap) in msg
    # exactly at the limit the size check passes (this request then fails a later check instead: keep / 4 lags < acf_len)
    rc = fn(h, V.data_ptr(), cap // (2 * 16), 16, dp, 2, 5.0, 0, stats.data_ptr(), stats.data_ptr(), 16, st)
    assert rc == _capi.LGS_E_INVALID and "acf_len" in _err(L, h)
    # the same argument errors as lgs_chain_diagnostics
    bad = np.array([4], dtype=np.int32)
    assert fn(h, V.data_ptr(), 3, 16, bad.ctypes.data_as(ctypes.POINTER(ctypes.c_int32)), 1, 5.0, 0, stats.data_ptr(),
              None, 0, st) == _capi.LGS_E_INVALID
    assert fn(h, V.data_ptr(), 3, 1, dp, 2, 5.0, 0, stats.data_ptr(), None, 0, st) == _capi.LGS_E_INVALID
    # a refused call leaves the entry point usable
    assert fn(h, V.data_ptr(), 3, 16, dp, 2, 5.0, 0, stats.data_ptr(), None, 0, st) == 0, _err(L, h)
    ref = torch.zeros_like(stats)
    assert L.lgs_chain_diagnostics(h, V.data_ptr(), 3, 16, dp, 2, 5.0, 0, ref.data_ptr(), None, 0, st) == 0
    torch.cuda.synchronize()
    assert np.max(np.abs(stats.cpu().numpy() - ref.cpu().numpy()) / np.abs(ref.cpu().numpy())) <= STAT_RTOL
