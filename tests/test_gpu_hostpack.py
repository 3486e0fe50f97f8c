"""Packed device-to-host transport of lgs_klein_sample_host (integer bases: lattice points cross PCIe as int16 and are
widened by the host pool).  Routes: the chain kernel packing int16 in its epilogue, finished groups of tiles shipped by the copy engine behind
cuStreamWaitValue32 while the kernel runs (n % 8 == 0; calls of several batches, LGS_HOST_ZEROCOPY=3) or stored by the kernel into mapped
pinned memory (single batch, LGS_HOST_ZEROCOPY=2), and chunked pack-kernel + copy (everything else, LGS_HOST_ZEROCOPY=0).  It is transport
only: the float64 array handed back must be bit-identical to the plain fp64 copy (LGS_HOST_PACK=0), including batches
that fall back to fp64 because a value leaves int16."""
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


def _with_pack(flag, make):
    """LGS_HOST_PACK is read when a handle is created"""
    old = os.environ.get("LGS_HOST_PACK")
    os.environ["LGS_HOST_PACK"] = str(flag)
    try:
        return make()
    finally:
        if old is None:
            del os.environ["LGS_HOST_PACK"]
        else:
            os.environ["LGS_HOST_PACK"] = old


@pytest.mark.parametrize("n,S", [(10, 10000), (64, 4097), (200, 70001), (1024, 3 * 18944 + 77)])
def test_packed_transport_is_bit_identical_to_the_fp64_copy(lgs, n, S):
    """several chunks, a ragged last chunk, n not a multiple of 8 (scalar tail of the pack kernel and of the widening)"""
    from lattice_gaussian_mcmc_b200.lattices import dd_lattice
    lat = dd_lattice(n, seed=300 + n)
    sigma = 1.5 * lat.max_gram_schmidt_norm
    a = _with_pack(1, lambda: lgs.KleinSampler(lat, sigma, seed=21))
    b = _with_pack(0, lambda: lgs.KleinSampler(lat, sigma, seed=21))
    Va = a._engine.klein_sample(S, 21, 5, "fast")[0]
    Vb = b._engine.klein_sample(S, 21, 5, "fast")[0]
    assert Va.dtype == np.float64 and Va.shape == (S, n)
    assert np.array_equal(Va, Vb)
    for route in ("0", "2", "3"):     # read per call: chunked pack-kernel copies / mapped-memory stores / copy engine
        os.environ["LGS_HOST_ZEROCOPY"] = route
        try:
            assert np.array_equal(a._engine.klein_sample(S, 21, 5, "fast")[0], Vb)
        finally:
            del os.environ["LGS_HOST_ZEROCOPY"]
    # a second call reuses the mapped buffers with a new epoch
    assert np.array_equal(a._engine.klein_sample(S // 2 + 1, 22, 0, "fast")[0], b._engine.klein_sample(S // 2 + 1, 22, 0, "fast")[0])
    assert np.array_equal(Va, np.rint(Va)) and np.abs(Va).max() < 32768
    # with coefficients requested the call takes the unpacked route: same points again, V = B x
    Vc, Xc, _ = a.sample_with_coefficients(257)
    assert np.array_equal(Vc, Xc.astype(np.float64) @ lat.basis.T)


def test_zero_copy_over_several_full_batches(lgs):
    """S larger than two chain-kernel batches: double-buffered mapped memory, the next launch queued while the pool
    drains the previous batch, ragged last batch"""
    from lattice_gaussian_mcmc_b200.lattices import dd_lattice
    lat = dd_lattice(64, seed=364)
    sigma = 1.5 * lat.max_gram_schmidt_norm
    a = _with_pack(1, lambda: lgs.KleinSampler(lat, sigma, seed=8))
    b = _with_pack(0, lambda: lgs.KleinSampler(lat, sigma, seed=8))
    S = 2 * a._engine.preferred_batch + 12345
    Va = a._engine.klein_sample(S, 8, 0, "fast")[0]
    assert np.array_equal(Va, b._engine.klein_sample(S, 8, 0, "fast")[0])
    # split invariance through the host route: the same ids in two calls
    Vh = a._engine.klein_sample(1000, 8, S - 1000, "fast")[0]
    assert np.array_equal(Vh, Va[-1000:])


def test_values_outside_int16_fall_back_to_the_fp64_chunk(lgs):
    from lattice_gaussian_mcmc_b200.lattices import identity_lattice
    big = lgs.SimpleLattice(1000.0 * np.eye(16))
    a = _with_pack(1, lambda: lgs.KleinSampler(big, 30000.0, seed=3))
    b = _with_pack(0, lambda: lgs.KleinSampler(big, 30000.0, seed=3))
    lat = identity_lattice(16)
    c = np.full(16, 32760.5)      # straddles the int16 bound: a few values per chunk leave the range
    a2 = _with_pack(1, lambda: lgs.KleinSampler(lat, 5.0, c, seed=3))
    b2 = _with_pack(0, lambda: lgs.KleinSampler(lat, 5.0, c, seed=3))
    for x, y in ((a, b), (a2, b2)):
        Vx = x._engine.klein_sample(200000, 3, 0, "fast")[0]
        Vy = y._engine.klein_sample(200000, 3, 0, "fast")[0]
        assert np.abs(Vy).max() > 32767
        assert np.array_equal(Vx, Vy)


def test_real_valued_basis_is_never_packed(lgs):
    rng = np.random.RandomState(9)
    B = np.eye(12) * 7 + rng.uniform(-1, 1, (12, 12))
    lat = lgs.SimpleLattice(B)
    a = _with_pack(1, lambda: lgs.KleinSampler(lat, 30.0, seed=4))
    b = _with_pack(0, lambda: lgs.KleinSampler(lat, 30.0, seed=4))
    Va = a._engine.klein_sample(50000, 4, 0, "fast")[0]
    assert not np.array_equal(Va, np.rint(Va))
    assert np.array_equal(Va, b._engine.klein_sample(50000, 4, 0, "fast")[0])


def test_imhk_kept_states_travel_packed_and_match_the_fp64_copy(lgs):
    """lgs_imhk_run_host: kept states [chains, keep, n] of an integer basis over the int16 transport (several chunks),
    and a centre that pushes some values out of int16 (those chunks fall back to fp64)"""
    from lattice_gaussian_mcmc_b200.lattices import dd_lattice, identity_lattice
    lat = dd_lattice(64, seed=77)
    sigma = 1.5 * lat.max_gram_schmidt_norm
    cases = [(lat, sigma, None, 4096, 500), (identity_lattice(16), 5.0, np.full(16, 32760.5), 8192, 48)]
    for L, sg, c, chains, keep in cases:
        mk = lambda: lgs.IMHKSampler(L, sg, c, burn_in=0, n_chains=chains, seed=13)   # noqa: E731
        a = _with_pack(1, mk)
        b = _with_pack(0, mk)
        Va = a.sample(keep, thin=1)
        Vb = b.sample(keep, thin=1)
        assert Va.shape == Vb.shape and Va.dtype == np.float64 and Va.size >= 1 << 22
        assert np.array_equal(Va, Vb)
        assert a.acceptance_rate == b.acceptance_rate


def test_pageable_destinations_take_the_pinned_bounce_ring_and_match_the_device_result(lgs):
    """fp64 points, int64 coefficients and log q of several chunks delivered to pageable numpy arrays (staged through
    the pinned ring, pieces of 32 MB, ragged last piece) == the same call with page-locked destinations (direct
    cudaMemcpyAsync) == the device-resident result of the same seed and sample ids."""
    import torch
    from lattice_gaussian_mcmc_b200 import _capi
    from lattice_gaussian_mcmc_b200.lattices import dd_lattice
    n, S = 200, 2 * 65536 + 4321            # 211 MB of points: several pieces per chunk
    lat = dd_lattice(n, seed=91)
    smp = _with_pack(0, lambda: lgs.KleinSampler(lat, 1.5 * lat.max_gram_schmidt_norm, seed=8))
    eng = smp._engine
    V, X, lq = eng.klein_sample(S, 8, 3, "fast", want_v=True, want_x=True, want_logq=True)
    Vd = torch.empty((S, n), dtype=torch.float64, device="cuda")
    Xd = torch.empty((S, n), dtype=torch.int64, device="cuda")
    qd = torch.empty(S, dtype=torch.float64, device="cuda")
    eng.klein_sample_dev(S, 8, 3, "fast", V=Vd.data_ptr(), X=Xd.data_ptr(), logq=qd.data_ptr(),
                         stream=torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    assert np.array_equal(V, Vd.cpu().numpy()) and np.array_equal(X, Xd.cpu().numpy())
    assert np.array_equal(lq, qd.cpu().numpy())
    Vp = torch.empty((S, n), dtype=torch.float64).pin_memory()
    Xp = torch.empty((S, n), dtype=torch.int64).pin_memory()
    _capi.check(eng._h, _capi.lib().lgs_klein_sample_host(eng._h, S, 8, 3, _capi.SAMPLEZ["fast"],
                                                          Vp.data_ptr(), Xp.data_ptr(), None))
    assert np.array_equal(V, Vp.numpy()) and np.array_equal(X, Xp.numpy())
    # IMHK kept states and coefficients of a real-valued basis (never packed): pageable, above the staging threshold
    rng = np.random.RandomState(2)
    B = np.eye(32) * 9 + rng.uniform(-1, 1, (32, 32))
    im = lgs.IMHKSampler(lgs.SimpleLattice(B), 40.0, burn_in=0, n_chains=4096, seed=5)
    im._engine.imhk_reset(4096, seed=5)
    Vk, Xk = im._engine.imhk_run(0, 80, 1, "fast", "reference", want_v=True, want_x=True)
    assert Vk.nbytes >= 8 << 20 and Xk.nbytes >= 8 << 20
    np.testing.assert_allclose(Vk, Xk.astype(np.float64) @ B.T, rtol=0, atol=1e-9)
    assert len(np.unique(Xk[:, -1, 0])) > 3


def test_sample_into_a_caller_owned_array(lgs):
    """KleinSampler.sample(num, out=...) fills the caller's array with exactly what a fresh call returns"""
    from lattice_gaussian_mcmc_b200.lattices import dd_lattice
    lat = dd_lattice(72, seed=5)
    sigma = 1.5 * lat.max_gram_schmidt_norm
    a = lgs.KleinSampler(lat, sigma, seed=3)
    b = lgs.KleinSampler(lat, sigma, seed=3)
    buf = np.full((5000, 72), np.nan)
    for _ in range(2):                       # the second round continues the sample ids in both
        got = a.sample(5000, out=buf)
        assert got is buf
        assert np.array_equal(buf, b.sample(5000))
    for bad in (np.empty((5000, 71)), np.empty((5000, 72), np.float32), np.empty((72, 5000)).T):
        with pytest.raises(ValueError):
            a.sample(5000, out=bad)
