"""Host-side setup (numpy): lattice constructors, the NTRU trapdoor generator, QR precompute -- no GPU needed."""
import numpy as np
import pytest

import lattice_gaussian_mcmc_b200  # noqa: F401  (registers the package)
from lattice_gaussian_mcmc_b200 import ntru
from lattice_gaussian_mcmc_b200._engine import precompute_qr
from lattice_gaussian_mcmc_b200.lattices import (SimpleLattice, column_gram_schmidt_norms, dd_lattice,
                                                 identity_lattice, qary_lattice, row_gram_schmidt_norms)


def test_simple_lattice_duck_type():
    lat = SimpleLattice([[4, 1], [1, 3]])
    assert lat.dimension == 2 and lat.basis.dtype == np.float64
    # row Gram-Schmidt as src/lattices/simple.py:39-58: ||b*_0|| = |(4,1)|, ||b*_1|| = det / ||b*_0||
    gs = row_gram_schmidt_norms(lat.basis)
    assert np.allclose(gs, [np.sqrt(17), 11 / np.sqrt(17)])
    assert lat.min_gram_schmidt_norm == pytest.approx(11 / np.sqrt(17))
    assert lat.smoothing_parameter() > 0
    assert identity_lattice(10).basis.shape == (10, 10)
    with pytest.raises(ValueError):
        SimpleLattice(np.zeros((2, 3)))


def test_qary_shape_c3():
    lat = qary_lattice(n=8, m=16, q=97, seed=0)
    B = lat.basis
    assert B.shape == (16, 16) and np.allclose(np.triu(B), B)
    assert np.allclose(np.diag(B)[:8], 97) and np.allclose(np.diag(B)[8:], 1)
    assert abs(np.linalg.det(B) - 97.0 ** 8) < 1e-3 * 97.0 ** 8


def test_precompute_qr_matches_reference_convention():
    B = dd_lattice(12, seed=3).basis
    c = np.linspace(-1, 1, 12)
    Q, R, t = precompute_qr(B, c)
    assert np.all(np.diag(R) > 0) and np.allclose(np.tril(R, -1), 0)
    assert np.allclose(Q @ R, B) and np.allclose(t, Q.T @ c)
    assert np.allclose(np.abs(np.diag(R)), column_gram_schmidt_norms(B))


@pytest.mark.parametrize("n", [2, 8, 32, 64])
def test_ntru_solve_and_gs_profile(n):
    q = 12289
    f, g, F, G = ntru.ntru_keygen(n, q, seed=1)
    fG = ntru.poly_mul([int(x) for x in f], [int(x) for x in G])
    gF = ntru.poly_mul([int(x) for x in g], [int(x) for x in F])
    d = [a - b for a, b in zip(fG, gF)]
    assert d[0] == q and not any(d[1:])
    rows = ntru.ntru_basis_rows(f, g, F, G)
    # rows are lattice vectors of the NTRU lattice {(u, v): u h = v}: determinant q^n, short Gram-Schmidt profile
    gs = row_gram_schmidt_norms(rows)
    assert abs(np.sum(np.log(gs)) - n * np.log(q)) < 1e-6
    assert gs.max() <= 1.17 * np.sqrt(q) + 1e-9
    lat = ntru.ntru_lattice(n, q, seed=1)
    assert np.array_equal(lat.basis, rows.T) and lat.dimension == 2 * n
    assert lat.max_gram_schmidt_norm == pytest.approx(gs.max())


def test_poly_mul_negacyclic():
    rng = np.random.RandomState(0)
    a, b = rng.randint(-50, 50, 16), rng.randint(-50, 50, 16)
    want = ntru.negacyclic(a).T @ b   # column convention of the rotation matrix
    got = ntru.poly_mul([int(x) for x in a], [int(x) for x in b])
    assert np.array_equal(np.array(got, dtype=np.float64), want)


def _recorded():
    import json
    import os
    return json.load(open(os.path.join(os.path.dirname(__file__), "golden", "klein_lll_recorded.json")))["cases"]


@pytest.mark.parametrize("n", [16, 32, 64])
def test_lll_stand_in_reproduces_the_reference_recorded_basis_profile(n):
    """The reference reduced its scaling-run bases with Sage's LLL and recorded the Gram-Schmidt norms and
    sigma = 1.5 max||b*|| (results/Klein_LLL_n=*.json).  The host LLL stand-in must land on the same profile."""
    from lattice_gaussian_mcmc_b200.lattices import lll_reduce, random_int_basis, row_gram_schmidt_norms
    rec = _recorded()[str(n)]
    M = random_int_basis(n)
    L = lll_reduce(M)
    gs = row_gram_schmidt_norms(L)
    np.testing.assert_allclose(gs, np.array(rec["gs_norms"]), rtol=1e-10)
    assert abs(1.5 * gs.max() - rec["sigma"]) < 1e-9 * rec["sigma"]
    # same lattice: L = U M with U unimodular
    U = np.linalg.solve(M.T, L.T).T
    assert np.abs(U - np.rint(U)).max() < 1e-6
    assert abs(abs(np.linalg.det(np.rint(U))) - 1.0) < 1e-6 or n > 32   # det of a 64x64 integer matrix in fp64 is noisy
    assert abs(np.linalg.slogdet(L)[1] - np.linalg.slogdet(M)[1]) < 1e-9
    # (delta, eta) = (0.99, 0.501)-reduced
    R = np.linalg.qr(L.T, mode="r")
    R = R * np.sign(np.diag(R))[:, None]
    mu = (R / np.diag(R)[:, None]).T
    assert np.all(np.abs(np.tril(mu, -1)) <= 0.5011)
    d = np.diag(R) ** 2
    assert all(d[k] >= (0.99 - mu[k, k - 1] ** 2) * d[k - 1] * (1 - 1e-12) for k in range(1, n))


@pytest.mark.parametrize("threads", [0, 3, 8])
def test_host_widening_of_the_int16_transport(threads):
    """lgs_host_widen_i16 (csrc/lgs_hostpack.cpp): the host half of the packed lattice-point transport -- unaligned
    destinations, lengths that are not a multiple of the vector width, the threaded split (no GPU involved)."""
    from lattice_gaussian_mcmc_b200 import _capi
    L = _capi.lib()
    rng = np.random.RandomState(threads)
    for count in (0, 1, 7, 8, 63, 1000, (1 << 16) + 5, 300001):
        for off in (0, 1, 3):
            src = rng.randint(-32768, 32768, count + 8).astype(np.int16)
            dst = np.full(count + off + 4, np.nan)
            rc = L.lgs_host_widen_i16(src.ctypes.data, dst[off:].ctypes.data, count, threads)
            assert rc == 0
            assert np.array_equal(dst[off:off + count], src[:count].astype(np.float64))
            assert np.all(np.isnan(dst[:off])) and np.all(np.isnan(dst[off + count:]))      # nothing outside the range
    assert L.lgs_host_widen_i16(None, None, 5, 1) == _capi.LGS_E_INVALID


def test_host_result_cache_is_bounded_and_evicts_the_oldest_mapping():
    """_HostArrayCache (the recycled host memory behind KleinSampler.sample's result arrays): a dropped array backs the
    next result of the same size; at most two mappings are retained and a third size unmaps the oldest; mappings
    beyond LGS_HOST_CACHE_MB are not retained at all."""
    import gc
    from lattice_gaussian_mcmc_b200._engine import _HostArrayCache
    c = _HostArrayCache()
    assert c.max_arrays == 2 and c.limit == 4096 << 20
    mb = 1 << 20
    a = c.empty((64 * mb // 8,))
    assert a.flags.writeable and a.flags.c_contiguous and a.ctypes.data % (2 << 20) == 0
    addr = a.ctypes.data
    a[:3] = 1.0
    del a
    gc.collect()
    assert [e[0] for e in c.free] == [64 * mb] and c.cached == 64 * mb
    b = c.empty((64 * mb // 8,))                    # same size: the retained mapping, already faulted in
    assert b.ctypes.data == addr and c.cached == 0 and not c.free
    v = b[5:10]                                     # a view keeps the mapping out of the cache
    del b
    gc.collect()
    assert not c.free
    del v
    gc.collect()
    assert len(c.free) == 1
    for size in (72, 80):                           # varying batch sizes: never more than two, oldest goes first
        x = c.empty((size * mb // 8,))
        del x
        gc.collect()
    assert [e[0] // mb for e in c.free] == [72, 80] and c.cached == 152 * mb
    c.limit = 100 * mb                              # total budget: the 96 MB mapping fits only alone
    x = c.empty((96 * mb // 8,))
    del x
    gc.collect()
    assert [e[0] // mb for e in c.free] == [96] and c.cached == 96 * mb
    x = c.empty((128 * mb // 8,))                   # larger than the whole budget: not retained, nothing evicted
    del x
    gc.collect()
    assert [e[0] // mb for e in c.free] == [96]
    assert isinstance(c.empty((8,)), np.ndarray)    # small results are plain arrays


def test_reference_named_lattice_classes():
    """IdentityLattice / QaryLattice / NTRULattice: the constructor names and signatures of the reference's README
    quick-start (README.md:37-77; src/lattices/identity.py:33-52, qary.py:34-113, ntru.py:71,269-310,482-595)"""
    import lattice_gaussian_mcmc_b200 as lgs
    z = lgs.IdentityLattice(dimension=10)
    assert z.dimension == 10 and np.array_equal(z.get_basis(), np.eye(10)) and z.min_gram_schmidt_norm == 1.0
    assert z.get_gram_schmidt()[2] == [1.0] * 10 and z.first_minimum() == 1.0
    with pytest.raises(ValueError):
        lgs.IdentityLattice(0)
    A = np.random.RandomState(1).randint(0, 97, size=(3, 5))
    qa = lgs.QaryLattice(A, 97)
    assert qa.dimension == 8 and qa.q == 97 and (qa.n, qa.m) == (3, 5)
    B = qa.basis
    assert np.array_equal(B[:3, :3], 97 * np.eye(3)) and np.array_equal(B[:3, 3:], A) and np.array_equal(B[3:, 3:], np.eye(5))
    x = np.random.RandomState(2).randint(-9, 10, size=8)
    v = B @ x
    assert np.all((v[:3] - A @ v[3:]) % 97 == 0)                   # membership: x1 = A x2 (mod q)
    assert lgs.QaryLattice.from_lwe_instance(A, 97).lattice_type == "dual"
    lwe = lgs.QaryLattice.from_lwe_instance(256, 512, 3329, alpha=3.2 / np.sqrt(3329))      # README.md:71
    assert lwe.dimension == 512 and lwe.alpha == pytest.approx(3.2 / np.sqrt(3329))
    assert np.array_equal(lwe.basis, lgs.lattices.qary_lattice(256, 512, 3329, seed=0).basis)
    for bad in (lambda: lgs.QaryLattice(A, 1), lambda: lgs.QaryLattice(A, 97, "other"), lambda: lgs.QaryLattice(A + 0.5, 97)):
        with pytest.raises(ValueError):
            bad()


@pytest.mark.parametrize("n", [16, 64])
def test_ntru_public_key_and_public_basis(n):
    """h = g / f mod q exactly; [[q I, 0], [H, I]] spans the lattice of the trapdoor basis [[g, -f], [G, -F]]"""
    import lattice_gaussian_mcmc_b200 as lgs
    from lattice_gaussian_mcmc_b200 import ntru
    q = 12289
    L = lgs.NTRULattice(n=n, q=q, sigma=4.05 * np.sqrt(512 / n), seed=3)
    with pytest.raises(ValueError):
        L.get_basis()                                              # ntru.py:489: keys first
    assert L.generate_basis() is L
    assert L.verify_basis() == {"ntru_equation": True, "public_key": True, "gram_schmidt_bound": True}
    f, g = L.get_private_key()
    h = L.get_public_key()
    assert h.min() >= 0 and h.max() < q
    hf = ntru.poly_mul([int(c) for c in h], [int(c) for c in f])
    assert all((a - int(b)) % q == 0 for a, b in zip(hf, g))
    finv = ntru.poly_inv_mod_q(f, q)
    one = [c % q for c in ntru.poly_mul(finv, [int(c) for c in f])]
    assert one == [1] + [0] * (n - 1)
    with pytest.raises(ValueError):
        ntru.poly_inv_mod_q([q] + [0] * (n - 1), q)
    T, P = L.get_basis(), L.get_public_basis()                    # rows
    assert np.array_equal(T.T, L.basis) and T.shape == (2 * n, 2 * n)
    assert np.array_equal(P[:n, :n], q * np.eye(n)) and not P[:n, n:].any() and np.array_equal(P[n:, n:], np.eye(n))
    # every trapdoor row (u, v) satisfies u = v H (mod q) -> T = U P with U integer; equal volumes -> U unimodular
    H = P[n:, :n]
    assert np.all((T[:, :n] - T[:, n:] @ H) % q == 0)
    assert abs(np.linalg.slogdet(T)[1] - n * np.log(q)) < 1e-6
    pub = L.public_lattice()
    gs = np.sort(pub.get_gram_schmidt_norms())
    assert np.allclose(gs[:n], 1.0) and np.allclose(gs[n:], q)
    assert L.get_max_gram_schmidt_norm()[0] <= 1.17 * np.sqrt(q)
    with pytest.raises(ValueError):
        lgs.NTRULattice(n=48)
