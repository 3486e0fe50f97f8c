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
