"""Host-side NTRU trapdoor bases (numpy + Python big integers), the lattice of BASELINE config C4.

The reference builds the same block basis  [[rot(g), -rot(f)], [rot(G), -rot(F)]]  with negacyclic rot()
(src/lattices/ntru.py:482-537) from f, g ~ D_{Z, sigma_fg}^n (ntru.py:186-214) and (F, G) with f G - g F = q
(ntru.py:269-310), but it needs SageMath and its _ntru_solve treats the field norm as an integer (ntru.py:342-356,
SURVEY.md section 8f-1), so it cannot produce a basis here.  This module follows the same formulas with an exact
NTRUSolve: the tower-of-rings recursion of Pornin & Prest ("More efficient algorithms for the NTRU key generation
using the field norm", PKC 2019) -- field norm down to degree 1, extended gcd, lift, and Babai size-reduction of
(F, G) against (f, g) at every level.  All of it is setup that runs once on the host.
"""
from __future__ import annotations

import math

import numpy as np


# ------------------------------------------------------------------ arithmetic in Z[x]/(x^n + 1), big integers
def _bitlen(v):
    return max((abs(int(c)).bit_length() for c in v), default=0)


def poly_mul(a, b):
    """a * b mod x^n + 1 by Kronecker substitution (one big-integer product)."""
    n = len(a)
    k = _bitlen(a) + _bitlen(b) + n.bit_length() + 2
    A = sum(int(c) << (k * i) for i, c in enumerate(a))
    B = sum(int(c) << (k * i) for i, c in enumerate(b))
    C = A * B
    mask, half = (1 << k) - 1, 1 << (k - 1)
    out = [0] * (2 * n)
    for i in range(2 * n - 1):
        r = C & mask
        if r >= half:
            r -= 1 << k
        out[i] = r
        C = (C - r) >> k
    return [out[i] - out[i + n] for i in range(n)]


def galois_conjugate(a):
    return [(-c if i & 1 else c) for i, c in enumerate(a)]


def field_norm(a):
    """N(a) = a(x) a(-x) as an element of Z[y]/(y^(n/2) + 1), y = x^2."""
    ae, ao = a[0::2], a[1::2]
    ae2, ao2 = poly_mul(ae, ae), poly_mul(ao, ao)
    m = len(ae)
    # ae^2 - y * ao^2, with y * p = (-p[m-1], p[0], ..., p[m-2])
    shifted = [-ao2[m - 1]] + ao2[:m - 1]
    return [ae2[i] - shifted[i] for i in range(m)]


def lift(a):
    out = [0] * (2 * len(a))
    out[0::2] = a
    return out


def _xgcd(a, b):
    x0, x1, y0, y1 = 1, 0, 0, 1
    while b:
        q, a, b = a // b, b, a % b
        x0, x1 = x1, x0 - q * x1
        y0, y1 = y1, y0 - q * y1
    return a, x0, y0


def _fft(a):
    """evaluation of a real polynomial at the n odd 2n-th roots of unity"""
    n = len(a)
    w = np.exp(1j * np.pi * np.arange(n) / n)
    return np.fft.fft(np.asarray(a, dtype=np.float64) * w)


def _ifft(A):
    n = len(A)
    w = np.exp(-1j * np.pi * np.arange(n) / n)
    return (np.fft.ifft(A) * w).real


def _reduce(f, g, F, G):
    """Babai round-off of (F, G) against (f, g): (F, G) -= k (f, g), k = round((F f* + G g*) / (f f* + g g*))."""
    size = max(53, _bitlen(f), _bitlen(g))
    fa = _fft([c >> (size - 53) for c in f])
    ga = _fft([c >> (size - 53) for c in g])
    den = fa * np.conj(fa) + ga * np.conj(ga)
    while True:
        Size = max(53, _bitlen(F), _bitlen(G))
        if Size < size:
            break
        Fa = _fft([c >> (Size - 53) for c in F])
        Ga = _fft([c >> (Size - 53) for c in G])
        k = np.rint(_ifft((Fa * np.conj(fa) + Ga * np.conj(ga)) / den))
        if not np.any(k):
            break
        k = [int(v) for v in k]
        fk, gk = poly_mul(f, k), poly_mul(g, k)
        sh = Size - size
        F = [F[i] - (fk[i] << sh) for i in range(len(F))]
        G = [G[i] - (gk[i] << sh) for i in range(len(G))]
    return F, G


def ntru_solve(f, g, q):
    """(F, G) in Z[x]/(x^n+1) with f G - g F = q; raises ValueError when the resultants are not coprime."""
    n = len(f)
    if n == 1:
        d, u, v = _xgcd(int(f[0]), int(g[0]))
        if d not in (1, -1):
            raise ValueError("gcd(Res(f), Res(g)) != 1")
        return [-q * v * d], [q * u * d]
    Fp, Gp = ntru_solve(field_norm(f), field_norm(g), q)
    F = poly_mul(lift(Fp), galois_conjugate(g))
    G = poly_mul(lift(Gp), galois_conjugate(f))
    return _reduce(f, g, F, G)


# ------------------------------------------------------------------ key generation and the block basis
def negacyclic(a) -> np.ndarray:
    """rot(a): row i is x^i * a mod x^n + 1 (ntru.py:482-510)."""
    a = np.asarray(a, dtype=np.float64)
    n = a.size
    M = np.empty((n, n))
    row = a.copy()
    for i in range(n):
        M[i] = row
        row = np.concatenate(([-row[-1]], row[:-1]))
    return M


def sample_fg(n: int, sigma_fg: float, rng: np.random.RandomState):
    """coefficients ~ D_{Z, sigma_fg} by a truncated table (ntru.py:186-214 samples them one at a time)"""
    tail = int(math.ceil(12 * sigma_fg))
    sup = np.arange(-tail, tail + 1)
    p = np.exp(-sup.astype(np.float64) ** 2 / (2 * sigma_fg ** 2))
    p /= p.sum()
    return rng.choice(sup, size=n, p=p), rng.choice(sup, size=n, p=p)


def ntru_keygen(n: int = 512, q: int = 12289, sigma_fg: float | None = None, seed: int = 0, max_gs: float = 1.17,
                max_tries: int = 200):
    """f, g, F, G with f G - g F = q and max ||b*_i|| <= max_gs * sqrt(q) (the Falcon acceptance test)."""
    if sigma_fg is None:
        sigma_fg = 1.17 * math.sqrt(q / (2.0 * n))
    rng = np.random.RandomState(seed)
    for _ in range(max_tries):
        f, g = sample_fg(n, sigma_fg, rng)
        # cheap GS bound first (Falcon spec, algorithm 5): both halves of the profile peak at b*_0 and b*_n
        n0 = math.sqrt(float(np.sum(f * f) + np.sum(g * g)))
        fa, ga = _fft(f), _fft(g)
        den = (fa * np.conj(fa) + ga * np.conj(ga)).real
        if np.min(den) < 1e-9:
            continue
        n1 = math.sqrt(float(np.sum((q * q) / den) / n))   # ||(q f*/(ff*+gg*), q g*/(ff*+gg*))||
        if max(n0, n1) > max_gs * math.sqrt(q):
            continue
        try:
            F, G = ntru_solve([int(c) for c in f], [int(c) for c in g], q)
        except ValueError:
            continue
        return (np.array(f, dtype=np.int64), np.array(g, dtype=np.int64),
                np.array(F, dtype=np.int64), np.array(G, dtype=np.int64))
    raise RuntimeError("no NTRU key found")


def ntru_basis_rows(f, g, F, G) -> np.ndarray:
    """[[rot(g), -rot(f)], [rot(G), -rot(F)]]: ROWS are lattice vectors (ntru.py:512-537)."""
    return np.block([[negacyclic(g), -negacyclic(f)], [negacyclic(G), -negacyclic(F)]])


def ntru_lattice(n: int = 512, q: int = 12289, sigma_fg: float | None = None, seed: int = 0):
    """SimpleLattice over the trapdoor basis, handed to the samplers with the short vectors as COLUMNS (the samplers
    read `basis` column-wise, klein.py:59,218), so that the conditional widths are sigma / ||b*_i|| with
    ||b*_i|| <= 1.17 sqrt(q)."""
    from .lattices import SimpleLattice
    f, g, F, G = ntru_keygen(n, q, sigma_fg, seed)
    rows = ntru_basis_rows(f, g, F, G)
    lat = SimpleLattice(np.ascontiguousarray(rows.T), name=f"NTRULattice(n={n}, q={q})")
    lat.f, lat.g, lat.F, lat.G, lat.q = f, g, F, G, q
    # the reference's SimpleLattice computes Gram-Schmidt over rows; the vectors here are the columns
    from .lattices import column_gram_schmidt_norms
    lat._gs_norms = column_gram_schmidt_norms(lat.basis)
    return lat
