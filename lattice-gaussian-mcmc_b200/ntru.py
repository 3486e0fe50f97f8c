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


# ------------------------------------------------------------------ public key and public basis
def poly_inv_mod_q(f, q: int):
    """f^-1 in Z_q[x]/(x^n + 1), n a power of two, exactly: f(x) f(-x) = N(f)(x^2), so f^-1 = f(-x) * lift(N(f)^-1),
    with N(f)^-1 from the same rule one level down (degree 1: a modular inverse).  ValueError if f is not invertible
    (the reference's _is_invertible / _invert_poly_Rq, ntru.py:224-267, ask Sage for the same thing)."""
    f = [int(c) % q for c in f]
    if len(f) == 1:
        try:
            return [pow(f[0], -1, q)]
        except ValueError:
            raise ValueError("f is not invertible mod q") from None
    inv_norm = poly_inv_mod_q(field_norm(f), q)
    return [c % q for c in poly_mul(galois_conjugate(f), lift(inv_norm))]


def ntru_public_key(f, g, q: int):
    """h = g * f^-1 mod q, coefficients in [0, q) (ntru.py:564-578 get_public_key; ntru_clean.py:100-108)."""
    return np.array([c % q for c in poly_mul([int(c) for c in g], poly_inv_mod_q(f, q))], dtype=np.int64)


def ntru_public_basis_rows(h, q: int) -> np.ndarray:
    """[[q I, 0], [H, I]] (ntru_clean.py:116-134), ROWS are lattice vectors.  H = rot(-h mod q), negacyclic like the
    trapdoor blocks: with h = g / f the trapdoor rows (g, -f), (G, -F) of ntru_basis_rows satisfy u = -h v (mod q), so
    this is a basis of the SAME lattice -- the one anybody can write down from the public key.  Its Gram-Schmidt norms
    are q (n times) and ~1 (n times): under one sigma the sampler meets sigma_i = sigma / q (a three-entry window) and
    sigma_i ~ sigma (the 1001-entry truncated window) in one lattice."""
    h = np.asarray(h, dtype=np.int64)
    n = h.size
    H = negacyclic((-h) % q)
    return np.block([[q * np.eye(n), np.zeros((n, n))], [H, np.eye(n)]])


class NTRULattice:
    """The reference's NTRULattice surface that its README and experiment scripts touch (src/lattices/ntru.py:46-112,
    269-310, 482-595, 724-747; README.md:66-77), without SageMath:

        ntru = NTRULattice(n=512, q=12289, sigma=4.05); ntru.generate_basis(); KleinSampler(ntru, sigma=...)

    `sigma` is the width of the coefficients of f and g; a value of sqrt(q)/2 or more is read as the reference's
    default "1.17 sqrt(q)" (a Gram-Schmidt target, ntru.py:71) and selects the Falcon width 1.17 sqrt(q / 2n).
    After generate_basis() the object is the lattice duck type the samplers read: `basis` holds the trapdoor vectors
    as COLUMNS (klein.py:59,218 read columns); get_basis() returns the reference's row form [[g, -f], [G, -F]];
    public_lattice() is the same lattice under the public basis [[q I, 0], [H, I]]."""

    def __init__(self, n: int, q: int = 12289, sigma: float | None = None, seed: int = 0):
        if n < 2 or n & (n - 1):
            raise ValueError(f"n must be a power of 2, got {n}")
        self.n, self.q, self.seed = int(n), int(q), int(seed)
        self.sigma = float(sigma) if sigma is not None else 1.17 * math.sqrt(q)
        self.dimension = 2 * self.n
        self.name = f"NTRULattice(n={n}, q={q})"
        self.f = self.g = self.F = self.G = self.h = None
        self._basis = self._gs_norms = None

    # -- key and basis generation (ntru.py:269-310, 482-503)
    def generate_keys(self, max_attempts: int = 200) -> bool:
        sigma_fg = None if self.sigma >= math.sqrt(self.q) / 2 else self.sigma
        # f must also be invertible mod q for the public key to exist: draw again (other seed) until it is
        for attempt in range(8):
            f, g, F, G = ntru_keygen(self.n, self.q, sigma_fg, self.seed + 7919 * attempt, max_tries=max_attempts)
            try:
                h = ntru_public_key(f, g, self.q)
            except ValueError:
                continue
            self.f, self.g, self.F, self.G, self.h = f, g, F, G, h
            return True
        return False

    def generate_basis(self):
        from .lattices import column_gram_schmidt_norms
        if self.f is None and not self.generate_keys():
            raise RuntimeError("no NTRU key found")
        rows = ntru_basis_rows(self.f, self.g, self.F, self.G)
        self._basis = np.ascontiguousarray(rows.T)
        self._gs_norms = column_gram_schmidt_norms(self._basis)
        return self

    def _need_basis(self):
        if self._basis is None:
            raise ValueError("Must generate keys before constructing basis")     # ntru.py:489

    @property
    def basis(self) -> np.ndarray:
        self._need_basis()
        return self._basis

    def get_basis(self) -> np.ndarray:
        """the reference's orientation (ntru.py:539-558): ROWS (g, -f), (G, -F) and their rotations"""
        self._need_basis()
        return self._basis.T.copy()

    def get_dimension(self) -> int:
        return self.dimension

    def get_public_key(self) -> np.ndarray:
        self._need_basis()
        return self.h.copy()

    def get_private_key(self):
        self._need_basis()
        return self.f.copy(), self.g.copy()

    def get_public_basis(self) -> np.ndarray:
        """[[q I, 0], [H, I]], rows (ntru_clean.py:116-134)"""
        self._need_basis()
        return ntru_public_basis_rows(self.h, self.q)

    def public_lattice(self):
        """the same lattice under its public basis, as the samplers read it (columns)"""
        from .lattices import SimpleLattice, column_gram_schmidt_norms
        lat = SimpleLattice(np.ascontiguousarray(self.get_public_basis().T), name=f"NTRUPublic(n={self.n}, q={self.q})")
        lat._gs_norms = column_gram_schmidt_norms(lat.basis)
        lat.q, lat.h = self.q, self.h.copy()
        return lat

    # -- Gram-Schmidt data (ntru.py:639-677, 724-747)
    def gram_schmidt_norms(self) -> np.ndarray:
        self._need_basis()
        return self._gs_norms.copy()

    get_gram_schmidt_norms = gram_schmidt_norms

    @property
    def min_gram_schmidt_norm(self) -> float:
        return float(np.min(self.gram_schmidt_norms()))

    @property
    def max_gram_schmidt_norm(self) -> float:
        return float(np.max(self.gram_schmidt_norms()))

    def get_max_gram_schmidt_norm(self):
        """(max ||b*_i||, the Falcon bound 1.17 sqrt(q))  (ntru.py:724-747)"""
        return self.max_gram_schmidt_norm, 1.17 * math.sqrt(self.q)

    def smoothing_parameter(self, epsilon: float = 0.01) -> float:
        return float(self.max_gram_schmidt_norm * np.sqrt(np.log(2 * self.dimension * (1 + 1 / epsilon)) / np.pi))

    def verify_basis(self) -> dict:
        """ntru.py:749-801: f G - g F = q, h f = g (mod q), the Falcon Gram-Schmidt bound"""
        self._need_basis()
        fi, gi, Fi, Gi = ([int(c) for c in p] for p in (self.f, self.g, self.F, self.G))
        eq = [a - b for a, b in zip(poly_mul(fi, Gi), poly_mul(gi, Fi))]
        hf = [c % self.q for c in poly_mul([int(c) for c in self.h], fi)]
        return {"ntru_equation": eq == [self.q] + [0] * (self.n - 1),
                "public_key": hf == [c % self.q for c in gi],
                "gram_schmidt_bound": self.max_gram_schmidt_norm <= 1.17 * math.sqrt(self.q) * (1 + 1e-9)}

    def __repr__(self) -> str:
        return self.name + ("" if self._basis is not None else " (no basis yet)")
