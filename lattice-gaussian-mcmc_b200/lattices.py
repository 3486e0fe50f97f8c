"""Host-side lattice objects (numpy only).

The reference's lattice classes need SageMath and three of them are not valid UTF-8, so they
cannot be imported; what the samplers actually consume is a four-attribute duck type:
`basis` (columns are basis vectors as klein.py:59,218 use it), `dimension`, `min_gram_schmidt_norm`,
`smoothing_parameter()`, plus `name` for __repr__.  `SimpleLattice` supplies exactly that, following the formulas of
the reference's src/lattices/simple.py:39-82 (row Gram-Schmidt) and src/lattices/base.py (smoothing bound).
The constructors below restate the reference's *formulas* for the shapes BASELINE.json names:

  identity_lattice(n)            src/lattices/identity.py:33-52      B = I_n
  qary_lattice(A, q)             src/lattices/qary.py:131-164        [[q I, A], [0, I]]  (columns generate {x: x1 = A x2 mod q})
  dd_lattice(n, seed)            SURVEY.md appendix A.5              randint(-3,4) + 2n I   (the dimension-sweep family)
  random_int_lattice(n, seed)    klein_scaling_analysis.py:83-86     randint(0,51), seed 0+n  (input of the LLL runs)
  ntru_lattice(...)              src/lattices/ntru.py:482-537        [[rot(g), -rot(f)], [rot(G), -rot(F)]], see ntru.py here

IdentityLattice / QaryLattice (here) and NTRULattice (ntru.py) carry the reference's class names and the constructor
signatures its README quick-start uses, so that the quick-start runs against this package with the imports swapped.

Everything here is host setup that runs once per sampler; nothing in this file is on the sampling path.
"""
from __future__ import annotations

import numpy as np


def row_gram_schmidt_norms(basis: np.ndarray) -> np.ndarray:
    """||b*_i|| of the ROWS of `basis` (the reference's SimpleLattice convention, simple.py:39-58), via QR."""
    b = np.asarray(basis, dtype=np.float64)
    r = np.linalg.qr(b.T, mode="r")
    return np.abs(np.diag(r))


def column_gram_schmidt_norms(basis: np.ndarray) -> np.ndarray:
    """||b*_i|| of the COLUMNS of `basis`: |R_ii| of the factor the samplers use (klein.py:59)."""
    r = np.linalg.qr(np.asarray(basis, dtype=np.float64), mode="r")
    return np.abs(np.diag(r))


class SimpleLattice:
    """The lattice duck type the samplers read (src/lattices/simple.py:17-82, src/lattices/base.py:16-120)."""

    def __init__(self, basis, name: str = "SimpleLattice"):
        b = np.array(basis, dtype=np.float64)
        if b.ndim != 2 or b.shape[0] != b.shape[1]:
            raise ValueError("basis must be a square matrix")
        self._basis = np.ascontiguousarray(b)
        self.dimension = b.shape[0]
        self.name = name
        self._gs_norms = row_gram_schmidt_norms(self._basis)

    @property
    def basis(self) -> np.ndarray:
        return self._basis

    def get_basis(self) -> np.ndarray:
        return self._basis

    def get_dimension(self) -> int:
        return self.dimension

    @property
    def min_gram_schmidt_norm(self) -> float:
        return float(np.min(self._gs_norms))

    @property
    def max_gram_schmidt_norm(self) -> float:
        return float(np.max(self._gs_norms))

    def get_gram_schmidt_norms(self):
        return [float(x) for x in self._gs_norms]

    def smoothing_parameter(self, epsilon: float = 0.01) -> float:
        """eta_eps(L) <= max||b*_i|| sqrt(ln(2n(1+1/eps))/pi)  (the bound lattices/base.py uses)."""
        n = self.dimension
        return float(np.max(self._gs_norms) * np.sqrt(np.log(2 * n * (1 + 1 / epsilon)) / np.pi))

    def __repr__(self):
        return f"{self.name}(n={self.dimension})"


def identity_lattice(n: int) -> SimpleLattice:
    return SimpleLattice(np.eye(n), name=f"IdentityLattice(Z^{n})")


class IdentityLattice(SimpleLattice):
    """Z^n under the reference's class name and signature (src/lattices/identity.py:17-52; README.md:37-48:
    `IdentityLattice(dimension=10)`), with the closed forms of identity.py:74-90,177-209."""

    def __init__(self, dimension: int):
        if dimension <= 0:
            raise ValueError(f"Dimension must be positive, got {dimension}")
        super().__init__(np.eye(int(dimension)), name=f"IdentityLattice(Z^{int(dimension)})")

    def get_gram_schmidt(self):
        n = self.dimension
        return np.eye(n), np.eye(n), [1.0] * n

    def volume(self) -> float:
        return 1.0

    def first_minimum(self) -> float:
        return 1.0

    def is_self_dual(self) -> bool:
        return True


class QaryLattice(SimpleLattice):
    """q-ary lattice under the reference's class name (src/lattices/qary.py:21-164).  The samplers need a square
    basis, so both lattice types are given by the full-rank column basis [[q I_k, A], [0, I_m]] of
    {(x1, x2): x1 = A x2 (mod q)}, A of shape (k, m) (qary.py:131-152 writes the primal one as [[q I, -A^T], [0, I]]:
    the same construction for -A^T).  `from_lwe_instance(A, q)` is the code's signature (qary.py:100-113);
    `from_lwe_instance(n, m, q, alpha=...)` is the README's (README.md:71: n secret dimensions, m in total): a
    uniform A of shape (m - n, n) from `seed`."""

    def __init__(self, A, q: int, lattice_type: str = "primal"):
        if not isinstance(q, (int, np.integer)) or q <= 1:
            raise ValueError(f"Modulus q must be an integer > 1, got {q}")
        if lattice_type not in ("primal", "dual"):
            raise ValueError(f"Unknown lattice type: {lattice_type}")
        A = np.asarray(A)
        if A.ndim != 2 or not np.all(A == np.round(A)):
            raise ValueError("A must be an integer matrix")
        self.A = A.astype(np.int64) % int(q)
        self.q = int(q)
        self.lattice_type = lattice_type
        self.n, self.m = self.A.shape
        super().__init__(qary_basis(self.A, self.q), name=f"QaryLattice({lattice_type}, {self.n}x{self.m}, q={self.q})")

    @classmethod
    def from_random_matrix(cls, n: int, m: int, q: int, lattice_type: str = "primal", seed: int = 0):
        return cls(np.random.RandomState(seed).randint(0, q, size=(n, m)), q, lattice_type)

    @classmethod
    def from_lwe_instance(cls, A, q_or_m, q: int | None = None, alpha: float | None = None, seed: int = 0):
        if q is None:                                    # qary.py:100-113: (A, q)
            return cls(A, int(q_or_m), lattice_type="dual")
        n, m = int(A), int(q_or_m)                       # README.md:71: (n, m, q, alpha=...)
        if not 0 < n < m:
            raise ValueError("need 0 < n < m")
        lat = cls(np.random.RandomState(seed).randint(0, q, size=(m - n, n)), int(q), lattice_type="dual")
        lat.alpha = alpha
        return lat

    def volume(self) -> float:
        return float(self.q) ** self.n

    def get_determinant(self) -> float:
        return self.volume()


def dd_basis(n: int, seed: int) -> np.ndarray:
    rng = np.random.RandomState(seed)
    return (rng.randint(-3, 4, size=(n, n)) + 2 * n * np.eye(n)).astype(np.float64)


def dd_lattice(n: int, seed: int | None = None) -> SimpleLattice:
    return SimpleLattice(dd_basis(n, n if seed is None else seed), name=f"DiagDominant(n={n})")


def random_int_basis(n: int, seed: int | None = None) -> np.ndarray:
    """The unreduced input of the reference's scaling runs (klein_scaling_analysis.py:83-86)."""
    np_rng = np.random.RandomState(0 + n if seed is None else seed)
    return np_rng.randint(0, 51, size=(n, n)).astype(np.float64)


def qary_basis(A: np.ndarray, q: int) -> np.ndarray:
    """[[q I_k, A], [0, I_m]] with A of shape (k, m): columns generate {(x1, x2): x1 = A x2 (mod q)}."""
    A = np.asarray(A, dtype=np.float64)
    k, m = A.shape
    top = np.hstack([q * np.eye(k), A])
    bot = np.hstack([np.zeros((m, k)), np.eye(m)])
    return np.vstack([top, bot])


def qary_lattice(n: int = 256, m: int = 512, q: int = 3329, seed: int = 0) -> SimpleLattice:
    """BASELINE config C3: q-ary LWE lattice, n = 256 secret dims, m = 512 total dims, q = 3329."""
    A = np.random.RandomState(seed).randint(0, q, size=(m - n, n))
    return SimpleLattice(qary_basis(A, q), name=f"QaryLattice(n={n}, m={m}, q={q})")


def lattice_from_rows(rows: np.ndarray, name: str = "SimpleLattice") -> SimpleLattice:
    """Wrap a row-basis the way the reference's scripts do: the matrix is handed to the sampler unchanged
    (klein_scaling_analysis.py:175), i.e. the sampler then treats its columns as the basis."""
    return SimpleLattice(rows, name=name)


# ---------------------------------------------------------------------------------------------------------
# LLL.  The reference reduces through Sage's Matrix.LLL() (fplll; src/lattices/reduction.py:103,
# klein_scaling_analysis.py:95-108), which is not vendored and not installable here.  The reduction is host C++
# behind the C ABI (csrc/lgs_reduce.cpp, reduction.py); with (delta, eta) = (0.99, 0.501) it lands on the
# Gram-Schmidt profiles the reference recorded for its n = 16 / 32 / 64 runs (tests/test_host_setup.py).
# ---------------------------------------------------------------------------------------------------------
def lll_reduce(rows: np.ndarray, delta: float = 0.99, eta: float = 0.501) -> np.ndarray:
    """LLL-reduce the ROWS of an integer matrix; returns the reduced row basis as float64."""
    from .reduction import lll
    red, _, _ = lll(rows, delta=delta, eta=eta, with_transform=False)
    return red.astype(np.float64)


def lll_lattice(n: int, seed: int | None = None, delta: float = 0.99) -> SimpleLattice:
    """BASELINE config C2 / the n <= 64 points of C5: randint(0, 51) matrix with seed 0 + n
    (klein_scaling_analysis.py:83-86), LLL-reduced by rows, handed to the sampler unchanged (ibid. :175)."""
    return lattice_from_rows(lll_reduce(random_int_basis(n, seed), delta), name=f"LLL-reduced random lattice (n={n})")
