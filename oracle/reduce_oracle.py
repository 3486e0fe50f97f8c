"""TEST INFRASTRUCTURE -- CPU checkers for the host basis reduction (csrc/lgs_reduce.cpp).  Only tests/ import this.

Parity status: the reduction's parity target is Sage's Matrix.LLL()/BKZ() (fplll), called at
/root/reference/src/lattices/reduction.py:103,275 and klein_scaling_analysis.py:95-108 -- a third-party dependency
that is absent from /root/reference and un-pinned (requirements.txt names no fpylll version).  What pins it: the
Gram-Schmidt profiles the reference RECORDED from that LLL for its n = 16 / 32 / 64 scaling runs
(results/Klein_LLL_n=*.json -> tests/golden/klein_lll_recorded.json).  `lll_textbook` below (the published
Lenstra-Lenstra-Lovasz algorithm with float Gram-Schmidt on an exact integer basis) reproduces those profiles, and
the C++ reducer is compared with it basis-for-basis.  BKZ output of the reference is recorded nowhere: BKZ parity is
UNPINNED; it is checked through properties (exactly verified LLL-reducedness, same lattice, block-wise shortest
vectors against brute force).
"""
from __future__ import annotations

from fractions import Fraction
from itertools import product

import numpy as np


def lll_textbook(rows: np.ndarray, delta: float = 0.99, eta: float = 0.501) -> np.ndarray:
    """LLL-reduce the ROWS of an integer matrix (Lenstra-Lenstra-Lovasz with parameters (delta, eta)).
    The basis is kept in exact Python integers; Gram-Schmidt data in float64 (adequate for the n <= 64,
    |entries| <= 50 inputs of the reference's scaling runs).  Returns the reduced row basis as float64."""
    B = np.array(np.rint(np.asarray(rows, dtype=np.float64)), dtype=object)
    n = B.shape[0]
    Bf = np.array(B, dtype=np.float64)
    Bs = np.zeros((n, B.shape[1]))
    mu = np.zeros((n, n))
    nrm = np.zeros(n)

    def gso_row(k):
        v = Bf[k].copy()
        if k:
            mu[k, :k] = (Bs[:k] @ Bf[k]) / nrm[:k]
            v -= mu[k, :k] @ Bs[:k]
        Bs[k] = v
        nrm[k] = v @ v

    gso_row(0)
    k = 1
    while k < n:
        gso_row(k)
        for j in range(k - 1, -1, -1):          # size reduction
            if abs(mu[k, j]) > eta:
                q = int(np.rint(mu[k, j]))
                B[k] = B[k] - q * B[j]
                Bf[k] = np.array(B[k], dtype=np.float64)
                mu[k, :j] -= q * mu[j, :j]
                mu[k, j] -= q
        gso_row(k)
        if nrm[k] >= (delta - mu[k, k - 1] ** 2) * nrm[k - 1]:   # Lovasz condition
            k += 1
        else:
            B[[k, k - 1]] = B[[k - 1, k]]
            Bf[[k, k - 1]] = Bf[[k - 1, k]]
            k = max(k - 1, 1)
            if k == 1:
                gso_row(0)
    return np.array(B, dtype=np.float64)


def exact_gso(rows):
    """Exact rational Gram-Schmidt of integer rows: (mu [n][n] Fractions, r [n] squared norms)."""
    B = [[int(x) for x in row] for row in np.asarray(rows)]
    n = len(B)
    G = [[sum(a * b for a, b in zip(B[i], B[j])) for j in range(i + 1)] for i in range(n)]
    r = [[Fraction(0)] * n for _ in range(n)]
    mu = [[Fraction(0)] * n for _ in range(n)]
    for i in range(n):
        for j in range(i + 1):
            a = Fraction(G[i][j])
            for l in range(j):
                a -= mu[j][l] * r[i][l]
            r[i][j] = a
            if j < i:
                mu[i][j] = a / r[j][j]
    return mu, [r[i][i] for i in range(n)]


def is_lll_reduced_exact(rows, delta, eta) -> bool:
    """(delta, eta)-reducedness decided in exact rational arithmetic: |mu_ij| <= eta and
    delta ||b*_{k-1}||^2 <= ||b*_k||^2 + mu_{k,k-1}^2 ||b*_{k-1}||^2."""
    mu, r = exact_gso(rows)
    n = len(r)
    d, e = Fraction(delta).limit_denominator(10 ** 9), Fraction(eta).limit_denominator(10 ** 9)
    for i in range(n):
        if r[i] <= 0:
            return False
        for j in range(i):
            if abs(mu[i][j]) > e:
                return False
        if i and d * r[i - 1] > r[i] + mu[i][i - 1] ** 2 * r[i - 1]:
            return False
    return True


def exact_det(M) -> int:
    """Determinant of an integer matrix by fraction-free (Bareiss) elimination."""
    A = [[int(x) for x in row] for row in np.asarray(M)]
    n = len(A)
    sign, prev = 1, 1
    for k in range(n - 1):
        if A[k][k] == 0:
            for i in range(k + 1, n):
                if A[i][k]:
                    A[k], A[i] = A[i], A[k]
                    sign = -sign
                    break
            else:
                return 0
        for i in range(k + 1, n):
            for j in range(k + 1, n):
                A[i][j] = (A[i][j] * A[k][k] - A[i][k] * A[k][j]) // prev
        prev = A[k][k]
    return sign * A[n - 1][n - 1]


def shortest_projected_norm2(rows, j, k, radius=3):
    """Brute force: min over nonzero integer coefficient vectors in [-radius, radius]^(k-j+1) of the squared norm of
    sum_a c_a b_{j+a} projected orthogonally to rows < j (small blocks only)."""
    mu, r = exact_gso(rows)
    d = k - j + 1
    best = None
    for c in product(range(-radius, radius + 1), repeat=d):
        if not any(c):
            continue
        tot = Fraction(0)
        for a in range(d):
            w = Fraction(c[a]) + sum(Fraction(c[l]) * mu[j + l][j + a] for l in range(a + 1, d))
            tot += w * w * r[j + a]
        if best is None or tot < best:
            best = tot
    return best


def shortest_vector_norm2(rows) -> int:
    """lambda_1^2 of the lattice spanned by the integer ROWS, by an independent Fincke-Pohst / Schnorr-Euchner search:
    the basis is first LLL-reduced with `lll_textbook` (so the search tree is small), the enumeration runs on float64
    Gram-Schmidt data with a 1e-9 relative safety margin on the radius, and every candidate's norm is recomputed in exact
    integers.  For n <= ~24 (pure Python).  Used to cross-check lgs_bkz_reduce with beta = n (a full-block BKZ tour must
    end on a shortest vector up to its delta slack) -- the only external check BKZ has on this image (fplll is absent)."""
    B = np.array(np.rint(lll_textbook(rows)), dtype=object)
    n = B.shape[0]
    Bf = np.array(B, dtype=np.float64)
    Bs = np.zeros_like(Bf)
    mu = np.zeros((n, n))
    r = np.zeros(n)
    for k in range(n):
        v = Bf[k].copy()
        for j in range(k):
            mu[k, j] = (Bs[j] @ Bf[k]) / r[j]
            v -= mu[k, j] * Bs[j]
        Bs[k] = v
        r[k] = v @ v
    best = int(sum(int(x) * int(x) for x in B[0]))       # the first reduced vector bounds the search
    x = [0] * n

    def descend(k, partial):          # coordinates k+1 .. n-1 fixed, `partial` = their contribution to the squared norm
        nonlocal best
        c = -sum(x[j] * mu[j, k] for j in range(k + 1, n))
        bound = best * (1 + 1e-9) - partial
        if bound < 0:
            return
        rad = (bound / r[k]) ** 0.5
        lo, hi = int(np.ceil(c - rad)), int(np.floor(c + rad))
        for xi in sorted(range(lo, hi + 1), key=lambda t: abs(t - c)):
            x[k] = xi
            p2 = partial + (xi - c) ** 2 * r[k]
            if k == 0:
                if any(x):
                    vec = [sum(x[i] * int(B[i][col]) for i in range(n)) for col in range(B.shape[1])]
                    nrm = sum(v * v for v in vec)
                    if nrm < best:
                        best = nrm
            else:
                descend(k - 1, p2)
        x[k] = 0

    descend(n - 1, 0.0)
    return best
