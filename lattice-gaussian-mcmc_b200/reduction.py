"""Basis reduction before the sampling path: host-side mirror of the reference's `LatticeReduction`
(src/lattices/reduction.py:23-697) over the C-ABI entry points `lgs_lll_reduce` / `lgs_bkz_reduce` / `lgs_gso_norms`
(csrc/lgs_reduce.cpp, host C++).

The reference drives Sage's `Matrix.LLL(delta)` / `Matrix.BKZ(block_size)` (fplll), which is neither vendored nor
installable here; the algorithms are re-implemented (L2-style LLL on an exact integer basis, Schnorr-Euchner BKZ).
Conventions follow the reference: ROWS are basis vectors; every reducer returns `(reduced_basis, U, diagnostics)`
with `reduced = U @ basis` (reduction.py:104 `B.solve_left(B_reduced)`) as float64 arrays (reduction.py:133
`Matrix(RDF, ...)`).  Plotting and report files (reduction.py:539-579, 672-697) are out of scope.
"""
from __future__ import annotations

import math
import time
from typing import Any, Dict, List, Optional, Tuple

import numpy as np

from . import _capi


class ReductionError(RuntimeError):
    pass


_ERR = {
    _capi.LGS_E_INVALID: "invalid argument (m < n, linearly dependent rows, or delta / eta out of range)",
    _capi.LGS_E_RANGE: "an integer left the exact range of the reducer (|entry| >= 2^56)",
    _capi.LGS_E_PRECISION: "floating-point Gram-Schmidt data not accurate enough at the highest precision tier",
}


def _as_int_rows(basis) -> np.ndarray:
    A = np.asarray(basis, dtype=np.float64)
    if A.ndim != 2:
        raise ValueError("basis must be a 2-D array (rows = basis vectors)")
    Ai = np.rint(A)
    if not np.array_equal(Ai, A):
        raise ValueError("the reducer works on integer bases (scale rational bases to integers first)")
    return np.ascontiguousarray(Ai.astype(np.int64))


def _call(rc: int):
    if rc != _capi.LGS_OK:
        raise ReductionError(f"liblgs_b200 error {rc}: {_ERR.get(rc, 'unknown')}")


def lll(basis, delta: float = 0.99, eta: float = 0.501, precision: int = 0, with_transform: bool = True):
    """(delta, eta)-LLL-reduce the rows of an integer matrix.  Returns (reduced int64 [n,m], U int64 [n,n] or None,
    stats dict).  (0.99, 0.501) reproduces Sage's `LLL(delta=0.99)` Gram-Schmidt profiles recorded by the reference
    (tests/golden/klein_lll_recorded.json)."""
    A = _as_int_rows(basis)
    n, m = A.shape
    U = np.zeros((n, n), dtype=np.int64) if with_transform else None
    st = np.zeros(5, dtype=np.int64)
    _call(_capi.lib().lgs_lll_reduce(A.ctypes.data, n, m, float(delta), float(eta), int(precision),
                                     U.ctypes.data if U is not None else None, st.ctypes.data))
    return A, U, _stats(st)


def bkz(basis, block_size: int, delta: float = 0.99, max_tours: int = 0, prune: int = -1, precision: int = 0,
        with_transform: bool = True):
    """BKZ-`block_size` reduce the rows of an integer matrix (LLL first).  max_tours <= 0: until no block improves."""
    A = _as_int_rows(basis)
    n, m = A.shape
    U = np.zeros((n, n), dtype=np.int64) if with_transform else None
    st = np.zeros(5, dtype=np.int64)
    _call(_capi.lib().lgs_bkz_reduce(A.ctypes.data, n, m, int(block_size), float(delta), int(max_tours), int(prune),
                                     int(precision), U.ctypes.data if U is not None else None, st.ctypes.data))
    return A, U, _stats(st)


def gso_norms(basis) -> np.ndarray:
    """||b*_i|| of the rows, Gram-Schmidt in double-double on exact inner products (reduction.py:386-391)."""
    A = _as_int_rows(basis)
    n, m = A.shape
    out = np.zeros(n, dtype=np.float64)
    _call(_capi.lib().lgs_gso_norms(A.ctypes.data, n, m, out.ctypes.data))
    return out


def _stats(st) -> Dict[str, int]:
    return {"swaps": int(st[0]), "size_reduction_passes": int(st[1]), "bkz_insertions": int(st[2]),
            "enumeration_nodes": int(st[3]), "float_bits": int(st[4])}


def _float_gs_norms(B: np.ndarray) -> np.ndarray:
    """Row Gram-Schmidt norms of a real matrix = |diag R| of QR(B^T)."""
    R = np.linalg.qr(np.asarray(B, dtype=np.float64).T, mode="r")
    return np.abs(np.diag(R))


class LatticeReduction:
    """Same surface as the reference class (reduction.py:23): lll_reduce, lll_with_removals, bkz_reduce,
    hermite_factor, orthogonality_defect, basis_quality_profile, sampling_reduce, gram_schmidt_improve,
    estimate_reduction_cost, compare_bases; `stats` as reduction.py:43-47."""

    def __init__(self, log_dir: Optional[str] = None):
        self.log_dir = log_dir
        self.stats = {"reductions_performed": 0, "total_time": 0.0, "quality_improvements": []}

    # ---- 1. LLL (reduction.py:68-133) ----
    def lll_reduce(self, basis, delta: float = 0.75, track_progress: bool = False):
        start = time.time()
        initial = self.basis_quality_profile(basis)
        red, U, st = lll(basis, delta=delta)
        final = self.basis_quality_profile(red)
        dt = time.time() - start
        diag = {"time": dt, "initial_quality": initial, "final_quality": final,
                "improvement_factor": initial["hermite_factor"] / final["hermite_factor"], "delta": delta,
                "reducer": st}
        if track_progress:  # the reference logs (swap index, potential) per swap; here: the totals of the run
            diag["progress"] = [{"swaps": st["swaps"], "log_potential": final["log_potential"]}]
        self._account(dt, diag["improvement_factor"])
        return red.astype(np.float64), U.astype(np.float64), diag

    # ---- LLL with removals (reduction.py:188-234) ----
    def lll_with_removals(self, basis, target_dim: Optional[int] = None, removal_threshold: float = 0.99):
        red, _, _ = self.lll_reduce(basis)
        n = red.shape[0]
        if target_dim is None:
            target_dim = max(1, n // 2)
        # the reference reads fpylll's r[i,i] (squared norms) into `gs_norms` (reduction.py:215) -- kept as is
        gs_sq = gso_norms(red) ** 2
        threshold = removal_threshold * gs_sq.max()
        removed: List[int] = []
        for i in range(n - 1, -1, -1):
            if len(removed) >= n - target_dim:
                break
            if gs_sq[i] > threshold:
                removed.append(i)
        keep = [i for i in range(n) if i not in removed]
        return red[keep, :], removed

    # ---- 2. BKZ (reduction.py:238-318) ----
    def bkz_reduce(self, basis, block_size: int, strategy: str = "default", progressive: bool = False):
        start = time.time()
        initial = self.basis_quality_profile(basis)
        n = np.asarray(basis).shape[0]
        sizes = self._progressive_block_sizes(n, block_size) if progressive else [block_size]
        cur = _as_int_rows(basis)
        Utot = np.eye(n, dtype=object)
        reducer_stats = []
        for beta in sizes:
            cur, U, st = bkz(cur, min(int(beta), n))
            Utot = U.astype(object) @ Utot
            reducer_stats.append(st)
        final = self.basis_quality_profile(cur)
        dt = time.time() - start
        diag = {"time": dt, "block_size": block_size, "block_sizes_used": sizes, "strategy": strategy,
                "initial_quality": initial, "final_quality": final,
                "improvement_factor": initial["hermite_factor"] / final["hermite_factor"], "reducer": reducer_stats}
        self._account(dt, diag["improvement_factor"])
        return cur.astype(np.float64), np.array(Utot, dtype=np.float64), diag

    @staticmethod
    def _progressive_block_sizes(n: int, target: int) -> List[int]:
        """reduction.py:302-312."""
        sizes = []
        beta = min(20, n)
        while beta < target:
            sizes.append(beta)
            beta = min(beta + 10, target)
        sizes.append(target)
        return sizes

    # ---- 3. quality (reduction.py:322-405) ----
    def hermite_factor(self, basis) -> float:
        B = np.asarray(basis, dtype=np.float64)
        n = B.shape[0]
        logdet = float(np.sum(np.log(_float_gs_norms(B))))   # |det| = prod ||b*_i||, overflow-free
        return float(np.linalg.norm(B[0]) / math.exp(logdet / n))

    def orthogonality_defect(self, basis) -> float:
        B = np.asarray(basis, dtype=np.float64)
        logdef = float(np.sum(np.log(np.linalg.norm(B, axis=1))) - np.sum(np.log(_float_gs_norms(B))))
        return float(math.exp(logdef)) if logdef < 700 else float("inf")

    def basis_quality_profile(self, basis) -> Dict[str, Any]:
        B = np.asarray(basis, dtype=np.float64)
        try:
            gs = gso_norms(B)
        except ValueError:                     # real-valued basis: float Gram-Schmidt
            gs = _float_gs_norms(B)
        gs_list = [float(x) for x in gs]
        return {"dimension": B.shape[0], "hermite_factor": self.hermite_factor(B),
                "orthogonality_defect": self.orthogonality_defect(B), "gs_norms": gs_list,
                "max_gs_norm": max(gs_list), "min_gs_norm": min(gs_list), "gs_ratio": max(gs_list) / min(gs_list),
                "log_potential": float(sum(math.log(x) for x in gs_list))}

    # ---- 4. sampling-oriented (reduction.py:409-535) ----
    def sampling_reduce(self, basis, target_sigma: float, max_iterations: int = 10,
                        lll_deltas=(0.75, 0.85, 0.95, 0.99), bkz_blocks=(20, 30, 40)):
        B = _as_int_rows(basis)
        n = B.shape[0]
        best, best_max = B.astype(np.float64), float("inf")
        log: List[Dict[str, Any]] = []
        for it in range(max_iterations):
            cands = [("LLL", d, self.lll_reduce(B, delta=d)[0]) for d in lll_deltas]
            cands += [("BKZ", b, self.bkz_reduce(B, block_size=b)[0]) for b in bkz_blocks if b < n]
            for method, param, cand in cands:
                mx = float(gso_norms(cand).max())
                if mx < target_sigma * math.sqrt(n) and mx < best_max:
                    # the reference logs best/new AFTER the assignment, i.e. always 1.0 (reduction.py:457-465)
                    best, best_max = cand, mx
                    log.append({"iteration": it, "method": method, "parameter": param, "max_gs_norm": mx,
                                "improvement": best_max / mx})
            B = _as_int_rows(self.gram_schmidt_improve(best))
            if len(log) > 1 and log[-1]["improvement"] < 1.01:
                break
        info = {"iterations": len(log), "final_max_norm": best_max,
                "suitable_for_sigma": best_max < target_sigma * math.sqrt(n), "optimization_log": log,
                "final_profile": self.basis_quality_profile(best)}
        return np.asarray(best, dtype=np.float64), info

    def gram_schmidt_improve(self, basis, max_swaps: int = 100):
        """Greedy adjacent swaps where ||b*_i|| > ||b*_{i+1}|| by the largest ratio (reduction.py:491-535)."""
        B = np.array(basis, dtype=np.float64)
        n = B.shape[0]
        for _ in range(max_swaps):
            r = _float_gs_norms(B) ** 2
            ratios = r[:-1] / r[1:]
            i = int(np.argmax(ratios)) if n > 1 else 0
            if n > 1 and ratios[i] > 1.01:
                B[[i, i + 1]] = B[[i + 1, i]]
            else:
                break
        return B

    # ---- 5. analysis (reduction.py:581-670) ----
    def estimate_reduction_cost(self, basis, block_size: int, model: str = "bkz2") -> Dict[str, float]:
        n = np.asarray(basis).shape[0]
        if model == "bkz2":
            ops = (n / block_size) * 2 ** (0.292 * block_size)
        elif model == "classical":
            ops = 2 ** (0.292 * block_size + 16.4)
        else:
            ops = 2 ** (0.265 * block_size)
        secs = ops / 1e9
        return {"model": model, "block_size": block_size, "dimension": n, "operations": ops,
                "log2_operations": float(math.log2(ops)), "estimated_seconds": secs, "estimated_hours": secs / 3600}

    def compare_bases(self, basis1, basis2, labels: Tuple[str, str] = ("Before", "After")) -> Dict:
        p1, p2 = self.basis_quality_profile(basis1), self.basis_quality_profile(basis2)
        out: Dict[str, Any] = {"labels": labels}
        for key in ("hermite_factor", "orthogonality_defect", "max_gs_norm", "gs_ratio"):
            out[key] = {labels[0]: p1[key], labels[1]: p2[key], "improvement": p1[key] / p2[key]}
        return out

    def _account(self, dt: float, factor: float):
        self.stats["reductions_performed"] += 1
        self.stats["total_time"] += dt
        self.stats["quality_improvements"].append(factor)


def optimal_delta_for_sampling(n: int, target_sigma: float) -> float:
    """reduction.py:701-721."""
    if n < 50:
        return 0.99
    if n < 100:
        return 0.95
    if n < 200:
        return 0.85
    return 0.75


def reduction_strategy_for_lattice(lattice_type: str, n: int) -> Dict[str, Any]:
    """reduction.py:724-764."""
    strategies = {
        "random": {"algorithm": "LLL", "delta": 0.99, "use_bkz": n < 100, "bkz_block": min(40, n // 2)},
        "qary": {"algorithm": "BKZ", "delta": 0.95, "use_bkz": True, "bkz_block": min(50, n // 2), "progressive": True},
        "ntru": {"algorithm": "BKZ", "delta": 0.99, "use_bkz": True, "bkz_block": min(60, n // 3), "progressive": True},
        "ideal": {"algorithm": "LLL", "delta": 0.85, "use_bkz": n < 200, "bkz_block": min(30, n // 4)},
    }
    return strategies.get(lattice_type, strategies["random"])
