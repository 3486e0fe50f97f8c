"""ctypes front-end of the CPU oracle (oracle/lgs_oracle.c).

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
--impl reference legs.  The product package never imports this module.

Each function names the reference code it restates (paths relative to the reference tree):
  precompute_qr      src/samplers/klein.py:56-79   (_precompute_qr_stable)
  klein_replay       src/samplers/klein.py:181-220 (sample_single) x S, 222-271 (compute_log_density)
  imhk_replay        src/samplers/imhk.py:126-177  (_initialize_chain, step), 196-229 (sample)
  uniform_stream     numpy legacy RandomState(seed).random_sample(N) -- the "recorded uniform stream"
  conditional_counts src/samplers/klein.py:189-195 (conditional means) + 101-170 (the table) applied to sampler OUTPUT
"""
from __future__ import annotations

import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "liblgs_oracle.so")
_lib = None

_f64p = ctypes.POINTER(ctypes.c_double)
_i64p = ctypes.POINTER(ctypes.c_int64)
_u8p = ctypes.POINTER(ctypes.c_uint8)


def build(force: bool = False) -> str:
    """Compile the oracle with the recipe in oracle/Makefile (gcc, no GPU needed)."""
    src = os.path.join(_HERE, "lgs_oracle.c")
    stale = (not os.path.exists(_LIB_PATH)) or os.path.getmtime(_LIB_PATH) < os.path.getmtime(src)
    if force or stale:
        subprocess.check_call(["make", "-C", _HERE, "-B", "liblgs_oracle.so"], stdout=subprocess.DEVNULL)
    return _LIB_PATH


def lib():
    global _lib
    if _lib is None:
        build()
        _lib = ctypes.CDLL(_LIB_PATH)
        _lib.lgs_oracle_np_sum.restype = ctypes.c_double
    return _lib


def _p(a, typ):
    return None if a is None else a.ctypes.data_as(typ)


def precompute_qr(basis: np.ndarray, center: np.ndarray):
    """klein.py:56-79: Householder QR of the column basis, rows of R flipped to a positive diagonal."""
    basis = np.ascontiguousarray(basis, dtype=np.float64)
    Q, R = np.linalg.qr(basis, mode="reduced")
    for i in range(basis.shape[0]):
        if R[i, i] < 0:
            R[i, :] *= -1
            Q[:, i] *= -1
    t = Q.T @ np.asarray(center, dtype=np.float64)
    return np.ascontiguousarray(Q), np.ascontiguousarray(R), np.ascontiguousarray(t)


def uniform_stream(seed: int, count: int) -> np.ndarray:
    out = np.empty(int(count), dtype=np.float64)
    lib().lgs_oracle_uniform_stream(ctypes.c_uint32(seed), ctypes.c_int64(count), _p(out, _f64p))
    return out


def np_sum(a: np.ndarray) -> float:
    a = np.ascontiguousarray(a, dtype=np.float64)
    return float(lib().lgs_oracle_np_sum(_p(a, _f64p), ctypes.c_int64(a.size)))


def samplez(mean: float, sigma: float, u: float, precision: int = 10):
    out = ctypes.c_int64(0)
    used = lib().lgs_oracle_samplez(ctypes.c_double(mean), ctypes.c_double(sigma), ctypes.c_int(precision),
                                    ctypes.c_double(u), ctypes.byref(out))
    return int(out.value), bool(used)


def table(mean: float, sigma: float, precision: int = 10):
    logp = np.empty(1002, dtype=np.float64)
    lower = ctypes.c_int64(0)
    ln = lib().lgs_oracle_table(ctypes.c_double(mean), ctypes.c_double(sigma), ctypes.c_int(precision),
                                ctypes.byref(lower), _p(logp, _f64p))
    return np.arange(lower.value, lower.value + ln), logp[:ln].copy()


def uniforms_per_sample(R: np.ndarray, sigma: float) -> int:
    R = np.ascontiguousarray(R, dtype=np.float64)
    return int(lib().lgs_oracle_uniforms_per_sample(ctypes.c_int(R.shape[0]), _p(R, _f64p), ctypes.c_double(sigma)))


class Problem:
    """Everything the sampler derives from (basis, sigma, center) on the host."""

    def __init__(self, basis, sigma, center=None, precision=10):
        self.B = np.ascontiguousarray(basis, dtype=np.float64)
        self.n = self.B.shape[0]
        self.sigma = float(sigma)
        self.center = np.zeros(self.n) if center is None else np.ascontiguousarray(center, dtype=np.float64)
        self.precision = int(precision)
        self.Q, self.R, self.t = precompute_qr(self.B, self.center)
        self.ups = uniforms_per_sample(self.R, self.sigma)

    def _args(self):
        return (ctypes.c_int(self.n), _p(self.R, _f64p), _p(self.B, _f64p), _p(self.t, _f64p),
                _p(self.center, _f64p), ctypes.c_double(self.sigma), ctypes.c_int(self.precision))


def klein_replay(prob: Problem, U: np.ndarray, S: int, want_v=True, want_logq=True, exact_logz=True, threads=0):
    """S samples of sample_single(); sample s consumes U[s*ups:(s+1)*ups]."""
    U = np.ascontiguousarray(U, dtype=np.float64)
    assert U.size >= S * prob.ups, "uniform stream too short"
    X = np.zeros((S, prob.n), dtype=np.int64)
    V = np.zeros((S, prob.n), dtype=np.float64) if want_v else None
    logq = np.zeros(S, dtype=np.float64) if want_logq else None
    rc = lib().lgs_oracle_klein_replay(*prob._args(), ctypes.c_int64(S), _p(U, _f64p), _p(X, _i64p),
                                       _p(V, _f64p), _p(logq, _f64p), ctypes.c_int(int(exact_logz)),
                                       ctypes.c_int(threads))
    if rc != 0:
        raise RuntimeError("oracle: a draw took the no-uniform fallback; stream layout undefined")
    return X, V, logq


def conditional_counts(prob: Problem, X: np.ndarray, K: int = 6, FB: int = 4, threads=0, width=None):
    """Per coordinate, per bin of frac(m_i) and per clipped offset x_i - round(m_i): observed counts in the sampler
    output X [S, n] and the counts the reference's own table (klein.py:101-170) expects at the recomputed conditional
    means (klein.py:189-195).  `width` (int32 [n], optional) coarsens the offsets of a coordinate into bins of that many
    integers.  Returns (obs int64 [n, FB, 2K+1], exp float64 [n, FB, 2K+1])."""
    X = np.ascontiguousarray(X, dtype=np.int64)
    if width is not None:
        width = np.ascontiguousarray(width, dtype=np.int32)
        assert width.shape == (prob.n,)
    S = X.shape[0]
    obs = np.zeros((prob.n, FB, 2 * K + 1), dtype=np.int64)
    exp = np.zeros((prob.n, FB, 2 * K + 1), dtype=np.float64)
    rc = lib().lgs_oracle_conditional_counts(*prob._args(), ctypes.c_int64(S), _p(X, _i64p), ctypes.c_int(K),
                                             ctypes.c_int(FB), _p(width, ctypes.POINTER(ctypes.c_int32)), _p(obs, _i64p),
                                             _p(exp, _f64p), ctypes.c_int(threads))
    if rc != 0:
        raise MemoryError("oracle: conditional_counts could not allocate its accumulators")
    return obs, exp


def klein_seeded(prob: Problem, S: int, seed: int, chains: int = 1, want_v=True, threads=0):
    X = np.zeros((S, prob.n), dtype=np.int64)
    V = np.zeros((S, prob.n), dtype=np.float64) if want_v else None
    lib().lgs_oracle_klein_seeded(*prob._args(), ctypes.c_int64(S), ctypes.c_uint32(seed), ctypes.c_int(chains),
                                  _p(X, _i64p), _p(V, _f64p), ctypes.c_int(threads))
    return X, V


def imhk_stream_len(prob: Problem, steps: int) -> int:
    return prob.ups + steps * (prob.ups + 1)


def imhk_replay(prob: Problem, U: np.ndarray, chains: int, steps: int, thin: int = 1, want_v=True,
                exact_logz=True, threads=0):
    """`chains` fresh IMHKSampler objects, each: _initialize_chain() + `steps` step()s, state kept every `thin`."""
    U = np.ascontiguousarray(U, dtype=np.float64)
    per = imhk_stream_len(prob, steps)
    assert U.size >= chains * per, "uniform stream too short"
    n_keep = steps // thin
    Xk = np.zeros((chains, n_keep, prob.n), dtype=np.int64)
    Vk = np.zeros((chains, n_keep, prob.n), dtype=np.float64) if want_v else None
    acc = np.zeros((chains, steps), dtype=np.uint8)
    logw = np.zeros((chains, steps + 1), dtype=np.float64)
    tot = np.zeros(chains, dtype=np.int64)
    rc = lib().lgs_oracle_imhk_replay(*prob._args(), ctypes.c_int64(chains), ctypes.c_int64(steps),
                                      ctypes.c_int64(thin), _p(U, _f64p), _p(Xk, _i64p), _p(Vk, _f64p),
                                      _p(acc, _u8p), _p(logw, _f64p), _p(tot, _i64p),
                                      ctypes.c_int(int(exact_logz)), ctypes.c_int(threads))
    if rc != 0:
        raise RuntimeError("oracle: a draw took the no-uniform fallback; stream layout undefined")
    return {"X": Xk, "V": Vk, "accept": acc, "logw": logw, "accepted": tot}


def imhk_seeded(prob: Problem, chains: int, steps: int, seed: int, exact_logz=False, threads=0):
    Xl = np.zeros((chains, prob.n), dtype=np.int64)
    tot = np.zeros(chains, dtype=np.int64)
    lib().lgs_oracle_imhk_seeded(*prob._args(), ctypes.c_int64(chains), ctypes.c_int64(steps),
                                 ctypes.c_uint32(seed), _p(Xl, _i64p), _p(tot, _i64p),
                                 ctypes.c_int(int(exact_logz)), ctypes.c_int(threads))
    return Xl, tot
