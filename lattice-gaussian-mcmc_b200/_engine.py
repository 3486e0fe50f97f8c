"""One lgs_handle (include/lgs_b200.h) behind a small Python object.

Host forms take/return numpy arrays (what the reference-facing classes use); device forms take raw device pointers
(`tensor.data_ptr()` of torch CUDA tensors) and a CUDA stream handle, for callers that keep samples resident on the
GPU (bench.py, the multi-GPU driver).  No sampling happens in Python and nothing here falls back to the CPU.
"""
from __future__ import annotations

import ctypes
import mmap
import os
import threading
import weakref

import numpy as np

from . import _capi
from ._capi import PATH, SAMPLEZ, WEIGHT, LgsDesc, c_f64p, check, lib


def precompute_qr(basis: np.ndarray, center: np.ndarray):
    """RefinedKleinSampler._precompute_qr_stable (klein.py:56-79): LAPACK Householder QR of the column basis, rows of
    R (and columns of Q) flipped so that diag(R) > 0, center carried into the Q frame.  Host-side, once."""
    Q, R = np.linalg.qr(basis, mode="reduced")
    n = basis.shape[0]
    for i in range(n):
        if R[i, i] < 0:
            R[i, :] *= -1
            Q[:, i] *= -1
    t = Q.T @ center
    return Q, R, t


_HUGE = 2 << 20


class _Block:
    """Owner of one result buffer: exports the mapping through the buffer protocol (PEP 688), so that every numpy
    array and view made from it keeps it alive; when the last one dies the mapping goes back to the cache."""
    __slots__ = ("mm", "off", "nbytes", "__weakref__")

    def __init__(self, mm, off, nbytes):
        self.mm, self.off, self.nbytes = mm, off, nbytes

    def __buffer__(self, flags):
        return memoryview(self.mm)[self.off:self.off + self.nbytes]


class _HostArrayCache:
    """Recycles the host memory of large result arrays.  The reference returns a NEW array per call (klein.py:322);
    a new 1.4 GB array means 350 k page faults (half the cost of KleinSampler.sample(170496) in round 1).  Arrays
    handed out here are new objects over 2 MB-aligned anonymous mappings (MADV_HUGEPAGE); when the caller drops an
    array -- and every view of it -- its mapping is kept and backs a later result of the same size, already faulted in.
    What is retained is bounded twice: at most LGS_HOST_CACHE_ARRAYS mappings (default 2: a caller that drops each
    result before asking for the next needs one) and at most LGS_HOST_CACHE_MB in total (default 4096); when a
    mapping of a new size comes back and there is no room, the OLDEST retained ones are unmapped, so a caller that
    varies its batch size does not accumulate mappings nobody will ask for again.  With 8 ranks on one host the
    worst case is 8 x the per-process bound: lower LGS_HOST_CACHE_MB there if host memory is tight.
    LGS_HOST_CACHE_MB=0 disables the cache (np.empty)."""

    def __init__(self):
        self.free = []      # [(nbytes, mm, off)], oldest first
        self.cached = 0
        self.lock = threading.Lock()
        self.limit = int(os.environ.get("LGS_HOST_CACHE_MB", "4096")) << 20
        self.max_arrays = max(int(os.environ.get("LGS_HOST_CACHE_ARRAYS", "2")), 0)

    def _release(self, mm, off, nbytes):
        evict = []
        with self.lock:
            if nbytes <= self.limit and self.max_arrays > 0:
                while self.free and (self.cached + nbytes > self.limit or len(self.free) >= self.max_arrays):
                    old = self.free.pop(0)
                    self.cached -= old[0]
                    evict.append(old[1])
                self.free.append((nbytes, mm, off))
                self.cached += nbytes
            else:
                evict.append(mm)
        for m in evict:
            try:
                m.close()
            except BufferError:
                pass

    def empty(self, shape, dtype=np.float64):
        count = int(np.prod(shape))
        nbytes = count * np.dtype(dtype).itemsize
        if nbytes < (64 << 20) or self.limit <= 0:
            return np.empty(shape, dtype=dtype)
        hit = None
        with self.lock:
            for k in range(len(self.free) - 1, -1, -1):
                if self.free[k][0] == nbytes:
                    _, mm, off = self.free.pop(k)
                    hit = (mm, off)
                    self.cached -= nbytes
                    break
        try:
            if hit:
                mm, off = hit
            else:
                mm = mmap.mmap(-1, nbytes + _HUGE)
                addr = ctypes.addressof(ctypes.c_char.from_buffer(mm))
                off = (-addr) % _HUGE
                if hasattr(mmap, "MADV_HUGEPAGE") and os.environ.get("LGS_FRESH_HUGEPAGE", "1") != "0":
                    mm.madvise(mmap.MADV_HUGEPAGE)
            blk = _Block(mm, off, nbytes)
            weakref.finalize(blk, self._release, mm, off, nbytes)
            return np.frombuffer(blk, dtype=dtype, count=count).reshape(shape)
        except (OSError, ValueError, BufferError, TypeError):
            return np.empty(shape, dtype=dtype)


_host_cache = _HostArrayCache()


def fresh_array(shape, dtype=np.float64):
    """A new writeable C-contiguous array for a result the device pipeline fills (see _HostArrayCache)."""
    return _host_cache.empty(shape, dtype)


def _ptr(a, typ=ctypes.c_void_p):
    return None if a is None else a.ctypes.data_as(typ)


class Engine:
    def __init__(self, R, B, t, center, sigma, precision=10, max_range=1000, device=0, path="auto"):
        self.n = int(R.shape[0])
        # keep the uploads alive until lgs_create returns (it copies)
        R = np.ascontiguousarray(R, dtype=np.float64)
        B = np.ascontiguousarray(B, dtype=np.float64)
        t = np.ascontiguousarray(t, dtype=np.float64)
        center = np.ascontiguousarray(center, dtype=np.float64)
        desc = LgsDesc(self.n, _ptr(R, c_f64p), _ptr(B, c_f64p), _ptr(t, c_f64p), _ptr(center, c_f64p),
                       float(sigma), int(precision), int(max_range), int(device), PATH[path])
        h = ctypes.c_void_p()
        self._lib = lib()
        self._h = None
        check(None, self._lib.lgs_create(ctypes.byref(desc), ctypes.byref(h)))
        self._h = h
        self.device = int(device)
        self.ups = int(self._lib.lgs_uniforms_per_sample(self._h))
        self.preferred_batch = int(self._lib.lgs_preferred_batch(self._h))
        self.chains = 0

    @property
    def replay_on_chain(self) -> bool:
        """True when the replay entry points of this handle run on the throughput (chain) kernel (lgs_replay_path)"""
        return int(self._lib.lgs_replay_path(self._h)) == PATH["blocked"]

    def close(self):
        if getattr(self, "_h", None) is not None:
            self._lib.lgs_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ------------------------------------------------------------------ Klein, host buffers
    def klein_sample(self, S, seed, first_id, samplez="fast", want_v=True, want_x=False, want_logq=False, out=None):
        """`out`: a caller-owned C-contiguous float64 [S, n] array for the lattice points (reused between calls it
        saves the page faults of a fresh array: about 2x at n = 1024)."""
        S = int(S)
        if out is not None:
            if not (isinstance(out, np.ndarray) and out.dtype == np.float64 and out.shape == (S, self.n)
                    and out.flags.c_contiguous and out.flags.writeable):
                raise ValueError(f"out must be a writeable C-contiguous float64 array of shape ({S}, {self.n})")
            want_v = True
        V = out if out is not None else (fresh_array((S, self.n)) if want_v else None)
        X = np.empty((S, self.n), dtype=np.int64) if want_x else None
        lq = np.empty(S, dtype=np.float64) if want_logq else None
        check(self._h, self._lib.lgs_klein_sample_host(self._h, S, int(seed) & (2**64 - 1), int(first_id),
                                                       SAMPLEZ[samplez], _ptr(V), _ptr(X), _ptr(lq)))
        return V, X, lq

    def klein_replay(self, U, S, want_v=True, want_x=False, want_logq=False):
        S = int(S)
        U = np.ascontiguousarray(U, dtype=np.float64)
        if U.size < S * self.ups:
            raise ValueError(f"uniform stream too short: need {S * self.ups}, got {U.size}")
        V = np.empty((S, self.n), dtype=np.float64) if want_v else None
        X = np.empty((S, self.n), dtype=np.int64) if want_x else None
        lq = np.empty(S, dtype=np.float64) if want_logq else None
        fb = np.zeros(1, dtype=np.int32)
        check(self._h, self._lib.lgs_klein_replay_host(self._h, S, _ptr(U), _ptr(V), _ptr(X), _ptr(lq), _ptr(fb)))
        return V, X, lq, int(fb[0])

    def klein_logdensity(self, X):
        X = np.ascontiguousarray(X, dtype=np.int64).reshape(-1, self.n)
        out = np.empty(X.shape[0], dtype=np.float64)
        check(self._h, self._lib.lgs_klein_logdensity_host(self._h, X.shape[0], _ptr(X), _ptr(out)))
        return out

    # ------------------------------------------------------------------ IMHK, host buffers
    def imhk_reset(self, chains, seed, first_chain_id=0):
        check(self._h, self._lib.lgs_imhk_reset(self._h, int(chains), int(seed) & (2**64 - 1), int(first_chain_id)))
        self.chains = int(chains)

    def imhk_run(self, burn_in, keep, thin, samplez="fast", weight="reference", want_v=True, want_x=False):
        C = self.chains
        V = np.empty((C, keep, self.n), dtype=np.float64) if want_v else None
        X = np.empty((C, keep, self.n), dtype=np.int64) if want_x else None
        check(self._h, self._lib.lgs_imhk_run_host(self._h, int(burn_in), int(keep), int(thin), SAMPLEZ[samplez],
                                                   WEIGHT[weight], _ptr(V), _ptr(X)))
        return V, X

    def imhk_replay(self, U, stride, burn_in, keep, thin, weight="reference", want_v=True, want_x=False):
        C = self.chains
        steps = int(burn_in) + int(keep) * int(thin)
        U = np.ascontiguousarray(U, dtype=np.float64)
        if U.size < C * stride:
            raise ValueError("uniform stream too short")
        V = np.empty((C, keep, self.n), dtype=np.float64) if want_v else None
        X = np.empty((C, keep, self.n), dtype=np.int64) if want_x else None
        acc = np.empty((C, steps), dtype=np.uint8)
        lw = np.empty((C, steps + 1), dtype=np.float64)
        fb = np.zeros(1, dtype=np.int32)
        check(self._h, self._lib.lgs_imhk_replay_host(self._h, int(burn_in), int(keep), int(thin), WEIGHT[weight],
                                                      _ptr(U), int(stride), _ptr(V), _ptr(X), _ptr(acc), _ptr(lw),
                                                      _ptr(fb)))
        return {"V": V, "X": X, "accept": acc, "logw": lw, "fallbacks": int(fb[0])}

    def imhk_state(self):
        C = self.chains
        counts = np.empty((C, 2), dtype=np.int64)
        X = np.empty((C, self.n), dtype=np.int64)
        V = np.empty((C, self.n), dtype=np.float64)
        lw = np.empty(C, dtype=np.float64)
        check(self._h, self._lib.lgs_imhk_get_state_host(self._h, _ptr(counts), _ptr(X), _ptr(V), _ptr(lw)))
        return counts, X, V, lw

    def imhk_chain_counts(self):
        """int64 [chains, 2] = (accepted, total) per chain; nothing else is computed or copied"""
        counts = np.empty((self.chains, 2), dtype=np.int64)
        check(self._h, self._lib.lgs_imhk_get_state_host(self._h, _ptr(counts), None, None, None))
        return counts

    def imhk_steps_done(self) -> int:
        return int(self._lib.lgs_imhk_steps_done(self._h))

    def imhk_set_state(self, X, logw, counts=None, steps_done=0, samplez="fast"):
        """lgs_imhk_set_state_host: install a chain state taken with imhk_state() / imhk_steps_done()"""
        C = self.chains
        X = np.ascontiguousarray(X, dtype=np.int64).reshape(C, self.n)
        logw = np.ascontiguousarray(logw, dtype=np.float64).reshape(C)
        if counts is not None:
            counts = np.ascontiguousarray(counts, dtype=np.int64).reshape(C, 2)
        check(self._h, self._lib.lgs_imhk_set_state_host(self._h, SAMPLEZ[samplez], _ptr(X), _ptr(logw), _ptr(counts),
                                                         int(steps_done)))

    def imhk_counts(self):
        a, t = ctypes.c_int64(0), ctypes.c_int64(0)
        check(self._h, self._lib.lgs_imhk_counts_host(self._h, ctypes.byref(a), ctypes.byref(t)))
        return int(a.value), int(t.value)

    # ------------------------------------------------------------------ measurement hooks
    PROF_CLASSES = ("panel", "centers", "accum", "point", "other", "chain")

    def profile_enable(self, on=True):
        check(self._h, self._lib.lgs_profile_enable(self._h, int(bool(on))))

    def profile_read(self):
        ms = (ctypes.c_double * len(self.PROF_CLASSES))()
        cnt = (ctypes.c_int64 * len(self.PROF_CLASSES))()
        check(self._h, self._lib.lgs_profile_read(self._h, ms, cnt))
        return {k: (float(ms[i]), int(cnt[i])) for i, k in enumerate(self.PROF_CLASSES)}

    # ------------------------------------------------------------------ device-pointer forms
    def klein_sample_dev(self, S, seed, first_id, samplez="fast", V=0, X=0, logq=0, stream=0):
        check(self._h, self._lib.lgs_klein_sample(self._h, int(S), int(seed) & (2**64 - 1), int(first_id),
                                                  SAMPLEZ[samplez], V or None, X or None, logq or None,
                                                  stream or None))

    def klein_replay_dev(self, S, U, V=0, X=0, logq=0, fallback=0, stream=0):
        check(self._h, self._lib.lgs_klein_replay(self._h, int(S), U, V or None, X or None, logq or None,
                                                  fallback or None, stream or None))

    def imhk_run_dev(self, burn_in, keep, thin, samplez="fast", weight="reference", V=0, X=0, stream=0):
        check(self._h, self._lib.lgs_imhk_run(self._h, int(burn_in), int(keep), int(thin), SAMPLEZ[samplez],
                                              WEIGHT[weight], V or None, X or None, stream or None))

    def imhk_state_dev(self, counts=0, X=0, V=0, logw=0, stream=0):
        check(self._h, self._lib.lgs_imhk_get_state(self._h, counts or None, X or None, V or None, logw or None,
                                                    stream or None))

    DIAG_FIELDS = ("mean", "var", "tau_int", "window", "acf1")

    def chain_diagnostics_dev(self, V, chains, keep, dims, stats, c_window=5.0, max_lag=0, acf=0, acf_len=0, stream=0):
        """stats[chains][len(dims)][5] (device pointer) <- per-chain mean / var / tau_int / window / acf[1] of the
        tracked coordinates of V[chains][keep][n] (device pointer); see lgs_chain_diagnostics in include/lgs_b200.h"""
        dims = np.ascontiguousarray(dims, dtype=np.int32)
        check(self._h, self._lib.lgs_chain_diagnostics(self._h, V, int(chains), int(keep),
                                                       dims.ctypes.data_as(_capi.c_i32p), dims.size, float(c_window),
                                                       int(max_lag), stats, acf or None, int(acf_len), stream or None))

    def chain_diagnostics_static_dev(self, V, chains, keep, dims, stats, c_window=5.0, max_lag=0, acf=0, acf_len=0,
                                     stream=0):
        """chain_diagnostics_dev for callers that must not allocate device memory: same results, work space in the
        library's fixed 32 MiB device array (chains * len(dims) * keep <= LGS_DIAG_STATIC_MAX_DOUBLES), one call at a
        time per device; see lgs_chain_diagnostics_static in include/lgs_b200.h"""
        dims = np.ascontiguousarray(dims, dtype=np.int32)
        check(self._h, self._lib.lgs_chain_diagnostics_static(self._h, V, int(chains), int(keep),
                                                              dims.ctypes.data_as(_capi.c_i32p), dims.size,
                                                              float(c_window), int(max_lag), stats, acf or None,
                                                              int(acf_len), stream or None))

    def gelman_rubin_sums_dev(self, stats, chains, n_dims, sums, stream=0):
        check(self._h, self._lib.lgs_gelman_rubin_sums(self._h, stats, int(chains), int(n_dims), sums, stream or None))

    def attach_histogram(self, dims, lo, width, nbins, hist):
        """Count the tracked coordinates of every lattice point the chain kernel writes to a device-resident V into
        hist (device pointer to int64[len(dims)][nbins], accumulated) from inside the sampling kernel; hist = 0
        detaches.  See lgs_attach_histogram in include/lgs_b200.h."""
        dims = np.ascontiguousarray(dims if dims is not None else [], dtype=np.int32)
        check(self._h, self._lib.lgs_attach_histogram(self._h, dims.ctypes.data_as(_capi.c_i32p), dims.size,
                                                      float(lo), float(width), int(nbins), hist or None))

    def histogram_dev(self, V, S, dims, lo, width, nbins, hist, stream=0):
        dims = np.ascontiguousarray(dims, dtype=np.int32)
        check(self._h, self._lib.lgs_histogram(self._h, V, int(S), dims.ctypes.data_as(_capi.c_i32p), dims.size,
                                               float(lo), float(width), int(nbins), hist, stream or None))
