"""Drop-in for the reference's RefinedKleinSampler / KleinSampler (src/samplers/klein.py:21-353).

Same constructor, attributes and methods; the sampling itself is the CUDA path behind include/lgs_b200.h.

Host side (once, numpy, as the reference does it):   QR of `lattice.basis` (klein.py:56-79).
Device side (every call):                            sample_single / sample / compute_log_density.

Two modes (keyword-only, reference-preserving defaults for everything else):

  mode="native"  (default)  Philox4x32-10 counter RNG keyed by (seed, global sample id); output is distributed as the
                            reference's, but is not the same stream of points.
  mode="replay"             consumes the reference's uniform stream: by default it draws the uniforms from numpy's
                            global legacy RandomState, so `np.random.seed(s); KleinSampler(L, sigma, mode="replay")
                            .sample(N)` returns the very points the reference returns after `np.random.seed(s)`
                            (integer coefficients bit-exact); `uniforms=` supplies a recorded stream instead.
"""
from __future__ import annotations

import logging
from typing import Any, Dict, Optional

import numpy as np

from .._engine import Engine, precompute_qr
from .base import DiscreteGaussianSampler, lattice_basis

logger = logging.getLogger(__name__)


def native_seed(seed, mode: str) -> int:
    """The 64-bit Philox key of a native-mode sampler.  An explicit `seed` is used as is.  `seed=None` (the default)
    draws the key from numpy's GLOBAL legacy generator at construction -- the stream the reference itself samples
    from -- so that the reference's own worker pattern keeps its meaning on the drop-in:
        np.random.seed(seed + 1000 * chain_id); sampler = KleinSampler(...); sampler.sample(n)
    (experiments/dimension_scaling.py:877-884, convergence_study.py:559-562) gives different chains for different
    chain_id, the same chain for the same seed, and two samplers built one after the other differ.  Replay mode
    consumes the global stream draw by draw as the reference does and must not be disturbed: no key is drawn there."""
    if seed is not None:
        return int(seed)
    if mode != "native":
        return 0
    hi, lo = np.random.randint(0, 2 ** 32, size=2, dtype=np.uint64)
    return (int(hi) << 32) | int(lo)


class RefinedKleinSampler(DiscreteGaussianSampler):
    """Klein's randomized nearest-plane sampler (klein.py:21-24), batched on a B200."""

    def __init__(self, lattice, sigma: float, center: Optional[np.ndarray] = None, precision: int = 10,
                 use_log_space: bool = True, *, mode: str = "native", seed: Optional[int] = None, uniforms=None,
                 samplez: Optional[str] = None, device: int = 0, path: str = "auto", max_range: int = 1000,
                 first_sample_id: int = 0):
        super().__init__(lattice, sigma, center)
        if mode not in ("native", "replay"):
            raise ValueError("mode must be 'native' or 'replay'")
        self.precision = precision
        self.use_log_space = use_log_space
        self.mode = mode
        self.seed = native_seed(seed, mode)
        self.samplez = samplez or ("table" if mode == "replay" else "fast")
        self._basis = lattice_basis(lattice)
        self._precompute_qr_stable()
        self._sample_cache = {}          # kept for interface parity (klein.py:45-46); the GPU path has no cache
        self._max_cache_size = 10000
        self._validate_klein_parameters()
        self._log_2pi = np.log(2 * np.pi)
        self._log_sigma = np.log(self.sigma)
        self._engine_kwargs = dict(max_range=max_range, device=device, path=path)
        self._alt_engines = {}           # precision -> Engine, for adaptive_precision_sample
        self._engine = Engine(self.R, self._basis, self.center_transformed, self.center, self.sigma,
                              precision=precision, **self._engine_kwargs)
        # Philox counter of the next sample: sample i of this sampler is a function of (seed, first_sample_id + i) only,
        # so ranks / workers sharing a seed take disjoint ranges (rank r: first_sample_id = r * samples_per_rank)
        self._next_sample_id = int(first_sample_id)
        self._uniforms = None if uniforms is None else np.ascontiguousarray(uniforms, dtype=np.float64).ravel()
        self._upos = 0

    # ------------------------------------------------------------------ host precompute (klein.py:56-99)
    def _precompute_qr_stable(self):
        self.Q, self.R, self.center_transformed = precompute_qr(self._basis, self.center)
        self.P = np.arange(self.dimension)
        condition_number = np.linalg.cond(self.R)
        if condition_number > 1e10:
            logger.warning(f"Poorly conditioned basis: condition number = {condition_number:.2e}")
        self.R_diag = np.diag(self.R)
        self.log_R_diag = np.log(np.abs(self.R_diag) + 1e-300)

    def _validate_klein_parameters(self):
        min_gs_norm = getattr(self.lattice, "min_gram_schmidt_norm", None)
        if min_gs_norm is None:
            min_gs_norm = float(np.min(np.abs(self.R_diag)))
        klein_lower = min_gs_norm / np.sqrt(2 * np.log(self.dimension + 1))
        if self.sigma < klein_lower * 0.9:
            logger.warning(f"σ={self.sigma:.4f} is below Klein's requirement (minimum ≈ {klein_lower:.4f}). "
                           "Sampling may fail.")

    # ------------------------------------------------------------------ uniform stream for replay mode
    def set_uniform_stream(self, uniforms):
        self._uniforms = np.ascontiguousarray(uniforms, dtype=np.float64).ravel()
        self._upos = 0

    def _take_uniforms(self, count: int) -> np.ndarray:
        if self._uniforms is None:
            return np.random.random_sample(count)  # the global legacy stream the reference itself consumes
        if self._upos + count > self._uniforms.size:
            raise ValueError("recorded uniform stream exhausted")
        u = self._uniforms[self._upos:self._upos + count]
        self._upos += count
        return u

    # ------------------------------------------------------------------ sampling (klein.py:181-220, 304-337)
    def _draw(self, num_samples: int, want_v=True, want_x=False, want_logq=False, out=None):
        if out is not None and self.mode == "replay":
            raise ValueError("out= is a native-mode option")
        if self.mode == "replay":
            U = self._take_uniforms(num_samples * self._engine.ups)
            V, X, lq, fb = self._engine.klein_replay(U, num_samples, want_v, want_x, want_logq)
            if fb:
                logger.warning(f"{fb} draws took the reference's no-uniform fallback (klein.py:176-179); "
                               "the replay stream layout after that point is undefined")
        else:
            V, X, lq = self._engine.klein_sample(num_samples, self.seed, self._next_sample_id, self.samplez, want_v,
                                                 want_x, want_logq, out=out)
            self._next_sample_id += num_samples
        return V, X, lq

    @property
    def next_sample_id(self) -> int:
        """Philox counter the next native-mode sample will use (first_sample_id + samples drawn so far); assignable."""
        return self._next_sample_id

    @next_sample_id.setter
    def next_sample_id(self, value: int) -> None:
        if int(value) < 0:
            raise ValueError("next_sample_id must be non-negative")
        self._next_sample_id = int(value)

    def sample_single(self) -> np.ndarray:
        return self._draw(1)[0][0]

    def parallel_sample_batch(self, num_samples: int, batch_size: int = 100) -> np.ndarray:
        # `batch_size` is accepted for signature parity (klein.py:304); the device batches internally
        return self._draw(int(num_samples))[0]

    def sample(self, num_samples: int = 1, *, out: Optional[np.ndarray] = None) -> np.ndarray:
        """klein.py:322-337.  `out` (an addition to the reference signature): a float64 [num_samples, n] array to
        fill and return instead of a fresh one."""
        if out is not None:
            return self._draw(int(num_samples), out=out)[0]
        if num_samples == 1:
            return self.sample_single().reshape(1, -1)
        return self.parallel_sample_batch(num_samples)

    def sample_coefficients(self, num_samples: int = 1) -> np.ndarray:
        """base.py:98-118 without the O(n^3) lstsq per sample: the kernel already has the coefficients."""
        return self._draw(int(num_samples), want_v=False, want_x=True)[1]

    def sample_with_coefficients(self, num_samples: int = 1):
        """(lattice points, integer coefficients, Klein log-densities) of the same draws."""
        return self._draw(int(num_samples), True, True, True)

    def adaptive_precision_sample(self, target_distance: Optional[float] = None) -> np.ndarray:
        """klein.py:273-302: one draw with the table window (`precision`, in conditional sigmas) doubled (>= 20) for
        a target closer than sigma, halved (>= 5) for one further than 5 sigma.  The window is a per-handle constant on
        the device, so the draw comes from a second handle built for that precision (created on first use, kept)."""
        if target_distance is None:
            return self.sample_single()
        relative_distance = target_distance / self.sigma
        precision = self.precision
        if relative_distance < 1:
            precision = max(20, 2 * self.precision)
        elif relative_distance > 5:
            precision = max(5, self.precision // 2)
        if precision == self.precision:
            return self.sample_single()
        main = self._engine
        if precision not in self._alt_engines:
            self._alt_engines[precision] = Engine(self.R, self._basis, self.center_transformed, self.center, self.sigma,
                                                  precision=precision, **self._engine_kwargs)
        try:
            self._engine = self._alt_engines[precision]
            return self.sample_single()
        finally:
            self._engine = main

    # ------------------------------------------------------------------ density (klein.py:222-271)
    def compute_log_density(self, point: np.ndarray) -> float:
        point = np.asarray(point, dtype=np.float64)
        coeffs = np.linalg.solve(self._basis, point)
        coeffs_int = np.round(coeffs).astype(np.int64)
        reconstruction = self._basis @ coeffs_int
        if np.linalg.norm(reconstruction - point) > 1e-10:
            return -np.inf
        return float(self._engine.klein_logdensity(coeffs_int[None, :])[0])

    def compute_log_density_batch(self, coeffs: np.ndarray) -> np.ndarray:
        """Klein log-density of many points given by their integer coefficient vectors [S, n]."""
        return self._engine.klein_logdensity(coeffs)

    def diagnostic_info(self) -> Dict[str, Any]:
        return {
            'algorithm': 'Refined Klein',
            'sigma': self.sigma,
            'precision': self.precision,
            'use_log_space': self.use_log_space,
            'cache_size': len(self._sample_cache),
            'condition_number': np.linalg.cond(self.R),
            'min_R_diag': np.min(np.abs(self.R_diag)),
            'max_R_diag': np.max(np.abs(self.R_diag)),
            'min_conditional_sigma': self.sigma / np.max(np.abs(self.R_diag)),
            'max_conditional_sigma': self.sigma / np.min(np.abs(self.R_diag)),
            'backend': 'liblgs_b200 (sm_100a)',
            'mode': self.mode,
            'samplez': self.samplez,
        }
