"""Drop-in for the reference's IMHKSampler (src/samplers/imhk.py:19-318): Independent Metropolis-Hastings-Klein.

The chain state lives on the GPU (lgs_imhk_* in include/lgs_b200.h); burn-in and thinning happen inside the kernel and
only kept states come back.  Both call shapes are accepted:

    sampler.sample(num_samples, thin)                       the real signature (imhk.py:196); no burn-in is applied,
                                                            exactly like the reference
    sampler.sample(n_samples=N, burn_in=B, thin=T)          the README / src/core/base_sampler.py:45-76 shape
    sampler.acceptance_rate                                 accepted_proposals / total_proposals (0.0 before any step)

New keyword-only arguments keep reference behaviour by default: `n_chains=1`, `mode` in {"native", "replay"},
`seed`, `uniforms`, `weight_mode` in {"reference", "exact"}.  With `weight_mode="reference"` the importance weight is
the reference's (imhk.py:102-124), a constant up to rounding, so every proposal is accepted -- that is what the
reference does.  `weight_mode="exact"` uses the Wang-Ling weight
prod_i rho_{sigma_i, m_i(x)}(Z) and gives a genuine Metropolis-Hastings chain.
"""
from __future__ import annotations

import logging
import time
from typing import List, Optional, Tuple

import numpy as np

from .base import DiscreteGaussianSampler
from .klein import RefinedKleinSampler as KleinSampler, native_seed

logger = logging.getLogger(__name__)


class IMHKSampler(DiscreteGaussianSampler):
    def __init__(self, lattice, sigma: float, center: Optional[np.ndarray] = None, precision: int = 10,
                 burn_in: Optional[int] = None, *, n_chains: int = 1, mode: str = "native",
                 seed: Optional[int] = None,
                 uniforms=None, samplez: Optional[str] = None, weight_mode: str = "reference", device: int = 0,
                 path: str = "auto", first_chain_id: int = 0):
        super().__init__(lattice, sigma, center)
        if weight_mode not in ("reference", "exact"):
            raise ValueError("weight_mode must be 'reference' or 'exact'")
        seed = native_seed(seed, mode)          # None: from numpy's global legacy state (see klein.native_seed)
        self.seed = seed
        self.proposal_sampler = KleinSampler(lattice, sigma, center, precision, mode=mode, seed=seed,
                                             uniforms=uniforms, samplez=samplez, device=device, path=path)
        self.mode = mode
        self.weight_mode = weight_mode
        self.n_chains = int(n_chains)
        self.burn_in = burn_in if burn_in is not None else self._estimate_burn_in()
        self._engine = self.proposal_sampler._engine
        self._engine.imhk_reset(self.n_chains, seed, first_chain_id)
        self._initialized = False
        self._precompute_partition_bounds()
        logger.info(f"Initialized IMHK with burn-in={self.burn_in}")

    # ------------------------------------------------------------------ imhk.py:68-100
    def _estimate_burn_in(self) -> int:
        epsilon = 0.01
        min_gs = getattr(self.lattice, "min_gram_schmidt_norm", None)
        if min_gs is None:
            min_gs = float(np.min(np.abs(self.proposal_sampler.R_diag)))
        try:
            sigma_factor = (self.sigma / min_gs) ** self.dimension
            mixing_time = int(np.ceil(-np.log(epsilon) * self.dimension * sigma_factor))
        except (OverflowError, ValueError):
            # the reference raises OverflowError here for large n; its own cap is the intent
            return 10000
        return min(mixing_time * 2, 10000)

    def _precompute_partition_bounds(self):
        self.klein_log_partition = 0.0
        for i in range(self.dimension):
            sigma_i = self.sigma / abs(self.proposal_sampler.R[i, i])
            self.klein_log_partition += 0.5 * np.log(2 * np.pi) + np.log(sigma_i)

    # ------------------------------------------------------------------ imhk.py:102-124
    def _compute_importance_weight(self, lattice_point: np.ndarray) -> Tuple[float, float]:
        log_target = self.log_gaussian_weight(np.asarray(lattice_point, dtype=np.float64))
        log_proposal = self.proposal_sampler.compute_log_density(lattice_point)
        log_weight = log_target - log_proposal
        if log_weight > 700:
            return np.inf, log_weight
        elif log_weight < -700:
            return 0.0, log_weight
        return np.exp(log_weight), log_weight

    # ------------------------------------------------------------------ running the chains
    def _run(self, burn_in: int, keep: int, thin: int, want_v=True, want_x=False):
        eng = self._engine
        if self.mode == "replay":
            ups = eng.ups
            steps = burn_in + keep * thin
            stride = (0 if self._initialized else ups) + steps * (ups + 1)
            U = self.proposal_sampler._take_uniforms(self.n_chains * stride)
            out = eng.imhk_replay(U, stride, burn_in, keep, thin, self.weight_mode, want_v, want_x)
            V, X = out["V"], out["X"]
            self._last_accept = out["accept"]
            self._last_logw = out["logw"]
        else:
            V, X = eng.imhk_run(burn_in, keep, thin, self.proposal_sampler.samplez, self.weight_mode, want_v, want_x)
        self._initialized = True
        return V, X

    def _initialize_chain(self):
        if not self._initialized:
            self._run(0, 0, 1, want_v=False)

    def step(self) -> Tuple[np.ndarray, bool]:
        """One MH step of every chain (imhk.py:141-177); returns (state, accepted) -- of chain 0 when n_chains == 1,
        else arrays over chains."""
        # the accept bit from the per-chain counters alone (two 16-byte-per-chain copies, no kernel, no state copies)
        before = (self._engine.imhk_chain_counts()[:, 0] if self._initialized
                  else np.zeros(self.n_chains, np.int64))
        V, _ = self._run(0, 1, 1)
        accepted = (self._engine.imhk_chain_counts()[:, 0] - before) > 0
        if self.n_chains == 1:
            return V[0, 0].copy(), bool(accepted[0])
        return V[:, 0].copy(), accepted

    def sample_single(self) -> np.ndarray:
        """imhk.py:179-194: burn-in runs only while the chain has no state yet, then one step."""
        b = 0 if self._initialized else int(self.burn_in)
        V, _ = self._run(b, 1, 1)
        return V[0, 0] if self.n_chains == 1 else V[:, 0]

    def sample(self, num_samples: int = 1, thin: int = 1, *, n_samples: Optional[int] = None,
               burn_in: int = 0) -> np.ndarray:
        """imhk.py:196-229.  Returns float64 [num_samples, n] (n_chains == 1) or [n_chains, num_samples, n]."""
        if n_samples is not None:
            num_samples = n_samples
        start_time = time.time()
        V, _ = self._run(int(burn_in), int(num_samples), int(thin))
        self.stats.samples_generated += num_samples
        self.stats.time_elapsed += time.time() - start_time
        self.stats.acceptance_rate = self.acceptance_rate
        return V[0] if self.n_chains == 1 else V

    def sample_coefficients(self, num_samples: int = 1, thin: int = 1, burn_in: int = 0) -> np.ndarray:
        _, X = self._run(int(burn_in), int(num_samples), int(thin), want_v=False, want_x=True)
        return X[0] if self.n_chains == 1 else X

    def run_chain(self, num_steps: int, save_every: int = 1) -> List[np.ndarray]:
        """imhk.py:231-250: states after steps 0, save_every, 2*save_every, ... (the reference saves when
        i % save_every == 0, i.e. after step i+1)."""
        saved = []
        done = 0
        for i in range(0, num_steps, save_every):
            V, _ = self._run(i + 1 - done - 1, 1, 1)
            done = i + 1
            saved.append(V[0, 0].copy() if self.n_chains == 1 else V[:, 0].copy())
        if done < num_steps:
            self._run(num_steps - done, 0, 1, want_v=False)
        return saved

    # ------------------------------------------------------------------ state and statistics (imhk.py:54-61)
    @property
    def total_proposals(self) -> int:
        return self._engine.imhk_counts()[1]

    @property
    def accepted_proposals(self) -> int:
        return self._engine.imhk_counts()[0]

    @property
    def acceptance_rate(self) -> float:
        a, t = self._engine.imhk_counts()
        return a / t if t > 0 else 0.0

    @property
    def per_chain_acceptance(self) -> np.ndarray:
        c = self._engine.imhk_state()[0]
        return c[:, 0] / np.maximum(c[:, 1], 1)

    def _state(self, k):
        if not self._initialized:
            return None
        out = self._engine.imhk_state()[k]
        return out[0] if self.n_chains == 1 else out

    # The reference keeps the chain state in three plain attributes and its experiment scripts assign to them
    # (`sampler.current_state = v`, convergence_study.py:563, parameter_sensitivity.py:470,539,
    # cryptographic_experiments.py:344).  Here the state lives on the device as (integer coefficients, log weight) per
    # chain, so the setters install a consistent pair through lgs_imhk_set_state:
    #   current_state = v      x = round(B^-1 v) as imhk.py:130-133; [n] (all chains) or [n_chains, n]
    #   current_coeffs = x     the same from integer coefficients
    #   current_log_weight = w keeps the coefficients, replaces the weight
    # The weight installed with a new state is the reference's importance weight (imhk.py:102-124) of the lattice point
    # B x when v IS that lattice point and weight_mode == "reference"; otherwise (v off the lattice: zero mass under
    # the lattice Gaussian; or exact weights, which the host does not evaluate) it is -inf, for which the accept rule
    # (imhk.py:158-160) takes the next proposal with probability one.  Counters and the Philox position are kept.
    def _install(self, X, logw):
        X = np.asarray(X)
        if X.ndim == 1:
            X = np.broadcast_to(X, (self.n_chains, self.dimension))
        logw = np.broadcast_to(np.asarray(logw, dtype=np.float64), (self.n_chains,))
        if self._initialized:
            counts, steps = self._engine.imhk_chain_counts(), self._engine.imhk_steps_done()
        else:
            counts, steps = None, 0
        self._engine.imhk_set_state(np.ascontiguousarray(X, dtype=np.int64), np.ascontiguousarray(logw), counts,
                                    steps, self.proposal_sampler.samplez)
        self._initialized = True

    def _weights_of(self, X, on_lattice):
        B = self.proposal_sampler._basis
        X = np.atleast_2d(X)
        w = np.full(X.shape[0], -np.inf)
        if self.weight_mode == "reference":
            for c in range(X.shape[0]):
                if on_lattice[c]:
                    w[c] = self._compute_importance_weight(B @ X[c].astype(np.float64))[1]
        return w

    @property
    def current_state(self):
        return self._state(2)

    @current_state.setter
    def current_state(self, value):
        if value is None:
            raise ValueError("use a new sampler (or lgs_imhk_reset) to discard the chain state")
        B = self.proposal_sampler._basis
        v = np.asarray(value, dtype=np.float64)
        if v.shape not in ((self.dimension,), (self.n_chains, self.dimension)):
            raise ValueError(f"state must have shape ({self.dimension},) or ({self.n_chains}, {self.dimension})")
        v2 = np.atleast_2d(v)
        X = np.linalg.solve(B, v2.T).T.round().astype(np.int64)         # imhk.py:130-133
        scale = max(1.0, float(np.abs(v2).max()))
        on = np.max(np.abs(X.astype(np.float64) @ B.T - v2), axis=1) <= 1e-9 * scale
        w = self._weights_of(X, on)
        self._install(X if v.ndim == 2 else X[0], w if v.ndim == 2 else w[0])

    @property
    def current_coeffs(self):
        return self._state(1)

    @current_coeffs.setter
    def current_coeffs(self, value):
        X = np.asarray(value)
        if X.shape not in ((self.dimension,), (self.n_chains, self.dimension)) or not np.all(X == np.round(X)):
            raise ValueError("coefficients must be integers of shape (n,) or (n_chains, n)")
        X = X.astype(np.int64)
        w = self._weights_of(X, np.ones(np.atleast_2d(X).shape[0], dtype=bool))
        self._install(X, w if X.ndim == 2 else w[0])

    @property
    def current_log_weight(self):
        s = self._state(3)
        return None if s is None else (float(s) if self.n_chains == 1 else s)

    @current_log_weight.setter
    def current_log_weight(self, value):
        if not self._initialized:
            raise RuntimeError("the chains have no state yet: assign current_state or current_coeffs first")
        self._install(self._engine.imhk_state()[1], value)

    def get_state(self) -> dict:
        """Checkpoint: everything a new IMHKSampler on the same lattice / sigma needs to continue these chains
        (the reference's current_coeffs / current_log_weight / counters, imhk.py:54-61, for every chain, plus the
        Philox stream position)."""
        if not self._initialized:
            raise RuntimeError("the chains have no state yet")
        counts, X, _, lw = self._engine.imhk_state()
        return {"coeffs": X, "log_weight": lw, "counts": counts, "steps_done": self._engine.imhk_steps_done(),
                "n_chains": self.n_chains}

    def set_state(self, state: dict) -> None:
        """Resume from get_state(): with the same seed and first_chain_id the run continues the uninterrupted one."""
        if int(state["n_chains"]) != self.n_chains:
            raise ValueError("state was taken from a different number of chains")
        self._engine.imhk_set_state(state["coeffs"], state["log_weight"], state["counts"], state["steps_done"],
                                    self.proposal_sampler.samplez)
        self._initialized = True

    # ------------------------------------------------------------------ imhk.py:252-313
    def estimate_spectral_gap(self, num_samples: int = 1000) -> float:
        V, X, lq = self.proposal_sampler.sample_with_coefficients(num_samples)
        diff = V - self.center
        log_w = -np.einsum("ij,ij->i", diff, diff) / (2 * self.sigma ** 2) - lq
        log_w = log_w[log_w <= 700]
        if log_w.size == 0:
            logger.warning("All importance weights were infinite")
            return 0.0
        max_weight = float(np.exp(np.max(log_w))) if np.max(log_w) >= -700 else 0.0
        return 1.0 / max_weight if max_weight > 0 else 0.0

    def diagnose_convergence(self, num_samples: int = 1000) -> dict:
        old_stats = self.stats
        self.reset_stats()
        samples = self.sample(num_samples)
        flat = samples.reshape(-1, self.dimension)
        diagnostics = {
            'acceptance_rate': self.stats.acceptance_rate,
            'spectral_gap_estimate': self.estimate_spectral_gap(100),
            'empirical_mean': self.empirical_mean(flat),
            'empirical_std': np.std(flat, axis=0),
            'theoretical_std': self.sigma * np.ones(self.dimension),
            'samples_per_second': num_samples / max(self.stats.time_elapsed, 1e-12),
        }
        self.stats = old_stats
        return diagnostics

    def __repr__(self) -> str:
        return (f"IMHKSampler(lattice={getattr(self.lattice, 'name', '?')}, σ={self.sigma:.4f}, "
                f"burn_in={self.burn_in}, acceptance_rate={self.acceptance_rate:.2%})")
