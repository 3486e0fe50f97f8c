"""Host-side base class: the interface of the reference's DiscreteGaussianSampler (src/samplers/base.py:18-176).

Only bookkeeping lives here (constructor checks, the two Gaussian weight helpers, empirical statistics); everything
that samples is a CUDA kernel behind include/lgs_b200.h.
"""
from __future__ import annotations

import logging
from abc import ABC, abstractmethod
from dataclasses import dataclass
from typing import Any, Dict, Optional

import numpy as np

logger = logging.getLogger(__name__)


@dataclass
class SamplingStats:
    """Same fields as the reference's SamplingStats (base.py:18-28)."""
    samples_generated: int = 0
    time_elapsed: float = 0.0
    acceptance_rate: Optional[float] = None
    extra_stats: Dict[str, Any] = None

    def __post_init__(self):
        if self.extra_stats is None:
            self.extra_stats = {}


def lattice_basis(lattice) -> np.ndarray:
    """`lattice.basis` as the reference samplers read it (klein.py:59), falling back to `get_basis()` so that lattice
    classes which only expose the latter (identity / q-ary / NTRU in the reference) also work."""
    if hasattr(lattice, "basis"):
        b = lattice.basis
    elif hasattr(lattice, "get_basis"):
        b = lattice.get_basis()
    else:
        raise AttributeError("lattice has neither .basis nor .get_basis()")
    return np.ascontiguousarray(np.array(b, dtype=np.float64))


class DiscreteGaussianSampler(ABC):
    """D_{L,sigma,c}(x) proportional to exp(-||x-c||^2 / (2 sigma^2)) over a lattice L (base.py:31-41)."""

    def __init__(self, lattice, sigma: float, center: Optional[np.ndarray] = None):
        self.lattice = lattice
        self.sigma = sigma
        self.dimension = lattice.dimension
        if center is None:
            self.center = np.zeros(self.dimension)
        else:
            self.center = np.array(center, dtype=np.float64)
        if len(self.center) != self.dimension:
            raise ValueError(f"Center dimension {len(self.center)} != lattice dimension {self.dimension}")
        self.stats = SamplingStats()
        self._validate_parameters()
        logger.info(f"Initialized {self.__class__.__name__} with σ={sigma:.4f}")

    def _validate_parameters(self):
        # base.py:72-83: only sigma <= 0 is an error; below-smoothing is a warning
        if self.sigma <= 0:
            raise ValueError(f"Standard deviation must be positive, got {self.sigma}")
        sm = getattr(self.lattice, "smoothing_parameter", None)
        if sm is not None:
            eta = sm()
            if self.sigma < eta:
                logger.warning(f"σ={self.sigma:.4f} is below smoothing parameter η={float(eta):.4f}. "
                               "Sampling may be biased.")

    @abstractmethod
    def sample(self, num_samples: int = 1) -> np.ndarray:
        ...

    def gaussian_weight(self, point: np.ndarray) -> float:
        diff = point - self.center
        return np.exp(-np.dot(diff, diff) / (2 * self.sigma ** 2))

    def log_gaussian_weight(self, point: np.ndarray) -> float:
        diff = point - self.center
        return -np.dot(diff, diff) / (2 * self.sigma ** 2)

    def reset_stats(self):
        self.stats = SamplingStats()

    def get_stats(self) -> SamplingStats:
        return self.stats

    def empirical_mean(self, samples: np.ndarray) -> np.ndarray:
        return np.mean(samples, axis=0)

    def empirical_covariance(self, samples: np.ndarray) -> np.ndarray:
        return np.cov(samples.T)

    def theoretical_covariance(self) -> np.ndarray:
        return self.sigma ** 2 * np.eye(self.dimension)

    def __repr__(self) -> str:
        return (f"{self.__class__.__name__}(lattice={getattr(self.lattice, 'name', '?')}, "
                f"σ={self.sigma:.4f}, center={self.center})")
