"""B200-native batched lattice Gaussian sampler: drop-in for the Klein / IMHK sampler classes of
NickQrumpton/lattice-gaussian-mcmc (src/samplers/__init__.py:3-7).

    from lattice_gaussian_mcmc_b200.samplers import KleinSampler, IMHKSampler

Host code is Python (lattice construction, QR / Gram-Schmidt in numpy, as in the reference); the sampling itself
runs in hand-written sm_100a CUDA behind the C ABI declared in include/lgs_b200.h (liblgs_b200.so).
There is no CPU fallback: constructing a sampler without a CUDA device raises.
"""
from . import samplers  # noqa: F401
from .lattices import IdentityLattice, QaryLattice, SimpleLattice  # noqa: F401
from .ntru import NTRULattice  # noqa: F401
from .samplers import DiscreteGaussianSampler, IMHKSampler, KleinSampler, RefinedKleinSampler, SamplingStats  # noqa: F401

__all__ = ["samplers", "DiscreteGaussianSampler", "KleinSampler", "RefinedKleinSampler", "IMHKSampler",
           "SamplingStats", "SimpleLattice", "IdentityLattice", "QaryLattice", "NTRULattice"]
