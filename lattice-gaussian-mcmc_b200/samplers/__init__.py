"""Mirror of the reference's src/samplers/__init__.py:3-7 (same exported names, same alias)."""
from .base import DiscreteGaussianSampler, SamplingStats
from .imhk import IMHKSampler
from .klein import RefinedKleinSampler
from .klein import RefinedKleinSampler as KleinSampler

__all__ = ["DiscreteGaussianSampler", "KleinSampler", "RefinedKleinSampler", "IMHKSampler", "SamplingStats"]
