#!/usr/bin/env python
"""Mints tests/golden/surface_api.npz by running the REFERENCE's sampler classes (src/samplers/base.py:98-175,
klein.py:222-271,339-354, imhk.py:68-100,315-318) after np.random.seed(seed): the small methods around the hot path --
sample_coefficients, (log_)gaussian_weight, compute_log_density (lattice and non-lattice points), empirical_mean /
covariance, theoretical_covariance, diagnostic_info, get_stats, __repr__, the default burn-in estimate.

    python tests/golden/make_golden_surface.py
"""
import json
import os
import sys

sys.dont_write_bytecode = True
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from make_golden import IMHKSampler, NoCache, RefinedKleinSampler, ShimLattice, np  # noqa: E402  (imports the reference)

rng = np.random.RandomState(7)
n, seed, S, k = 8, 21, 60, 12
B = (rng.randint(-3, 4, (n, n)) + 2 * n * np.eye(n)).astype(np.float64)
center = rng.uniform(-5, 5, n)
lat = ShimLattice(B, "dd8-surface")
sigma = 1.5 * float(np.abs(np.diag(np.linalg.qr(B, mode="r"))).max())
np.random.seed(seed)
smp = RefinedKleinSampler(lat, sigma, center)
smp._sample_cache = NoCache()
V = smp.sample(S)                                   # S * n uniforms
X = smp.sample_coefficients(k)                      # k * n more
pts = V[:5]
off = pts[0] + 0.25                                 # not a lattice point
used = (S + k) * n
U = np.random.RandomState(seed).random_sample(used + 1)
assert np.random.random_sample() == U[used], "stream accounting is off"
info = smp.diagnostic_info()
st = smp.get_stats()
burn = {}
for nn, ratio in ((4, 1.5), (8, 1.5), (16, 1.2), (64, 1.05), (8, 0.9)):
    Bn = np.eye(nn) * 10.0
    im = IMHKSampler(ShimLattice(Bn, f"d{nn}"), 10.0 * ratio)
    burn[f"{nn}:{ratio}"] = int(im.burn_in)
im = IMHKSampler(lat, sigma, center, burn_in=17)
try:                                                # imhk.py:315-318 formats stats.acceptance_rate, None before sampling
    repr(im)
    repr_raises = False
except TypeError:
    repr_raises = True
np.random.seed(seed + 1)
im.proposal_sampler._sample_cache = NoCache()
im.sample(3)
out = {
    "basis": B, "center": center, "sigma": sigma, "seed": seed, "S": S, "k": k, "uniforms": U[:used],
    "V": V, "X": X, "points": pts, "off_point": off,
    "log_gaussian_weight": np.array([smp.log_gaussian_weight(p) for p in pts]),
    "gaussian_weight": np.array([smp.gaussian_weight(p) for p in pts]),
    "log_density": np.array([smp.compute_log_density(p) for p in pts]),
    "log_density_off": np.array(smp.compute_log_density(off)),
    "empirical_mean": smp.empirical_mean(V), "empirical_covariance": smp.empirical_covariance(V),
    "theoretical_covariance": smp.theoretical_covariance(),
    "diagnostic_info": json.dumps({a: (float(b) if not isinstance(b, (str, bool)) else b) for a, b in info.items()}),
    "stats": json.dumps({"samples_generated": st.samples_generated, "acceptance_rate": st.acceptance_rate}),
    "repr_klein": repr(smp), "repr_imhk": repr(im), "repr_imhk_raises_before_sampling": repr_raises,
    "imhk_uniforms": np.random.RandomState(seed + 1).random_sample(n + 3 * (n + 1)), "burn_in": json.dumps(burn),
    "imhk_klein_log_partition": im.klein_log_partition, "numpy_version": np.__version__,
}
path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "surface_api.npz")
np.savez_compressed(path, **out)
print(path, "\n", info, "\n", out["stats"], "\n", out["repr_klein"], "\n", out["repr_imhk"], "\n", burn, out["log_density_off"])
