"""Mints tests/golden/tvd_d1.npz from the reference's OWN compute_tvd / compute_acceptance_rate
(/root/reference/src/diagnostics/convergence_diag.py:15-72, mcmc_diag.py:107-117; importable in the build container):
integer-valued sample sets shaped like lattice points -> exact-value TVD (1-D and multivariate = mean of marginals),
histogram TVD with `bins`, acceptance rate of a boolean array.

    python tests/golden/make_golden_tvd.py
"""
import os
import sys

import numpy as np

sys.dont_write_bytecode = True
sys.path.insert(0, "/root/reference")
from src.diagnostics import convergence_diag as cd   # noqa: E402
from src.diagnostics import mcmc_diag as md           # noqa: E402

rng = np.random.RandomState(515)
cases = {}
for name, (S1, S2, n, s1, s2, shift) in {"same": (4000, 3000, 6, 5.0, 5.0, 0.0), "wider": (5000, 5000, 4, 5.0, 7.0, 0.0),
                                         "shifted": (2500, 4000, 3, 12.0, 12.0, 2.0), "one_dim": (3000, 3000, 1, 3.0, 3.5, 0.0)}.items():
    A = np.round(rng.normal(0.0, s1, (S1, n)))
    B = np.round(rng.normal(shift, s2, (S2, n)))
    if n == 1:
        A, B = A[:, 0], B[:, 0]
    cases[name] = (A, B, float(cd.compute_tvd(A, B)), float(cd.compute_tvd(A, B, bins=16)), float(cd.compute_tvd(A, B, bins=50)))
acc = rng.uniform(size=(7, 333)) < 0.37
out = os.path.join(os.path.dirname(os.path.abspath(__file__)), "tvd_d1.npz")
flat = {"names": np.array(list(cases)), "accepted": acc, "acceptance_rate": np.array([md.compute_acceptance_rate(a) for a in acc]),
        "acceptance_rate_all": np.array(md.compute_acceptance_rate(acc)), "numpy_version": np.__version__}
for k, (A, B, t0, t16, t50) in cases.items():
    flat[k + "_a"], flat[k + "_b"], flat[k + "_tvd"] = A, B, np.array([t0, t16, t50])
np.savez_compressed(out, **flat)
print(out, {k: v[2:] for k, v in cases.items()})
