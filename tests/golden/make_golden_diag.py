"""Mints tests/golden/diag_ar1.npz from the reference's OWN diagnostics functions
(/root/reference/src/diagnostics/convergence_diag.py, mcmc_diag.py; importable in the build container):
AR(1) integer-valued series of several correlation strengths -> acf, tau_int, ess, r_hat.

    python tests/golden/make_golden_diag.py
"""
import os
import sys

import numpy as np

sys.dont_write_bytecode = True
sys.path.insert(0, "/root/reference")
from src.diagnostics import convergence_diag as cd   # noqa: E402
from src.diagnostics import mcmc_diag as md           # noqa: E402

rng = np.random.RandomState(2026)
chains, keep = 12, 400
phis = [0.0, 0.3, 0.6, 0.9]
X = np.zeros((len(phis), chains, keep))
for a, phi in enumerate(phis):
    for c in range(chains):
        e = rng.normal(size=keep) * 7.0
        x = np.zeros(keep)
        x[0] = e[0]
        for t in range(1, keep):
            x[t] = phi * x[t - 1] + e[t]
        X[a, c] = np.round(x + 3.0 * a)           # integer-valued like lattice coordinates
acf = np.zeros((len(phis), chains, keep // 4 + 1))
tau = np.zeros((len(phis), chains))
tau_md = np.zeros((len(phis), chains))
ess = np.zeros((len(phis), chains))
rhat = np.zeros(len(phis))
for a in range(len(phis)):
    for c in range(chains):
        acf[a, c] = cd.compute_autocorrelation(X[a, c])
        tau[a, c] = cd.integrated_autocorrelation_time(X[a, c])
        tau_md[a, c] = md.integrated_autocorrelation_time(X[a, c])
        ess[a, c] = md.effective_sample_size(X[a, c])
    rhat[a] = cd.gelman_rubin_statistic([X[a, c] for c in range(chains)])
out = os.path.join(os.path.dirname(os.path.abspath(__file__)), "diag_ar1.npz")
np.savez_compressed(out, X=X, phis=np.array(phis), acf=acf, tau=tau, tau_md=tau_md, ess=ess, rhat=rhat,
                    numpy_version=np.__version__)
print(out, "tau range", tau.min(), tau.max(), "rhat", rhat, "max |tau - tau_md|", np.abs(tau - tau_md).max())
