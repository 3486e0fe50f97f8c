import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import lattice_gaussian_mcmc_b200 as lgs
lat = lgs.SimpleLattice(np.array([[4.0, 1.7], [0.0, 0.6]]))
sigma, center = 1.0, np.array([0.4, 0.25])
kw = dict(burn_in=0, n_chains=500, seed=17, weight_mode="reference", samplez="fast")
a = lgs.IMHKSampler(lat, sigma, center, path="ordered", **kw)
b = lgs.IMHKSampler(lat, sigma, center, path="blocked", **kw)
for step in range(6):
    Xa = a.sample_coefficients(1, thin=1)[:, 0]
    Xb = b.sample_coefficients(1, thin=1)[:, 0]
    bad = np.where(np.any(Xa != Xb, axis=1))[0]
    print("step", step, "mismatching chains", bad[:10], len(bad), "acc", a.accepted_proposals, b.accepted_proposals)
    if len(bad):
        k = bad[0]
        print("  ordered", Xa[k], "blocked", Xb[k], "lw", a.current_log_weight[k], b.current_log_weight[k], a.klein_log_partition)
