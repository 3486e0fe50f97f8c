#!/usr/bin/env python
"""Mints tests/golden/m6_imhk_diag.npz by running the REFERENCE's IMHKSampler.estimate_spectral_gap /
diagnose_convergence (src/samplers/imhk.py:252-313) after np.random.seed(seed), through the duck-typed lattice of
make_golden.py.  Recorded: inputs, the uniform stream the calls consumed, the spectral-gap estimate, and the
deterministic entries of the diagnose_convergence dict (everything except samples_per_second).

    python tests/golden/make_golden_m6.py
"""
import os
import sys

sys.dont_write_bytecode = True
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from make_golden import IMHKSampler, NoCache, ShimLattice, np  # noqa: E402  (imports the reference)

rng = np.random.RandomState(99)
out = {}
for name, n, sigma_f, seed in (("dd8", 8, 1.5, 11), ("dd24", 24, 1.2, 12)):
    B = (rng.randint(-3, 4, (n, n)) + 2 * n * np.eye(n)).astype(np.float64)
    lat = ShimLattice(B, name)
    R = np.abs(np.diag(np.linalg.qr(B, mode="r")))
    sigma = sigma_f * float(R.max())
    k, num = 40, 25
    np.random.seed(seed)
    smp = IMHKSampler(lat, sigma, burn_in=5)
    smp.proposal_sampler._sample_cache = NoCache()
    gap = smp.estimate_spectral_gap(k)
    diag = smp.diagnose_convergence(num)
    used = k * n + (n + num * (n + 1)) + 100 * n
    st = np.random.get_state()
    U = np.random.RandomState(seed).random_sample(used + 1)
    probe = np.random.random_sample()                       # the reference's stream position == `used`
    assert probe == U[used], "stream accounting is off"
    out.update({f"{name}_basis": B, f"{name}_sigma": sigma, f"{name}_seed": seed, f"{name}_k": k, f"{name}_num": num,
                f"{name}_uniforms": U[:used], f"{name}_gap": gap, f"{name}_acceptance_rate": diag["acceptance_rate"],
                f"{name}_gap2": diag["spectral_gap_estimate"], f"{name}_empirical_mean": diag["empirical_mean"],
                f"{name}_empirical_std": diag["empirical_std"], f"{name}_theoretical_std": diag["theoretical_std"],
                f"{name}_klein_log_partition": smp.klein_log_partition})
    print(name, "gap", gap, "exp(-log partition)", np.exp(-smp.klein_log_partition), "acc", diag["acceptance_rate"])
out["names"] = np.array(["dd8", "dd24"])
out["numpy_version"] = np.__version__
path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "m6_imhk_diag.npz")
np.savez_compressed(path, **out)
print(path)
