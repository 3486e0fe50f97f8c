"""BASELINE config C5: dimension sweep n = 16 .. 1024 -- Klein / IMHK throughput on one GPU, the CPU port of the
reference path on the host cores beside it, and the marginal TVD of GPU samples against the CPU port's.

    python tools/dimension_sweep.py > gpurun_out/dimension_sweep.json

Recipe: n <= 64 the reference's seed-(0+n) randint(0,51) matrix, LLL-reduced (klein_scaling_analysis.py
:83-108); above, diagonally dominant integer bases; two extra rows run the same LLL recipe at n = 128 / 256 through the
C++ reducer (lgs_lll_reduce).  sigma = 1.5 max||b*||, S = max(2^14, 2^26 / n) samples (chains) rounded to whole
kernel batches (at least one), thin = 10.
Development record, not the bench contract (bench.py)."""
import json
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import __graft_entry__ as g  # noqa: E402

g.build()
import lattice_gaussian_mcmc_b200 as lgs  # noqa: E402
from lattice_gaussian_mcmc_b200.lattices import dd_lattice, lll_lattice  # noqa: E402
from oracle import oracle  # noqa: E402  (checker / CPU baseline only)

PEAK = 37.142  # measured FP64 DMMA TFLOP/s, profiles/fp64_peak_r01.json


def tvd_ratio(Xa, Xb):
    S = min(len(Xa), len(Xb))
    worst = 0.0
    for d in range(Xa.shape[1]):
        lo = min(Xa[:, d].min(), Xb[:, d].min())
        hi = max(Xa[:, d].max(), Xb[:, d].max())
        a = np.bincount(Xa[:, d] - lo, minlength=hi - lo + 1) / len(Xa)
        b = np.bincount(Xb[:, d] - lo, minlength=hi - lo + 1) / len(Xb)
        bins = int(((a + b) > 0).sum())
        worst = max(worst, 0.5 * np.abs(a - b).sum() / np.sqrt(bins / S))
    return worst


def timed(fn, reps):
    fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for r in range(reps):
        fn(r)
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


rows = []
cores = os.cpu_count()
cases = [(n, "lll" if n <= 64 else "dd") for n in (16, 32, 64, 128, 256, 512, 1024)] + [(128, "lll"), (256, "lll")]
for n, kind in cases:
    t0 = time.time()
    lat = lll_lattice(n) if kind == "lll" else dd_lattice(n)
    t_setup = time.time() - t0
    sigma = 1.5 * lat.max_gram_schmidt_norm
    smp = lgs.KleinSampler(lat, sigma, seed=n + 1)
    eng = smp._engine
    S = max(1 << 14, (1 << 26) // n)
    S = max(1, round(S / eng.preferred_batch)) * eng.preferred_batch   # whole batches of the kernel that serves this n
    V = torch.empty((S, n), dtype=torch.float64, device="cuda")
    st = torch.cuda.current_stream().cuda_stream
    ms = timed(lambda r=0: eng.klein_sample_dev(S, n + 1, (r + 1) * S, "fast", V=V.data_ptr(), stream=st), 4)
    klein = S / ms * 1e3
    # IMHK: S chains, thin 10, the kept state written as a lattice point
    thin = 10
    im = lgs.IMHKSampler(lat, sigma, burn_in=0, n_chains=S, seed=n + 2)
    ie = im._engine
    ie.imhk_reset(S, seed=n + 2)
    Vk = torch.empty((S, 1, n), dtype=torch.float64, device="cuda")
    ms_i = timed(lambda r=0: ie.imhk_run_dev(0, 1, thin, "fast", "reference", V=Vk.data_ptr(), stream=st), 3)
    imhk = S * thin / ms_i * 1e3
    # marginals against the CPU port (same lattice, sigma), and the port's own rate on the host cores
    So = min(20000, max(600, 600 * 1024 // n))
    prob = oracle.Problem(lat.basis, sigma)
    t0 = time.time()
    Xs, _ = oracle.klein_seeded(prob, So, seed=n, chains=cores, threads=cores)
    cpu_rate = So / (time.time() - t0)
    _, X, _ = smp.sample_with_coefficients(So)
    step = max(1, n // 16)
    rows.append({
        "n": n, "basis": "LLL-reduced randint(0,51), seed 0+n" if kind == "lll" else "diagonally dominant integer",
        "host_setup_s": round(t_setup, 3), "sigma": sigma, "samples_per_call": S,
        "klein_samples_per_s": klein, "klein_tflops": (3 * n * n - n) * klein * 1e-12,
        "klein_frac_of_fp64_peak": (3 * n * n - n) * klein * 1e-12 / PEAK,
        "imhk_chain_steps_per_s": imhk,
        "imhk_frac_of_fp64_peak": (n * (n - 1) + 6 * n + 2 * n * n / thin) * imhk * 1e-12 / PEAK,
        "imhk_acceptance": float(ie.imhk_counts()[0]) / max(1, ie.imhk_counts()[1]),
        "cpu_port_samples_per_s": cpu_rate, "cpu_threads": cores, "oracle_samples": So,
        "marginal_tvd_over_noise": tvd_ratio(Xs[:, ::step], X[:, ::step]),
    })
    print(json.dumps(rows[-1]), file=sys.stderr)
    del V, Vk, smp, im
    torch.cuda.empty_cache()
print(json.dumps({"config": "C5 dimension sweep, one B200, device-resident outputs", "fp64_peak_tflops": PEAK,
                  "rows": rows}, indent=1))
