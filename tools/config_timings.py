"""BASELINE configs C1-C3 at their FULL sizes on one GPU (C4 is bench.py, C5 is tools/dimension_sweep.py), the CPU
port of the reference path timed beside each on a bounded sample.  Development record, not the bench contract.

    python tools/config_timings.py > gpurun_out/config_timings.json

C1  IdentityLattice Z^10, sigma = 5: KleinSampler.sample(10 000) through the Python mirror (host array returned).
C2  LLL-reduced randint(0, 51) lattice n = 64 (seed 0 + 64), sigma = 1.5 max||b*||: IMHK, 4096 chains, burn_in = 1000,
    thin = 10, 10 000 kept states per chain; kept states stay on the device in blocks of 250 and are reduced to
    marginal histograms there (the [4096, 10000, 64] float64 array would be 21 GB).
C3  q-ary LWE lattice n = 256, m = 512, q = 3329, sigma = 1.5 q: 2^20 Klein samples, device-resident."""
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
from lattice_gaussian_mcmc_b200.lattices import identity_lattice, lll_lattice, qary_lattice  # noqa: E402
from oracle import oracle  # noqa: E402  (checker / CPU baseline only)

cores = os.cpu_count()
out = {}
st = torch.cuda.current_stream().cuda_stream

# ---------------------------------------------------------------------------------------------------- C1
lat = identity_lattice(10)
smp = lgs.KleinSampler(lat, 5.0, seed=0)
smp.sample(10000)
t0 = time.perf_counter()
reps = 20
for _ in range(reps):
    pts = smp.sample(10000)
dt = (time.perf_counter() - t0) / reps
prob = oracle.Problem(lat.basis, 5.0)
t0 = time.perf_counter()
oracle.klein_seeded(prob, 10000, seed=0, chains=1, threads=1)
dt_cpu = time.perf_counter() - t0
out["C1"] = {"config": "Z^10, sigma=5, KleinSampler.sample(10000), host float64 array returned",
             "ms_per_call": dt * 1e3, "samples_per_s": 10000 / dt, "mean_abs": float(np.abs(pts.mean(0)).max()),
             "std": float(pts.std()), "cpu_port_1_thread_samples_per_s": 10000 / dt_cpu}
print(json.dumps(out["C1"]), file=sys.stderr)

# ---------------------------------------------------------------------------------------------------- C2
lat = lll_lattice(64)
sigma = 1.5 * lat.max_gram_schmidt_norm
chains, burn_in, thin, keep, block = 4096, 1000, 10, 10000, 250
im = lgs.IMHKSampler(lat, sigma, burn_in=0, n_chains=chains, seed=64)
eng = im._engine
V = torch.empty((chains, block, 64), dtype=torch.float64, device="cuda")
dims = np.array([0, 21, 42, 63], dtype=np.int32)
hist = torch.zeros((4, 64), dtype=torch.int64, device="cuda")


def c2_run():
    eng.imhk_reset(chains, seed=64)
    hist.zero_()
    eng.imhk_run_dev(burn_in, 0, 1, "fast", "reference", stream=st)
    for _ in range(keep // block):
        eng.imhk_run_dev(0, block, thin, "fast", "reference", V=V.data_ptr(), stream=st)
        eng.histogram_dev(V.data_ptr(), chains * block, dims, -8 * sigma, sigma / 4, 64, hist.data_ptr(),
                          stream=st)


c2_run()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
c2_run()
e1.record()
torch.cuda.synchronize()
ms = e0.elapsed_time(e1)
steps = chains * (burn_in + keep * thin)
acc, tot = (int(x) for x in eng.imhk_counts())
prob = oracle.Problem(lat.basis, sigma)
t0 = time.perf_counter()
cpu_steps = cores * 4000
oracle.imhk_seeded(prob, cores, 4000, seed=1, threads=cores)
dt_cpu = time.perf_counter() - t0
out["C2"] = {"config": "IMHK, LLL lattice n=64, 4096 chains, burn_in=1000, thin=10, 10000 kept states per chain, "
                       "marginal histograms on the device",
             "seconds": ms / 1e3, "chain_steps": steps, "chain_steps_per_s": steps / ms * 1e3,
             "kept_states": int(hist[0].sum().item()), "acceptance_rate": acc / max(tot, 1),
             "cpu_port_chain_steps_per_s": cpu_steps / dt_cpu,
             "cpu_threads": cores}
print(json.dumps(out["C2"]), file=sys.stderr)
del V

# ---------------------------------------------------------------------------------------------------- C3
lat = qary_lattice(256, 512, 3329, seed=0)
sigma = 1.5 * 3329
smp = lgs.KleinSampler(lat, sigma, seed=3)
eng = smp._engine
S = 1 << 20
V = torch.empty((S, 512), dtype=torch.float64, device="cuda")
eng.klein_sample_dev(S, 3, 0, "fast", V=V.data_ptr(), stream=st)
torch.cuda.synchronize()
e0.record()
for r in range(3):
    eng.klein_sample_dev(S, 3, (r + 1) * S, "fast", V=V.data_ptr(), stream=st)
e1.record()
torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / 3
prob = oracle.Problem(lat.basis, sigma)
t0 = time.perf_counter()
oracle.klein_seeded(prob, 8192, seed=3, chains=cores, threads=cores)
dt_cpu = time.perf_counter() - t0
out["C3"] = {"config": "q-ary LWE lattice n=256 m=512 q=3329, sigma=1.5q, 2^20 Klein samples, device-resident",
             "ms": ms, "samples_per_s": S / ms * 1e3, "fp64_tflops": (3 * 512 * 512 - 512) * S / ms * 1e-9,
             "cpu_port_samples_per_s": 8192 / dt_cpu, "cpu_threads": cores}
print(json.dumps(out["C3"]), file=sys.stderr)
print(json.dumps(out, indent=1))
