"""Development timing of lgs_klein_sample_host (packed / zero-copy transport) on one GPU; not the bench contract.
usage: python tools/quick_e2e.py [n] [batches] [reps]"""
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import __graft_entry__ as g  # noqa: E402

g.build() if not os.environ.get("LGS_B200_LIB") else None
import lattice_gaussian_mcmc_b200 as lgs  # noqa: E402
from lattice_gaussian_mcmc_b200._capi import check  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
nb = float(sys.argv[2]) if len(sys.argv) > 2 else 1
reps = int(sys.argv[3]) if len(sys.argv) > 3 else 6

from lattice_gaussian_mcmc_b200.ntru import ntru_lattice  # noqa: E402
lat = ntru_lattice(n // 2, 12289, seed=0)     # the bench lattice at n = 1024
smp = lgs.KleinSampler(lat, 165.74, seed=1, path="blocked")
eng = smp._engine
S = int(eng.preferred_batch * nb)
Vh = torch.empty((S, n), dtype=torch.float64).pin_memory()
for r in range(2):
    check(eng._h, eng._lib.lgs_klein_sample_host(eng._h, S, 1, r * S, 1, Vh.data_ptr(), None, None))
eng.profile_enable(True)
t0 = time.perf_counter()
for r in range(reps):
    check(eng._h, eng._lib.lgs_klein_sample_host(eng._h, S, 1, r * S, 1, Vh.data_ptr(), None, None))
dt = (time.perf_counter() - t0) / reps
prof = eng.profile_read()
eng.profile_enable(False)
ms, cnt = prof["chain"]
print(f"n={n} S={S} host call {dt*1e3:.2f} ms = {S/dt/1e6:.2f} M samples/s | chain kernel {ms/max(cnt,1):.2f} ms x {cnt/reps:.1f} per call"
      f" | env pack={os.environ.get('LGS_HOST_PACK')} zc={os.environ.get('LGS_HOST_ZEROCOPY')} threads={os.environ.get('LGS_HOST_THREADS')}")

# the user-facing call: KleinSampler.sample(S) returns a FRESH pageable float64 array (klein.py:322)
for huge in ("1", "0"):
    os.environ["LGS_FRESH_HUGEPAGE"] = huge
    for _ in range(2):
        smp.sample(S)
    t0 = time.perf_counter()
    for r in range(3):
        out = smp.sample(S)
    dt = (time.perf_counter() - t0) / 3
    print(f"KleinSampler.sample({S}) -> {out.shape} {out.dtype}: {dt*1e3:.2f} ms = {S/dt/1e6:.2f} M samples/s "
          f"(fresh pageable array per call, MADV_HUGEPAGE mapping = {huge})")
    del out
buf = np.empty((S, n))
for _ in range(2):
    smp.sample(S, out=buf)
t0 = time.perf_counter()
for r in range(3):
    smp.sample(S, out=buf)
dt = (time.perf_counter() - t0) / 3
print(f"KleinSampler.sample({S}, out=buf): {dt*1e3:.2f} ms = {S/dt/1e6:.2f} M samples/s (caller-owned pageable array, reused)")
