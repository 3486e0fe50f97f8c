"""Development timing of the blocked Klein path on one GPU (not the bench contract; see bench.py).
usage: python tools/quick_klein.py [n] [log2 S] [reps]"""
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import __graft_entry__ as g  # noqa: E402

g.build() if not os.environ.get("LGS_B200_LIB") else None
import lattice_gaussian_mcmc_b200 as lgs  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
_a = int(sys.argv[2]) if len(sys.argv) > 2 else 17
S = (1 << _a) if _a <= 40 else _a
reps = int(sys.argv[3]) if len(sys.argv) > 3 else 5

rng = np.random.RandomState(0)
diag = rng.uniform(95, 130, n)
R = np.triu(rng.uniform(-0.5, 0.5, (n, n)) * diag[None, :], 1) + np.diag(diag)
lat = lgs.SimpleLattice(R)
t0 = time.time()
smp = lgs.KleinSampler(lat, 165.74, seed=1, path="blocked")
print(f"setup {time.time() - t0:.2f}s  sigma_i in [{165.74 / diag.max():.3f}, {165.74 / diag.min():.3f}]")
eng = smp._engine
V = torch.empty((S, n), dtype=torch.float64, device="cuda")
st = torch.cuda.current_stream().cuda_stream
for want_v in (True, False):
    for _ in range(2):
        eng.klein_sample_dev(S, 1, 0, "fast", V=V.data_ptr() if want_v else 0, stream=st)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for r in range(reps):
        eng.klein_sample_dev(S, 1, r * S, "fast", V=V.data_ptr() if want_v else 0, stream=st)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / reps
    flop = (3 * n * n - n) if want_v else n * (n - 1)
    print(f"n={n} S={S} want_v={want_v}: {ms:.2f} ms/batch  {S / ms * 1e3:.3e} samples/s  "
          f"{flop * S / ms * 1e-9:.2f} TFLOP/s algorithmic")

eng.profile_enable(True)
for r in range(reps):
    eng.klein_sample_dev(S, 1, r * S, "fast", V=V.data_ptr(), stream=st)
prof = eng.profile_read()
eng.profile_enable(False)
print("  per-batch ms by class:", {k: round(v[0] / reps, 3) for k, v in prof.items()}, "lib:",
      os.environ.get("LGS_B200_LIB", "default"))
