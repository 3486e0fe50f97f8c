"""Development timing of the blocked IMHK path on one GPU (not the bench contract; see bench.py).
usage: python tools/quick_imhk.py [n] [chains or log2] [thin] [reps] [weight] [kept states per call]"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import __graft_entry__ as g  # noqa: E402

g.build() if not os.environ.get("LGS_B200_LIB") else None
import lattice_gaussian_mcmc_b200 as lgs  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
_a = int(sys.argv[2]) if len(sys.argv) > 2 else 16
C = (1 << _a) if _a <= 40 else _a
thin = int(sys.argv[3]) if len(sys.argv) > 3 else 10
reps = int(sys.argv[4]) if len(sys.argv) > 4 else 3
weight = sys.argv[5] if len(sys.argv) > 5 else "reference"
keep = int(sys.argv[6]) if len(sys.argv) > 6 else 1

rng = np.random.RandomState(0)
diag = rng.uniform(95, 130, n)
R = np.triu(rng.uniform(-0.5, 0.5, (n, n)) * diag[None, :], 1) + np.diag(diag)
lat = lgs.SimpleLattice(R)
smp = lgs.KleinSampler(lat, 165.74, seed=1, path="blocked")
eng = smp._engine
eng.imhk_reset(C, seed=3)
V = torch.empty((C, keep, n), dtype=torch.float64, device="cuda")
st = torch.cuda.current_stream().cuda_stream
for _ in range(2):
    eng.imhk_run_dev(0, keep, thin, "fast", weight, V=V.data_ptr(), stream=st)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for r in range(reps):
    eng.imhk_run_dev(0, keep, thin, "fast", weight, V=V.data_ptr(), stream=st)
e1.record()
torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / reps / keep
flop = n * (n - 1) + 6 * n + 2 * n * n / thin
print(f"n={n} chains={C} thin={thin} weight={weight} keep={keep}: {ms:.2f} ms per {thin} steps  {C * thin / ms * 1e3:.3e} chain-steps/s  "
      f"{flop * C * thin / ms * 1e-9:.2f} TFLOP/s algorithmic  acc={eng.imhk_counts()}")
eng.profile_enable(True)
for r in range(reps):
    eng.imhk_run_dev(0, keep, thin, "fast", weight, V=V.data_ptr(), stream=st)
prof = eng.profile_read()
eng.profile_enable(False)
print("  ms per MH step by class:", {k: round(v[0] / reps / thin, 3) for k, v in prof.items()},
      "launches/step", sum(v[1] for v in prof.values()) / reps / thin)
