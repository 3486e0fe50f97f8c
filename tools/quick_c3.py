"""BASELINE config C3 alone (q-ary LWE n = 256, m = 512, q = 3329, sigma = 1.5 q; half the coordinates have wide
truncated windows): the workload for profiling the WIDE chain-kernel instantiation.   python tools/quick_c3.py [reps]"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import __graft_entry__ as g  # noqa: E402

g.build()
import lattice_gaussian_mcmc_b200 as lgs  # noqa: E402
from lattice_gaussian_mcmc_b200.lattices import qary_lattice  # noqa: E402

reps = int(sys.argv[1]) if len(sys.argv) > 1 else 3
lat = qary_lattice(256, 512, 3329, seed=0)
smp = lgs.KleinSampler(lat, 1.5 * 3329, seed=3)
eng = smp._engine
S = 1 << 20
V = torch.empty((S, 512), dtype=torch.float64, device="cuda")
st = torch.cuda.current_stream().cuda_stream
eng.klein_sample_dev(S, 3, 0, "fast", V=V.data_ptr(), stream=st)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for r in range(reps):
    eng.klein_sample_dev(S, 3, (r + 1) * S, "fast", V=V.data_ptr(), stream=st)
e1.record()
torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / reps
print(f"C3: {ms:.2f} ms per 2^20 samples = {S / ms / 1e3:.2f} M samples/s = {(3 * 512 * 512 - 512) * S / ms * 1e-9:.2f} TFLOP/s")
