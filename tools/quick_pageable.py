"""fp64 / int64 deliveries of lgs_klein_sample_host to pageable host memory: pinned bounce ring (default) against the
driver's own pageable copy (LGS_HOST_BOUNCE=0).  Development record.   python tools/quick_pageable.py"""
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import __graft_entry__ as g  # noqa: E402

g.build()
os.environ["LGS_HOST_PACK"] = "0"
import lattice_gaussian_mcmc_b200 as lgs  # noqa: E402
from lattice_gaussian_mcmc_b200.lattices import dd_lattice  # noqa: E402

for n in (1024, 256):
    lat = dd_lattice(n)
    smp = lgs.KleinSampler(lat, 1.5 * lat.max_gram_schmidt_norm, seed=1)
    eng = smp._engine
    S = 2 * eng.preferred_batch if n == 1024 else 8 * eng.preferred_batch
    for want in (dict(want_v=True), dict(want_v=True, want_x=True, want_logq=True)):
        eng.klein_sample(S, 1, 0, "fast", **want)
        t0 = time.time()
        reps = 3
        for r in range(reps):
            eng.klein_sample(S, 1, r * S, "fast", **want)
        dt = (time.time() - t0) / reps
        print(f"n={n} S={S} outputs={sorted(want)} bounce={os.environ.get('LGS_HOST_BOUNCE', '1')}: "
              f"{S / dt / 1e6:.3f} M samples/s, {dt * 1e3:.1f} ms per call", flush=True)
