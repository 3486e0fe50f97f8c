#!/usr/bin/env python
"""Extract the reference's RECORDED outputs for the LLL-reduced scaling runs into a small fixture.

    python tests/golden/make_lll_recorded.py         # writes tests/golden/klein_lll_recorded.json

Source: /root/reference/results/Klein_LLL_n={16,32,64}.json, written by klein_scaling_analysis.py:288-348
(random integer matrix, seed 0+n, entries 0..50 -> Sage LLL -> sigma = 1.5 max||b*|| -> 50 000 Klein samples ->
per-coordinate statistics of the recovered integer coefficients, klein_scaling_analysis.py:190-230).  These are data
the reference published about this path; they pin (i) the host LLL stand-in (Gram-Schmidt profile of the reduced
basis, sigma) and (ii) the sampler's per-coordinate spread on that basis."""
import json
import os

REF = os.environ.get("LGS_REFERENCE", "/root/reference")
OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "klein_lll_recorded.json")

out = {"source": "results/Klein_LLL_n=*.json of NickQrumpton/lattice-gaussian-mcmc (klein_scaling_analysis.py:288-348)",
       "n_samples": 50000, "cases": {}}
for n in (16, 32, 64):
    d = json.load(open(os.path.join(REF, "results", f"Klein_LLL_n={n}.json")))
    out["cases"][str(n)] = {
        "seed": d["seed"], "sigma": d["sigma"], "max_GS_norm": d["max_GS_norm"],
        "gs_norms": d["additional_info"]["gs_norms"],
        "all_stds": d["quality_metrics"]["all_stds"],
        "all_means": d["quality_metrics"]["all_means"],
    }
json.dump(out, open(OUT, "w"), indent=1)
print(OUT)
