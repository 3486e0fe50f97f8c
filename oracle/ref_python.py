"""The reference's OWN numpy sampler (src/samplers/{base,klein,imhk}.py) staged for the GPU box and timed there.

TEST / MEASUREMENT INFRASTRUCTURE ONLY (same rule as the rest of oracle/): used by bench.py's `cpu_baseline` and
`--impl reference` legs and by tests/; the product package never imports it.

`/root/reference` exists in the build container only.  stage() -- called by __graft_entry__.build() -- copies the three
sampler modules the hot path consists of, byte for byte, into oracle/_ref/src/samplers/ (git-ignored, NOT
gpurun-ignored: it travels to the GPU box with the snapshot, like the built .so files; nothing from the reference is
committed).  The package __init__ files written next to them are empty: the reference's own src/__init__.py pulls in
Sage-only modules.

time_klein() / time_imhk() run the reference the way its own scaling experiment does
(experiments/dimension_scaling.py:841-845, 877-884; experiments/scripts/benchmark_performance.py:240-266): a
multiprocessing.Pool of independent samplers, worker i seeded np.random.seed(seed + 1000 * i), one warm-up sample
excluded, wall clock by time.perf_counter(), OMP_NUM_THREADS=1.
"""
from __future__ import annotations

import os
import shutil
import sys
import time

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
REF_ROOT = os.environ.get("LGS_REFERENCE", "/root/reference")
STAGE = os.path.join(_HERE, "_ref")
FILES = ("base.py", "klein.py", "imhk.py")


def stage(force: bool = False) -> bool:
    """Copy the reference sampler modules into oracle/_ref when the reference tree is mounted.  Returns True when the
    staged copy exists afterwards."""
    src_dir = os.path.join(REF_ROOT, "src", "samplers")
    dst_dir = os.path.join(STAGE, "src", "samplers")
    if os.path.isdir(src_dir):
        os.makedirs(dst_dir, exist_ok=True)
        for f in FILES:
            s, d = os.path.join(src_dir, f), os.path.join(dst_dir, f)
            if force or not os.path.exists(d) or os.path.getmtime(d) < os.path.getmtime(s):
                shutil.copyfile(s, d)
        for init in (os.path.join(STAGE, "src", "__init__.py"), os.path.join(dst_dir, "__init__.py")):
            if not os.path.exists(init):
                open(init, "w").close()
    return available()


def available() -> bool:
    return all(os.path.exists(os.path.join(STAGE, "src", "samplers", f)) for f in FILES)


class ShimLattice:
    """The duck type the reference samplers consume: basis (columns = basis vectors, as
    klein.py:59,218 uses it), dimension, min_gram_schmidt_norm, smoothing_parameter(), name."""

    def __init__(self, basis, min_gs=None, name="shim"):
        self.basis = np.array(basis, dtype=np.float64)
        self.dimension = self.basis.shape[0]
        self.name = name
        if min_gs is None:   # |R_ii| of the column basis -- only used for a warning and the burn-in estimate
            min_gs = float(np.min(np.abs(np.diag(np.linalg.qr(self.basis, mode="r")))))
        self.min_gram_schmidt_norm = float(min_gs)

    def smoothing_parameter(self):
        return 0.0


def _import_reference():
    sys.dont_write_bytecode = True
    if STAGE not in sys.path:
        sys.path.insert(0, STAGE)
    # the staged package is named `src` like the reference's; make sure no other `src` shadows it
    for k in [k for k in sys.modules if k == "src" or k.startswith("src.")]:
        if not getattr(sys.modules[k], "__file__", "").startswith(STAGE):
            del sys.modules[k]
    from src.samplers.klein import RefinedKleinSampler
    from src.samplers.imhk import IMHKSampler
    return RefinedKleinSampler, IMHKSampler


def _worker(args):
    kind, basis, sigma, count, seed, wid, thin = args
    os.environ["OMP_NUM_THREADS"] = "1"
    import logging
    logging.disable(logging.CRITICAL)
    Klein, IMHK = _import_reference()
    lat = ShimLattice(basis)
    np.random.seed(seed + 1000 * wid)                      # dimension_scaling.py:879
    if kind == "klein":
        smp = Klein(lat, sigma)
        smp.sample(1)                                      # warm-up, excluded (benchmark_performance.py:240-266)
        t0 = time.perf_counter()
        out = smp.sample(count)
        dt = time.perf_counter() - t0
        units = count
    else:
        smp = IMHK(lat, sigma, burn_in=0)
        smp.sample(1)                                      # _initialize_chain + one step
        t0 = time.perf_counter()
        out = smp.sample(count, thin=thin)
        dt = time.perf_counter() - t0
        units = count * thin
    return units, dt, float(np.asarray(out).std())


def _run(kind, basis, sigma, procs, count, seed, thin=1):
    import multiprocessing as mp
    basis = np.ascontiguousarray(basis, dtype=np.float64)
    jobs = [(kind, basis, float(sigma), int(count), int(seed), i, int(thin)) for i in range(procs)]
    t0 = time.perf_counter()
    if procs == 1:
        res = [_worker(jobs[0])]
    else:
        # dimension_scaling.py:841-845 uses Pool(n) (fork); here the parent may hold CUDA / OpenMP threads, so the
        # workers are spawned: same processes, same seeding, and the timed region lies inside the worker
        ctx = mp.get_context("spawn")
        saved = {k: os.environ.get(k) for k in ("OMP_NUM_THREADS", "OPENBLAS_NUM_THREADS", "MKL_NUM_THREADS")}
        for k in saved:                                    # inherited by the workers BEFORE they import numpy
            os.environ[k] = "1"
        try:
            with ctx.Pool(procs) as pool:
                res = pool.map(_worker, jobs)
        finally:
            for k, v in saved.items():
                if v is None:
                    os.environ.pop(k, None)
                else:
                    os.environ[k] = v
    wall = time.perf_counter() - t0
    units = sum(r[0] for r in res)
    per_core = float(np.mean([r[0] / r[1] for r in res]))
    slowest = max(r[1] for r in res)
    return {"value": units / slowest, "per_core": per_core, "cores": procs, "procs": procs,
            "units_per_proc": res[0][0], "sampling_s": slowest, "wall_s": wall,
            "versions": {"numpy": np.__version__, "scipy": __import__("scipy").__version__,
                         "python": sys.version.split()[0]}}


def time_klein(basis, sigma, procs=None, samples_per_proc=32, seed=0):
    """RefinedKleinSampler.sample(k) (klein.py:324-337) in `procs` independently seeded processes.
    value = samples/s of the box (all processes, by the slowest process' sampling time); per_core = mean rate of one."""
    procs = procs or (os.cpu_count() or 1)
    out = _run("klein", basis, sigma, procs, samples_per_proc, seed)
    out["unit"] = "samples/s"
    return out


def time_imhk(basis, sigma, procs=None, steps_per_proc=16, thin=1, seed=0):
    """IMHKSampler.sample(k, thin) (imhk.py:196-229): value = chain-steps/s of the box."""
    procs = procs or (os.cpu_count() or 1)
    out = _run("imhk", basis, sigma, procs, steps_per_proc, seed, thin=thin)
    out["unit"] = "chain-steps/s"
    return out
