"""Development: what bounds the end-to-end (host-buffer) rate when several ranks share one host.  Run under torchrun;
every rank times, with a barrier before each leg: (a) the pool widening a resident int16 buffer into the 4-batch pinned
float64 buffer, 4 batches back to back (the host's sustained ceiling for this delivery), (b) lgs_klein_sample_host, 4
batches per call, for several pool sizes (LGS_HOST_THREADS is read when a handle first needs its pool).
usage: torchrun --nproc-per-node N tools/e2e_ranks.py"""
import os
import sys
import time

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
if world > 1:
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
import __graft_entry__ as g  # noqa: E402
if rank == 0:
    g.build()
if world > 1:
    dist.barrier()
import lattice_gaussian_mcmc_b200 as lgs  # noqa: E402
from lattice_gaussian_mcmc_b200._capi import check  # noqa: E402
from lattice_gaussian_mcmc_b200.ntru import ntru_lattice  # noqa: E402

n = 1024
lat = ntru_lattice(512, 12289, seed=0)
cpus = len(os.sched_getaffinity(0))


def sync():
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
        torch.cuda.synchronize()


def tmax(dt):
    t = torch.tensor([dt], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


for threads in sorted({max(1, cpus // world), max(1, cpus // world - 1), max(1, cpus // world + 2)}, reverse=True):
    os.environ["LGS_HOST_THREADS"] = str(threads)
    smp = lgs.KleinSampler(lat, 165.74, seed=1, device=local, path="blocked")
    eng = smp._engine
    S = eng.preferred_batch
    S4 = 4 * S
    Vh = torch.empty((S4, n), dtype=torch.float64).pin_memory()
    src16 = np.ascontiguousarray(np.tile(np.random.RandomState(1).randint(-3000, 3000, (1024, n)).astype(np.int16), (S4 // 1024 + 1, 1))[:S4])
    dst = Vh.numpy()
    best = 1e9
    for r in range(3):
        sync()
        t0 = time.perf_counter()
        check(None, eng._lib.lgs_host_widen_i16(src16.ctypes.data, dst.ctypes.data, S4 * n, threads))
        best = min(best, tmax(time.perf_counter() - t0))
    peak = world * S4 * n * 8 / best / 1e9
    for r in range(2):
        check(eng._h, eng._lib.lgs_klein_sample_host(eng._h, S4, 1, (r * world + rank) * S4, 1, Vh.data_ptr(), None, None))
    sync()
    t0 = time.perf_counter()
    reps = 3
    for r in range(reps):
        check(eng._h, eng._lib.lgs_klein_sample_host(eng._h, S4, 1, ((r + 2) * world + rank) * S4, 1, Vh.data_ptr(), None, None))
    torch.cuda.synchronize()
    dt = tmax(time.perf_counter() - t0)
    rate = world * S4 * reps / dt
    if rank == 0:
        print(f"ranks {world} cpus {cpus} pool threads/rank {threads}: sustained widen {peak:.1f} GB/s = {peak / (8 * n) * 1e3:.2f} M samples/s "
              f"ceiling | e2e {rate / 1e6:.2f} M samples/s = {rate * n * 8 / 1e9:.1f} GB/s ({rate * n * 8 / 1e9 / peak * 100:.0f} % of it)", flush=True)
    del smp, eng, Vh, src16, dst
if world > 1:
    dist.destroy_process_group()
