#!/usr/bin/env python
"""bench.py -- Klein samples/s and IMHK chain-steps/s on the NTRU-512 trapdoor lattice (1024-dim), BASELINE config C4.

    python bench.py --gpus N --steps K --warmup W            # our arm (one process per GPU; torchrun for N > 1)
    python bench.py --impl reference --gpus N --steps K ...  # the reference's CPU algorithm (oracle port), rank 0 only

One "step" of the Klein leg = one batch of 170 496 samples per GPU (lgs_preferred_batch: 148 SMs x 9 tiles x 128)
through the hot path (lgs_klein_sample: blocked fp64 back-substitution + SampleZ + lattice points, one persistent
kernel), output resident in HBM.  One step of the IMHK leg = 10 MH steps of as many chains per GPU (lgs_imhk_run,
thin = 10, the kept state written as a lattice point), followed by the NCCL all-reduce of the accept counters and
marginal histograms.  Independent samples / chains are sharded over ranks by global id
(Philox key = seed, counter = global id): no data-path collective, weak scaling.

Prints ONE JSON line (rank 0).  `value` = Klein samples/s over all GPUs; the IMHK figure rides in "imhk".
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

N_RING, Q, SIGMA = 512, 12289, 165.74          # NTRU n = 512, q = 12289; sampler sigma = Falcon-512's
# config.workload of BOTH arms (the reference arm times the CPU implementation of the same path on this workload)
WORKLOAD = (f"NTRU n={N_RING} q={Q} trapdoor basis (1024-dim, f,g ~ D_Z,4.05 = BASELINE's key-generation sigma, exact "
            f"NTRUSolve, seed 0), sampler sigma={SIGMA}, Klein sampling + IMHK")
DIM = 2 * N_RING
KLEIN_FLOP = 3 * DIM * DIM - DIM                # back-substitution n(n-1) + lattice point 2n^2   
NCU_TRAFFIC_BYTES = 37.94e9                     # dram read 33.81 GB + write 4.13 GB per launch of the chain kernel at
                                                # S = 170 496 (profiles/r01j_ncu_full_summary.json; r01c: 37.72 GB)


def imhk_flop(thin):                             # proposal back-substitution + two quadratic forms + kept point / thin
    return DIM * (DIM - 1) + 6 * DIM + 2 * DIM * DIM / thin


def fp64_peak():
    """Measured FP64 DMMA peak of this pool's B200 (tools/fp64_peak.cu): MEASURED_PEAKS.json carries no FP64 figure."""
    for p in (os.path.join(ROOT, "profiles", "fp64_peak_r01.json"),):
        if os.path.exists(p):
            return float(json.load(open(p))["fp64_peak_tflops"]), "measured: tools/fp64_peak.cu (DMMA m8n8k4), profiles/fp64_peak_r01.json"
    return 37.2, "fallback: 148 SMs x 64 DFMA/clk x 1.965 GHz"


def build_lattice():
    """NTRU-512 trapdoor basis, seed 0 (about 3 s of host big-integer work; every rank builds the same one)."""
    from lattice_gaussian_mcmc_b200.ntru import ntru_lattice
    return ntru_lattice(N_RING, Q, seed=0)


class ClockSampler:
    """nvidia-smi clocks / throttle reasons while the timed regions run (B200_PROFILING.md recipe)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, device):
        self.rows, self.proc, self.device = [], None, device

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "100", "-i", str(self.device)], stdout=subprocess.PIPE, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append((time.perf_counter(), line.strip()))

    def stop(self, windows):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        sm, smax, reasons = [], None, set()
        for ts, line in self.rows:
            if not any(a <= ts <= b for a, b in windows):
                continue
            f = [x.strip() for x in line.split(",")]
            try:
                sm.append(float(f[1])); smax = float(f[2])
            except Exception:
                continue
            for name, val in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[5:9]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": smax, "reasons": sorted(reasons),
                "samples": len(sm)}


def cpu_baseline(lat, seconds=12.0):
    """The oracle port (oracle/lgs_oracle.c, OpenMP over independent seeded chains exactly like the reference's
    multiprocessing pattern, dimension_scaling.py:877-884) on this box's host cores, bounded sample."""
    from oracle import oracle
    cores = os.cpu_count() or 1
    prob = oracle.Problem(lat.basis, SIGMA)
    t0 = time.perf_counter()
    oracle.klein_seeded(prob, cores, seed=1, chains=cores, want_v=True, threads=cores)   # one sample per core
    per = max(time.perf_counter() - t0, 1e-3)
    S = int(max(cores, min(seconds / per, 4096) * cores))
    t0 = time.perf_counter()
    oracle.klein_seeded(prob, S, seed=2, chains=cores, want_v=True, threads=cores)
    dt = time.perf_counter() - t0
    return {"value": S / dt, "unit": "samples/s", "cores": cores, "kind": "port",
            "sample": f"{S} Klein samples of the same NTRU-512 lattice, {cores} OpenMP threads, {dt:.1f} s; "
                      "C restatement of klein.py:101-220 (the Python reference itself measured 2.7 samples/s/core at "
                      "n=1024 in the survey container, BASELINE.md section 2)"}


def run_reference(args, rank, world):
    if rank != 0:
        return
    import __graft_entry__ as g
    g.build()
    from oracle import oracle
    lat = build_lattice()
    cores = os.cpu_count() or 1
    prob = oracle.Problem(lat.basis, SIGMA)
    S = 2048 * cores                                                   # bounded sample per step (~3 s on 16 cores)
    for w in range(args.warmup):
        oracle.klein_seeded(prob, S // 8, seed=100 + w, chains=cores, want_v=True, threads=cores)
    t0 = time.perf_counter()
    for k in range(args.steps):
        oracle.klein_seeded(prob, S, seed=200 + k, chains=cores, want_v=True, threads=cores)
    dt = time.perf_counter() - t0
    val = S * args.steps / dt
    steps_imhk = 4
    t1 = time.perf_counter()
    oracle.imhk_seeded(prob, cores, steps_imhk, seed=5, threads=cores)
    dti = time.perf_counter() - t1
    imhk_val = cores * (steps_imhk + 1) / dti
    line = {"impl": "reference", "metric": "klein_samples_per_s", "value": val, "unit": "samples/s",
            "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": WORKLOAD, "samples_per_step": S,
                       "sample_of": "the same lattice, sigma and sampler as the GPU arm; a bounded sample per step"},
            "cpu_baseline": {"value": val, "unit": "samples/s", "cores": cores, "kind": "port",
                             "sample": f"{S} samples per step x {args.steps} steps, {cores} OpenMP threads"},
            "imhk": {"value": imhk_val, "unit": "chain-steps/s", "sample": f"{cores} chains x {steps_imhk} steps"},
            "e2e": {"value": val, "unit": "samples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=8)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours")
    ap.add_argument("--samples", type=int, default=0, help="Klein samples per GPU per step (0: lgs_preferred_batch)")
    ap.add_argument("--chains", type=int, default=0, help="IMHK chains per GPU (0: lgs_preferred_batch)")
    ap.add_argument("--thin", type=int, default=10)
    ap.add_argument("--skip-imhk", action="store_true")
    ap.add_argument("--skip-cpu", action="store_true")
    ap.add_argument("--skip-e2e", action="store_true")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        return run_reference(args, rank, world)

    import torch
    import torch.distributed as dist
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the sampling path has no CPU fallback")
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    import __graft_entry__ as g
    if rank == 0:
        g.build()
    if world > 1:
        dist.barrier()
    import lattice_gaussian_mcmc_b200 as lgs

    W = max(args.warmup, 3)
    K = args.steps
    n = DIM
    lat = build_lattice()
    smp = lgs.KleinSampler(lat, SIGMA, seed=2026, device=local, path="blocked")
    eng = smp._engine
    S = args.samples or eng.preferred_batch
    C = args.chains or eng.preferred_batch
    stream = torch.cuda.current_stream().cuda_stream
    V = torch.empty((S, n), dtype=torch.float64, device="cuda")
    clocks = ClockSampler(local)
    if rank == 0:
        clocks.start()
    windows = []

    def sync_all():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    def timed(fn, steps):
        """K calls bracketed by barrier + synchronize, CUDA events on the launching stream, max over ranks."""
        sync_all()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.perf_counter()
        e0.record()
        for k in range(steps):
            fn(k)
        e1.record()
        sync_all()
        windows.append((t0, time.perf_counter()))
        ms = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms.item())

    # ---------------------------------------------------------------- Klein leg (device resident)
    def klein_step(k, base=0):
        gid = ((base + k) * world + rank) * S           # global sample ids: disjoint per (step, rank)
        eng.klein_sample_dev(S, smp.seed, gid, "fast", V=V.data_ptr(), stream=stream)

    for w in range(W):
        klein_step(w, base=10_000)
    ms_klein = timed(klein_step, K)
    klein_rate = world * S * K / (ms_klein * 1e-3)

    # per-kernel-class device time of the same K steps (events around every launch; separate pass so that the
    # bracketing does not sit inside the figure above)
    eng.profile_enable(True)
    for k in range(K):
        klein_step(k, base=20_000)
    prof = eng.profile_read()
    eng.profile_enable(False)
    launches_per_step = sum(c for _, c in prof.values()) // K
    peak, peak_src = fp64_peak()
    if prof["chain"][1] > 0:        # the persistent chain kernel does the whole sample: (3 n^2 - n) flop per sample
        pt_ms, pt_cnt = prof["chain"]
        pt_flop = float(KLEIN_FLOP) * S
        kernel_name = ("klein_chain_kernel (persistent, warp-specialised: centers + SampleZ panels + lattice points; "
                       "fp64 DMMA fed by cp.async.bulk)")
    else:                           # LGS_CHAIN=0: the multi-kernel v1 path, dominant kernel = the lattice-point contraction
        pt_ms, pt_cnt = prof["point"]
        pt_flop = 2.0 * n * n * S
        kernel_name = "contract_kernel<64, MODE_POINT> (V = X B^T, fp64 DMMA)"
    pt_tf = pt_flop / (pt_ms / pt_cnt * 1e-3) * 1e-12
    share = {k: round(v[0] / max(sum(x[0] for x in prof.values()), 1e-9), 4) for k, v in prof.items()}

    # ---------------------------------------------------------------- e2e: the C ABI with HOST buffers
    e2e = None
    if not args.skip_e2e:
        S2 = S
        Vh = torch.empty((S2, n), dtype=torch.float64).pin_memory()
        import ctypes
        from lattice_gaussian_mcmc_b200._capi import check
        def host_step(k, base=30_000):
            gid = ((base + k) * world + rank) * S2
            check(eng._h, eng._lib.lgs_klein_sample_host(eng._h, S2, smp.seed, gid, 1, Vh.data_ptr(), None, None))
        for w in range(2):
            host_step(w, base=40_000)
        sync_all()
        t0 = time.perf_counter()
        for k in range(K):
            host_step(k)
        torch.cuda.synchronize()
        dt = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device="cuda")
        windows.append((t0, time.perf_counter()))
        if world > 1:
            dist.all_reduce(dt, op=dist.ReduceOp.MAX)
        transport = int(eng._lib.lgs_host_transport(eng._h))
        tname = {0: "fp64 copies", 1: "int16 pack kernel + chunked copies, widened to float64 by the host pool",
                 2: "int16 written by the chain kernel into mapped pinned memory (zero-copy), widened to float64 by "
                    "the host pool tile by tile while the kernel runs"}[transport]
        d2h = S2 * n * 8 if transport == 0 else S2 * n * 2 + 4 * ((S2 + 127) // 128)
        e2e = {"value": world * S2 * K / float(dt.item()), "unit": "samples/s", "h2d_bytes_per_step": 16,
               "d2h_bytes_per_step": d2h, "host_bytes_delivered_per_step": S2 * n * 8, "transport": tname,
               "host_threads_per_rank": int(os.environ.get("LGS_HOST_THREADS", 0)) or max(
                   1, len(os.sched_getaffinity(0)) // int(os.environ.get("LOCAL_WORLD_SIZE", "1"))),
               "note": "lgs_klein_sample_host: the sampler's only inputs are (seed, first id) scalars; every step "
                       "delivers its S x n float64 lattice points (the reference's return array) into a pinned host "
                       "buffer; for this integer basis they cross PCIe as int16, bit-identical after widening"}
        del Vh

    # ---------------------------------------------------------------- IMHK leg
    imhk = None
    if not args.skip_imhk:
        thin = args.thin
        eng.imhk_reset(C, seed=77, first_chain_id=rank * C)
        Vk = torch.empty((C, 1, n), dtype=torch.float64, device="cuda")
        counts = torch.zeros((C, 2), dtype=torch.int64, device="cuda")
        hist = torch.zeros((4, 64), dtype=torch.int64, device="cuda")
        dims = np.array([0, 1, n // 2, n - 1], dtype=np.int32)

        def imhk_step(k):
            eng.imhk_run_dev(0, 1, thin, "fast", "reference", V=Vk.data_ptr(), stream=stream)
            # statistics of src/diagnostics (acceptance mean, marginal histograms): accumulate per rank ...
            eng.histogram_dev(Vk.data_ptr(), C, dims, -8 * SIGMA, SIGMA / 4, 64, hist.data_ptr(), stream=stream)

        for w in range(2):
            imhk_step(w)
        hist.zero_()

        from lattice_gaussian_mcmc_b200 import distributed as D

        def imhk_all(k):
            imhk_step(k)
            if k == K - 1:                                # ... and all-reduce once at the end of the run (NCCL)
                eng.imhk_state_dev(counts=counts.data_ptr(), stream=stream)
                imhk_all.tot, imhk_all.hist = D.allreduce_stats(counts, hist)
        ms_imhk = timed(imhk_all, K)
        tot = imhk_all.tot.cpu().numpy()
        rate = world * C * thin * K / (ms_imhk * 1e-3)
        # per-class device time of the same work (separate pass, events around every launch)
        eng.profile_enable(True)
        for k in range(2):
            imhk_step(k)
        iprof = eng.profile_read()
        eng.profile_enable(False)
        itot = max(sum(v[0] for v in iprof.values()), 1e-9)
        imhk = {"value": rate, "unit": "chain-steps/s", "chains_per_gpu": C, "thin": thin, "ms_per_step": ms_imhk / K,
                "acceptance_rate": D.acceptance_rate(imhk_all.tot.cpu()),
                "flop_per_chain_step": imhk_flop(thin),
                "fp64_tflops_per_gpu": rate / world * imhk_flop(thin) * 1e-12,
                "fp64_frac_of_peak": rate / world * imhk_flop(thin) * 1e-12 / peak,
                "share_of_step": {k: round(v[0] / itot, 4) for k, v in iprof.items()},
                "gpu_launches_per_step": sum(v[1] for v in iprof.values()) // 2,
                "hist_total": int(imhk_all.hist.sum().item())}

    clk = clocks.stop(windows) if rank == 0 else None
    cpu = None
    if rank == 0 and world == 1 and not args.skip_cpu:
        cpu = cpu_baseline(lat)

    if rank == 0:
        gs = np.abs(np.diag(smp.R))
        line = {
            "metric": "klein_samples_per_s", "value": klein_rate, "unit": "samples/s", "n_gpus": world, "steps": K,
            "warmup": W, "ms_per_step": ms_klein / K, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": WORKLOAD,
                       "samples_per_gpu_per_step": S, "cond_sigma_range": [float(SIGMA / gs.max()), float(SIGMA / gs.min())],
                       "l2_policy": "each step writes S x n fp64 = %.2f GB of fresh output (> 126 MB L2) and streams "
                                    "its %.2f GB tiled coefficient workspace" % (S * n * 8 / 1e9, S * n * 8.25 / 1e9),
                       "parallelism": f"independent sample ranges per GPU x{world}, no data-path collective"},
            "fp64_tflops_per_gpu": klein_rate / world * KLEIN_FLOP * 1e-12,
            "fp64_frac_of_peak": klein_rate / world * KLEIN_FLOP * 1e-12 / peak,
            "roofline": {"bound": "tensor", "kernel": kernel_name,
                         "achieved": pt_tf, "peak": peak, "unit": "TFLOP/s", "frac": pt_tf / peak,
                         "peak_source": peak_src,
                         # dram__bytes_read.sum + dram__bytes_write.sum of this kernel per launch from the ncu
                         # --set full capture of this command (profiles/r01c_ncu_full_summary.json)
                         "traffic": NCU_TRAFFIC_BYTES if (S == 170496 and n == 1024 and prof["chain"][1] > 0) else None,
                         "algorithmic_flop_per_launch": pt_flop, "avg_launch_ms": pt_ms / pt_cnt,
                         "share_of_step": share},
            "gpu_launches": int(launches_per_step * K),
            "e2e": e2e, "imhk": imhk, "cpu_baseline": cpu, "clocks": clk,
        }
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
