#!/usr/bin/env python
"""bench.py -- Klein samples/s and IMHK chain-steps/s on the NTRU-512 trapdoor lattice (1024-dim), BASELINE config C4.

    python bench.py --gpus N --steps K --warmup W            # our arm (one process per GPU; torchrun for N > 1)
    python bench.py --impl reference --gpus N --steps K ...  # the reference's CPU algorithm (oracle port), rank 0 only

One "step" of the Klein leg = one batch of 170 496 samples per GPU (lgs_preferred_batch: 148 SMs x 9 tiles x 128)
through the hot path (lgs_klein_sample: blocked fp64 back-substitution + SampleZ + lattice points, one persistent
kernel), output resident in HBM.  One step of the IMHK leg = one lgs_imhk_run call for 4 kept states at thin = 10 of as
many chains per GPU (40 MH steps, the kept states written as lattice points, marginal histograms counted by the kernel's
epilogue), followed at the end of the run by the NCCL all-reduce of the accept counters and histograms.  Independent samples / chains are sharded over ranks by global id
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
SIGMA_REJECT = 55.0                            # exact-weight IMHK leg that rejects (acceptance ~0.5-0.9 on this lattice)
# config.workload of BOTH arms (the reference arm times the CPU implementation of the same path on this workload)
WORKLOAD = (f"NTRU n={N_RING} q={Q} trapdoor basis (1024-dim, f,g ~ D_Z,4.05 = BASELINE's key-generation sigma, exact "
            f"NTRUSolve, seed 0), sampler sigma={SIGMA}, Klein sampling + IMHK")
DIM = 2 * N_RING
KLEIN_FLOP = 3 * DIM * DIM - DIM                # back-substitution n(n-1) + lattice point 2n^2   


def ncu_traffic():
    """dram__bytes_read.sum + dram__bytes_write.sum per launch of the Klein chain kernel, from the NEWEST committed
    `ncu --set full` summary of this command (profiles/*ncu_full_summary.json, written by tools/ncu_summary.py).
    Returns (bytes or None, source file)."""
    import glob
    import re
    def val(txt):
        m = re.match(r"([0-9.eE+-]+)\s*([KMGT]?)byte", txt.strip())
        return float(m.group(1)) * {"": 1, "K": 1e3, "M": 1e6, "G": 1e9, "T": 1e12}[m.group(2)] if m else None
    files = sorted(glob.glob(os.path.join(ROOT, "profiles", "r*_ncu_full_summary.json")),
                   key=lambda f: (re.sub(r"^r(\d+)([a-z]*)_.*", lambda m: f"{int(m.group(1)):04d}{m.group(2):>3}",
                                         os.path.basename(f))))
    for f in reversed(files):
        try:
            for k in json.load(open(f)).get("kernels", []):
                name = k.get("Kernel Name", "")
                if "klein_chain_kernel<0, 0, 0, 0" in name and k.get("launch__grid_size") == "148":
                    r, w = val(k["dram__bytes_read.sum"]), val(k["dram__bytes_write.sum"])
                    if r is not None and w is not None:
                        return r + w, os.path.relpath(f, ROOT)
        except Exception:
            continue
    return None, None


_T0 = time.perf_counter()


def phase(name):
    """wall-clock trace of the run on stderr (the JSON line on stdout stays alone)"""
    if int(os.environ.get("RANK", "0")) == 0:
        print(f"[bench +{time.perf_counter() - _T0:7.1f} s] {name}", file=sys.stderr, flush=True)


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
                      "C restatement of klein.py:101-220 (value); the reference's own numpy code timed beside it in "
                      "reference_python",
            "reference_python": reference_python_leg(lat)}


def reference_python_leg(lat, samples_per_proc=24, imhk_steps_per_proc=8):
    """The reference's OWN src/samplers/klein.py + imhk.py (staged into oracle/_ref by build(), see
    oracle/ref_python.py) on this box's host cores, the way its scaling experiment runs them: one independently seeded
    process per core (experiments/dimension_scaling.py:841-845, 877-884), warm-up sample excluded.  About 0.35 s per
    1024-dim sample per core => ~10 s for the Klein leg."""
    from oracle import ref_python
    if not ref_python.available():
        return {"unavailable": "oracle/_ref not staged (build() stages it where /root/reference is mounted)"}
    procs = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)
    k = ref_python.time_klein(lat.basis, SIGMA, procs=procs, samples_per_proc=samples_per_proc, seed=0)
    m = ref_python.time_imhk(lat.basis, SIGMA, procs=procs, steps_per_proc=imhk_steps_per_proc, thin=1, seed=0)
    return {"value": k["value"], "unit": "samples/s", "cores": procs, "procs": procs, "per_core": k["per_core"],
            "sample": f"RefinedKleinSampler.sample({samples_per_proc}) in each of {procs} processes "
                      f"(np.random.seed(0 + 1000 i)), {k['sampling_s']:.1f} s; same lattice and sigma",
            "imhk": {"value": m["value"], "unit": "chain-steps/s", "per_core": m["per_core"],
                     "sample": f"IMHKSampler.sample({imhk_steps_per_proc}) per process, {m['sampling_s']:.1f} s"},
            "versions": k["versions"]}


def host_info():
    """what the host offers the ranks: CPUs, NUMA nodes, this process' affinity"""
    info = {"nproc": os.cpu_count(), "affinity": sorted(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else None}
    try:
        nodes = sorted(d for d in os.listdir("/sys/devices/system/node") if d.startswith("node"))
        info["numa_nodes"] = {d: open(f"/sys/devices/system/node/{d}/cpulist").read().strip() for d in nodes}
    except Exception:
        info["numa_nodes"] = None
    try:
        for line in open("/proc/cpuinfo"):
            if line.startswith("model name"):
                info["cpu_model"] = line.split(":", 1)[1].strip()
                break
        info["thp"] = open("/sys/kernel/mm/transparent_hugepage/enabled").read().strip()
    except Exception:
        pass
    a = info["affinity"]
    if a is not None:
        info["affinity"] = f"{len(a)} cpus: {a[0]}-{a[-1]}" if a == list(range(a[0], a[-1] + 1)) else a
    return info


def run_reference(args, rank, world):
    if rank != 0:
        return
    import __graft_entry__ as g
    g.build_oracle()               # the checker only: this arm never loads the product library
    from oracle import oracle
    phase("oracle built")
    lat = build_lattice()
    phase("lattice (NTRUSolve)")
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
    phase("port: klein steps")
    steps_imhk = 4
    t1 = time.perf_counter()
    oracle.imhk_seeded(prob, cores, steps_imhk, seed=5, threads=cores)
    dti = time.perf_counter() - t1
    imhk_val = cores * (steps_imhk + 1) / dti
    phase("port: imhk")
    refpy = reference_python_leg(lat)
    phase("reference python legs")
    line = {"impl": "reference", "metric": "klein_samples_per_s", "value": val, "unit": "samples/s",
            "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            # `config` names the workload and is the same dict in both arms; how each arm sizes a step is `step_shape`
            "config": {"workload": WORKLOAD},
            "step_shape": {"samples_per_step": S,
                           "sample_of": "the same lattice, sigma and sampler as the GPU arm; a bounded sample per step"},
            "cpu_baseline": {"value": val, "unit": "samples/s", "cores": cores, "kind": "port",
                             "sample": f"{S} samples per step x {args.steps} steps, {cores} OpenMP threads",
                             "reference_python": refpy},
            "host": host_info(),
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
    ap.add_argument("--check-invariance", action="store_true",
                    help="(always on) every rank hashes the samples of ids [0, 4096); see `invariance` in the line")
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
    phase("built / loaded the library")
    lat = build_lattice()
    phase("lattice (NTRUSolve)")
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

    phase("sampler created")
    for w in range(W):
        klein_step(w, base=10_000)
    ms_klein = timed(klein_step, K)
    phase("klein leg")
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
    traffic, traffic_src = ncu_traffic()
    try:
        hbm_peak = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
    except Exception:
        hbm_peak = None
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
        import ctypes
        from lattice_gaussian_mcmc_b200._capi import check
        BATCHES = 4                                       # batches per host call: the streamed int16 route
        S2 = BATCHES * S
        Vh = torch.empty((S2, n), dtype=torch.float64).pin_memory()

        def host_step(k, base=30_000):
            gid = ((base + k) * world + rank) * S2
            check(eng._h, eng._lib.lgs_klein_sample_host(eng._h, S2, smp.seed, gid, 1, Vh.data_ptr(), None, None))

        def wall(fn, steps):
            sync_all()
            t0 = time.perf_counter()
            for k in range(steps):
                fn(k)
            torch.cuda.synchronize()
            dt = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device="cuda")
            windows.append((t0, time.perf_counter()))
            if world > 1:
                dist.all_reduce(dt, op=dist.ReduceOp.MAX)
            return float(dt.item())

        phase("pinned host buffer")
        for w in range(2):
            host_step(w, base=40_000)
        Ke = max(2, (K + BATCHES - 1) // BATCHES)         # the same number of batches as K single-batch steps
        dt = wall(host_step, Ke)
        phase("e2e leg")
        transport = int(eng._lib.lgs_host_transport(eng._h))
        tname = {0: "fp64 copies", 1: "int16 pack kernel + chunked copies, widened to float64 by the host pool",
                 2: "int16 written by the chain kernel's lattice-point epilogue (HBM), shipped group by group by the "
                    "copy engine behind cuStreamWaitValue32 while the kernel runs, widened to float64 by the host pool"
                 }[transport]
        # the library's own pool size (csrc/lgs_hostpack.cpp hostpool_default_threads): the CPUs a rank may use, half
        # again as many when four or more ranks share the host
        lws = int(os.environ.get("LOCAL_WORLD_SIZE", "1"))
        threads = int(os.environ.get("LGS_HOST_THREADS", 0)) or max(1, len(os.sched_getaffinity(0)) // lws)
        if not os.environ.get("LGS_HOST_THREADS") and lws >= 4:
            threads += threads // 2
        # the host's own ceiling for this delivery: the pool widening a RESIDENT int16 buffer into the same float64
        # array with non-temporal stores, every rank at once (what the transport does minus the GPU and PCIe)
        # (a source of distinct, touched pages: np.zeros would map every page to the kernel's zero page and the "ceiling"
        # would have no read traffic at all -- it read 194 GB/s where the same pool sustains ~130 GB/s on real data)
        src16 = np.tile(np.random.RandomState(1).randint(-3000, 3000, (1024, n)).astype(np.int16), ((S + 1023) // 1024, 1))[:S]
        src16 = np.ascontiguousarray(src16)
        dst64 = Vh.numpy()[:S]
        best = 1e9
        for r in range(4):
            sync_all()
            t0 = time.perf_counter()
            check(None, eng._lib.lgs_host_widen_i16(src16.ctypes.data, dst64.ctypes.data, S * n, threads))
            best = min(best, time.perf_counter() - t0)
        tb = torch.tensor([best], dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(tb, op=dist.ReduceOp.MAX)
        store_gbs = world * S * n * 8 / float(tb.item()) / 1e9
        e2e_rate = world * S2 * Ke / dt
        d2h = S * n * 8 if transport == 0 else S * n * 2 + 4 * ((S + 127) // 128)
        e2e = {"value": e2e_rate, "unit": "samples/s", "h2d_bytes_per_step": 16,
               "d2h_bytes_per_step": d2h, "host_bytes_delivered_per_step": S * n * 8,
               "call_shape": f"lgs_klein_sample_host into a pinned, reused float64 [{S2}, {n}] buffer, {BATCHES} batches "
                             f"of {S} per call (a 'step' of this leg = one batch), {Ke} calls timed",
               "transport": tname, "host_threads_per_rank": threads,
               "host_store_gbs_peak": store_gbs,
               "host_store_peak_how": "lgs_host_widen_i16 on a resident int16 buffer of distinct pages into the same pinned "
                                      "float64 array, all ranks at once with their pool threads, best of 4 (the delivery "
                                      "also takes the DMA write of the int16, 2 more bytes per value)",
               "frac_of_host_peak": e2e_rate * n * 8 / 1e9 / store_gbs,
               "note": "the sampler's only inputs are (seed, first id) scalars; every step delivers its S x n float64 "
                       "lattice points (the reference's return array) into host memory; for this integer basis they "
                       "cross PCIe as int16, bit-identical after widening"}
        phase("host store peak")
        del Vh, src16, dst64
        # the reference-shaped call: KleinSampler.sample(S) returns a FRESH float64 array every time (klein.py:322)
        def fresh_step(k):
            smp.next_sample_id = ((50_000 + k) * world + rank) * S   # public: disjoint Philox ranges per rank and step
            out = smp.sample(S)
            assert out.shape == (S, n)
        for w in range(2):
            fresh_step(w)
        Kf = max(2, min(K, 6))
        dtf = wall(fresh_step, Kf)
        e2e["fresh_array"] = {"value": world * S * Kf / dtf, "unit": "samples/s",
                              "call_shape": f"KleinSampler.sample({S}) -> a fresh float64 array per call (the reference's "
                                            "return contract; 2 MB-aligned anonymous mapping + MADV_HUGEPAGE, first "
                                            "touched by the widening pool)", "calls": Kf}
        e2e["host"] = host_info()
        phase("e2e fresh-array leg")

    # ---------------------------------------------------------------- IMHK leg
    imhk = imhk_exact = None
    if not args.skip_imhk:
        from lattice_gaussian_mcmc_b200 import distributed as D
        thin = args.thin
        KEEP = 4          # kept states per lgs_imhk_run call, as IMHKSampler.sample(num_samples, thin) asks for many per call
        Vk = torch.empty((C, KEEP, n), dtype=torch.float64, device="cuda")
        dims = np.array([0, 1, n // 2, n - 1], dtype=np.int32)

        def imhk_leg(weight, steps, eng=eng, sigma=SIGMA):
            eng.imhk_reset(C, seed=77, first_chain_id=rank * C)
            counts = torch.zeros((C, 2), dtype=torch.int64, device="cuda")
            hist = torch.zeros((4, 64), dtype=torch.int64, device="cuda")
            res = {}

            # statistics of src/diagnostics (acceptance mean, marginal histograms): accumulated per rank by the lattice-point
            # epilogue of the sampling kernel itself (lgs_attach_histogram: no separate pass over V) ...
            eng.attach_histogram(dims, -8 * sigma, sigma / 4, 64, hist.data_ptr())

            def imhk_step(k):
                eng.imhk_run_dev(0, KEEP, thin, "fast", weight, V=Vk.data_ptr(), stream=stream)

            for w in range(W):
                imhk_step(w)
            hist.zero_()

            def imhk_all(k):
                imhk_step(k)
                if k == steps - 1:                            # ... and all-reduce once at the end of the run (NCCL)
                    eng.imhk_state_dev(counts=counts.data_ptr(), stream=stream)
                    res["tot"], res["hist"] = D.allreduce_stats(counts, hist)
            ms = timed(imhk_all, steps)
            rate = world * C * thin * KEEP * steps / (ms * 1e-3)
            # per-class device time of the same work (separate pass, events around every launch)
            eng.profile_enable(True)
            for k in range(2):
                imhk_step(k)
            iprof = eng.profile_read()
            eng.profile_enable(False)
            itot = max(sum(v[0] for v in iprof.values()), 1e-9)
            eng.attach_histogram(None, 0.0, 1.0, 1, 0)
            return {"value": rate, "unit": "chain-steps/s", "weight_mode": weight, "chains_per_gpu": C, "thin": thin,
                    "kept_states_per_call": KEEP,
                    "ms_per_step": ms / steps, "acceptance_rate": D.acceptance_rate(res["tot"].cpu()),
                    "flop_per_chain_step": imhk_flop(thin),
                    "fp64_tflops_per_gpu": rate / world * imhk_flop(thin) * 1e-12,
                    "fp64_frac_of_peak": rate / world * imhk_flop(thin) * 1e-12 / peak,
                    "share_of_step": {k: round(v[0] / itot, 4) for k, v in iprof.items()},
                    "gpu_launches_per_step": sum(v[1] for v in iprof.values()) // 2,
                    "hist_total": int(res["hist"].sum().item())}

        imhk = imhk_leg("reference", K)
        phase("imhk leg")
        # the Wang-Ling weights (LGS_WEIGHT_EXACT): normalisers accumulated, exp + accept draw per step.  At this
        # lattice's sigma_i >= 1.29 the normalisers are constant to ~1e-13, so the acceptance rate stays 1 - O(1e-13);
        # tools/quick_imhk.py runs the same mode where it rejects (sigma = 60: sigma_i ~ 0.5)
        imhk_exact = imhk_leg("exact", max(3, K // 2))
        phase("imhk exact leg")
        # ... and the same mode where it does reject: the same lattice at sigma = SIGMA_REJECT (sigma_i ~ 0.4-0.55, the
        # window masses now depend on the centers), i.e. accept / reject, the deferred commit of accepted states and the
        # exp per step all do work.  sigma_i < 1 takes the generic FP64 SampleZ (samplez_fast), not the FP32-screened one.
        smp_r = lgs.KleinSampler(lat, SIGMA_REJECT, seed=2027, device=local, path="blocked")
        imhk_exact["rejecting"] = imhk_leg("exact", max(3, K // 2), eng=smp_r._engine, sigma=SIGMA_REJECT)
        imhk_exact["rejecting"]["sigma"] = SIGMA_REJECT
        imhk_exact["rejecting"]["cond_sigma_range"] = [SIGMA_REJECT / lat.max_gram_schmidt_norm,
                                                       SIGMA_REJECT / lat.min_gram_schmidt_norm]
        del smp_r
        phase("imhk exact leg, rejecting sigma")

    # ---------------------------------------------------------------- replay through the same kernel (parity mode)
    replay = None
    if not args.skip_imhk:
        Sr = S // 3                                           # one round of tiles: the table sampler is ~40x slower
        U = torch.rand((Sr, eng.ups), dtype=torch.float64, device="cuda")
        if eng.replay_on_chain:
            eng.klein_replay_dev(Sr, U.data_ptr(), V=V.data_ptr(), stream=stream)
            ms_r = timed(lambda k: eng.klein_replay_dev(Sr, U.data_ptr(), V=V.data_ptr(), stream=stream), 2)
            replay = {"value": world * Sr * 2 / (ms_r * 1e-3), "unit": "samples/s", "samples_per_gpu_per_step": Sr,
                      "kernel": "klein_chain_kernel<0,1,0,0,REPLAY> (recorded uniforms + the reference's table SampleZ "
                                "on DMMA-contracted means; lgs_klein_replay on a LGS_PATH_BLOCKED handle)"}
        del U

    # ---------------------------------------------------------------- rank-count invariance (SURVEY 8e "test this")
    # Every rank samples ids [0, 4096) in full AND its own 1/world slice of them; a 64-bit hash of the integer
    # coefficients, additive over samples, must (a) be the same on every rank and (b) equal the all-reduced sum of the
    # slice hashes.  The N = 1, 2, 4, 8 lines of a scaling run therefore carry the same `hash`.
    def coeff_hash(first, count):
        Xi = torch.empty((count, n), dtype=torch.int64, device="cuda")
        eng.klein_sample_dev(count, 424242, first, "fast", X=Xi.data_ptr(), stream=stream)
        gid = torch.arange(first, first + count, dtype=torch.int64, device="cuda")[:, None] * n + \
            torch.arange(n, dtype=torch.int64, device="cuda")[None, :]
        wgt = gid * -7046029254386353131 + 0x2545F4914F6CDD1D        # odd multiplier: int64 arithmetic wraps
        return (Xi * wgt).sum()
    NI = 4096
    h_full = coeff_hash(0, NI)
    part = NI // world
    h_part = coeff_hash(rank * part, part if rank < world - 1 else NI - rank * part)
    h_sum, h_min, h_max = h_part.clone(), h_full.clone(), h_full.clone()
    if world > 1:
        dist.all_reduce(h_sum, op=dist.ReduceOp.SUM)
        dist.all_reduce(h_min, op=dist.ReduceOp.MIN)
        dist.all_reduce(h_max, op=dist.ReduceOp.MAX)
    invariance = {"ids": f"[0, {NI}) under seed 424242", "hash": f"{int(h_full.item()) & (2**64 - 1):016x}",
                  "all_ranks_equal": bool(h_min.item() == h_max.item() == h_full.item()),
                  "sum_of_rank_slices_equals_full": bool(h_sum.item() == h_full.item())}
    if not (invariance["all_ranks_equal"] and invariance["sum_of_rank_slices_equals_full"]):
        raise SystemExit(f"rank-count invariance violated: {invariance}")

    phase("replay + invariance")
    clk = clocks.stop(windows) if rank == 0 else None
    cpu = None
    if rank == 0 and world == 1 and not args.skip_cpu:
        cpu = cpu_baseline(lat)
        phase("cpu baseline (port + reference python)")

    if rank == 0:
        gs = np.abs(np.diag(smp.R))
        line = {
            "metric": "klein_samples_per_s", "value": klein_rate, "unit": "samples/s", "n_gpus": world, "steps": K,
            "warmup": W, "ms_per_step": ms_klein / K, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            # `config` names the workload and is the same dict in both arms; how each arm sizes a step is `step_shape`
            "config": {"workload": WORKLOAD},
            "step_shape": {"samples_per_gpu_per_step": S,
                           "cond_sigma_range": [float(SIGMA / gs.max()), float(SIGMA / gs.min())],
                           "l2_policy": "each step writes S x n fp64 = %.2f GB of fresh output (> 126 MB L2) and streams "
                                        "its %.2f GB tiled coefficient workspace" % (S * n * 8 / 1e9, S * n * 8.25 / 1e9),
                           "parallelism": f"independent sample ranges per GPU x{world}, no data-path collective"},
            "fp64_tflops_per_gpu": klein_rate / world * KLEIN_FLOP * 1e-12,
            "fp64_frac_of_peak": klein_rate / world * KLEIN_FLOP * 1e-12 / peak,
            "roofline": {"bound": "tensor", "kernel": kernel_name,
                         "achieved": pt_tf, "peak": peak, "unit": "TFLOP/s", "frac": pt_tf / peak,
                         "peak_source": peak_src,
                         # dram__bytes_read.sum + dram__bytes_write.sum of this kernel per launch from the newest
                         # committed ncu --set full capture of this command
                         "traffic": traffic if (S == 170496 and n == 1024 and prof["chain"][1] > 0) else None,
                         "traffic_source": traffic_src,
                         "traffic_measured_in_run": False,   # read from the committed ncu capture, not measured by this run
                         "traffic_frac_of_hbm_peak": (traffic / (pt_ms / pt_cnt * 1e-3) / 1e9 / hbm_peak
                                                      if traffic and hbm_peak else None),
                         "algorithmic_flop_per_launch": pt_flop, "avg_launch_ms": pt_ms / pt_cnt,
                         "share_of_step": share},
            "gpu_launches": int(launches_per_step * K),
            "e2e": e2e, "imhk": imhk, "imhk_exact": imhk_exact, "replay": replay, "invariance": invariance,
            "cpu_baseline": cpu, "clocks": clk,
        }
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
