"""Chain diagnostics on device-resident IMHK output: the callers' next step on the sampler output
(src/diagnostics/convergence_diag.py, mcmc_diag.py) without moving the chains to the host.

`ChainDiagnostics(sampler_or_engine).run(V, dims)` takes the kept states as a CUDA tensor `[chains, keep, n]` (what
`lgs_imhk_run` writes) and returns, per chain and tracked coordinate, what the reference computes per series:

  reference function                                   here
  compute_autocorrelation   (convergence_diag.py:75)   `acf`      [chains, dims, acf_len]   (optional)
  integrated_autocorrelation_time  (:115)              `tau_int`  [chains, dims]
  effective_sample_size     (mcmc_diag.py:59)          `ess`      [chains, dims]; multivariate = min over dims (:98-105)
  gelman_rubin_statistic    (convergence_diag.py:176)  `r_hat`    [dims], chains = the reference's list of chains
  np.mean / np.var(ddof=1) of a chain                  `mean`, `var`

Across ranks only `3 * dims` doubles travel (one SUM all-reduce of the chain-mean / chain-variance sums).
"""
from __future__ import annotations

import numpy as np
import torch
import torch.distributed as dist


def r_hat_from_sums(sums: torch.Tensor, chains: int, keep: int) -> torch.Tensor:
    """gelman_rubin_statistic (convergence_diag.py:176-213) from per-coordinate sums over chains
    `[sum mean_c, sum mean_c^2, sum var_c]`: B = n/(m-1) sum (mean_c - mean)^2, W = mean var_c,
    R = sqrt(((n-1)/n W + B/n) / W)."""
    s = sums.to(torch.float64).reshape(-1, 3)
    m, n = float(chains), float(keep)
    overall = s[:, 0] / m
    between = n / (m - 1.0) * (s[:, 1] - m * overall * overall)
    within = s[:, 2] / m
    pooled = (n - 1.0) / n * within + between / n
    return torch.sqrt(pooled / within)


class ChainDiagnostics:
    def __init__(self, sampler_or_engine):
        self.eng = getattr(sampler_or_engine, "_engine", sampler_or_engine)

    def run(self, V: torch.Tensor, dims, c_window: float = 5.0, max_lag: int = 0, acf_len: int = 0, group=None):
        if not (V.is_cuda and V.dtype == torch.float64 and V.dim() == 3 and V.is_contiguous()):
            raise ValueError("V must be a contiguous CUDA float64 tensor [chains, keep, n]")
        chains, keep, n = V.shape
        if n != self.eng.n:
            raise ValueError("last dimension of V must be the lattice dimension")
        dims = np.ascontiguousarray(dims, dtype=np.int32)
        st = torch.cuda.current_stream(V.device).cuda_stream
        stats = torch.empty((chains, dims.size, len(self.eng.DIAG_FIELDS)), dtype=torch.float64, device=V.device)
        acf = torch.empty((chains, dims.size, acf_len), dtype=torch.float64, device=V.device) if acf_len > 0 else None
        self.eng.chain_diagnostics_dev(V.data_ptr(), chains, keep, dims, stats.data_ptr(), c_window, max_lag,
                                       acf.data_ptr() if acf is not None else 0, acf_len, st)
        sums = torch.zeros((dims.size, 3), dtype=torch.float64, device=V.device)
        self.eng.gelman_rubin_sums_dev(stats.data_ptr(), chains, dims.size, sums.data_ptr(), st)
        total_chains = chains
        if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
            buf = torch.cat([sums.reshape(-1), torch.tensor([float(chains)], dtype=torch.float64, device=V.device)])
            dist.all_reduce(buf, op=dist.ReduceOp.SUM, group=group)
            sums = buf[:-1].reshape(dims.size, 3)
            total_chains = int(round(float(buf[-1])))
        out = {k: stats[:, :, i] for i, k in enumerate(self.eng.DIAG_FIELDS)}
        out["ess"] = keep / out["tau_int"]                      # mcmc_diag.py:76-78
        out["ess_min"] = out["ess"].min(dim=1).values           # mcmc_diag.py:98-105
        out["r_hat"] = r_hat_from_sums(sums, total_chains, keep) if total_chains > 1 else None
        out["acf"] = acf
        out["gr_sums"] = sums
        return out


# ---------------------------------------------------------------------------------------------------------
# D1: total-variation distance between two sample sets / acceptance rate, with the reference's semantics
# ---------------------------------------------------------------------------------------------------------
def compute_tvd(samples1, samples2, bins=None) -> float:
    """Host drop-in for compute_tvd (src/diagnostics/convergence_diag.py:15-72) on numpy arrays: exact-value counting
    in 1-D (or a shared-range histogram when `bins` is given), the MEAN of the per-coordinate marginal TVDs for
    [S, n] inputs."""
    a, b = np.asarray(samples1), np.asarray(samples2)
    if a.ndim == 1 and b.ndim == 1:
        if bins is None:
            values, inv = np.unique(np.concatenate([a, b]), return_inverse=True)
            p1 = np.bincount(inv[:len(a)], minlength=len(values)) / len(a)
            p2 = np.bincount(inv[len(a):], minlength=len(values)) / len(b)
            return float(0.5 * np.abs(p1 - p2).sum())
        lo, hi = min(a.min(), b.min()), max(a.max(), b.max())
        h1, _ = np.histogram(a, bins=bins, range=(lo, hi))
        h2, _ = np.histogram(b, bins=bins, range=(lo, hi))
        return float(0.5 * np.abs(h1 / h1.sum() - h2 / h2.sum()).sum())
    return float(np.mean([compute_tvd(a[:, i], b[:, i], bins) for i in range(a.shape[1])]))


def compute_acceptance_rate(accepted) -> float:
    """mcmc_diag.py:107-117: the mean of a boolean array of accept / reject decisions."""
    return float(np.mean(np.asarray(accepted)))


def device_tvd(sampler_or_engine, V1: torch.Tensor, V2: torch.Tensor, dims=None, group=None) -> float:
    """compute_tvd (bins=None) of two device-resident sets of INTEGER-valued lattice points `[S1, n]`, `[S2, n]` (CUDA
    float64, e.g. what lgs_klein_sample writes) without moving them to the host: exact-value counting is a histogram
    with unit bins centred on the integers, accumulated by `lgs_histogram`; with a process group the int64 bins are
    summed over ranks first (every rank holds its share of both sets).  Returns the mean marginal TVD over `dims`
    (default: every coordinate), as the reference does for multivariate input."""
    eng = getattr(sampler_or_engine, "_engine", sampler_or_engine)
    n = V1.shape[1]
    dims = np.arange(n, dtype=np.int32) if dims is None else np.ascontiguousarray(dims, dtype=np.int32)
    ext = torch.stack([torch.minimum(V1.min(), V2.min()), torch.maximum(V1.max(), V2.max())])
    if group is not None or (dist.is_available() and dist.is_initialized()):
        ext[0] = -ext[0]
        dist.all_reduce(ext, op=dist.ReduceOp.MAX, group=group)   # shared binning on every rank
        ext[0] = -ext[0]
    lo, hi = (float(x) for x in ext.tolist())
    nbins = int(round(hi - lo)) + 1
    hist = torch.zeros((2, len(dims), nbins), dtype=torch.int64, device=V1.device)
    st = torch.cuda.current_stream().cuda_stream
    eng.histogram_dev(V1.data_ptr(), V1.shape[0], dims, lo - 0.5, 1.0, nbins, hist[0].data_ptr(), stream=st)
    eng.histogram_dev(V2.data_ptr(), V2.shape[0], dims, lo - 0.5, 1.0, nbins, hist[1].data_ptr(), stream=st)
    if group is not None or (dist.is_available() and dist.is_initialized()):
        dist.all_reduce(hist, op=dist.ReduceOp.SUM, group=group)
    h = hist.to(torch.float64)
    p = h / h.sum(dim=-1, keepdim=True)
    return float((0.5 * (p[0] - p[1]).abs().sum(dim=-1)).mean())
