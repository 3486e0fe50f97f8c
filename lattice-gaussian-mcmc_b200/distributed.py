"""Multi-GPU plumbing of the sampling path: one process per GPU, independent units sharded by global id.

The reference's only parallelism is `multiprocessing.Pool` over independent chains, each worker returning its whole
chain to the parent, which then averages acceptance rates / TVD curves (experiments/dimension_scaling.py:841-884,
experiments/convergence_study.py:500-522,543).  Here every rank keeps its samples on its GPU and only the
statistics cross ranks:

  * `shard_range`   contiguous range of global sample / chain ids per rank.  The Philox counter is the GLOBAL id, so
                    the union of the ranks' outputs is the same set of samples for any world size.
  * `allreduce_stats`  one SUM all-reduce of the int64 `[accepted, total]` pair and of the marginal histogram bins
                    (NCCL over NVLink for CUDA tensors, gloo for the CPU tests) -- a few KB.
  * `acceptance_rate`, `marginal_tvd`  what src/diagnostics/mcmc_diag.py:107-117 (`compute_acceptance_rate`) and
                    src/diagnostics/convergence_diag.py:15-72 (`compute_tvd`, multivariate branch = mean of the
                    per-coordinate marginal TVDs over a shared binning) compute, from the reduced integers.

No sample ever crosses a rank boundary; there is no data-path collective.
"""
from __future__ import annotations

from typing import Optional, Tuple

import torch
import torch.distributed as dist


def world() -> Tuple[int, int]:
    """(rank, world_size) of the default process group, (0, 1) when torch.distributed is not initialised."""
    if dist.is_available() and dist.is_initialized():
        return dist.get_rank(), dist.get_world_size()
    return 0, 1


def shard_range(total: int, rank: int, world_size: int) -> Tuple[int, int]:
    """(first global id, count) of `rank`'s contiguous share of `total` independent units; the first
    `total % world_size` ranks take one extra unit."""
    if world_size < 1 or not (0 <= rank < world_size):
        raise ValueError("need 0 <= rank < world_size")
    if total < 0:
        raise ValueError("total must be non-negative")
    base, extra = divmod(int(total), world_size)
    start = rank * base + min(rank, extra)
    return start, base + (1 if rank < extra else 0)


def allreduce_stats(counts: torch.Tensor, hist: Optional[torch.Tensor] = None, group=None):
    """SUM all-reduce of the acceptance counters (`[..., 2]` int64: accepted, total -- per chain or already summed)
    and, optionally, of histogram bins (`[dims, bins]` int64).  Returns (accepted_total [2], hist).  A no-op
    without an initialised process group."""
    tot = counts.reshape(-1, 2).sum(dim=0)
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        if hist is not None:  # one message: counters ride with the bins
            buf = torch.cat([tot, hist.reshape(-1)])
            dist.all_reduce(buf, op=dist.ReduceOp.SUM, group=group)
            tot = buf[:2]
            hist = buf[2:].reshape(hist.shape)
        else:
            dist.all_reduce(tot, op=dist.ReduceOp.SUM, group=group)
    return tot, hist


def acceptance_rate(accepted_total: torch.Tensor) -> float:
    """accepted / total (imhk.py:221-224; 0.0 before any step)."""
    a, t = int(accepted_total[0]), int(accepted_total[1])
    return a / t if t > 0 else 0.0


def marginal_tvd(hist_a: torch.Tensor, hist_b: torch.Tensor) -> float:
    """Mean over tracked coordinates of the marginal total-variation distance 0.5 * sum_bins |p - q| between two
    sets of histograms taken on the same binning (convergence_diag.py:15-72)."""
    if hist_a.shape != hist_b.shape:
        raise ValueError("histograms must share their binning")
    a = hist_a.to(torch.float64)
    b = hist_b.to(torch.float64)
    pa = a / a.sum(dim=-1, keepdim=True).clamp_min(1.0)
    pb = b / b.sum(dim=-1, keepdim=True).clamp_min(1.0)
    return float((0.5 * (pa - pb).abs().sum(dim=-1)).mean())


def host_histogram(values: torch.Tensor, dims, lo: float, width: float, nbins: int) -> torch.Tensor:
    """The binning rule of lgs_histogram (include/lgs_b200.h) restated on the host for checks: counts of
    floor((v - lo) / width) clamped to [0, nbins) for each tracked coordinate.  int64 [len(dims), nbins]."""
    out = torch.zeros((len(dims), nbins), dtype=torch.int64)
    for r, d in enumerate(dims):
        idx = torch.floor((values[:, d].to(torch.float64) - lo) / width).to(torch.int64).clamp_(0, nbins - 1)
        out[r] = torch.bincount(idx, minlength=nbins)
    return out
