"""world_size-2 gloo test of the multi-GPU host logic (lattice-gaussian-mcmc_b200/distributed.py): id-range sharding,
the one all-reduce of acceptance counters + marginal histograms, and the statistics formed from the reduced integers
(src/diagnostics/mcmc_diag.py:107-117, convergence_diag.py:15-72).  Runs on CPU; the sampling itself needs a GPU and
is covered by the `gpu` tests (same Philox key => same samples for any split)."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def _unit_values(ids, n):
    """Deterministic stand-in for 'the sample with global id g': a function of g only, like the Philox-keyed
    sampler, so any partition of the ids produces the same multiset."""
    g = ids.to(torch.float64).unsqueeze(1)
    k = torch.arange(n, dtype=torch.float64).unsqueeze(0)
    return 40.0 * torch.sin(0.37 * g + 1.3 * k) + 3.0 * torch.cos(0.011 * g * (k + 1.0))


def _accept_bits(ids):
    return ((ids * 2654435761) % 97 < 71).to(torch.int64)


def _worker(rank, world_size, port, total, n, out_dir):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world_size)
    from lattice_gaussian_mcmc_b200 import distributed as D
    assert D.world() == (rank, world_size)
    start, count = D.shard_range(total, rank, world_size)
    ids = torch.arange(start, start + count, dtype=torch.int64)
    vals = _unit_values(ids, n)
    dims = [0, 1, n // 2, n - 1]
    hist = D.host_histogram(vals, dims, -48.0, 1.5, 64)
    acc = _accept_bits(ids)
    counts = torch.stack([acc, torch.ones_like(acc)], dim=1)          # per-unit (accepted, total)
    tot, hist = D.allreduce_stats(counts, hist)
    np.savez(os.path.join(out_dir, f"rank{rank}.npz"), tot=tot.numpy(), hist=hist.numpy(), start=start, count=count)
    dist.barrier()
    dist.destroy_process_group()


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


@pytest.mark.parametrize("total", [1000, 1001])
def test_two_rank_allreduce_matches_single_process(tmp_path, total):
    from lattice_gaussian_mcmc_b200 import distributed as D
    n, world_size = 12, 2
    mp.spawn(_worker, args=(world_size, _free_port(), total, n, str(tmp_path)), nprocs=world_size, join=True)
    res = [np.load(tmp_path / f"rank{r}.npz") for r in range(world_size)]
    # the shards tile [0, total) with no gap or overlap
    assert res[0]["start"] == 0 and res[0]["start"] + res[0]["count"] == res[1]["start"]
    assert res[1]["start"] + res[1]["count"] == total
    # every rank holds the same reduced statistics, equal to the single-process ones
    ids = torch.arange(total, dtype=torch.int64)
    ref_hist = D.host_histogram(_unit_values(ids, n), [0, 1, n // 2, n - 1], -48.0, 1.5, 64)
    ref_acc = int(_accept_bits(ids).sum())
    for r in res:
        assert np.array_equal(r["hist"], ref_hist.numpy())
        assert r["tot"].tolist() == [ref_acc, total]
    assert D.acceptance_rate(torch.tensor(res[0]["tot"])) == ref_acc / total
    assert D.marginal_tvd(torch.tensor(res[0]["hist"]), ref_hist) == 0.0


def test_shard_range_properties():
    from lattice_gaussian_mcmc_b200 import distributed as D
    for total in (0, 1, 7, 8, 1 << 20, (1 << 20) + 5):
        for w in (1, 2, 4, 8):
            nxt = 0
            for r in range(w):
                s, c = D.shard_range(total, r, w)
                assert s == nxt and c in (total // w, total // w + 1)
                nxt = s + c
            assert nxt == total
    with pytest.raises(ValueError):
        D.shard_range(10, 2, 2)


def test_marginal_tvd_and_acceptance_formulas():
    from lattice_gaussian_mcmc_b200 import distributed as D
    a = torch.tensor([[10, 0, 0], [5, 5, 0]])
    b = torch.tensor([[0, 10, 0], [5, 5, 0]])
    assert D.marginal_tvd(a, b) == pytest.approx(0.5)          # (1.0 + 0.0) / 2
    assert D.acceptance_rate(torch.tensor([0, 0])) == 0.0
    assert D.acceptance_rate(torch.tensor([3, 4])) == 0.75
    # single process: all-reduce is the identity
    tot, h = D.allreduce_stats(torch.tensor([[1, 2], [3, 4]]), a)
    assert tot.tolist() == [4, 6] and torch.equal(h, a)


def _gr_worker(rank, world_size, port, out_dir):
    """each rank holds half of the chains: per-chain mean / variance sums -> one all-reduce -> R-hat"""
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world_size)
    from lattice_gaussian_mcmc_b200 import distributed as D
    from lattice_gaussian_mcmc_b200.diagnostics import r_hat_from_sums
    g = np.load(os.path.join(ROOT, "tests", "golden", "diag_ar1.npz"))
    X = torch.from_numpy(g["X"])                       # [series kinds = "coordinates", chains, keep]
    A, C, K = X.shape
    start, count = D.shard_range(C, rank, world_size)
    mine = X[:, start:start + count]
    means = mine.mean(dim=2)
    sums = torch.stack([means.sum(dim=1), (means * means).sum(dim=1), mine.var(dim=2, unbiased=True).sum(dim=1)], dim=1)
    buf = torch.cat([sums.reshape(-1), torch.tensor([float(count)], dtype=torch.float64)])
    dist.all_reduce(buf, op=dist.ReduceOp.SUM)
    rhat = r_hat_from_sums(buf[:-1].reshape(A, 3), int(buf[-1]), K)
    np.save(os.path.join(out_dir, f"rhat{rank}.npy"), rhat.numpy())
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_gelman_rubin_matches_reference_golden(tmp_path):
    """convergence_diag.py:176-213 from all-reduced per-rank sums == the reference's value on all chains"""
    mp.spawn(_gr_worker, args=(2, _free_port(), str(tmp_path)), nprocs=2, join=True)
    g = np.load(os.path.join(ROOT, "tests", "golden", "diag_ar1.npz"))
    for r in range(2):
        assert np.max(np.abs(np.load(tmp_path / f"rhat{r}.npy") - g["rhat"])) < 1e-12
