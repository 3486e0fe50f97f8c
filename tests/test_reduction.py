"""Host basis reduction: lgs_lll_reduce / lgs_bkz_reduce / lgs_gso_norms and the LatticeReduction mirror
of src/lattices/reduction.py.  CPU only (the reducer is host C++ inside liblgs_b200.so).

Pinning: the reference's LLL is Sage/fplll (absent); the Gram-Schmidt profiles it recorded for the n = 16 / 32 / 64
scaling runs are the golden data (tests/golden/klein_lll_recorded.json).  BKZ is checked through properties."""
import json
import os

import numpy as np
import pytest

from oracle import reduce_oracle as ro

HERE = os.path.dirname(__file__)


def _recorded():
    return json.load(open(os.path.join(HERE, "golden", "klein_lll_recorded.json")))["cases"]


def _qary_rows(k, m, q, seed=0):
    A = np.random.RandomState(seed).randint(0, q, (m - k, k))
    B = np.zeros((m, m), np.int64)
    B[:k, :k] = q * np.eye(k, dtype=np.int64)
    B[k:, :k] = A
    B[k:, k:] = np.eye(m - k, dtype=np.int64)
    return B


@pytest.mark.parametrize("n", [16, 32, 64])
@pytest.mark.parametrize("precision", [53, 64, 106])
def test_lll_lands_on_the_reference_recorded_profile_in_every_precision_tier(n, precision):
    """Sage's LLL(delta=0.99) output recorded by the reference (results/Klein_LLL_n=*.json gs_norms, sigma)."""
    from lattice_gaussian_mcmc_b200.lattices import random_int_basis
    from lattice_gaussian_mcmc_b200.reduction import gso_norms, lll
    rec = _recorded()[str(n)]
    M = np.rint(random_int_basis(n)).astype(np.int64)
    red, U, st = lll(M, delta=0.99, eta=0.501, precision=precision)
    gs = gso_norms(red)
    np.testing.assert_allclose(gs, np.array(rec["gs_norms"]), rtol=1e-10)
    assert abs(1.5 * gs.max() - rec["sigma"]) < 1e-9 * rec["sigma"]
    assert st["float_bits"] == precision
    assert np.array_equal(U.astype(object) @ M.astype(object), red.astype(object))     # reduced = U * input, exactly
    # the published algorithm run step by step in Python gives the same basis, row for row
    assert np.array_equal(red.astype(np.float64), ro.lll_textbook(M.astype(np.float64)))


@pytest.mark.parametrize("delta", [0.75, 0.99])
def test_lll_output_is_reduced_in_exact_arithmetic_and_unimodular(delta):
    from lattice_gaussian_mcmc_b200.reduction import lll
    rng = np.random.RandomState(7)
    for n, m in ((6, 6), (12, 12), (10, 14), (20, 20)):
        M = rng.randint(-200, 201, (n, m))
        red, U, _ = lll(M, delta=delta)
        assert ro.is_lll_reduced_exact(red, delta, 0.501)
        assert abs(ro.exact_det(U)) == 1
        assert np.array_equal(U.astype(object) @ M.astype(object), red.astype(object))


def test_lll_large_entries_take_the_int128_inner_products():
    """Entries around 2^40: the 53-bit double inner products no longer apply; the result must still be exact."""
    from lattice_gaussian_mcmc_b200.reduction import lll
    rng = np.random.RandomState(3)
    n = 8
    M = np.eye(n, dtype=np.int64)
    M[:, 0] = rng.randint(1 << 39, 1 << 40, n)     # knapsack-type basis
    M[0, 0] |= 1
    red, U, _ = lll(M, delta=0.99)
    assert ro.is_lll_reduced_exact(red, 0.99, 0.501)
    assert abs(ro.exact_det(U)) == 1
    assert np.abs(red).max() < 1 << 12


def test_qary_128_reduces_in_extended_precision_and_keeps_the_determinant():
    from lattice_gaussian_mcmc_b200.reduction import gso_norms, lll
    M = _qary_rows(64, 128, 3329)
    red, U, st = lll(M, delta=0.99, with_transform=False)
    gs = gso_norms(red)
    assert abs(np.log(gs).sum() - 64 * np.log(3329)) < 1e-8            # det = q^64
    assert st["float_bits"] >= 64 and U is None
    assert np.linalg.norm(red[0].astype(float)) < 1.03 ** 128 * 3329 ** 0.5   # root Hermite factor of an LLL basis
    assert gs.max() < 3329


def test_bkz_blocks_start_with_a_shortest_projected_vector():
    """After BKZ-beta has run to its fixed point every block's first Gram-Schmidt norm is within the 0.99 slack of
    the shortest vector of the projected block (brute force in exact arithmetic)."""
    from lattice_gaussian_mcmc_b200.reduction import bkz
    rng = np.random.RandomState(11)
    n, beta = 10, 4
    M = rng.randint(-30, 31, (n, n))
    red, U, st = bkz(M, beta, delta=0.99)
    assert ro.is_lll_reduced_exact(red, 0.99, 0.501)
    assert abs(ro.exact_det(U)) == 1
    assert np.array_equal(U.astype(object) @ M.astype(object), red.astype(object))
    _, r = ro.exact_gso(red)
    for j in range(n - 1):
        k = min(j + beta - 1, n - 1)
        assert float(ro.shortest_projected_norm2(red, j, k, radius=2)) >= 0.99 * float(r[j]) * (1 - 1e-9)


def test_bkz_with_full_block_finds_the_shortest_vector_and_never_loses_to_lll():
    from lattice_gaussian_mcmc_b200.reduction import bkz, lll
    rng = np.random.RandomState(5)
    n = 7
    M = rng.randint(-40, 41, (n, n))
    red, _, st = bkz(M, n)
    lam2 = float(ro.shortest_projected_norm2(red, 0, n - 1, radius=2))
    assert 0.99 * float(red[0] @ red[0]) <= lam2 <= float(red[0] @ red[0])
    # a harder instance: BKZ-20 on the n = 64 scaling-run lattice improves on LLL
    from lattice_gaussian_mcmc_b200.lattices import random_int_basis
    M64 = np.rint(random_int_basis(64)).astype(np.int64)
    l, _, _ = lll(M64, with_transform=False)
    b, _, st = bkz(M64, 20, with_transform=False)
    assert st["bkz_insertions"] > 0 and st["enumeration_nodes"] > 0
    assert np.linalg.norm(b[0].astype(float)) <= np.linalg.norm(l[0].astype(float))
    from lattice_gaussian_mcmc_b200.reduction import gso_norms
    assert gso_norms(b).max() <= gso_norms(l).max() * (1 + 1e-12)
    assert abs(np.log(gso_norms(b)).sum() - np.log(gso_norms(l)).sum()) < 1e-9


@pytest.mark.parametrize("n,seed", [(n, s) for n in (8, 12, 16, 20, 24) for s in range(4)])
def test_full_block_bkz_ends_on_a_shortest_vector_of_an_independent_search(n, seed):
    """BKZ parity is unpinned (no fplll on this image, no BKZ basis recorded by the reference): the one external check
    that exists -- lgs_bkz_reduce with beta = n must return a first vector as short as the shortest vector found by the
    oracle's own enumeration (oracle/reduce_oracle.py shortest_vector_norm2), up to the delta = 0.99 slack of its
    insertion rule -- on 20 random lattices of dimension 8 to 24 (src/lattices/reduction.py:238-300 is the call it stands in for)."""
    from lattice_gaussian_mcmc_b200.reduction import bkz
    rng = np.random.RandomState(1000 * n + seed)
    M = rng.randint(-60, 61, (n, n))
    while abs(np.linalg.det(M.astype(float))) < 0.5:
        M = rng.randint(-60, 61, (n, n))
    red, U, _ = bkz(M, n, delta=0.99)
    assert np.array_equal(U.astype(object) @ M.astype(object), red.astype(object))
    lam2 = ro.shortest_vector_norm2(M)
    b1 = int(sum(int(v) * int(v) for v in red[0]))
    assert lam2 <= b1 and 0.99 * b1 <= lam2, (lam2, b1)


def test_pruned_enumeration_still_returns_a_reduced_basis_of_the_same_lattice():
    from lattice_gaussian_mcmc_b200.lattices import random_int_basis
    from lattice_gaussian_mcmc_b200.reduction import bkz
    M = np.rint(random_int_basis(32)).astype(np.int64)
    red, U, _ = bkz(M, 24, prune=1, max_tours=3)
    assert np.array_equal(U.astype(object) @ M.astype(object), red.astype(object))
    assert ro.is_lll_reduced_exact(red[:12], 0.99, 0.501)    # exact check on a prefix (a prefix of a reduced basis is reduced)


def test_error_codes():
    from lattice_gaussian_mcmc_b200 import _capi
    from lattice_gaussian_mcmc_b200.reduction import ReductionError, bkz, lll
    L = _capi.lib()
    dep = np.array([[1, 2, 3], [2, 4, 6], [0, 1, 0]], dtype=np.int64)
    with pytest.raises(ReductionError, match="-1"):
        lll(dep)
    with pytest.raises(ReductionError, match="-1"):
        lll(np.eye(3), delta=0.2)
    with pytest.raises(ReductionError, match="-1"):
        lll(np.ones((3, 2)))                       # m < n
    with pytest.raises(ReductionError, match="-6"):
        lll(np.eye(2) * float(1 << 60))
    with pytest.raises(ReductionError, match="-1"):
        bkz(np.eye(3), 1)
    with pytest.raises(ValueError):
        lll(np.eye(2) * 0.5)
    assert L.lgs_lll_reduce(None, 3, 3, 0.99, 0.501, 0, None, None) == _capi.LGS_E_INVALID
    assert L.lgs_gso_norms(dep.ctypes.data, 3, 3, np.zeros(3).ctypes.data) == _capi.LGS_E_INVALID


def test_lattice_reduction_class_mirrors_the_reference_surface():
    """Return shapes, diagnostics keys and formulas of src/lattices/reduction.py:68-133, 238-405, 581-764."""
    from lattice_gaussian_mcmc_b200.reduction import (LatticeReduction, optimal_delta_for_sampling,
                                                      reduction_strategy_for_lattice)
    rng = np.random.RandomState(1)
    M = rng.randint(0, 51, (12, 12)).astype(np.float64)
    lr = LatticeReduction()
    red, U, diag = lr.lll_reduce(M, delta=0.75)
    assert red.dtype == np.float64 and U.dtype == np.float64
    np.testing.assert_array_equal(U @ M, red)
    assert set(diag) >= {"time", "initial_quality", "final_quality", "improvement_factor", "delta"}
    prof = diag["final_quality"]
    assert set(prof) == {"dimension", "hermite_factor", "orthogonality_defect", "gs_norms", "max_gs_norm",
                         "min_gs_norm", "gs_ratio", "log_potential"}
    # hermite factor = ||b_1|| / det^(1/n); orthogonality defect = prod ||b_i|| / det (reduction.py:322-371)
    det = abs(np.linalg.det(red))
    assert abs(lr.hermite_factor(red) - np.linalg.norm(red[0]) / det ** (1 / 12)) < 1e-9
    assert abs(lr.orthogonality_defect(red) / (np.prod(np.linalg.norm(red, axis=1)) / det) - 1) < 1e-9
    assert diag["improvement_factor"] >= 1.0
    assert lr.stats["reductions_performed"] == 1

    red2, U2, d2 = lr.bkz_reduce(M, block_size=8)
    np.testing.assert_array_equal(U2 @ M, red2)
    assert d2["block_sizes_used"] == [8] and d2["strategy"] == "default"
    assert lr._progressive_block_sizes(100, 45) == [20, 30, 40, 45]
    _, _, d3 = lr.bkz_reduce(M, block_size=12, progressive=True)
    assert d3["block_sizes_used"] == [12]

    kept, removed = lr.lll_with_removals(M, target_dim=9)
    assert kept.shape == (12 - len(removed), 12) and len(removed) <= 3

    best, info = lr.sampling_reduce(M, target_sigma=200.0, max_iterations=2, bkz_blocks=(8,))
    assert info["suitable_for_sigma"] and info["final_max_norm"] == max(info["final_profile"]["gs_norms"])
    cmp_ = lr.compare_bases(M, red)
    assert cmp_["max_gs_norm"]["improvement"] >= 1.0
    cost = lr.estimate_reduction_cost(M, 20)
    assert abs(cost["log2_operations"] - np.log2((12 / 20) * 2 ** (0.292 * 20))) < 1e-12
    assert optimal_delta_for_sampling(64, 1.0) == 0.95 and optimal_delta_for_sampling(300, 1.0) == 0.75
    assert reduction_strategy_for_lattice("qary", 128)["bkz_block"] == 50
    assert reduction_strategy_for_lattice("nonsense", 64) == reduction_strategy_for_lattice("random", 64)
