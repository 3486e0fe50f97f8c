"""The persistent chain kernel (lgs_chain.cu) against the ordered path, and against the multi-kernel blocked path it
replaced (LGS_CHAIN=0), on the cases its own machinery adds: the FP32 SampleZ screen and its FP64 fallback over the
sigma_i regimes, look-ahead over many coordinate pairs, partial tiles, IMHK launches split at kept steps and at the
64-step cap, the flush-only state read-out.

Same Philox key per (sample, draw) and the same map u -> x everywhere, so integer coefficients must be equal sample
for sample (see test_gpu_blocked.py for why a mismatch has probability ~1e-12 per draw).
"""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def lgs():
    import __graft_entry__ as g
    g.build()
    import lattice_gaussian_mcmc_b200 as m
    return m


def _with_chain(flag, make):
    """the choice between the chain kernel and the v1 kernels is read when a handle is created"""
    old = os.environ.get("LGS_CHAIN")
    os.environ["LGS_CHAIN"] = str(flag)
    try:
        return make()
    finally:
        if old is None:
            del os.environ["LGS_CHAIN"]
        else:
            os.environ["LGS_CHAIN"] = old


@pytest.mark.parametrize("n,S", [(64, 129), (200, 1000), (520, 300)])
def test_chain_equals_v1_blocked_path(lgs, n, S):
    from lattice_gaussian_mcmc_b200.lattices import dd_lattice
    lat = dd_lattice(n, seed=200 + n)
    sigma = 1.5 * lat.max_gram_schmidt_norm
    center = np.random.RandomState(n).uniform(-40, 40, n)
    a = _with_chain(1, lambda: lgs.KleinSampler(lat, sigma, center, seed=11, path="blocked", samplez="fast"))
    b = _with_chain(0, lambda: lgs.KleinSampler(lat, sigma, center, seed=11, path="blocked", samplez="fast"))
    assert a._engine.preferred_batch % 128 == 0 and a._engine.preferred_batch != b._engine.preferred_batch
    Va, Xa, la = a.sample_with_coefficients(S)
    Vb, Xb, lb = b.sample_with_coefficients(S)
    assert int(np.sum(np.any(Xa != Xb, axis=1))) == 0
    assert np.array_equal(Va, Vb)
    assert np.max(np.abs(la - lb) / np.abs(la)) < 1e-10
    # lattice points only (the kernel variant without the log-density accumulation) give the same points
    assert np.array_equal(a._engine.klein_sample(S, 11, 0, "fast")[0], Va)


@pytest.mark.parametrize("sigma", [0.8, 1.2, 5.0, 14.0, 20.0, 60.0])
def test_chain_sigma_regimes_equal_ordered(lgs, sigma):
    """sigma_i < 1 (summed normaliser), 1 <= sigma_i < 1.45 (theta-series term), the FP32-screened regime, beyond its
    sigma_i <= 16 eligibility (FP64 smooth), and the truncated 1001-entry window (sigma_i > 50)"""
    n = 72
    rng = np.random.RandomState(5)
    B = np.triu(rng.uniform(-0.4, 0.4, (n, n)), 1) + np.eye(n)
    lat = lgs.SimpleLattice(B)
    center = rng.uniform(-3, 3, n)
    a = lgs.KleinSampler(lat, sigma, center, seed=4, path="ordered", samplez="fast")
    b = lgs.KleinSampler(lat, sigma, center, seed=4, path="blocked", samplez="fast")
    S = 3000
    Va, Xa, la = a.sample_with_coefficients(S)
    Vb, Xb, lb = b.sample_with_coefficients(S)
    assert int(np.sum(np.any(Xa != Xb, axis=1))) == 0
    ref = Xb.astype(np.float64) @ B.T
    assert np.max(np.abs(Vb - ref)) <= 1e-12 * np.max(np.abs(ref))
    assert np.max(np.abs(la - lb) / np.abs(la)) < 1e-10


def test_chain_half_integer_centers_take_the_fp64_path(lgs):
    """means on exact .5 ties (Z^n, c = 1/2): rint's half-to-even decides c0, in FP32 and in FP64 alike"""
    from lattice_gaussian_mcmc_b200.lattices import identity_lattice
    lat = identity_lattice(66)
    center = np.full(66, 0.5)
    a = lgs.KleinSampler(lat, 3.0, center, seed=9, path="ordered", samplez="fast")
    b = lgs.KleinSampler(lat, 3.0, center, seed=9, path="blocked", samplez="fast")
    Xa = a.sample_with_coefficients(4000)[1]
    Xb = b.sample_with_coefficients(4000)[1]
    assert np.array_equal(Xa, Xb)
    assert abs(Xb.mean() - 0.5) < 0.05


@pytest.mark.parametrize("weight_mode", ["reference", "exact"])
def test_chain_imhk_long_burn_in_split_over_launches(lgs, weight_mode):
    """burn-in longer than the 64-step launch cap, kept states every 3 steps: launches end at every kept step and the
    chains continue across them exactly like the per-thread ordered chains"""
    from lattice_gaussian_mcmc_b200.lattices import dd_lattice
    n, chains = 40, 200
    lat = dd_lattice(n, seed=77)
    sigma = (0.5 if weight_mode == "exact" else 1.5) * lat.max_gram_schmidt_norm
    center = np.random.RandomState(3).uniform(-10, 10, n)
    kw = dict(burn_in=0, n_chains=chains, seed=23, weight_mode=weight_mode, samplez="fast")
    a = lgs.IMHKSampler(lat, sigma, center, path="ordered", **kw)
    b = lgs.IMHKSampler(lat, sigma, center, path="blocked", **kw)
    Va = a.sample(n_samples=3, burn_in=150, thin=3)
    Vb = b.sample(n_samples=3, burn_in=150, thin=3)
    assert np.array_equal(Va, Vb)
    assert a.total_proposals == b.total_proposals == chains * 159
    assert a.accepted_proposals == b.accepted_proposals
    if weight_mode == "exact":
        assert 0.05 < b.acceptance_rate < 0.9995, b.acceptance_rate   # rejections exercise the deferred commit
    assert np.array_equal(a.current_coeffs, b.current_coeffs)          # flush-only launch
    assert np.array_equal(a.per_chain_acceptance, b.per_chain_acceptance)
    # a state read-out does not disturb the chains
    Va2 = a.sample(n_samples=1, thin=1)
    Vb2 = b.sample(n_samples=1, thin=1)
    assert np.array_equal(Va2, Vb2)


def test_chain_imhk_equals_v1_at_full_dimension(lgs):
    from lattice_gaussian_mcmc_b200.lattices import dd_lattice
    lat = dd_lattice(1024, seed=2)
    sigma = 1.5 * lat.max_gram_schmidt_norm
    kw = dict(burn_in=0, n_chains=300, seed=5, path="blocked")
    a = _with_chain(1, lambda: lgs.IMHKSampler(lat, sigma, **kw))
    b = _with_chain(0, lambda: lgs.IMHKSampler(lat, sigma, **kw))
    Va = a.sample(n_samples=2, burn_in=2, thin=2)
    Vb = b.sample(n_samples=2, burn_in=2, thin=2)
    assert np.array_equal(Va, Vb)
    assert np.array_equal(a.current_coeffs, b.current_coeffs)
    assert a.total_proposals == b.total_proposals == 300 * 6


def test_imhk_checkpoint_resume_reproduces_the_uninterrupted_run(lgs):
    """get_state -> destroy -> create -> set_state -> continue == one uninterrupted run (imhk.py:59-61: the reference's
    chain state is three assignable attributes; with counter-based Philox the resume point is (x, log w, counts, step))"""
    from lattice_gaussian_mcmc_b200.lattices import dd_lattice
    lat = dd_lattice(96, seed=3)
    for weight, sigma in (("reference", 1.5 * lat.max_gram_schmidt_norm), ("exact", 0.45 * lat.max_gram_schmidt_norm)):
        chains = 300
        whole = lgs.IMHKSampler(lat, sigma, burn_in=0, n_chains=chains, seed=9, weight_mode=weight, first_chain_id=1000)
        Xw = whole.sample_coefficients(6, thin=3, burn_in=4)                  # 22 steps
        first = lgs.IMHKSampler(lat, sigma, burn_in=0, n_chains=chains, seed=9, weight_mode=weight, first_chain_id=1000)
        Xa = first.sample_coefficients(2, thin=3, burn_in=4)                  # 10 steps
        state = first.get_state()
        assert state["steps_done"] == 10
        del first
        second = lgs.IMHKSampler(lat, sigma, burn_in=0, n_chains=chains, seed=9, weight_mode=weight, first_chain_id=1000)
        second.set_state(state)
        assert np.array_equal(second.current_coeffs, state["coeffs"])
        Xb = second.sample_coefficients(4, thin=3)                           # 12 more
        assert np.array_equal(np.concatenate([Xa, Xb], axis=1), Xw)
        assert second.total_proposals == whole.total_proposals == chains * 22
        assert second.accepted_proposals == whole.accepted_proposals
        if weight == "exact":
            assert 0.05 < whole.acceptance_rate < 0.999
        assert np.allclose(second.current_log_weight, whole.current_log_weight, rtol=0, atol=0)
    # the ordered kernels take a state too (table SampleZ, native RNG)
    lat = dd_lattice(12, seed=1)
    kw = dict(burn_in=0, n_chains=40, seed=4, samplez="table", path="ordered")
    whole = lgs.IMHKSampler(lat, 30.0, **kw)
    Xw = whole.sample_coefficients(5, thin=2)
    first = lgs.IMHKSampler(lat, 30.0, **kw)
    Xa = first.sample_coefficients(2, thin=2)
    st = first.get_state()
    second = lgs.IMHKSampler(lat, 30.0, **kw)
    second.set_state(st)
    Xb = second.sample_coefficients(3, thin=2)
    assert np.array_equal(np.concatenate([Xa, Xb], axis=1), Xw)


def test_adaptive_precision_sample_uses_the_requested_window(lgs):
    """klein.py:273-302: a target further than 5 sigma halves the window (precision 10 -> 5): draws beyond 5 sigma_i
    can no longer occur; without a target the constructor's window is used"""
    lat = lgs.SimpleLattice(np.eye(1))
    smp = lgs.KleinSampler(lat, 1.0, seed=3, samplez="table", path="ordered")
    assert smp.adaptive_precision_sample().shape == (1,)
    x = np.array([smp.adaptive_precision_sample(target_distance=100.0)[0] for _ in range(50)])
    assert np.all(np.abs(x) <= 5)
    assert 5 in smp._alt_engines and smp._engine is not smp._alt_engines[5]
    y = smp.adaptive_precision_sample(target_distance=0.5)
    assert 20 in smp._alt_engines and y.shape == (1,)


@pytest.mark.parametrize("weight", ["reference", "exact"])
def test_imhk_step_parallel_equals_step_by_step(lgs, weight):
    """Few chains: the proposals of many steps are drawn as one batch and the accept / reject recursion runs afterwards
    (lgs_chain.cu, ChainArgs::tiles_per_step).  Same Philox keys, same expressions: kept states, lattice points,
    counters and log-weights equal the step-by-step launches bit for bit -- also across calls, across the 32-kept-
    states-per-launch cap, and when the two modes alternate on one chain set."""
    import os
    from lattice_gaussian_mcmc_b200.lattices import dd_lattice
    lat = dd_lattice(96, seed=3)
    sigma = (1.5 if weight == "reference" else 0.45) * lat.max_gram_schmidt_norm
    chains = 333

    def run(mode, calls):
        os.environ["LGS_IMHK_STEPPAR"] = mode
        try:
            im = lgs.IMHKSampler(lat, sigma, burn_in=0, n_chains=chains, seed=21, weight_mode=weight, first_chain_id=7)
            outs = []
            for (burn, keep, thin, what) in calls:
                if isinstance(what, str) and what in "01":
                    os.environ["LGS_IMHK_STEPPAR"] = what
                    what = "x"
                if keep == 0:
                    im._run(burn, 0, 1, want_v=False)
                elif what == "v":
                    outs.append(im.sample(n_samples=keep, burn_in=burn, thin=thin))
                else:
                    outs.append(im.sample_coefficients(keep, thin=thin, burn_in=burn))
            return outs, im.total_proposals, im.accepted_proposals, im.current_log_weight, im.current_coeffs
        finally:
            os.environ.pop("LGS_IMHK_STEPPAR", None)

    calls = [(3, 2, 2, "x"), (0, 37, 3, "v"), (5, 0, 1, "x"), (0, 1, 1, "x"), (7, 3, 5, "v")]
    seq = run("0", calls)
    par = run("1", calls)
    for a, b in zip(seq[0], par[0]):
        assert a.shape == b.shape and np.array_equal(a, b)
    assert seq[1] == par[1] == chains * (3 + 4 + 111 + 5 + 1 + 7 + 15)
    assert seq[2] == par[2]
    assert np.array_equal(seq[3], par[3]) and np.array_equal(seq[4], par[4])
    if weight == "exact":
        assert 0.05 < seq[2] / seq[1] < 0.999
    # alternate the modes on one chain set
    mixed = run("0", [(3, 2, 2, "x"), (0, 37, 3, "1"), (5, 0, 1, "0"), (0, 1, 1, "1"), (7, 3, 5, "0")])
    for a, b in zip(seq[0], mixed[0]):
        if a.shape[-1] == 96 and a.dtype == np.float64:     # lattice points of the reference run
            a = np.rint(np.linalg.solve(lat.basis, a.reshape(-1, 96).T).T).astype(np.int64).reshape(a.shape)
        assert np.array_equal(a, b)
    assert np.array_equal(seq[4], mixed[4]) and seq[2] == mixed[2]


def test_imhk_config_c2_shape_runs_step_parallel(lgs):
    """BASELINE config 2's shape (4096 chains, n = 64, burn-in + thinning, kept states) picks the step-parallel path by
    itself and equals the step-by-step chains"""
    import os
    lat = lgs.lattices.lll_lattice(64)
    sigma = 1.5 * lat.max_gram_schmidt_norm
    a = lgs.IMHKSampler(lat, sigma, burn_in=0, n_chains=4096, seed=64).sample_coefficients(5, thin=10, burn_in=100)
    os.environ["LGS_IMHK_STEPPAR"] = "0"
    try:
        b = lgs.IMHKSampler(lat, sigma, burn_in=0, n_chains=4096, seed=64).sample_coefficients(5, thin=10, burn_in=100)
    finally:
        os.environ.pop("LGS_IMHK_STEPPAR", None)
    assert np.array_equal(a, b)
