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


def test_imhk_step_index_cannot_leave_the_32_bit_philox_tag(lgs):
    """the step index is the Philox tag of a step's draws (tag 0 = initial draw): a resume point or a run that would
    take a chain beyond 2^32 - 2 steps is refused instead of wrapping onto streams already used"""
    from lattice_gaussian_mcmc_b200 import _capi
    from lattice_gaussian_mcmc_b200.lattices import dd_lattice
    lat = dd_lattice(96, seed=3)
    top = 2 ** 32 - 2
    for kw in (dict(), dict(samplez="table", path="ordered")):
        im = lgs.IMHKSampler(lat, 1.5 * lat.max_gram_schmidt_norm, burn_in=0, n_chains=8, seed=2, **kw)
        im.sample(1)
        state = im.get_state()
        assert state["steps_done"] == 1
        with pytest.raises(_capi.LgsError) as ei:
            im.set_state(dict(state, steps_done=top + 1))
        assert ei.value.code == _capi.LGS_E_INVALID and "2^32" in str(ei.value)
        im.set_state(dict(state, steps_done=top - 3))
        assert im._engine.imhk_steps_done() == top - 3
        im.sample(1, thin=2)                                   # two steps fit ...
        assert im._engine.imhk_steps_done() == top - 1
        with pytest.raises(_capi.LgsError) as ei:
            im.sample(1, thin=2)                               # ... two more do not
        assert ei.value.code == _capi.LGS_E_INVALID and "2^32" in str(ei.value)
        with pytest.raises(_capi.LgsError):                    # nor does a product that overflows 64 bits
            im._engine.imhk_run_dev(0, 2 ** 40, 2 ** 40, im.proposal_sampler.samplez, "reference")
        im.sample(1)                                           # the last step that fits
        assert im._engine.imhk_steps_done() == top


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


class _env:
    """environment switches the chain kernel reads per launch (development A/B knobs)"""

    def __init__(self, **kv):
        self.kv = {k: str(v) for k, v in kv.items()}

    def __enter__(self):
        self.old = {k: os.environ.get(k) for k in self.kv}
        os.environ.update(self.kv)

    def __exit__(self, *a):
        for k, v in self.old.items():
            if v is None:
                os.environ.pop(k, None)
            else:
                os.environ[k] = v


@pytest.mark.parametrize("n", [40, 832])
def test_scheduling_choices_do_not_change_a_single_coefficient(lgs, n):
    """Tiles per panel warpgroup (1 / 2 / 3), staggered start, look-ahead (none / every pair / on a dry ring) and the
    global lattice-point queue only decide WHEN a tile's pairs are drawn and contracted; the Philox key of every draw is
    (sample id, draw, step), so coefficients and lattice points must be identical under all of them (log-densities too,
    to rounding where the summation order of the means changes) --
    Klein batches with and without lattice points, a full batch (every slot of every SM gets a tile, which the staggered
    start requires) plus a ragged tile.  n = 832 has 13 pairs: above the 12 the staggered start needs."""
    import torch
    from lattice_gaussian_mcmc_b200.lattices import dd_lattice
    lat = dd_lattice(n, seed=7 + n)
    sigma = 1.5 * lat.max_gram_schmidt_norm
    smp = lgs.KleinSampler(lat, sigma, seed=21, path="blocked")
    eng = smp._engine
    S = 148 * 128 * 9 + 77
    st = torch.cuda.current_stream().cuda_stream
    V = torch.empty((S, n), dtype=torch.float64, device="cuda")
    X = torch.empty((S, n), dtype=torch.int64, device="cuda")
    X2 = torch.empty((S, n), dtype=torch.int64, device="cuda")
    lq = torch.empty(S, dtype=torch.float64, device="cuda")
    ref = None
    for tpw in (1, 2, 3):
        for stagger, la, gq in ((0, 0, 1), (100, 0, 1), (100, 1, 0), (50, 2, 1)):
            with _env(LGS_CHAIN_TPW=tpw, LGS_CHAIN_STAGGER=stagger, LGS_CHAIN_LOOKAHEAD=la, LGS_CHAIN_GQ=gq):
                eng.klein_sample_dev(S, 21, 0, "fast", V=V.data_ptr(), X=X.data_ptr(), logq=lq.data_ptr(), stream=st)
                eng.klein_sample_dev(S, 21, 0, "fast", X=X2.data_ptr(), stream=st)            # proposal-only launch
                torch.cuda.synchronize()
            if ref is None:
                ref = (V.clone(), X.clone(), lq.clone())
                B = torch.from_numpy(np.ascontiguousarray(lat.basis.T)).cuda()
                assert torch.equal(V[:4096], X[:4096].double() @ B)
                assert float(X.double().std()) > 0.5
            assert torch.equal(X, ref[1]), (tpw, stagger, la, gq)
            assert torch.equal(X2, ref[1]), (tpw, stagger, la, gq)
            assert torch.equal(V, ref[0]), (tpw, stagger, la, gq)
            # the split centers items of the look-ahead modes sum the conditional means in another order: the integers
            # are the same, the log-density (a continuous function of the means) to rounding
            if la == 0:
                assert torch.equal(lq, ref[2]), (tpw, stagger, la, gq)
            else:
                assert torch.allclose(lq, ref[2], rtol=1e-12, atol=0.0), (tpw, stagger, la, gq)


def test_imhk_scheduling_and_kept_states_in_one_launch_equal_one_launch_per_kept_state(lgs):
    """IMHK with exact weights (real rejections) at n = 832, every slot of every SM holding a tile: the kept states and
    the accept counters of a call for several kept states are the same whether the kept states are snapshots inside
    launches of 4 * thin steps (ChainArgs::keep_thin) or each ends a launch of its own, and under every slot geometry /
    staggered start."""
    import torch
    from lattice_gaussian_mcmc_b200.lattices import dd_lattice
    n = 832
    lat = dd_lattice(n, seed=5)
    sigma = 0.42 * lat.max_gram_schmidt_norm              # sigma_i >= 0.42, mostly < 1: the window masses move with the centers
    smp = lgs.KleinSampler(lat, sigma, seed=3, path="blocked")
    eng = smp._engine
    C, keep, thin = 148 * 128 * 9 + 50, 6, 3
    st = torch.cuda.current_stream().cuda_stream
    V = torch.empty((C, keep, n), dtype=torch.float64, device="cuda")
    X = torch.empty((C, keep, n), dtype=torch.int64, device="cuda")
    ref = None
    for kin, tpw, stagger in ((0, 0, -1), (1, 0, -1), (1, 1, 0), (1, 2, 100), (1, 3, 100), (1, 3, 0)):
        env = dict(LGS_IMHK_KEEP_IN_LAUNCH=kin, LGS_CHAIN_TPW=tpw, LGS_IMHK_STEPPAR=0)
        if stagger >= 0:
            env["LGS_CHAIN_STAGGER"] = stagger
        with _env(**env):
            eng.imhk_reset(C, seed=9)
            eng.imhk_run_dev(4, keep, thin, "fast", "exact", V=V.data_ptr(), X=X.data_ptr(), stream=st)
            torch.cuda.synchronize()
            acc, tot = eng.imhk_counts()
        if ref is None:
            ref = (V.clone(), X.clone(), acc, tot)
            assert tot == C * (4 + keep * thin) and 0.2 * tot < acc < 0.995 * tot, (acc, tot)    # it does reject
            B = torch.from_numpy(np.ascontiguousarray(lat.basis.T)).cuda()
            assert torch.equal(V[:512].reshape(-1, n), X[:512].reshape(-1, n).double() @ B)
        assert torch.equal(X, ref[1]), (kin, tpw, stagger)
        assert torch.equal(V, ref[0]), (kin, tpw, stagger)
        assert (acc, tot) == ref[2:], (kin, tpw, stagger)
