"""What the reference's own scripts do to a sampler, done to the drop-in (boundary items):

  * the worker pattern of experiments/dimension_scaling.py:877-884 and convergence_study.py:559-562 --
    `np.random.seed(seed + 1000 * i); sampler = Sampler(lattice, sigma); sampler.sample(n)` -- must give different
    chains for different i and the same chain for the same seed: with `seed=None` (the default) the Philox key is drawn
    from numpy's global legacy generator at construction;
  * `sampler.current_state = v` followed by `step()` (convergence_study.py:563-571, parameter_sensitivity.py:470,539);
  * `first_sample_id=`: ranks sharing a seed take disjoint Philox ranges through the public API.
"""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def lgs():
    import __graft_entry__ as g
    g.build()
    import lattice_gaussian_mcmc_b200 as m
    return m


@pytest.fixture(scope="module")
def lat():
    from lattice_gaussian_mcmc_b200.lattices import dd_lattice
    return dd_lattice(32, seed=32)


@pytest.mark.parametrize("kind", ["klein", "imhk"])
def test_reference_worker_pattern_gives_distinct_reproducible_chains(lgs, lat, kind):
    sigma = 1.5 * lat.max_gram_schmidt_norm
    seed = 1234

    def worker(i):
        np.random.seed(seed + 1000 * i)                       # dimension_scaling.py:877-884, verbatim shape
        if kind == "klein":
            sampler = lgs.KleinSampler(lat, sigma)
        else:
            sampler = lgs.IMHKSampler(lat, sigma, burn_in=3)
        return sampler.sample(40)

    a0, a1, a0_again = worker(0), worker(1), worker(0)
    assert a0.shape == (40, 32)
    assert not np.array_equal(a0, a1)
    assert np.array_equal(a0, a0_again)
    # two samplers built one after the other without reseeding differ as well
    np.random.seed(7)
    mk = (lambda: lgs.KleinSampler(lat, sigma)) if kind == "klein" else (lambda: lgs.IMHKSampler(lat, sigma, burn_in=0))
    s1, s2 = mk(), mk()
    assert s1.seed != s2.seed and not np.array_equal(s1.sample(8), s2.sample(8))
    # an explicit seed is used as is and does not touch the global stream
    st = np.random.get_state()[1].copy()
    e1 = lgs.KleinSampler(lat, sigma, seed=5).sample(8)
    e2 = lgs.KleinSampler(lat, sigma, seed=5).sample(8)
    assert np.array_equal(e1, e2) and np.array_equal(np.random.get_state()[1], st)


def test_replay_mode_construction_leaves_the_global_stream_alone(lgs, lat):
    """replay mode consumes np.random draw by draw like the reference: building the sampler must not draw a key"""
    np.random.seed(99)
    want = np.random.random_sample(3)
    np.random.seed(99)
    lgs.KleinSampler(lat, 1.5 * lat.max_gram_schmidt_norm, mode="replay")
    lgs.IMHKSampler(lat, 1.5 * lat.max_gram_schmidt_norm, burn_in=0, mode="replay")
    assert np.array_equal(np.random.random_sample(3), want)


def test_first_sample_id_selects_a_range_of_the_same_stream(lgs, lat):
    sigma = 1.5 * lat.max_gram_schmidt_norm
    whole = lgs.KleinSampler(lat, sigma, seed=3).sample(300)
    part = lgs.KleinSampler(lat, sigma, seed=3, first_sample_id=100)
    assert part.next_sample_id == 100
    assert np.array_equal(part.sample(50), whole[100:150])
    assert part.next_sample_id == 150
    part.next_sample_id = 250
    assert np.array_equal(part.sample(50), whole[250:300])
    with pytest.raises(ValueError):
        part.next_sample_id = -1


@pytest.mark.parametrize("weight_mode", ["reference", "exact"])
def test_assigning_current_state_then_stepping(lgs, lat, weight_mode):
    """convergence_study.py:559-571: seed, construct, `current_state = randn * sigma * 5`, then step() in a loop"""
    n = 32
    sigma = (1.5 if weight_mode == "reference" else 0.5) * lat.max_gram_schmidt_norm
    np.random.seed(42)
    sampler = lgs.IMHKSampler(lat, sigma, weight_mode=weight_mode)
    assert sampler.current_state is None
    v = np.random.randn(n) * sigma * 5
    sampler.current_state = v
    x0 = np.linalg.solve(lat.basis, v).round().astype(np.int64)               # imhk.py:130-133
    assert np.array_equal(sampler.current_coeffs, x0)
    assert np.array_equal(sampler.current_state, lat.basis @ x0.astype(np.float64))
    assert sampler.current_log_weight == -np.inf                              # off the lattice: zero mass
    assert sampler.total_proposals == 0
    states, accepts = [], []
    for i in range(12):
        state, accepted = sampler.step()
        states.append(sampler.current_state.copy())
        accepts.append(accepted)
        assert np.array_equal(state, states[-1])
    assert accepts[0] is True                                                  # imhk.py:158-160: weight -inf always moves
    assert sampler.total_proposals == 12 and sampler.accepted_proposals == sum(accepts)
    assert np.isfinite(sampler.current_log_weight)
    X = np.linalg.solve(lat.basis, np.array(states).T).T
    assert np.max(np.abs(X - X.round())) < 1e-6                                # every state is a lattice point
    for k in range(1, 12):                                                     # a rejected step keeps the state
        assert accepts[k] or np.array_equal(states[k], states[k - 1])


def test_assigning_a_lattice_point_installs_the_reference_weight(lgs, lat):
    """parameter_sensitivity.py:470: `sampler.current_state = center.copy()` -- a lattice point (the origin): its
    weight is the reference's importance weight (imhk.py:102-124), counters and Philox position are kept"""
    n = 32
    sigma = 1.5 * lat.max_gram_schmidt_norm
    s = lgs.IMHKSampler(lat, sigma, burn_in=0, seed=8)
    s.sample(5)
    steps, acc = s.total_proposals, s.accepted_proposals
    s.current_state = np.zeros(n)
    assert np.array_equal(s.current_coeffs, np.zeros(n, dtype=np.int64))
    assert s.current_log_weight == pytest.approx(s._compute_importance_weight(np.zeros(n))[1], rel=1e-12, abs=1e-12)
    assert (s.total_proposals, s.accepted_proposals) == (steps, acc)
    assert s._engine.imhk_steps_done() == 5
    x = np.arange(n) % 3 - 1
    s.current_coeffs = x
    assert np.array_equal(s.current_state, lat.basis @ x.astype(np.float64))
    s.current_log_weight = -3.5
    assert s.current_log_weight == -3.5 and np.array_equal(s.current_coeffs, x)
    # several chains: one vector for all, or one row per chain
    m = lgs.IMHKSampler(lat, sigma, burn_in=0, n_chains=4, seed=8)
    m.current_state = np.zeros(n)
    assert m.current_coeffs.shape == (4, n) and not m.current_coeffs.any()
    rows = np.stack([lat.basis @ (np.full(n, c, dtype=np.float64)) for c in range(4)])
    m.current_state = rows
    assert np.array_equal(m.current_coeffs, np.repeat(np.arange(4)[:, None], n, axis=1))
    V, acc = m.step()
    assert V.shape == (4, n) and acc.shape == (4,) and m.total_proposals == 4
    with pytest.raises(ValueError):
        m.current_state = np.zeros(n + 1)
    with pytest.raises(ValueError):
        m.current_coeffs = np.full(n, 0.5)
