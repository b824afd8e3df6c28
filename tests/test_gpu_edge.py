"""Edge cases of the device path against the oracle: empty and ragged plates, ragged chain counts
(1 chain, counts straddling the warp / CTA widths), groups that are empty or longer than the
shared-memory chunk of the grouped-plate kernel, dense matrices of width 1 / 3 / 104 (the maximum) and
the typed rejection beyond it, degenerate resampling inputs (n = 1, zero weights, u = 0 wrap), models
without continuous sites, and the typed structural errors (301 / 302)."""
import math
import numpy as np
import pytest
import oracle
from fugue_b200.model import ModelBuilder, Normal, Gamma, Bernoulli, Poisson, Categorical, dense, sigmoid_clamped
from fugue_b200._ir import FugueError

pytestmark = pytest.mark.gpu


def _close(a, b, scale=0.0, rel=1e-10):
    a, b = np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64)
    both_inf = np.isinf(a) & np.isinf(b) & (np.sign(a) == np.sign(b))
    with np.errstate(invalid="ignore"):
        return (np.abs(a - b) <= rel * np.maximum(np.maximum(np.abs(a), np.abs(b)), scale)) | both_inf


def _check(m, Cn, rng, pos=("s",)):
    from fugue_b200 import api
    O, P = oracle.OracleProgram(m), api.Program(m)
    q = rng.normal(0, 0.5, (O.d_cont, Cn))
    for i, nm in enumerate(O.sites()):
        if nm in pos:
            q[i] = rng.uniform(0.5, 2.0, Cn)
    olp, oll, olf, og = O.logp(q, grad=True)
    lp, ll, lf = P.logp(q)
    assert _close(lp, olp).all() and _close(ll, oll).all() and _close(lf, olf).all()
    tot, g = P.logp_grad(q)
    assert _close(tot, (olp + oll) + olf).all()
    if O.d_cont:
        sc = np.abs(og).max(axis=0, keepdims=True) * 1e-4
        assert _close(g, og, scale=sc).all(), np.abs(g - og).max()
    P.close()


def _mixed_model(rng, N, G, D):
    m = ModelBuilder()
    a = m.sample("a", Normal(0.0, 2.0))
    s = m.sample("s", Gamma(2.0, 2.0))
    al = m.plate_sample("al", G, Normal(a, 1.0))
    gid = m.data(rng.integers(0, G, N).astype(np.uint32))
    m.plate_observe("y", Normal(al[gid] + a, s), m.data(rng.normal(size=N)))
    b = m.plate_sample("b", D, Normal(0.0, 1.0))
    X = m.data(rng.normal(size=(N, D)))
    m.plate_observe("k", Bernoulli(sigmoid_clamped(dense(X, b), 1e-10, 1 - 1e-10)), m.data((rng.random(N) < 0.5).astype(np.uint8)))
    m.plate_observe("c", Poisson(dense(m.data(rng.random((N, D)) + 0.1), m.plate_sample("th", D, Gamma(2.0, 1.0)))), m.data(rng.poisson(2.0, N).astype(np.uint32)))
    return m


@pytest.mark.parametrize("Cn", [1, 2, 31, 33, 127, 129, 257])
def test_ragged_chain_counts(Cn):
    rng = np.random.default_rng(Cn)
    from fugue_b200 import api
    m = _mixed_model(rng, 211, 7, 5)
    O, P = oracle.OracleProgram(m), api.Program(m)
    q = rng.normal(0, 0.4, (O.d_cont, Cn))
    for i, nm in enumerate(O.sites()):
        if nm == "s" or nm.startswith("th#"):
            q[i] = rng.uniform(0.5, 2.0, Cn)
    olp, oll, olf, og = O.logp(q, grad=True)
    tot, g = P.logp_grad(q)
    assert _close(tot, (olp + oll) + olf).all()
    assert _close(g, og, scale=np.abs(og).max(axis=0, keepdims=True) * 1e-4).all()


@pytest.mark.parametrize("N", [1, 31, 32, 33, 1023, 1025, 2049])
def test_ragged_plate_lengths(N):
    rng = np.random.default_rng(N)
    from fugue_b200 import api
    m = _mixed_model(rng, N, 3, 4)
    O, P = oracle.OracleProgram(m), api.Program(m)
    Cn = 19
    q = rng.normal(0, 0.4, (O.d_cont, Cn))
    for i, nm in enumerate(O.sites()):
        if nm == "s" or nm.startswith("th#"):
            q[i] = rng.uniform(0.5, 2.0, Cn)
    olp, oll, olf, og = O.logp(q, grad=True)
    tot, g = P.logp_grad(q)
    assert _close(tot, (olp + oll) + olf).all()
    assert _close(g, og, scale=np.abs(og).max(axis=0, keepdims=True) * 1e-4).all()


def test_empty_plates_and_empty_groups_and_giant_group():
    rng = np.random.default_rng(0)
    # (a) plates of length 0
    m = ModelBuilder()
    a = m.sample("a", Normal(0.0, 2.0))
    s = m.sample("s", Gamma(2.0, 2.0))
    m.plate_observe("y", Normal(a, s), m.data(np.zeros(0)))
    m.plate_sample("z", 0, Normal(a, 1.0))
    _check(m, 5, rng)
    # (b) groups without any element (gid skips 1 and 4) and one group longer than the 1024-element chunk
    G, big = 6, 2600
    gid = np.concatenate([np.zeros(3), np.full(big, 2), np.full(5, 3), np.full(2, 5)]).astype(np.uint32)
    rng.shuffle(gid)
    m = ModelBuilder()
    a = m.sample("a", Normal(0.0, 2.0))
    s = m.sample("s", Gamma(2.0, 2.0))
    al = m.plate_sample("al", G, Normal(a, 1.0))
    x = rng.normal(size=gid.size)
    m.plate_observe("y", Normal(al[m.data(gid)] + a * m.data(x), s), m.data(rng.normal(size=gid.size)))
    _check(m, 130, rng)


@pytest.mark.parametrize("K", [1, 3, 8, 33, 104])
def test_dense_widths(K):
    rng = np.random.default_rng(K)
    N = 301
    m = ModelBuilder()
    b = m.plate_sample("b", K, Normal(0.0, 1.0))
    X = m.data(rng.normal(size=(N, K)) / math.sqrt(K))
    m.plate_observe("k", Bernoulli(sigmoid_clamped(dense(X, b), 1e-10, 1 - 1e-10)), m.data((rng.random(N) < 0.5).astype(np.uint8)))
    _check(m, 37, rng)


def test_dense_width_beyond_the_tile_is_a_typed_rejection():
    from fugue_b200 import api
    rng = np.random.default_rng(1)
    N, K = 64, 105
    m = ModelBuilder()
    b = m.plate_sample("b", K, Normal(0.0, 1.0))
    m.plate_observe("k", Bernoulli(sigmoid_clamped(dense(m.data(rng.normal(size=(N, K))), b), 1e-10, 1 - 1e-10)), m.data(np.zeros(N, dtype=np.uint8)))
    P = api.Program(m)
    with pytest.raises(FugueError) as e:
        P.logp(np.zeros((K, 2)))
    assert e.value.code == 1000  # no CPU fallback: unsupported shapes fail loudly


def test_degenerate_resampling_inputs():
    from fugue_b200 import api
    for method in (0, 1, 2):
        assert np.array_equal(api.resample_indices(method, [1.0], [0.3]), [0])
    w = np.array([0.0, 0.0, 0.5, 0.0, 0.5, 0.0])
    for method, us in ((1, [0.25]), (2, [0.1, 0.9, 0.5, 0.5, 0.01, 0.99]), (0, [0.0, 0.5, 0.50001, 1.0, 0.2, 0.7])):
        assert np.array_equal(api.resample_indices(method, w, us), oracle.resample(method, w, us))
    # u = 0: the first threshold is 0, `cum < thr` fails at once and (i - 1) wraps to n - 1
    w = np.full(8, 0.125)
    assert np.array_equal(api.resample_indices(1, w, [0.0]), oracle.resample(1, w, [0.0]))
    # unnormalised and all-zero weights follow the sequential loop too
    for w in (np.array([3.0, 1.0, 4.0, 1.0, 5.0]), np.zeros(4)):
        assert np.array_equal(api.resample_indices(1, w, [0.6]), oracle.resample(1, w, [0.6]))


def test_model_without_continuous_sites():
    from fugue_b200 import api
    m = ModelBuilder()
    m.sample("b", Bernoulli(0.3))
    m.sample("k", Categorical(0.25))
    draws, lp, st = api.hmc_chain(m, n_samples=50, n_warmup=10, n_chains=64, seed=1)   # hmc.rs:828-846: prior draws
    assert draws.shape == (50, 0, 64)
    d, dd, lpj, st = api.adaptive_mcmc_chain(m, n_samples=400, n_warmup=100, n_chains=256, seed=2, thin=2)
    assert abs(dd[:, 0, :].mean() - 0.3) < 0.03 and abs((dd[:, 1, :] == 0).mean() - 0.25) < 0.03


def test_structural_errors_are_typed():
    from fugue_b200 import api
    m = ModelBuilder()
    m.sample("x", Normal(0.0, 1.0))
    m.sample("x", Normal(0.0, 1.0))
    with pytest.raises(FugueError) as e:
        api.Program(m)
    assert e.value.code == 301  # AddressConflict
    m1 = ModelBuilder()
    m1.sample("x", Normal(0.0, 1.0))
    m1.plate_observe("y", Normal(0.0, 1.0), m1.data(np.zeros(3)))
    m2 = ModelBuilder()
    m2.sample("x", Normal(0.0, 1.0))
    m2.plate_observe("y", Normal(0.0, 1.0), m2.data(np.zeros(4)))  # the structure depends on the run
    P = api.Program(m1)
    with pytest.raises(FugueError) as e:
        P.check_structure(m2)
    assert e.value.code == 302  # UnexpectedModelStructure


def test_poisson_dense_rate_ranges():
    """Identity-link Poisson plate (distribution.rs:1237-1258) at the edges of the fast path: rates
    below / above the table-log range (2^-1021, 512), the lambda > 700 && k == 0 special case,
    zero / negative / NaN / infinite rates (-inf, no gradient: NaN, which HMC reads as divergent), one chain per regime."""
    from fugue_b200 import api
    rng = np.random.default_rng(5)
    N, D = 300, 3
    m = ModelBuilder()
    th = m.plate_sample("th", D, Normal(0.0, 1e6))
    X = rng.random((N, D)) + 0.5
    k = rng.poisson(2.0, N).astype(np.uint32)
    k[::7] = 0
    m.plate_observe("c", Poisson(dense(m.data(X), th)), m.data(k))
    O, P = oracle.OracleProgram(m), api.Program(m)
    scales = [1.0, 100.0, 300.0, 600.0, 2000.0, 1e-5, 1e-200, 1e-305, 1e150, 1e305]
    q = np.stack([np.full(D, s) for s in scales], axis=1)
    extra = np.stack([np.array([1.0, -0.2, 0.1]), np.array([-1.0, -1.0, -1.0]), np.array([0.0, 0.0, 0.0]),
                      np.array([np.nan, 1.0, 1.0]), np.array([np.inf, 1.0, 1.0])], axis=1)
    q = np.concatenate([q, extra], axis=1)
    # all-zero observations: every row takes the k == 0 rule
    olp, oll, olf, og = O.logp(q, grad=True)
    lp, ll, lf = P.logp(q)
    assert _close(ll, oll).all(), (ll, oll)
    tot, g = P.logp_grad(q)
    assert _close(tot, (olp + oll) + olf).all()
    fin = np.isfinite(oll)
    assert fin[:5].all() and not fin[-4:].any()
    fin &= np.isfinite(olp)  # theta = 1e305 overflows the prior's z^2: -inf prior, no gradient (NaN on both sides)
    assert np.isnan(g[:, ~np.isfinite(olp + oll)]).any(axis=0).all()
    assert _close(g[:, fin], og[:, fin], scale=np.abs(og[:, fin]).max(axis=0, keepdims=True) * 1e-6).all()
    P.close()
    m0 = ModelBuilder()
    th0 = m0.plate_sample("th", D, Normal(0.0, 1e6))
    m0.plate_observe("c", Poisson(dense(m0.data(X), th0)), m0.data(np.zeros(N, dtype=np.uint32)))
    O0, P0 = oracle.OracleProgram(m0), api.Program(m0)
    _, oll0, _, og0 = O0.logp(q[:, :10], grad=True)
    _, ll0, _ = P0.logp(q[:, :10])
    assert _close(ll0, oll0).all(), (ll0, oll0)
    _, g0 = P0.logp_grad(q[:, :10])
    f0 = np.isfinite(og0).all(axis=0)  # theta = 1e305: -inf prior => no gradient on either side
    assert f0[:9].all() and np.array_equal(np.isnan(g0).any(axis=0), ~f0)
    assert _close(g0[:, f0], og0[:, f0], scale=np.abs(og0[:, f0]).max(axis=0, keepdims=True) * 1e-6).all()
    P0.close()
