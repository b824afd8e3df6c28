"""GPU SMC: bit-exact resampling indices (given the same weights and uniforms), weight numerics
against the reference's golden values, tempered-SMC evidence / moments within Monte-Carlo error
(templates: /root/reference/tests/f_smc_smc.rs, src/core/numerical.rs tests)."""
import json, math, os
import numpy as np
import pytest
import oracle
from fugue_b200 import models
from fugue_b200.model import ModelBuilder, Normal, Beta, Bernoulli

pytestmark = pytest.mark.gpu
G = os.path.join(os.path.dirname(__file__), "golden")


def weight_sets(n, rng):
    out = {}
    out["uniform"] = np.full(n, 1.0 / n)
    w = rng.random(n); out["random"] = w / w.sum()
    w = np.zeros(n); w[n // 3] = 1.0; out["one_hot"] = w
    w = 0.5 ** np.arange(1, n + 1, dtype=np.float64); w[-1] = w[-2] if n > 1 else 1.0; out["geometric"] = w
    w = np.sort(rng.random(n)); out["sorted_up"] = w / w.sum()
    out["sorted_down"] = out["sorted_up"][::-1].copy()
    w = rng.random(n) * 1e-310; w[rng.integers(0, n)] = 1.0; out["denormal_heavy"] = w
    w = rng.exponential(size=n) ** 8; out["heavy_tail"] = w / w.sum()
    w = np.where(rng.random(n) < 0.5, 0.0, rng.random(n)); w[0] = 0.0; out["zeros"] = w / max(w.sum(), 1e-300)
    w = np.full(n, 2.0 ** -54); w[0] = 0.75; out["ties"] = w        # exact round-to-even ties against ulp(0.75)
    w = rng.random(n); out["unnormalised_sum_gt_1"] = w * 3.0 / w.sum()
    return out


@pytest.mark.parametrize("n", [1, 2, 7, 1000, 4097, 1 << 17])
def test_exact_prefix_sums_bitwise(n):
    from fugue_b200 import api
    rng = np.random.default_rng(n)
    for name, w in weight_sets(n, rng).items():
        S = api.exact_prefix_sums(w)
        ref = np.zeros(n + 1)
        c = 0.0
        for i in range(n):
            c = c + w[i]
            ref[i + 1] = c
        assert np.array_equal(S.view(np.uint64), ref.view(np.uint64)), (name, n, np.flatnonzero(S != ref)[:5])


@pytest.mark.parametrize("n", [1, 2, 1000, 1 << 16])
@pytest.mark.parametrize("method", [0, 1, 2])
def test_resampling_indices_bit_exact(n, method):
    from fugue_b200 import api
    rng = np.random.default_rng(100 * n + method)
    if method == 0 and n > 4096:
        n = 4096  # the reference's multinomial is O(n^2)
    for name, w in weight_sets(n, rng).items():
        us = rng.random(1 if method == 1 else n)
        got = api.resample_indices(method, w, us)
        ref = oracle.resample(method, w, us)
        assert np.array_equal(got, ref), (name, n, method, np.flatnonzero(got != ref)[:5])
    # u == 0 edge: i - 1 wraps to n - 1 in release builds
    if method == 1:
        w = np.full(n, 1.0 / n)
        assert np.array_equal(api.resample_indices(1, w, [0.0]), oracle.resample(1, w, [0.0]))


def test_resampling_large_population_bit_exact():
    from fugue_b200 import api
    n = 1 << 22
    rng = np.random.default_rng(5)
    lw = rng.normal(size=n) * 3.0
    w = np.exp(lw - lw.max()); w /= w.sum()
    got = api.resample_indices(1, w, [0.37])
    ref = oracle.resample(1, w, [0.37])
    assert np.array_equal(got, ref)


def test_weight_numerics_golden():
    from fugue_b200 import api
    N = json.load(open(os.path.join(G, "numerics.json")))
    dec = lambda v: [float(x) for x in v]
    for c in N["log_sum_exp"]:
        got, exp = api.log_sum_exp(dec(c["x"])), float(c["expected"])
        assert (got == exp) if (c.get("exact") or math.isinf(exp)) else abs(got - exp) < 1e-9
    for c in N["normalize_log_probs"]:
        got = api.normalize_log_probs(c["x"])
        assert np.allclose(got, c["expected"], atol=1e-9, rtol=0) and abs(got.sum() - 1.0) < 1e-9
    rng = np.random.default_rng(2)
    x = rng.normal(size=100000) * 5
    assert abs(api.log_sum_exp(x) - oracle.log_sum_exp(x)) < 1e-10 * abs(oracle.log_sum_exp(x))
    w = rng.random(5000); w /= w.sum()
    assert abs(api.effective_sample_size(w) - oracle.ess(w)) < 1e-10 * oracle.ess(w)


@pytest.mark.parametrize("n", [1, 2, 3, 1000, 4097, 1 << 18])
def test_next_beta_matches_oracle(n):
    """next_beta (smc.rs:588-622): bisection result vs the sequential oracle. The device sums in a
    different order (and uses sum e^2 for lse(2 lv)), so the ESS curve differs in the last bits and
    beta agrees to ~1e-12 relative, not bitwise."""
    from fugue_b200 import api
    rng = np.random.default_rng(n)
    for trial in range(4):
        ll = rng.normal(size=n) * (5.0 if trial % 2 else 50.0) - 30.0
        lw = np.full(n, -math.log(n)) if trial < 2 else np.log(rng.dirichlet(np.ones(n)))
        beta = [0.0, 0.3, 0.0, 0.9][trial]
        target = 0.5 * n
        got, ref = api.next_beta(beta, lw, ll, target), oracle.next_beta(beta, lw, ll, target)
        assert abs(got - ref) <= 1e-10 * abs(ref), (n, trial, got, ref)
    # degenerate inputs: all -inf log-likelihoods (lse non-finite => ESS := n => jump to 1.0), NaN
    ll = np.full(n, -np.inf)
    lw = np.full(n, -math.log(n))
    assert api.next_beta(0.0, lw, ll, 0.5 * n) == oracle.next_beta(0.0, lw, ll, 0.5 * n)
    ll = rng.normal(size=n); ll[0] = np.nan
    assert api.next_beta(0.0, lw, ll, 0.5 * n) == oracle.next_beta(0.0, lw, ll, 0.5 * n)
    # flat likelihood: a full jump keeps the ESS
    assert api.next_beta(0.2, lw, np.zeros(n), 0.5 * n) == 1.0


def test_importance_sampling_default_does_not_square_prior():
    """tests/f_smc_smc.rs: p ~ Beta(2,2) [prior], 27 coin flips with 19 heads... the reference checks
    the IS posterior mean of a Beta-Bernoulli model: Beta(1,1) prior, 19 heads / 10 tails ->
    Beta(20,11), mean 20/31 +- 0.03; default config (rejuvenation_steps = 0) never resamples."""
    from fugue_b200 import api
    m = ModelBuilder()
    p = m.sample("p", Beta(1.0, 1.0))
    y = np.array([1] * 19 + [0] * 10, dtype=np.uint8)
    m.plate_observe("y", Bernoulli(p), m.data(y))
    r = api.adaptive_smc(m, 20000, api.SMCConfig(), seed=4)
    assert r.n_temper_steps == 1
    mean = float((r.weight * r.values[0]).sum())
    assert abs(mean - 20.0 / 31.0) < 0.03
    assert abs(r.weight.sum() - 1.0) < 1e-9
    # log evidence of the Beta-Bernoulli model: log B(20,11) - log B(1,1)
    le = math.lgamma(20) + math.lgamma(11) - math.lgamma(31)
    assert abs(r.log_evidence - le) < 0.1


def test_tempered_smc_evidence_and_moments():
    """tests/f_smc_smc.rs / examples/smc_inference.rs: mu ~ N(0,2), y_i ~ N(mu,1) for y = [1.0, 1.2, 1.4]
    (conjugate): tempered SMC log Z within 0.2 of the closed form, posterior mean within 0.06;
    rejuvenation keeps the weights uniform after each intermediate resample."""
    from fugue_b200 import api
    ys = np.array([1.0, 1.2, 1.4])
    m = ModelBuilder()
    mu = m.sample("mu", Normal(0.0, 2.0))
    m.plate_observe("y", Normal(mu, 1.0), m.data(ys))
    # closed forms
    n, s0sq = len(ys), 4.0
    post_var = 1.0 / (1.0 / s0sq + n)
    post_mean = post_var * ys.sum()
    cov = np.eye(n) + s0sq * np.ones((n, n))
    sign, logdet = np.linalg.slogdet(cov)
    log_z = -0.5 * (n * math.log(2 * math.pi) + logdet + ys @ np.linalg.solve(cov, ys))
    r = api.adaptive_smc(m, 20000, api.SMCConfig(rejuvenation_steps=3), seed=11)
    assert r.n_temper_steps >= 2
    mean = float((r.weight * r.values[0]).sum())
    var = float((r.weight * (r.values[0] - mean) ** 2).sum())
    assert abs(r.log_evidence - log_z) < 0.2, (r.log_evidence, log_z)
    assert abs(mean - post_mean) < 0.06, (mean, post_mean)
    assert abs(var - post_var) < 0.1 * post_var + 0.02
    ess = 1.0 / (r.weight ** 2).sum()
    assert ess > 0.3 * 20000
    for method in (0, 2):
        r2 = api.adaptive_smc(m, 4096, api.SMCConfig(resampling_method=method, rejuvenation_steps=2), seed=3)
        assert abs(float((r2.weight * r2.values[0]).sum()) - post_mean) < 0.1


def test_mixture_model_smc_runs_and_orders_components():
    """BASELINE configs[3] at reduced P: gaussian mixture via tempered SMC with discrete z sites."""
    from fugue_b200 import api
    m, dat = models.gaussian_mixture(n_obs=80)
    r = api.adaptive_smc(m, 1 << 14, api.SMCConfig(rejuvenation_steps=2), seed=8)
    assert np.isfinite(r.log_evidence)
    assert r.dvalues.shape == (80, 1 << 14) and set(np.unique(r.dvalues)) <= {0, 1}
    assert abs(r.weight.sum() - 1.0) < 1e-9
    s = r.sites
    sig1 = r.values[s.index("sigma1")]
    assert (sig1 > 0).all()


def test_standalone_prior_particles_resample_rejuvenate():
    """The SMC surface around adaptive_smc on resident particles: smc_prior_particles (smc.rs:764-790,
    log_weight = log-likelihood), normalize_particles / effective_sample_size (smc.rs:230-233, 719-754),
    resample_particles (smc.rs:326-349; indices bit-exact vs the sequential oracle on the same weights
    and uniform, ancestors copied, weights reset to 1/n) and rejuvenate_particles (smc.rs:698-713;
    pi_1-invariant moves leave the SIR posterior in place and the weights untouched)."""
    from fugue_b200 import api
    y = np.array([1.5, 2.0, 2.5, 1.8, 2.2])
    m = ModelBuilder()
    mu = m.sample("mu", Normal(0.0, 2.0))
    m.plate_observe("y", Normal(mu, 1.0), m.data(y))
    n = 1 << 16
    S = api.smc_prior_particles(m, n, seed=9)
    log_norm, ess = S.normalize()
    r0 = S.result()
    v0 = r0.values[0].copy()
    ll = np.array([(-0.5 * (y - v) ** 2 - 0.5 * math.log(2 * math.pi)).sum() for v in v0])
    mx = ll.max()
    lse = mx + math.log(np.exp(ll - mx).sum())
    assert abs(log_norm - lse) < 1e-10 * abs(lse)
    w_ref = np.exp(ll - lse)
    assert np.allclose(r0.weight, w_ref, rtol=1e-9, atol=1e-300)
    assert abs(ess - 1.0 / (r0.weight ** 2).sum()) < 1e-9 * ess
    assert abs(S.effective_sample_size() - ess) == 0.0
    for method, us in ((1, [0.37]), (2, np.random.default_rng(1).random(n)), (0, np.random.default_rng(2).random(n))):
        S2 = api.smc_prior_particles(m, n, seed=9)
        S2.normalize()
        w_dev = S2.result().weight
        idx = S2.resample_particles(method, us)
        assert np.array_equal(idx, oracle.resample(method, w_dev, us))
        r1 = S2.result()
        assert np.array_equal(r1.values[0], v0[idx.astype(np.int64)])
        assert np.all(r1.weight == 1.0 / n) or np.allclose(r1.weight, 1.0 / n, rtol=1e-12)
        S2.close()
    idx = S.resample_particles(api.ResamplingMethod.Systematic)  # session stream uniform
    assert idx.shape == (n,) and (np.diff(idx.astype(np.int64)) >= 0).all()
    api.rejuvenate_particles(S, 1.0, 8)
    r2 = S.result()
    x = r2.values[0]
    assert abs(x.mean() - 1.90476) < 0.02 and abs(x.var() - 0.190476) < 0.02
    assert np.allclose(r2.weight, 1.0 / n, rtol=1e-12)
    assert len(np.unique(x)) > n // 4  # the moves diversified the resampled population
    S.close()


@pytest.mark.parametrize("vw", [1, 2, 3, 8])
@pytest.mark.parametrize("method", [1, 2])
def test_virtual_shards_resample_matches_single_population(vw, method):
    """The sharded resample's device logic on one GPU (fc_smc_resample_virtual_shards): the population is cut into
    vw virtual ranks processed one after the other -- exact scan continuing from the previous shard's end value,
    slot routing from the (carry, end) pairs, offset gathers -- and must reproduce the sequential reference loop
    over the whole population bit for bit (the 2-GPU run of the real thing: tests/test_gpu_smc_sharded.py)."""
    from fugue_b200 import api, models
    n, seed = 100003, 3
    S = api.SmcSession(models.gaussian_mixture(n_obs=12)[0], n, api.SMCConfig(rejuvenation_steps=1), seed)
    S.reweight(0.04)
    w = S.weights()[1]
    before = S.result()
    u = 0.0 if vw == 3 else 0.37
    idx = S.resample_virtual_shards(vw, method, u).astype(np.int64)
    us = [u] if method == 1 else oracle.philox_uniforms(seed, 0x8000000000000000, 0x1000000000000000 | 1, n)
    anc = oracle.resample(method, w, us).astype(np.int64)
    assert np.array_equal(idx, anc)
    after = S.result()
    assert np.array_equal(after.values, before.values[:, anc]) and np.array_equal(after.dvalues, before.dvalues[:, anc])
    assert np.allclose(after.weight, 1.0 / n, rtol=1e-12)
    S.close()
