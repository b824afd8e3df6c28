"""Mean-field VI as a batched scoring client of fc_logp (fugue_b200/vi.py; reference
src/inference/vi.rs:210-922). CPU: the guide families (for_support, log_prob, update clamps, the typed
discrete-latent error). GPU: ELBO identities on a conjugate model -- with q = the exact posterior the
integrand is constant, so the ELBO equals the log evidence to rounding -- and the optimiser."""
import math
import numpy as np
import pytest
from fugue_b200 import vi
from fugue_b200._ir import FugueError
from fugue_b200.model import ModelBuilder, Normal, Gamma, Beta, Bernoulli

Y = np.array([1.5, 2.0, 2.5, 1.8, 2.2])


def _conjugate():
    m = ModelBuilder()
    mu = m.sample("mu", Normal(0.0, 2.0))
    m.plate_observe("y", Normal(mu, 1.0), m.data(Y))
    return m


def _log_evidence():
    n = len(Y)
    cov = np.eye(n) + 4.0 * np.ones((n, n))
    _, logdet = np.linalg.slogdet(cov)
    return -0.5 * (Y @ np.linalg.solve(cov, Y)) - 0.5 * logdet - 0.5 * n * math.log(2 * math.pi)


def test_for_support_and_log_prob():
    """vi.rs:245-292, 378-404; reference test variational_param_sampling_and_log_prob."""
    from scipy import stats
    n = vi.VariationalParam.for_support(vi.REAL, -30.0)
    assert (n.kind, n.a) == ("Normal", -30.0) and abs(n.b - math.log(3.0)) < 1e-15
    assert abs(vi.VariationalParam.for_support(vi.REAL, 0.2).b - math.log(0.1)) < 1e-15     # floor 0.1
    p = vi.VariationalParam.for_support(vi.POSITIVE, 2.5)
    assert p.kind == "LogNormal" and abs(p.a - math.log(2.5)) < 1e-15 and abs(p.b - math.log(0.5)) < 1e-15
    assert vi.VariationalParam.for_support(vi.POSITIVE, -1.0).a == 0.0                        # unsafe init -> 1.0
    u = vi.VariationalParam.for_support(vi.UNIT, 0.25)
    assert u.kind == "Beta" and abs(math.exp(u.a) - 0.5) < 1e-15 and abs(math.exp(u.b) - 1.5) < 1e-15
    x = np.array([0.1, 0.7, 2.3])
    assert np.allclose(n.log_prob(x), stats.norm(-30.0, 3.0).logpdf(x), rtol=1e-13)
    assert np.allclose(p.log_prob(x), stats.lognorm(s=0.5, scale=2.5).logpdf(x), rtol=1e-12)
    xb = np.array([0.1, 0.5, 0.93])
    assert np.allclose(u.log_prob(xb), stats.beta(0.5, 1.5).logpdf(xb), rtol=1e-12)
    assert u.log_prob(np.array([1.5]))[0] == -math.inf and p.log_prob(np.array([-1.0]))[0] == -math.inf
    rng = vi._rng(1)
    assert ((u.sample(rng, 100) > 0) & (u.sample(rng, 100) < 1)).all() and (p.sample(rng, 100) > 0).all()
    bad = vi.VariationalParam("Normal", math.nan, 0.0)
    assert np.isnan(bad.sample(rng, 3)).all()


def test_update_clamps_and_common_random_numbers():
    p = vi.VariationalParam("Normal", 0.0, 0.0)
    p.apply_update(vi.LOCATION, 5e6); p.apply_update(vi.SCALE, 100.0)
    assert p.a == vi.MU_ABS_MAX and p.b == vi.LOG_SCALE_MAX
    b = vi.VariationalParam("Beta", 0.0, 0.0)
    b.apply_update(vi.LOCATION, -100.0); b.apply_update(vi.SCALE, 3.0)
    assert b.a == vi.LOG_SCALE_MIN and b.b == 3.0
    q = vi.VariationalParam("Normal", 1.0, math.log(0.5))
    z0, z1 = q.sample(vi._rng(7), 5), q.shifted(vi.LOCATION, 0.25).sample(vi._rng(7), 5)
    assert np.allclose(z1 - z0, 0.25, rtol=0, atol=1e-15)                                    # CRN: same base normals


def test_guide_from_values_and_discrete_error():
    """reference tests meanfield_from_trace_continuous_ok / discrete rejection (vi.rs:577-606)."""
    g = vi.MeanFieldGuide.from_values({"b": 2.0, "a": -0.5})
    assert sorted(g.params) == ["a", "b"] and g.params["b"].b == math.log(0.2)
    addrs, z, lq = g.sample(vi._rng(0), 8)
    assert addrs == ["a", "b"] and z.shape == (2, 8) and np.isfinite(lq).all()
    with pytest.raises(FugueError) as e:
        vi.MeanFieldGuide.from_values({"k": 3}, {"k": "u64"})
    assert "UnsupportedDiscreteLatent" in str(e.value)


@pytest.mark.gpu
def test_elbo_at_exact_posterior_equals_log_evidence():
    """q = posterior N(1.90476.., sqrt(0.190476..)): log p(y, mu) - log q(mu) = log p(y) for EVERY draw."""
    m = _conjugate()
    prec = 1.0 / 4.0 + len(Y)
    g = vi.MeanFieldGuide()
    g.params["mu"] = vi.VariationalParam("Normal", Y.sum() / prec, 0.5 * math.log(1.0 / prec))
    e = vi.elbo_with_guide(m, g, 4096, seed=3)
    assert abs(e - _log_evidence()) < 1e-9
    # any other guide is a strict lower bound
    g2 = vi.MeanFieldGuide()
    g2.params["mu"] = vi.VariationalParam("Normal", 1.0, math.log(0.8))
    assert vi.elbo_with_guide(m, g2, 20000, seed=4) < _log_evidence() - 0.05
    # a guide latent the model does not sample is ignored (vi.rs:662-668); a missing one is typed
    g.params["unused"] = vi.VariationalParam("Normal", 0.0, 0.0)
    assert abs(vi.elbo_with_guide(m, g, 1024, seed=5) - _log_evidence()) < 1e-9
    with pytest.raises(FugueError) as ex:
        vi.elbo_with_guide(m, vi.MeanFieldGuide(), 8)
    assert ex.value.code == 500


@pytest.mark.gpu
def test_elbo_gradient_and_optimizer_recover_the_posterior():
    m = _conjugate()
    prec = 1.0 / 4.0 + len(Y)
    g = vi.MeanFieldGuide()
    g.add_latent("mu", vi.REAL, 0.5)
    # with common random numbers z = mu_q + sigma_q * eps the integrand is quadratic in mu_q and log q does
    # not depend on it: the central difference is exactly sum(y) - prec * mean(z) on the draws used
    grad = vi.elbo_gradient_fd(11, m, g, "mu", vi.LOCATION, 0.01, 512)
    zbar = float(np.mean(g.params["mu"].sample(vi._rng(11), 512)))
    assert abs(grad - (Y.sum() - prec * zbar)) < 1e-6 * abs(grad)
    assert vi.elbo_gradient_fd(11, m, g, "absent", vi.LOCATION, 0.01, 8) == 0.0
    res = vi.optimize_meanfield_vi_with_config(m, g, vi.VIConfig(n_iterations=600, n_samples_per_iter=64, base_learning_rate=0.1, convergence_window=0), seed=2)
    q = res.guide.params["mu"]
    assert abs(q.a - Y.sum() / prec) < 0.05
    assert abs(math.exp(q.b) - math.sqrt(1.0 / prec)) < 0.08
    assert res.iterations == 600 and len(res.elbo_history) == res.iterations and not res.converged
    early = vi.optimize_meanfield_vi_with_config(m, g, vi.VIConfig(n_iterations=600, n_samples_per_iter=64, base_learning_rate=0.05), seed=2)
    assert early.converged and early.iterations < 600  # the ELBO-plateau stop of vi.rs:806-817
    assert np.mean(res.elbo_history[-10:]) > np.mean(res.elbo_history[:5])
    assert np.mean(res.elbo_history[-10:]) <= _log_evidence() + 0.05


@pytest.mark.gpu
def test_mixed_support_model_and_estimate_elbo():
    """reference tests elbo_computation_is_finite / optimize_vi_updates_parameters_and_is_stable:
    Real + Positive + Unit factors stay finite and inside their clamps; estimate_elbo (q = prior)."""
    m = ModelBuilder()
    mu = m.sample("mu", Normal(0.0, 2.0))
    s = m.sample("s", Gamma(2.0, 2.0))
    p = m.sample("p", Beta(2.0, 2.0))
    m.plate_observe("y", Normal(mu, s), m.data(Y))
    m.plate_observe("f", Bernoulli(p), m.data(np.array([1, 1, 0, 1], dtype=np.uint8)))
    g = vi.MeanFieldGuide()
    g.add_latent("mu", vi.REAL, 1.0); g.add_latent("s", vi.POSITIVE, 1.0); g.add_latent("p", vi.UNIT, 0.5)
    assert math.isfinite(vi.elbo_with_guide(m, g, 256, seed=1))
    out = vi.optimize_meanfield_vi(m, g, 60, 32, 0.05, seed=3)
    for a, q in out.params.items():
        assert math.isfinite(q.a) and math.isfinite(q.b) and vi.LOG_SCALE_MIN <= q.b <= vi.LOG_SCALE_MAX
    assert vi.elbo_with_guide(m, out, 4096, seed=9) > vi.elbo_with_guide(m, g, 4096, seed=9)
    # q = prior on the conjugate model: E[(y - mu)^2] = y^2 + 4
    want = float(np.sum(-0.5 * (Y ** 2 + 4.0) - 0.5 * math.log(2 * math.pi)))
    got = vi.estimate_elbo(_conjugate(), 1 << 18, seed=5)
    assert abs(got - want) < 0.05
