"""GPU single-site MH (adaptive_mcmc_chain): posterior moments within Monte-Carlo error on the
reference's own templates (tests/f_mcmc_proposals.rs, tests/analytical_validation.rs) and the
DiminishingAdaptation trajectory against the CPU restatement."""
import math
import numpy as np
import pytest
import oracle
from fugue_b200 import models
from fugue_b200.model import ModelBuilder, Normal, Gamma, Beta, Bernoulli, Poisson, Categorical, Uniform

pytestmark = pytest.mark.gpu


def test_log_space_walk_targets_gamma():
    """fg02_log_space_walk: x ~ Gamma(3,2) sampled by MH must have mean 1.5 (a missing Jacobian
    would target pi(x)/x with mean 1.0); the site is auto-detected as positive (mh.rs:351)."""
    from fugue_b200 import api
    m = ModelBuilder()
    m.sample("x", Gamma(3.0, 2.0))
    draws, dd, lp, st = api.adaptive_mcmc_chain(m, n_samples=600, n_warmup=400, n_chains=512, seed=1, thin=4)
    x = draws[:, 0, :]
    assert (x > 0).all()
    assert abs(x.mean() - 1.5) < 0.05
    assert abs(x.var() - 0.75) < 0.1


def test_conjugate_normal_mh():
    """analytical_validation.rs template: mu ~ N(0,2), y ~ N(mu,1) x5 -> mean 1.90476, var 0.190476."""
    from fugue_b200 import api
    m = ModelBuilder()
    mu = m.sample("mu", Normal(0.0, 2.0))
    m.plate_observe("y", Normal(mu, 1.0), m.data(np.array([1.5, 2.0, 2.5, 1.8, 2.2])))
    draws, dd, lp, st = api.adaptive_mcmc_chain(m, n_samples=500, n_warmup=500, n_chains=512, seed=2, thin=5)
    x = draws[:, 0, :]
    assert abs(x.mean() - 1.90476) < 0.03
    assert abs(x.var() - 0.190476) < 0.03
    rate = st["n_accept"] / st["n_steps"]
    assert 0.25 < rate.mean() < 0.65  # adapted towards 0.44


def test_beta_bernoulli_and_discrete_sites():
    """Beta(1,1) prior + 19/29 heads -> Beta(20,11); plus a Bernoulli latent (Flip proposal) and a
    2-way Categorical (PriorResample) whose posterior is its prior (no likelihood attached)."""
    from fugue_b200 import api
    m = ModelBuilder()
    p = m.sample("p", Beta(1.0, 1.0))
    m.sample("b", Bernoulli(0.3))
    m.sample("k", Categorical(0.25))
    m.plate_observe("y", Bernoulli(p), m.data(np.array([1] * 19 + [0] * 10, dtype=np.uint8)))
    draws, dd, lp, st = api.adaptive_mcmc_chain(m, n_samples=900, n_warmup=600, n_chains=512, seed=3, thin=3)
    x = draws[:, 0, :]
    assert (x > 0).all() and (x < 1).all()
    assert abs(x.mean() - 20.0 / 31.0) < 0.02
    P = api.Program(m)
    # discrete rows are in address order: "b", "k"
    assert abs(dd[:, 0, :].mean() - 0.3) < 0.03
    assert abs((dd[:, 1, :] == 0).mean() - 0.25) < 0.03


def test_poisson_rate_with_reflect_override():
    from fugue_b200 import api
    m = ModelBuilder()
    lam = m.sample("lam", Uniform(0.0, 20.0))
    y = np.array([3, 5, 4, 6, 2, 4, 5, 3], dtype=np.uint32)
    m.plate_observe("y", Poisson(lam), m.data(y))
    S = api.MhSession(m, n_chains=512, n_warmup=500, seed=4, overrides={"lam": api.SiteProposal.Reflect(0.0, 20.0)})
    S.step(500)
    xs = []
    for _ in range(60):
        S.step(5)
        xs.append(S.get()["q"][0])
    x = np.stack(xs)
    assert (x >= 0).all() and (x <= 20).all()
    # posterior Gamma(sum y + 1, n) truncated at 20 (negligible mass beyond)
    assert abs(x.mean() - (y.sum() + 1) / len(y)) < 0.1


def test_diminishing_adaptation_matches_oracle():
    """One chain, one site: replay the accept/reject sequence the device produced through the CPU
    restatement of DiminishingAdaptation::update (mcmc_utils.rs:88-150)."""
    from fugue_b200 import api
    m = ModelBuilder()
    m.sample("x", Normal(0.0, 1.0))
    S = api.MhSession(m, n_chains=4, n_warmup=200, seed=9)
    acc_hist = []
    prev = np.zeros(4, dtype=np.int64)
    scales = []
    for _ in range(200):
        S.step(1)
        g = S.get(scales=True)
        acc_hist.append(g["n_accept"] - prev)
        prev = g["n_accept"].copy()
        scales.append(g["scales"][0].copy())
    acc_hist, scales = np.array(acc_hist), np.array(scales)
    for c in range(4):
        ref = oracle.dim_adapt(acc_hist[:, c].astype(np.uint8))
        assert np.allclose(scales[:, c], ref, rtol=1e-12, atol=0), c
    S.step(50)  # frozen after warmup (mh.rs:989-1011)
    assert np.array_equal(S.get(scales=True)["scales"][0], scales[-1])


def test_poisson_glm_small_mh_recovers_rates():
    """BASELINE configs[4] shape at reduced size: theta_k ~ Gamma(2,1), y ~ Poisson(X theta)."""
    from fugue_b200 import api
    m, dat = models.poisson_glm(N=4000, D=3, seed=7)
    draws, dd, lp, st = api.adaptive_mcmc_chain(m, n_samples=1500, n_warmup=1500, n_chains=128, seed=5, thin=10)
    th = draws[50:].mean(axis=(0, 2))
    P = api.Program(m)
    true = dat["theta"]
    order = [int(s.split("#")[1]) for s in P.sites]
    assert np.allclose(th, true[order], atol=0.15), (th, true[order])
