"""The five BASELINE.json workloads as op-list builders + deterministic synthetic data
(SURVEY.md 8d). Each builder restates the model of the cited reference example."""
import numpy as np
from .model import (ModelBuilder, Normal, Gamma, Beta, Bernoulli, Categorical, Poisson, dense, select2,
                    sigmoid_clamped, clamp)


def linear_regression(n_obs=3, seed=12345):
    """C1: examples/linear_regression.rs:32-47 (only the first 3 observations are used, :41-43)."""
    rng = np.random.Generator(np.random.PCG64(seed))
    x = np.arange(n_obs, dtype=np.float64) / 19.0 * 10.0
    y = 2.0 + 1.5 * x + 0.5 * rng.standard_normal(n_obs)
    m = ModelBuilder()
    intercept = m.sample("intercept", Normal(0.0, 10.0))
    slope = m.sample("slope", Normal(0.0, 10.0))
    sigma = m.sample("sigma", Gamma(1.0, 1.0))
    xc, yc = m.data(x), m.data(y)
    m.plate_observe("y", Normal(intercept + slope * xc, sigma), yc)
    return m, dict(x=x, y=y)


def logistic_data(N, D, seed=42):
    rng = np.random.Generator(np.random.PCG64(seed))
    X = rng.standard_normal((N, D))
    X[:, 0] = 1.0
    beta = np.zeros(D)
    k = min(10, D)
    beta[:k] = 0.25 * rng.standard_normal(k)
    p = 1.0 / (1.0 + np.exp(-(X @ beta)))
    y = (rng.random(N) < p).astype(np.uint8)
    return X, y, beta


def logistic_regression(N=1_000_000, D=100, seed=42, data=None):
    """C2: examples/classification.rs:104-135."""
    X, y, beta = data if data is not None else logistic_data(N, D, seed)
    m = ModelBuilder()
    b = m.plate_sample("beta", X.shape[1], Normal(0.0, 2.0))
    Xc, yc = m.data(X), m.data(y)
    m.plate_observe("y", Bernoulli(sigmoid_clamped(dense(Xc, b), 1e-10, 1.0 - 1e-10)), yc)
    return m, dict(X=X, y=y, beta=beta)


def hierarchical(G=10_000, per_group=15, seed=123):
    """C3: examples/hierarchical_models.rs:7-33 (varying intercepts, shared slope)."""
    rng = np.random.Generator(np.random.PCG64(seed))
    N = G * per_group
    gid = np.repeat(np.arange(G, dtype=np.uint32), per_group)
    x = rng.uniform(-2.0, 2.0, N)
    alpha = rng.standard_normal(G)
    y = alpha[gid] + 0.6 * x + 0.8 * rng.standard_normal(N)
    m = ModelBuilder()
    mu_alpha = m.sample("mu_alpha", Normal(0.0, 5.0))
    sigma_alpha = m.sample("sigma_alpha", Gamma(1.0, 1.0))
    beta = m.sample("beta", Normal(0.0, 2.0))
    sigma_y = m.sample("sigma_y", Gamma(1.0, 1.0))
    alphas = m.plate_sample("alpha", G, Normal(mu_alpha, sigma_alpha))
    xc, yc, gc = m.data(x), m.data(y), m.data(gid)
    m.plate_observe("y", Normal(alphas[gc] + beta * xc, sigma_y), yc)
    return m, dict(x=x, y=y, gid=gid, alpha=alpha)


def hierarchical_inits(truth, sites, n_chains, jitter=0.5, seed=0):
    """Over-dispersed data-driven starting points for `hierarchical` (q0 [d][n_chains] in the sorted site
    order `sites`, as fc_hmc_create takes it): group means for the intercepts, their mean / spread for the
    hyper-parameters, slope 0, the overall spread of y for sigma_y; every chain jittered by N(0, jitter)
    (positive parameters on the log scale). From prior draws the intercepts sit ~5 prior sd away and
    relax over thousands of L = 16 transitions; a caller with data at hand starts here."""
    rng = np.random.Generator(np.random.PCG64(seed))
    y, gid = truth["y"], truth["gid"]
    G = int(gid.max()) + 1
    gm = np.bincount(gid, weights=y, minlength=G) / np.maximum(np.bincount(gid, minlength=G), 1)
    q0 = np.empty((len(sites), n_chains))
    for k, nm in enumerate(sites):
        z = jitter * rng.standard_normal(n_chains)
        if nm.startswith("alpha#"):
            q0[k] = gm[int(nm.split("#")[1])] + z
        elif nm == "mu_alpha":
            q0[k] = gm.mean() + z
        elif nm == "sigma_alpha":
            q0[k] = np.exp(np.log(gm.std()) + z)
        elif nm == "beta":
            q0[k] = z
        elif nm == "sigma_y":
            q0[k] = np.exp(np.log(y.std()) + z)
        else:
            raise KeyError(nm)
    return q0


def gaussian_mixture(n_obs=80, seed=42):
    """C4: examples/mixture_models.rs:77-111."""
    rng = np.random.Generator(np.random.PCG64(seed))
    comp = rng.random(n_obs) < 0.6
    x = np.where(comp, -1.5 + 0.8 * rng.standard_normal(n_obs), 2.0 + 1.2 * rng.standard_normal(n_obs))
    m = ModelBuilder()
    pi1 = m.sample("pi1", Beta(1.0, 1.0))
    mu1 = m.sample("mu1", Normal(0.0, 5.0))
    sigma1 = m.sample("sigma1", Gamma(1.0, 1.0))
    mu2 = m.sample("mu2", Normal(0.0, 5.0))
    sigma2 = m.sample("sigma2", Gamma(1.0, 1.0))
    # z#i / x#i are interleaved in the reference's plate; prior and likelihood accumulate
    # separately (trace.rs:172-176), so two consecutive plates give identical sums.
    z = m.plate_sample("z", n_obs, Categorical(clamp(pi1, 0.001, 0.999)))
    xc = m.data(x)
    m.plate_observe("x", Normal(select2(z, mu1, mu2), select2(z, sigma1, sigma2)), xc)
    return m, dict(x=x)


def poisson_glm(N=4_000_000, D=8, seed=7):
    """C5: builder-defined (no reference example; nearest tests/f_mcmc_proposals.rs fg41/fg02):
    theta#k ~ Gamma(2,1), y#i ~ Poisson(sum_k theta_k x_ik)."""
    rng = np.random.Generator(np.random.PCG64(seed))
    X = rng.random((N, D))
    theta = 0.5 + 0.25 * np.arange(D)
    y = rng.poisson(X @ theta).astype(np.uint32)
    m = ModelBuilder()
    th = m.plate_sample("theta", D, Gamma(2.0, 1.0))
    Xc, yc = m.data(X), m.data(y)
    m.plate_observe("y", Poisson(dense(Xc, th)), yc)
    return m, dict(X=X, y=y, theta=theta)
