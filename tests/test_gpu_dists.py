"""All 17 log-densities THROUGH THE DEVICE: the reference's own known-answer vectors
(tests/golden/dist_logprob.json, extracted from src/core/distribution.rs:2073-2355 and
tests/f_dist_distributions.rs:31-293) pushed through one-op programs via fc_logp, value + analytic
gradient against the oracle for every family as a prior and as a plate observation with
latent-dependent parameters, and goodness-of-fit of the 17 device samplers
(tests/f_tests_sampler_validation.rs pins the reference's by KS / chi-square / moments)."""
import json
import math
import os
import numpy as np
import pytest
import oracle
from fugue_b200 import api, model as M, _ir

pytestmark = pytest.mark.gpu
G = os.path.join(os.path.dirname(__file__), "golden")
CASES = json.load(open(os.path.join(G, "dist_logprob.json")))


def dec(v):
    if isinstance(v, str):
        return float(v)
    if isinstance(v, list):
        return [dec(x) for x in v]
    return v


def _one_op(c):
    d, a = c["dist"], dec(c["args"])
    cls = getattr(M, d)
    dist = cls(a[0]) if d == "Categorical" else cls(*a)
    m = M.ModelBuilder()
    m.sample("x", dist)
    return m, dist.value_type != "f64"


@pytest.mark.parametrize("i", range(len(CASES)))
def test_reference_known_answers_on_device(i):
    c = CASES[i]
    m, disc = _one_op(c)
    P = api.Program(m)
    x, exp = dec(c["x"]), dec(c["expected"])
    if disc:
        lp, ll, lf = P.logp(None, np.array([[int(x)]], dtype=np.int64))
    else:
        lp, ll, lf = P.logp(np.array([[float(x)]]))
    got = lp[0]
    assert ll[0] == 0.0 and lf[0] == 0.0
    if math.isinf(exp):
        assert got == exp, (c, got)
    else:
        assert abs(got - exp) < 1e-9, (c, got)  # the reference's own tolerance
        assert abs(got - exp) <= 1e-10 * max(1.0, abs(exp)) or abs(got - exp) < 1e-12, (c, got)
    # and the device agrees with the oracle on the same point
    O = oracle.OracleProgram(m)
    olp = O.logp(None if disc else np.array([[float(x)]]), np.array([[int(x)]], dtype=np.int64) if disc else None)[0][0]
    assert (got == olp) if math.isinf(olp) else abs(got - olp) <= 1e-10 * max(1.0, abs(olp))


def _family_model(name):
    """a ~ Gamma(2,1) (positive), a2 ~ Gamma(3,2), b ~ Normal(0,1), u ~ Beta(2,2); x ~ Family(params built from
    them) as a PRIOR site, and y#i ~ Family(same params) as a PLATE OBSERVATION over a data column."""
    rng = np.random.default_rng(5)
    m = M.ModelBuilder()
    a = m.sample("a", M.Gamma(2.0, 1.0))
    a2 = m.sample("a2", M.Gamma(3.0, 2.0))
    b = m.sample("b", M.Normal(0.0, 1.0))
    u = m.sample("u", M.Beta(2.0, 2.0))
    pos = rng.uniform(0.2, 4.0, 37)
    real = rng.normal(0.0, 2.0, 37)
    fam = {
        "Normal": (lambda: M.Normal(b, a), real), "Uniform": (lambda: M.Uniform(-6.0, 9.0), real),
        "LogNormal": (lambda: M.LogNormal(b, a), pos), "Exponential": (lambda: M.Exponential(a), pos),
        "Beta": (lambda: M.Beta(a, a2), rng.uniform(0.05, 0.95, 37)), "Gamma": (lambda: M.Gamma(a, a2), pos),
        "StudentT": (lambda: M.StudentT(a2, b, a), real), "Cauchy": (lambda: M.Cauchy(b, a), real),
        "Laplace": (lambda: M.Laplace(b, a), real), "Weibull": (lambda: M.Weibull(a, a2), pos),
        "ChiSquared": (lambda: M.ChiSquared(a2), pos), "InverseGamma": (lambda: M.InverseGamma(a, a2), pos),
        "Bernoulli": (lambda: M.Bernoulli(u), rng.integers(0, 2, 37).astype(np.uint8)),
        "Categorical": (lambda: M.Categorical(u), rng.integers(0, 2, 37).astype(np.int64)),
        "Binomial": (lambda: M.Binomial(20.0, u), rng.integers(0, 21, 37).astype(np.int64)),
        "Poisson": (lambda: M.Poisson(a), rng.integers(0, 9, 37).astype(np.int64)),
        "DiscreteUniform": (lambda: M.DiscreteUniform(-3.0, 8.0), rng.integers(-3, 9, 37).astype(np.int64)),
    }[name]
    mk, data = fam
    d = mk()
    m.sample("x", d)
    m.plate_observe("y", mk(), m.data(data))
    return m, d.value_type != "f64", data


FAMILIES = ["Normal", "Uniform", "LogNormal", "Exponential", "Bernoulli", "Categorical", "Beta", "Gamma", "Binomial", "Poisson",
            "StudentT", "Cauchy", "Laplace", "Weibull", "ChiSquared", "InverseGamma", "DiscreteUniform"]


@pytest.mark.parametrize("name", FAMILIES)
def test_family_as_prior_and_plate_observe_value_and_gradient(name):
    m, disc, data = _family_model(name)
    P, O = api.Program(m), oracle.OracleProgram(m)
    rng = np.random.default_rng(11)
    Cn = 64
    sites = P.sites
    q = np.zeros((P.d_cont, Cn))
    for s, nm in enumerate(sites):
        if nm in ("a", "a2"):
            q[s] = rng.uniform(0.4, 3.0, Cn)
        elif nm == "b":
            q[s] = rng.normal(0.0, 1.0, Cn)
        elif nm == "u":
            q[s] = rng.uniform(0.1, 0.9, Cn)
        else:  # the family's own f64 site: values inside its support (and a few outside: -inf on both sides)
            q[s] = rng.uniform(0.05, 0.95, Cn) if name == "Beta" else (rng.normal(0.0, 2.0, Cn) if name in ("Normal", "Uniform", "StudentT", "Cauchy", "Laplace") else rng.uniform(0.1, 5.0, Cn))
            q[s, :3] = [-1.0, 0.0, 1e3]
    qd = None
    if disc:
        lo, hi = {"Bernoulli": (0, 2), "Categorical": (0, 2), "Binomial": (0, 21), "Poisson": (0, 12), "DiscreteUniform": (-3, 9)}[name]
        qd = rng.integers(lo, hi, (P.d_disc, Cn))
        qd[0, :2] = [hi + 3, lo]  # one value outside the support
        if name in ("Bernoulli",):
            qd[0, 0] = 1
    lp, ll, lf = P.logp(q, qd)
    olp, oll, olf, og = O.logp(q, qd, grad=True)

    def close(a, b):
        fin = np.isfinite(b)
        assert np.array_equal(np.isfinite(a), fin) and np.array_equal(a[~fin], b[~fin])
        assert np.all(np.abs(a[fin] - b[fin]) <= 1e-10 * np.maximum(1.0, np.abs(b[fin])))

    close(lp, olp)
    close(ll, oll)
    tot, g = P.logp_grad(q, qd)
    ok = np.isfinite(olp + oll)
    assert ok.sum() >= Cn - 4
    scale = np.abs(og[:, ok]).max(axis=0, keepdims=True)
    assert np.all(np.abs(g[:, ok] - og[:, ok]) <= 1e-10 * np.maximum(np.abs(og[:, ok]), scale) + 1e-300), np.abs(g[:, ok] - og[:, ok]).max()


def test_k_way_categorical_prior_plate_and_validation():
    """Categorical::new(vec) (distribution.rs:679-713, 785-791): K-way constant probabilities as a prior site,
    as a plate of sites and as observations; constructor validation codes."""
    probs = [0.1, 0.2, 0.3, 0.4]
    m = M.ModelBuilder()
    m.sample("k", M.Categorical(probs))
    m.plate_sample("z", 5, M.Categorical(probs))
    m.plate_observe("o", M.Categorical([0.5, 0.0, 0.5]), m.data(np.array([0, 2, 2, 1], dtype=np.int64)))
    P, O = api.Program(m), oracle.OracleProgram(m)
    rng = np.random.default_rng(2)
    qd = rng.integers(0, 4, (P.d_disc, 32))
    qd[0, 0] = 4  # out of range -> -inf (distribution.rs:2313: index == len)
    lp, ll, lf = P.logp(None, qd)
    olp, oll, olf = O.logp(None, qd)
    assert lp[0] == -np.inf and np.all(ll == -np.inf)  # an observation with probability 0
    assert np.allclose(lp[1:], olp[1:], rtol=1e-12) and np.array_equal(ll, oll)
    assert abs(lp[1] - sum(math.log(probs[k]) for k in qd[:, 1])) < 1e-12
    for bad in ([0.5, 0.6], [], [0.5, float("nan"), 0.5], [1.5, -0.5]):
        with pytest.raises(_ir.FugueError) as e:
            M.Categorical(bad)
        assert e.value.code == 102
    # the library validates too (a caller that bypasses the Python constructor)
    mm = M.ModelBuilder()
    c = M.Categorical([0.5, 0.5])
    c.probs = np.array([0.7, 0.7])
    mm.sample("k", c)
    with pytest.raises(_ir.FugueError) as e:
        api.Program(mm)
    assert e.value.code == 102


# ---- goodness of fit of the device samplers (K0): prior populations of one-site programs ----
def _prior_draws(dist, n=200_000, seed=7, generic=0):
    api.debug_set("smc_generic", generic)
    try:
        m = M.ModelBuilder()
        m.sample("x", dist)
        S = api.SmcSession(m, n, api.SMCConfig(), seed)
        r = S.result()
        S.close()
    finally:
        api.debug_set("smc_generic", 0)
    return r.dvalues[0] if dist.value_type != "f64" else r.values[0]


CONT = [("Normal", (1.5, 2.0), lambda st: st.norm(1.5, 2.0)), ("Uniform", (-2.0, 3.0), lambda st: st.uniform(-2.0, 5.0)),
        ("LogNormal", (0.3, 0.6), lambda st: st.lognorm(0.6, scale=math.exp(0.3))), ("Exponential", (2.5,), lambda st: st.expon(scale=1 / 2.5)),
        ("Beta", (2.5, 4.0), lambda st: st.beta(2.5, 4.0)), ("Beta", (0.5, 0.7), lambda st: st.beta(0.5, 0.7)),
        ("Gamma", (3.0, 2.0), lambda st: st.gamma(3.0, scale=0.5)), ("Gamma", (0.4, 1.5), lambda st: st.gamma(0.4, scale=1 / 1.5)),
        ("StudentT", (5.0, 1.0, 2.0), lambda st: st.t(5.0, loc=1.0, scale=2.0)), ("Cauchy", (0.5, 1.5), lambda st: st.cauchy(0.5, 1.5)),
        ("Laplace", (-1.0, 0.8), lambda st: st.laplace(-1.0, 0.8)), ("Weibull", (1.7, 2.5), lambda st: st.weibull_min(1.7, scale=2.5)),
        ("ChiSquared", (4.0,), lambda st: st.chi2(4.0)), ("InverseGamma", (3.5, 2.0), lambda st: st.invgamma(3.5, scale=2.0)),
        # shape 1 takes the exponential short cut of DevRng::gamma (fc_prior.cuh)
        ("Gamma", (1.0, 1.5), lambda st: st.gamma(1.0, scale=1 / 1.5)), ("Beta", (1.0, 1.0), lambda st: st.uniform(0.0, 1.0)),
        ("ChiSquared", (2.0,), lambda st: st.chi2(2.0))]


@pytest.mark.parametrize("name,args,ref", CONT, ids=[f"{c[0]}{c[1]}" for c in CONT])
def test_continuous_samplers_ks(name, args, ref):
    import scipy.stats as st
    x = _prior_draws(getattr(M, name)(*args))
    assert np.all(np.isfinite(x))
    ks = st.kstest(x, ref(st).cdf)
    assert ks.statistic < 1.95 / math.sqrt(x.size), (name, ks)  # alpha ~ 1e-3 (the reference's KS gate is 0.05..0.01)


DISC = [("Bernoulli", (0.3,), lambda st, k: st.bernoulli(0.3).pmf(k), 2), ("Categorical", (0.35,), lambda st, k: np.array([0.35, 0.65])[k], 2),
        ("Categorical", ([0.1, 0.2, 0.3, 0.4],), lambda st, k: np.array([0.1, 0.2, 0.3, 0.4])[k], 4),
        ("Binomial", (20.0, 0.3), lambda st, k: st.binom(20, 0.3).pmf(k), 21), ("Binomial", (10000.0, 0.5), None, 0),
        ("Poisson", (3.5,), lambda st, k: st.poisson(3.5).pmf(k), 16), ("Poisson", (250.0,), None, 0),
        ("DiscreteUniform", (-3.0, 8.0), lambda st, k: np.full(len(k), 1 / 12.0), 12)]


@pytest.mark.parametrize("name,args,pmf,K", DISC, ids=[f"{c[0]}{c[1]}" for c in DISC])
@pytest.mark.parametrize("generic", [0, 1])
def test_discrete_samplers_chi_square_and_moments(name, args, pmf, K, generic):
    import scipy.stats as st
    if name == "Binomial" and args[0] > 255 or name == "Poisson" and args[0] > 100 or name == "DiscreteUniform":
        if not generic and name != "DiscreteUniform":
            pass  # values beyond one byte: the session stores 4-byte discrete sites (zw = 4) on either engine
    x = _prior_draws(getattr(M, name)(*args), generic=generic).astype(np.int64)
    n = x.size
    if pmf is None:  # large-mean families: moments against the closed forms (f_tests_sampler_validation.rs moment gates)
        mean, var = (args[0] * args[1], args[0] * args[1] * (1 - args[1])) if name == "Binomial" else (args[0], args[0])
        assert abs(x.mean() - mean) < 5 * math.sqrt(var / n), (x.mean(), mean)
        assert abs(x.var() - var) < 0.03 * var, (x.var(), var)
        return
    off = -3 if name == "DiscreteUniform" else 0
    ks = np.arange(K)
    cnt = np.bincount(np.clip(x - off, 0, K - 1), minlength=K).astype(float)
    p = np.asarray(pmf(st, ks + off if name != "DiscreteUniform" else ks), dtype=float)
    p[-1] += max(0.0, 1.0 - p.sum())  # tail mass into the last bin
    e = n * p
    keep = e > 5
    chi2 = ((cnt[keep] - e[keep]) ** 2 / e[keep]).sum()
    assert chi2 < st.chi2(keep.sum() - 1).ppf(0.9995), (name, chi2)
    assert x.min() >= off and (name in ("Poisson",) or x.max() <= off + K - 1)
