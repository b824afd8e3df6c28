"""Pins the CPU oracle against the reference's own known-answer tests (tests/golden/*.json,
extracted by tests/golden/make_golden.py) and independent closed forms."""
import json, math, os
import numpy as np
import pytest
import oracle
from fugue_b200 import _ir, models
from fugue_b200.model import ModelBuilder, Normal

G = os.path.join(os.path.dirname(__file__), "golden")


def dec(v):
    if isinstance(v, str):
        return float(v)
    if isinstance(v, list):
        return [dec(x) for x in v]
    return v


CASES = json.load(open(os.path.join(G, "dist_logprob.json")))


def eval_case(c):
    d, a, x = c["dist"], dec(c["args"]), dec(c["x"])
    did = _ir.DIST_IDS[d]
    if d == "Categorical":
        probs = oracle.f64(a[0])
        return oracle.lib().fo_categorical_lp(oracle._d(probs), probs.size, int(x))
    if d == "Binomial":
        return oracle.lib().fo_binomial_lp(int(a[0]), float(a[1]), int(x))
    if d == "DiscreteUniform":
        return oracle.lib().fo_discrete_uniform_lp(int(a[0]), int(a[1]), int(x))
    if d in ("Bernoulli", "Poisson"):
        return oracle.dist_logp(did, a, xi=int(x))
    return oracle.dist_logp(did, a, x=float(x))


@pytest.mark.parametrize("i", range(len(CASES)))
def test_dist_known_answers(i):
    c = CASES[i]
    got, exp = eval_case(c), dec(c["expected"])
    if math.isinf(exp) or c["exact"] and math.isinf(exp):
        assert got == exp, c
    else:
        assert abs(got - exp) < 1e-9, (c, got)  # the reference's own tolerance
        assert abs(got - exp) <= 1e-10 * max(1.0, abs(exp)) or abs(got - exp) < 1e-12, (c, got)


def test_numerics_golden():
    N = json.load(open(os.path.join(G, "numerics.json")))
    for c in N["log_sum_exp"]:
        got, exp = oracle.log_sum_exp(dec(c["x"])), dec(c["expected"])
        assert (got == exp) if (c.get("exact") or math.isinf(exp)) else abs(got - exp) < 1e-9
    for c in N["normalize_log_probs"]:
        got = oracle.normalize_log_probs(c["x"])
        assert np.allclose(got, c["expected"], atol=1e-9, rtol=0) and abs(got.sum() - 1.0) < 1e-9
    for c in N["log1p_exp"]:
        got = oracle.lib().fo_log1p_exp(c["x"])
        assert (got == c["expected"]) if c.get("exact") else abs(got - c["expected"]) < c.get("tol", 1e-9)


def test_whole_model_score_equals_closed_forms():
    """hmc.rs:975-989: mu ~ N(0,1); observe N(mu,1) at 1.0; total == sum of the two closed forms."""
    N = json.load(open(os.path.join(G, "numerics.json")))
    m = ModelBuilder()
    mu = m.sample("mu", Normal(0.0, 1.0))
    m.observe("y", Normal(mu, 1.0), 1.0)
    P = oracle.OracleProgram(m)
    for v in N["whole_model"]["mus"]:
        lp, ll, lf = P.logp([v])
        exp = (-0.5 * v * v - 0.5 * math.log(2 * math.pi)) + (-0.5 * (1.0 - v) ** 2 - 0.5 * math.log(2 * math.pi))
        assert abs(lp[0] + ll[0] + lf[0] - exp) < 1e-9


def test_gen_refs_formulas():
    """Independent stdlib re-derivation, as /root/reference/tests/gen_refs.py:129-167 does."""
    lg = math.lgamma
    for k, r, x in [(3.0, 2.0, 1.5), (0.5, 1.0, 0.2), (7.5, 0.3, 20.0)]:
        assert abs(oracle.dist_logp(_ir.FC_GAMMA, [k, r], x) - (k * math.log(r) + (k - 1) * math.log(x) - r * x - lg(k))) < 1e-12
    for a, b, x in [(2.0, 3.0, 0.5), (0.5, 0.5, 0.01)]:
        e = (a - 1) * math.log(x) + (b - 1) * math.log1p(-x) - (lg(a) + lg(b) - lg(a + b))
        assert abs(oracle.dist_logp(_ir.FC_BETA, [a, b], x) - e) < 1e-12
    for nu, loc, s, x in [(3.0, 1.0, 2.0, 2.5), (10.0, 2.0, 0.5, 1.0)]:
        z = (x - loc) / s
        e = lg((nu + 1) / 2) - lg(nu / 2) - 0.5 * math.log(nu * math.pi) - math.log(s) - 0.5 * (nu + 1) * math.log1p(z * z / nu)
        assert abs(oracle.dist_logp(_ir.FC_STUDENT_T, [nu, loc, s], x) - e) < 1e-12


CONT = [(_ir.FC_NORMAL, [0.3, 1.7], 0.9), (_ir.FC_UNIFORM, [-2.0, 2.0], 0.5), (_ir.FC_LOGNORMAL, [0.1, 0.8], 1.9),
        (_ir.FC_EXPONENTIAL, [2.0], 1.3), (_ir.FC_BETA, [2.5, 3.5], 0.3), (_ir.FC_GAMMA, [3.0, 2.0], 1.5),
        (_ir.FC_STUDENT_T, [4.0, 0.5, 1.5], 2.0), (_ir.FC_CAUCHY, [0.2, 1.1], 1.5), (_ir.FC_LAPLACE, [0.1, 0.7], 1.2),
        (_ir.FC_WEIBULL, [1.8, 2.2], 1.4), (_ir.FC_CHI_SQUARED, [5.0], 3.3), (_ir.FC_INVERSE_GAMMA, [3.0, 2.0], 0.9)]


@pytest.mark.parametrize("dist,P,x", CONT)
def test_oracle_analytic_derivatives_vs_central_differences(dist, P, x):
    g = oracle.dist_eval(dist, P, x)
    h = 1e-6
    fd = (oracle.dist_logp(dist, P, x + h) - oracle.dist_logp(dist, P, x - h)) / (2 * h)
    assert abs(g[1] - fd) < 1e-6 * max(1.0, abs(fd))
    for j in range(len(P)):
        Pp, Pm = list(P), list(P)
        Pp[j] += h; Pm[j] -= h
        fd = (oracle.dist_logp(dist, Pp, x) - oracle.dist_logp(dist, Pm, x)) / (2 * h)
        assert abs(g[2 + j] - fd) < 1e-6 * max(1.0, abs(fd)), (dist, j)


def test_discrete_param_derivatives():
    for dist, P, xi, j in [(_ir.FC_BERNOULLI, [0.3], 1, 0), (_ir.FC_BERNOULLI, [0.3], 0, 0), (_ir.FC_POISSON, [2.5], 4, 0),
                           (_ir.FC_BINOMIAL, [20.0, 0.3], 7, 1), (_ir.FC_CATEGORICAL, [0.35], 0, 0), (_ir.FC_CATEGORICAL, [0.35], 1, 0)]:
        g = oracle.dist_eval(dist, P, xi=xi)
        h = 1e-6
        Pp, Pm = list(P), list(P)
        Pp[j] += h; Pm[j] -= h
        fd = (oracle.dist_logp(dist, Pp, xi=xi) - oracle.dist_logp(dist, Pm, xi=xi)) / (2 * h)
        assert abs(g[2 + j] - fd) < 1e-6 * max(1.0, abs(fd))


def test_digamma():
    assert abs(oracle.lib().fo_digamma(1.0) + 0.5772156649015329) < 1e-14
    assert abs(oracle.lib().fo_digamma(0.5) + 1.9635100260214235) < 1e-14
    assert abs(oracle.lib().fo_digamma(25.0) - 3.198742512851974) < 1e-13  # scipy.special.digamma
    assert abs(oracle.lib().fo_digamma(3.7) - 1.1671535393615113) < 1e-13


@pytest.mark.parametrize("name", ["linear_regression", "hierarchical_small", "mixture", "logistic_small", "poisson_small"])
def test_model_analytic_grad_vs_fd(name):
    rng = np.random.default_rng(1)
    if name == "linear_regression":
        m, _ = models.linear_regression()
    elif name == "hierarchical_small":
        m, _ = models.hierarchical(G=7, per_group=4)
    elif name == "mixture":
        m, _ = models.gaussian_mixture(n_obs=12)
    elif name == "logistic_small":
        m, _ = models.logistic_regression(N=200, D=5)
    else:
        m, _ = models.poisson_glm(N=300, D=3)
    P = oracle.OracleProgram(m)
    sites = P.sites()
    Cn = 4
    q = rng.normal(0, 0.5, (P.d_cont, Cn))
    for s, nm in enumerate(sites):
        if nm.startswith("sigma") or nm.startswith("theta"):
            q[s] = rng.uniform(0.5, 2.0, Cn)
        if nm == "pi1":
            q[s] = rng.uniform(0.2, 0.8, Cn)
    qd = rng.integers(0, 2, (P.d_disc, Cn)) if P.d_disc else None
    lp, ll, lf, g = P.logp(q, qd, grad=True)
    assert np.all(np.isfinite(lp + ll + lf))
    gfd, ok = P.grad_fd(q, qd, h=1e-6)
    assert ok.all()
    assert np.allclose(g, gfd, rtol=2e-5, atol=2e-5), np.abs(g - gfd).max()
    assert sites == sorted(sites)  # BTreeMap order


def test_sorted_site_order_is_lexicographic():
    m, _ = models.logistic_regression(N=10, D=12)
    P = oracle.OracleProgram(m)
    s = P.sites()
    assert s[:4] == ["beta#0", "beta#1", "beta#10", "beta#11"] and s[4] == "beta#2"


def test_duplicate_address_conflict():
    m = ModelBuilder()
    m.sample("mu", Normal(0.0, 1.0))
    m.sample("mu", Normal(0.0, 1.0))
    with pytest.raises(_ir.FugueError) as e:
        oracle.OracleProgram(m)
    assert e.value.code == _ir.FC_ERR_ADDRESS_CONFLICT
