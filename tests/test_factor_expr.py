"""factor(f(latents)) -- FC_OP_FACTOR_EXPR: Model::Factor fed by a bind closure (src/core/model.rs:97-104),
the expression evaluated per chain with its analytic gradient. CPU: the oracle's forward-mode
restatement against closed forms and central differences. GPU: device (reverse sweep) == oracle to
1e-10 on the accumulators and every gradient component, samplers target the factor-defined density."""
import math
import numpy as np
import pytest
import oracle
from fugue_b200.model import ModelBuilder, Normal, Gamma, Bernoulli, exp, ln, sqrt, abs_, floor, sin, cos, tanh, pow_, min_, max_
from fugue_b200._ir import FugueError


def _model():
    m = ModelBuilder()
    x = m.sample("x", Normal(0.0, 1.0))
    s = m.sample("s", Gamma(2.0, 1.0))
    m.factor(-0.5 * pow_(x - 1.0, 2.0))
    m.factor(tanh(x * s) + ln(s) / (1.0 + exp(-x)) - sqrt(s) * abs_(x) + min_(x, s) + max_(0.25 * x, -s) + sin(x) * cos(s) + 0.01 * floor(3.0 * s))
    m.factor(-1.25)
    m.factor(x / s - s ** 2.0 + (-x) * 0.5)
    m.plate_observe("y", Normal(x, s), m.data(np.array([0.3, -0.2, 1.1])))
    return m


def _closed(sv, xv):
    return (-0.5 * (xv - 1) ** 2 + np.tanh(xv * sv) + np.log(sv) / (1 + np.exp(-xv)) - np.sqrt(sv) * np.abs(xv) + np.minimum(xv, sv)
            + np.maximum(0.25 * xv, -sv) + np.sin(xv) * np.cos(sv) + 0.01 * np.floor(3.0 * sv) - 1.25 + xv / sv - sv ** 2 - 0.5 * xv)


Q = np.array([[0.3, 0.7, 1.9, 0.45], [1.2, -0.8, 0.1, -2.5]])  # rows in address order: s, x


def test_oracle_factor_expr_value_and_gradient():
    O = oracle.OracleProgram(_model())
    assert O.sites() == ["s", "x"]
    lp, ll, lf, g = O.logp(Q, grad=True)
    assert np.allclose(lf, _closed(Q[0], Q[1]), rtol=1e-13, atol=1e-13)
    tot = lambda q: sum(O.logp(q))
    h = 1e-6
    for r in range(2):
        qp, qm = Q.copy(), Q.copy()
        qp[r] += h
        qm[r] -= h
        assert np.allclose(g[r], (tot(qp) - tot(qm)) / (2 * h), rtol=2e-6, atol=1e-6)


def test_nan_policy_and_validation():
    m = ModelBuilder()
    x = m.sample("x", Normal(0.0, 1.0))
    m.factor(ln(x), nan_is_neg_inf=True)   # the playground's policy: NaN -> -inf (dsl.rs:922-926)
    O = oracle.OracleProgram(m)
    lp, ll, lf = O.logp(np.array([[-1.0, 2.0]]))
    assert lf[0] == -math.inf and abs(lf[1] - math.log(2.0)) < 1e-15
    m2 = ModelBuilder()
    x2 = m2.sample("x", Normal(0.0, 1.0))
    m2.factor(ln(x2))                      # core factor(): the value is added as it is
    lf2 = oracle.OracleProgram(m2).logp(np.array([[-1.0]]))[2]
    assert math.isnan(lf2[0])
    m3 = ModelBuilder()
    b = m3.sample("b", Bernoulli(0.5))
    with pytest.raises(FugueError) as e:
        m3.factor(b * 2.0)                 # discrete latents cannot enter a differentiable expression
    assert e.value.code == 1000
    big = ModelBuilder()
    z = big.sample("z", Normal(0.0, 1.0))
    acc = z
    with pytest.raises(FugueError):
        for _ in range(70):
            acc = acc * 1.01 + 0.5
        big.factor(acc)


@pytest.mark.gpu
def test_device_matches_oracle_value_and_gradient():
    from fugue_b200 import api
    m = _model()
    O, P = oracle.OracleProgram(m), api.Program(m)
    rng = np.random.default_rng(0)
    q = np.vstack([rng.uniform(0.2, 3.0, 64), rng.normal(0, 1.5, 64)])
    olp, oll, olf, og = O.logp(q, grad=True)
    lp, ll, lf = P.logp(q)
    for a, b in ((lp, olp), (ll, oll), (lf, olf)):
        assert np.all(np.abs(a - b) <= 1e-10 * np.maximum(np.abs(a), np.abs(b)) + 1e-300)
    tot, g = P.logp_grad(q)
    assert np.all(np.abs(tot - ((olp + oll) + olf)) <= 1e-10 * np.abs(tot))
    assert np.all(np.abs(g - og) <= 1e-10 * np.maximum(np.abs(g), np.abs(og)) + 1e-13)


@pytest.mark.gpu
def test_nan_policy_on_device():
    from fugue_b200 import api
    m = ModelBuilder()
    x = m.sample("x", Normal(0.0, 1.0))
    m.factor(ln(x), nan_is_neg_inf=True)
    lf = api.Program(m).logp(np.array([[-1.0, 2.0]]))[2]
    assert lf[0] == -math.inf and abs(lf[1] - math.log(2.0)) < 1e-15


@pytest.mark.gpu
def test_samplers_target_factor_defined_density():
    """x ~ N(0,1), factor(-0.5 (x-1)^2) => N(0.5, 0.5) by HMC (analytic force through the factor) and
    by single-site MH; SMC evidence = log N(1 | 0, 2) + log sqrt(2 pi) (the factor is unnormalised)."""
    from fugue_b200 import api
    m = ModelBuilder()
    x = m.sample("x", Normal(0.0, 1.0))
    m.factor(-0.5 * pow_(x - 1.0, 2.0))
    d, lp, st = api.hmc_chain(m, n_samples=300, n_warmup=200, n_chains=256, seed=1)
    assert abs(d.mean() - 0.5) < 0.02 and abs(d.var() - 0.5) < 0.03
    d2, _, _, _ = api.adaptive_mcmc_chain(m, n_samples=600, n_warmup=400, n_chains=512, seed=2, thin=4)
    assert abs(d2.mean() - 0.5) < 0.03 and abs(d2.var() - 0.5) < 0.04
    r = api.adaptive_smc(m, 1 << 16, api.SMCConfig(rejuvenation_steps=2), seed=3)
    want = -0.25 - 0.5 * math.log(2.0)  # log E_{x~N(0,1)} exp(-0.5 (x-1)^2)
    assert abs(r.log_evidence - want) < 0.02
