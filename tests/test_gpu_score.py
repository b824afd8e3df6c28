"""GPU parity (through the C ABI): batched log-density and analytic gradient vs the CPU oracle.
Gate (BASELINE.json north_star): 1e-10 relative, f64."""
import numpy as np
import pytest
import oracle
from fugue_b200 import models
from fugue_b200.model import ModelBuilder, Normal, Gamma, Beta, Bernoulli, Poisson, StudentT, Exponential, dense, exp, sigmoid_clamped

pytestmark = pytest.mark.gpu
REL = 1e-10


def close(a, b, rel=REL, scale=None):
    a, b = np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64)
    both_inf = np.isinf(a) & np.isinf(b) & (np.sign(a) == np.sign(b))
    s = np.maximum(np.abs(a), np.abs(b)) if scale is None else np.maximum(np.maximum(np.abs(a), np.abs(b)), scale)
    with np.errstate(invalid="ignore"):
        ok = (np.abs(a - b) <= rel * s) | both_inf
    return ok


def init_q(P, rng, Cn, spread=0.5):
    q = rng.normal(0, spread, (P.d_cont, Cn))
    for s, nm in enumerate(P.sites()):
        if nm.startswith("sigma") or nm.startswith("theta"):
            q[s] = rng.uniform(0.5, 2.0, Cn)
        if nm == "pi1":
            q[s] = rng.uniform(0.2, 0.8, Cn)
    return q


def build(name):
    if name == "linear_regression":
        return models.linear_regression()[0]
    if name == "hierarchical":
        return models.hierarchical(G=41, per_group=7)[0]
    if name == "mixture":
        return models.gaussian_mixture(n_obs=80)[0]
    if name == "logistic":
        return models.logistic_regression(N=1000, D=3)[0]
    if name == "logistic_wide":
        return models.logistic_regression(N=2100, D=100)[0]
    if name == "logistic_odd":
        return models.logistic_regression(N=777, D=37)[0]
    if name == "poisson":
        return models.poisson_glm(N=1500, D=8)[0]
    raise KeyError(name)


@pytest.mark.parametrize("name,Cn", [("linear_regression", 1), ("linear_regression", 33), ("hierarchical", 64),
                                      ("mixture", 70), ("logistic", 64), ("logistic_wide", 200), ("logistic_odd", 5),
                                      ("poisson", 130)])
def test_logp_and_grad_match_oracle(name, Cn):
    from fugue_b200 import api
    rng = np.random.default_rng(7)
    m = build(name)
    O = oracle.OracleProgram(m)
    P = api.Program(m)
    assert P.sites == O.sites()
    assert (P.d_cont, P.d_disc) == (O.d_cont, O.d_disc)
    q = init_q(O, rng, Cn, spread=0.5 if "logistic" not in name else 0.3)
    qd = rng.integers(0, 2, (O.d_disc, Cn)) if O.d_disc else None
    lp, ll, lf, g = O.logp(q, qd, grad=True)
    dlp, dll, dlf = P.logp(q, qd)
    assert close(dlp, lp).all(), np.abs(dlp - lp).max()
    assert close(dll, ll).all(), (np.abs(dll - ll).max(), dll[:3], ll[:3])
    assert close(dlf, lf).all()
    tot, dg = P.logp_grad(q, qd)
    assert close(tot, lp + ll + lf).all()
    # gradient components: 1e-10 relative to the component, floored at 1e-10 x the chain's
    # largest component (cancelling sums cannot be better conditioned than their terms)
    scale = np.abs(g).max(axis=0, keepdims=True) * np.ones_like(g)
    bad = ~close(dg, g, scale=scale)
    assert not bad.any(), (np.argwhere(bad)[:5], dg[bad][:5], g[bad][:5])
    strict = close(dg, g)
    assert strict.mean() > 0.99


def test_prior_scale_points_with_active_clamp():
    """Prior draws beta ~ N(0,2) at D=100 put |eta| ~ 20: the sigmoid clamp (classification.rs:127)
    is active and the reference's ln(1-p) rounding is reproduced, not 'fixed' (SURVEY H4)."""
    from fugue_b200 import api
    rng = np.random.default_rng(3)
    m = models.logistic_regression(N=4000, D=100)[0]
    O, P = oracle.OracleProgram(m), api.Program(m)
    q = rng.normal(0, 2.0, (100, 64))
    lp, ll, lf, g = O.logp(q, grad=True)
    dlp, dll, dlf = P.logp(q)
    assert close(dll, ll).all(), np.abs((dll - ll) / ll).max()
    tot, dg = P.logp_grad(q)
    scale = np.abs(g).max(axis=0, keepdims=True) * np.ones_like(g)
    assert close(dg, g, scale=scale).all()


def test_invalid_parameters_give_neg_inf_not_nan():
    """sigma <= 0 reaches Normal::log_prob as -inf (distribution.rs:191-193); SURVEY H8."""
    from fugue_b200 import api
    m = models.linear_regression()[0]
    O, P = oracle.OracleProgram(m), api.Program(m)
    q = np.array([[0.1, 0.2], [-0.5, 0.0], [1.0, 1.0]])  # rows: intercept, sigma, slope
    lp, ll, lf = P.logp(q)
    olp, oll, olf = O.logp(q)
    assert np.isneginf(ll).all() and np.isneginf(oll).all()
    assert np.isneginf(lp).all() and np.isneginf(olp).all()  # Gamma prior at sigma <= 0


def test_generic_glm_links_and_dists():
    """Dense GLM plate through the generic path: Normal(X b + a, sigma), Poisson(exp(X b))."""
    from fugue_b200 import api
    rng = np.random.default_rng(11)
    N, D, Cn = 517, 6, 40
    X = rng.normal(size=(N, D))
    m = ModelBuilder()
    a = m.sample("a", Normal(0.0, 3.0))
    sigma = m.sample("sigma", Gamma(2.0, 2.0))
    b = m.plate_sample("b", D, Normal(0.0, 1.0))
    Xc = m.data(X)
    yc = m.data(rng.normal(size=N))
    m.plate_observe("y", Normal(a + dense(Xc, b), sigma), yc)
    kc = m.data(rng.poisson(2.0, N).astype(np.uint32))
    m.plate_observe("k", Poisson(exp(dense(Xc, b))), kc)
    O, P = oracle.OracleProgram(m), api.Program(m)
    q = rng.normal(0, 0.3, (O.d_cont, Cn))
    q[O.sites().index("sigma")] = rng.uniform(0.5, 2.0, Cn)
    lp, ll, lf, g = O.logp(q, grad=True)
    tot, dg = P.logp_grad(q)
    assert close(tot, lp + ll + lf).all()
    scale = np.abs(g).max(axis=0, keepdims=True) * np.ones_like(g)
    assert close(dg, g, scale=scale).all(), np.abs(dg - g).max()


def test_other_distributions_as_priors_and_observes():
    from fugue_b200 import api
    rng = np.random.default_rng(5)
    m = ModelBuilder()
    nu = m.sample("nu", Gamma(2.0, 0.1))
    loc = m.sample("loc", Normal(0.0, 2.0))
    sc = m.sample("sc", Exponential(1.0))
    pb = m.sample("pb", Beta(2.0, 3.0))
    yc = m.data(rng.standard_t(4, 200))
    m.plate_observe("y", StudentT(nu, loc, sc), yc)
    zc = m.data((rng.random(50) < 0.3))
    m.plate_observe("z", Bernoulli(pb), zc)
    m.observe("single", Normal(loc, 1.5), 0.7)
    m.factor(-1.25)
    O, P = oracle.OracleProgram(m), api.Program(m)
    Cn = 17
    q = np.zeros((4, Cn))
    s = O.sites()
    q[s.index("nu")] = rng.uniform(2, 9, Cn)
    q[s.index("loc")] = rng.normal(0, 1, Cn)
    q[s.index("sc")] = rng.uniform(0.5, 2, Cn)
    q[s.index("pb")] = rng.uniform(0.1, 0.9, Cn)
    lp, ll, lf, g = O.logp(q, grad=True)
    dlp, dll, dlf = P.logp(q)
    assert close(dlp, lp).all() and close(dll, ll).all() and close(dlf, lf).all()
    tot, dg = P.logp_grad(q)
    scale = np.abs(g).max(axis=0, keepdims=True) * np.ones_like(g)
    assert close(dg, g, scale=scale, rel=1e-9).all(), (dg - g)


def test_typed_errors():
    from fugue_b200 import api, _ir
    m = ModelBuilder()
    m.sample("mu", Normal(0.0, 1.0))
    m.sample("mu", Normal(0.0, 1.0))
    with pytest.raises(_ir.FugueError) as e:
        api.Program(m)
    assert e.value.code == _ir.FC_ERR_ADDRESS_CONFLICT
    m1 = models.linear_regression()[0]
    P = api.Program(m1)
    m2 = models.linear_regression(n_obs=4)[0]
    with pytest.raises(_ir.FugueError) as e:
        P.check_structure(m2)
    assert e.value.code == _ir.FC_ERR_UNEXPECTED_MODEL_STRUCTURE
    P.check_structure(models.linear_regression()[0])


def test_scaled_latent_terms_in_every_kernel_family():
    """FC_TERM_SCALED (c * latent) as a chain-constant part of scalar ops, generic / fast / CSR plates
    and of a dense GLM offset: value and gradient against the oracle."""
    from fugue_b200 import api
    rng = np.random.default_rng(3)
    N, G, D, Cn = 240, 12, 5, 37
    m = ModelBuilder()
    a = m.sample("a", Normal(0.0, 2.0))
    s = m.sample("s", Gamma(2.0, 2.0))
    m.sample("z", Normal(2.0 * a - 1.0, 0.5 * s + 0.1))                          # scalar op
    al = m.plate_sample("al", G, Normal(-0.5 * a, 1.0))                           # fast plate (sample), scaled prior mean
    gid = m.data(rng.integers(0, G, N).astype(np.uint32))
    xc = m.data(rng.normal(size=N))
    m.plate_observe("y", Normal(al[gid] + 1.5 * a, 0.25 * s + 0.2), m.data(rng.normal(size=N)))   # CSR plate, scaled sigma
    b = m.plate_sample("b", D, Normal(0.0, 1.0))
    X = m.data(rng.normal(size=(N, D)))
    m.plate_observe("k", Bernoulli(sigmoid_clamped(dense(X, b) + 0.7 * a, 1e-10, 1 - 1e-10)), m.data((rng.random(N) < 0.5).astype(np.uint8)))
    m.plate_observe("w", Normal(3.0 * a + a * xc, s), m.data(rng.normal(size=N)))  # LATENT_X + scaled latent
    O, P = oracle.OracleProgram(m), api.Program(m)
    q = rng.normal(0, 0.5, (O.d_cont, Cn))
    q[O.sites().index("s")] = rng.uniform(0.5, 2.0, Cn)
    olp, oll, olf, og = O.logp(q, grad=True)
    lp, ll, lf = P.logp(q)
    assert close(lp, olp).all() and close(ll, oll).all()
    tot, g = P.logp_grad(q)
    assert close(tot, (olp + oll) + olf).all()
    scale = np.abs(og).max(axis=0, keepdims=True) * np.ones_like(og) * 1e-4
    assert close(g, og, scale=scale).all(), np.abs(g - og).max()
