"""Text front-end (fugue_b200/dsl.py): the reference's own DSL tests
(/root/reference/crates/fugue-wasm/src/dsl.rs:1142-1329) restated against the op-list back end.
CPU tests cover parsing, static errors, lowering to plate ops and address parity; GPU tests run the
lowered models through the device samplers (coin posterior 9/14, regression slope sanity)."""
import math
import numpy as np
import pytest
from fugue_b200 import dsl, model as M
from fugue_b200._ir import FugueError

COIN = """
    let p <- sample(addr!("p"), Beta(2.0, 2.0));
    for i in 0..data.len() {
        observe(addr!("flip", i), Bernoulli(p), data[i]);
    }
    pure(p)
"""
REGRESSION = """
    let a <- sample(addr!("a"), Normal(0.0, 2.5));
    let b <- sample(addr!("b"), Normal(0.0, 2.5));
    for i in 0..x.len() {
        observe(addr!("y", i), Normal(a * x[i] + b, 0.8), y[i]);
    }
    pure(a)
"""
REG_DATA = '{"x": [-2, -1, 0, 1, 2], "y": [-2.1, -1.3, -0.4, 0.5, 1.2]}'

KIND_SAMPLE, KIND_OBSERVE, KIND_FACTOR, KIND_PLATE_SAMPLE, KIND_PLATE_OBSERVE, KIND_FACTOR_EXPR = 0, 1, 2, 3, 4, 5


def _kinds(cm):
    return [(op["kind"], op["name"], op["plate_len"]) for op in cm.builder.ops]


# ------------------------------------------------------------------------------------------------
# CPU: parse + lower
# ------------------------------------------------------------------------------------------------
def test_coin_lowers_to_one_plate_observe():
    """dsl.rs:1156-1187: the `for` body becomes ONE plate op with addresses flip#i."""
    cm = dsl.compile(COIN, "[1,0,1,1,0,1,1,0,1,1]")
    assert _kinds(cm) == [(KIND_SAMPLE, "p", 0), (KIND_PLATE_OBSERVE, "flip", 10)]
    assert cm.data_summary() == "data[10] "
    assert cm.take_warnings() == []
    assert M.addr("flip", 3) == "flip#3"


def test_regression_named_arrays_lower_to_affine_plate():
    """dsl.rs:1189-1213."""
    cm = dsl.compile(REGRESSION, REG_DATA)
    assert _kinds(cm) == [(KIND_SAMPLE, "a", 0), (KIND_SAMPLE, "b", 0), (KIND_PLATE_OBSERVE, "y", 5)]
    assert cm.data_summary() == "x[5] y[5] "


def test_indexed_sample_addresses():
    """dsl.rs:1215-1236: z#0..z#2 -- a plate of latents."""
    cm = dsl.compile('for i in 0..3 { let z <- sample(addr!("z", i), Normal(0.0, 1.0)); } pure(0.0)', "")
    assert _kinds(cm) == [(KIND_PLATE_SAMPLE, "z", 3)]


def test_rust_ish_sugar_parses():
    """dsl.rs:1238-1246: `Normal::new(..).unwrap()`."""
    cm = dsl.compile('let mu <- sample(addr!("mu"), Normal::new(0.0, 1.0).unwrap());'
                     'observe(addr!("y"), Normal::new(mu, 1.0).unwrap(), 0.5); pure(mu)', "")
    assert _kinds(cm) == [(KIND_SAMPLE, "mu", 0), (KIND_OBSERVE, "y", 0)]
    assert cm.builder.ops[1]["value"] == 0.5


def test_static_errors_are_caught():
    """dsl.rs:1248-1269."""
    for bad in ("pure(nope)",
                'let x <- sample(addr!("x"), Nope(1.0)); pure(x)',
                'let x <- sample(addr!("x"), Normal(1.0)); pure(x)',
                "let x = 1.0;"):
        with pytest.raises(dsl.DslError):
            dsl.compile(bad, "")
    with pytest.raises(dsl.DslError) as e:
        dsl.compile('let x <- sample(addr!("x") Normal(0,1)); pure(x)', "")
    assert "line" in str(e.value)


def test_line_numbers_in_errors():
    src = 'let a <- sample(addr!("a"), Normal(0.0, 1.0));\n\nlet b <- sample(addr!("b") Normal(0.0, 1.0));\npure(a)'
    with pytest.raises(dsl.DslError) as e:
        dsl.compile(src, "")
    assert str(e.value).startswith("line 3:")


def test_invalid_params_become_neg_inf_factor_and_warning():
    """dsl.rs:1271-1291: sigma=-1 must not crash; total weight is -inf and a warning is kept."""
    cm = dsl.compile('let mu <- sample(addr!("mu"), Normal(0.0, -1.0)); pure(mu)', "")
    ks = _kinds(cm)
    assert (KIND_FACTOR, "", 0) in ks
    fac = [op for op in cm.builder.ops if op["kind"] == KIND_FACTOR][0]
    assert fac["value"] == -math.inf
    w = cm.take_warnings()
    assert w and "sigma" in w[0]
    assert cm.take_warnings() == []  # take_ drains


def test_discrete_sites_sample_and_observe():
    """dsl.rs:1311-1328."""
    cm = dsl.compile('let k <- sample(addr!("k"), Poisson(4.0)); let z <- sample(addr!("z"), Categorical(0.3, 0.7));'
                     'observe(addr!("n"), Binomial(10, 0.5), 7); observe(addr!("flag"), Bernoulli(0.5), true); pure(k + z)', "")
    assert [k for k, _, _ in _kinds(cm)] == [KIND_SAMPLE, KIND_SAMPLE, KIND_OBSERVE, KIND_OBSERVE]
    assert cm.builder.ops[2]["value"] == 7.0 and cm.builder.ops[3]["value"] == 1.0
    assert cm.take_warnings() == []
    r = cm.return_value({"k": np.array([3.0, 5.0]), "z": np.array([1.0, 0.0])})
    assert np.array_equal(r, [4.0, 5.0])


def test_factor_and_math_functions():
    """dsl.rs:1293-1309: factor(-0.5 * pow(x - 1.0, 2.0)) is a log-weight computed from a latent -> a
    FACTOR_EXPR op (expression nodes evaluated per chain on the device); pure(sigmoid(x)) in (0,1)."""
    cm = dsl.compile('let x <- sample(addr!("x"), Normal(0.0, 1.0)); factor(-0.5 * pow(x - 1.0, 2.0)); pure(exp(x) / (1.0 + exp(x)))', "")
    assert [k for k, _, _ in _kinds(cm)] == [KIND_SAMPLE, KIND_FACTOR_EXPR]
    nodes = cm.builder.ops[1]["expr"].nodes
    assert nodes[-1][0] == 4 and any(n[0] == 15 for n in nodes)  # MUL at the root, a POW inside
    assert cm.builder.ops[1]["value"] == 1.0  # the playground's NaN -> -inf policy (dsl.rs:922-926)
    r = cm.return_value({"x": np.linspace(-3, 3, 7)})
    assert ((r > 0) & (r < 1)).all()
    assert abs(r[3] - 0.5) < 1e-15


def test_constant_factor_folds():
    cm = dsl.compile('let x <- sample(addr!("x"), Normal(0.0, 1.0)); factor(-0.5 * pow(2.0, 2.0)); pure(x)', "")
    fac = [op for op in cm.builder.ops if op["kind"] == KIND_FACTOR][0]
    assert fac["value"] == -2.0


def test_symbolic_let_and_unrolled_latent_factors():
    """A non-affine `let` stays symbolic (factor / pure only); a latent-dependent factor inside a loop
    is unrolled into one scalar factor per iteration (the data element is a constant there)."""
    cm = dsl.compile('let x <- sample(addr!("x"), Normal(0.0, 1.0)); let y = x * x; factor(-y);'
                     'for i in 0..d.len() { factor(-0.5 * pow(x - d[i], 2.0)); } pure(y)', '{"d": [1.0, 2.0]}')
    assert [k for k, _, _ in _kinds(cm)] == [KIND_SAMPLE, KIND_FACTOR_EXPR, KIND_FACTOR_EXPR, KIND_FACTOR_EXPR]
    assert np.array_equal(cm.return_value({"x": np.array([2.0, 3.0])}), [4.0, 9.0])
    with pytest.raises(FugueError) as e:  # ... but not as a distribution parameter
        dsl.compile('let x <- sample(addr!("x"), Normal(0.0, 1.0)); let y = x * x; observe(addr!("o"), Normal(y, 1.0), 0.3); pure(x)', "")
    assert e.value.code == 1000


def test_non_affine_latent_expression_is_a_typed_rejection():
    """No CPU fallback: what the op list cannot express is FC_ERR_UNSUPPORTED_OP (1000)."""
    with pytest.raises(FugueError) as e:
        dsl.compile('let x <- sample(addr!("x"), Normal(0.0, 1.0)); let y <- sample(addr!("y"), Normal(0.0, 1.0));'
                    'observe(addr!("o"), Normal(x * y, 1.0), 0.3); pure(x)', "")
    assert e.value.code == 1000


def test_constant_folding_and_comments():
    src = """
        // prior scale computed from constants
        let s = sqrt(4.0) * 1.25;
        let m <- sample(addr!("m"), Normal(0.0, s));
        pure(m)
    """
    cm = dsl.compile(src, "")
    d = cm.builder.ops[0]["dist"]
    assert isinstance(d, M.Normal)


def test_data_binding_forms():
    """dsl.rs:1055-1107: bare array binds `data`; object binds named arrays; empty string = no data."""
    assert dsl.compile("pure(data.len())", "[1,2,3]").return_value({}) == 3.0
    assert dsl.compile("pure(u.len() + v.len())", '{"u": [1], "v": [2, 3]}').return_value({}) == 3.0
    with pytest.raises(dsl.DslError):
        dsl.compile("pure(0.0)", '{"u": 3}')


# ------------------------------------------------------------------------------------------------
# GPU: run the lowered models through the device samplers
# ------------------------------------------------------------------------------------------------
@pytest.mark.gpu
def test_coin_posterior_mean_on_device():
    """dsl.rs:1176-1186: Beta(2+7, 2+3) posterior mean 9/14 through adaptive_mcmc_chain."""
    from fugue_b200 import api
    cm = dsl.compile(COIN, "[1,0,1,1,0,1,1,0,1,1]")
    draws, dd, lp, st = api.adaptive_mcmc_chain(cm, n_samples=500, n_warmup=500, n_chains=512, seed=11, thin=4)
    p = draws[:, 0, :]
    assert abs(p.mean() - 9.0 / 14.0) < 0.01
    assert abs(p.var() - (9.0 * 5.0) / (14.0 ** 2 * 15.0)) < 0.005
    r = cm.return_value({"p": p})
    assert r.shape == p.shape


@pytest.mark.gpu
def test_coin_density_matches_closed_form():
    """Scoring parity: log Beta(2,2)(p) + 7 ln p + 3 ln(1-p) at fixed points, 1e-10 relative."""
    from fugue_b200 import api
    cm = dsl.compile(COIN, "[1,0,1,1,0,1,1,0,1,1]")
    P = api.Program(cm)
    ps = np.array([[0.1, 0.35, 0.5, 0.9]])
    lp, ll, lf = P.logp(ps)
    assert np.allclose(lp, np.log(6.0 * ps[0] * (1 - ps[0])), rtol=1e-10, atol=0)
    assert np.allclose(ll, 7 * np.log(ps[0]) + 3 * np.log1p(-ps[0]), rtol=1e-10, atol=0)
    assert (lf == 0).all()


@pytest.mark.gpu
def test_regression_slope_on_device():
    """dsl.rs:1189-1213: slope sanity 0.4 < mean(a) < 1.4 (exact posterior mean ~0.83)."""
    from fugue_b200 import api
    cm = dsl.compile(REGRESSION, REG_DATA)
    draws, _, _ = api.hmc_chain(cm, n_samples=300, n_warmup=300, n_chains=256, seed=3)
    a = draws[:, 0, :]
    # conjugate: slope prior N(0,2.5), x centred so a and b decouple: precision 1/6.25 + sum(x^2)/0.64
    prec = 1 / 6.25 + 10 / 0.64
    mean = (np.dot([-2, -1, 0, 1, 2], [-2.1, -1.3, -0.4, 0.5, 1.2]) / 0.64) / prec
    assert abs(a.mean() - mean) < 0.02
    assert 0.4 < a.mean() < 1.4


@pytest.mark.gpu
def test_invalid_params_total_weight_is_neg_inf_on_device():
    from fugue_b200 import api
    cm = dsl.compile('let mu <- sample(addr!("mu"), Normal(0.0, -1.0)); pure(mu)', "")
    P = api.Program(cm)
    lp, ll, lf = P.logp(np.array([[0.3]]))
    assert (lp[0] + ll[0]) + lf[0] == -math.inf


@pytest.mark.gpu
def test_latent_factor_model_on_device():
    """x ~ N(0,1) with factor(-0.5 (x-1)^2): the target is N(0.5, 0.5) (dsl.rs:1293-1309 run through
    inference instead of one prior draw)."""
    from fugue_b200 import api
    cm = dsl.compile('let x <- sample(addr!("x"), Normal(0.0, 1.0)); factor(-0.5 * pow(x - 1.0, 2.0)); pure(exp(x) / (1.0 + exp(x)))', "")
    draws, _, _ = api.hmc_chain(cm, n_samples=300, n_warmup=200, n_chains=256, seed=5)
    x = draws[:, 0, :]
    assert abs(x.mean() - 0.5) < 0.02 and abs(x.var() - 0.5) < 0.03
    r = cm.return_value({"x": x})
    assert ((r > 0) & (r < 1)).all()


GRAMMAR_EXAMPLE = """
    let p <- sample(addr!("p"), Beta(2.0, 2.0));
    let mu = 2.0 * p - 1.0;
    for i in 0..y.len() {
        observe(addr!("y", i), Normal(mu, 0.8), y[i]);
    }
    factor(-0.5 * mu * mu);
    pure(p)
"""


def test_grammar_doc_example_lowers_with_a_scaled_latent():
    """The module's own grammar example (dsl.rs:11-19): `let mu = 2.0 * p - 1.0` is an AFFINE function of
    a latent and stays a parameter expression (a SCALED term + a constant), usable by the plate and by
    the latent-dependent factor."""
    import oracle
    cm = dsl.compile(GRAMMAR_EXAMPLE, '{"y": [0.1, 0.2, 0.3]}')
    assert [k for k, _, _ in _kinds(cm)] == [KIND_SAMPLE, KIND_PLATE_OBSERVE, KIND_FACTOR_EXPR]
    assert cm.builder.ops[1]["dist"].params[0].terms == [(8, 0, 0, 0, 2.0), (1, 0, 0, 0, -1.0)]  # SCALED(2, p), CONST(-1)
    O = oracle.OracleProgram(cm.builder)
    q = np.array([[0.3, 0.8]])
    lp, ll, lf, g = O.logp(q, grad=True)
    mu, y = 2 * q[0] - 1, np.array([0.1, 0.2, 0.3])
    assert np.allclose(ll, [(-0.5 * ((y - m) / 0.8) ** 2 - math.log(0.8) - 0.5 * math.log(2 * math.pi)).sum() for m in mu], rtol=1e-14)
    assert np.allclose(lf, -0.5 * mu * mu, rtol=1e-14)
    tot = lambda qq: sum(O.logp(qq))
    assert np.allclose(g[0], (tot(q + 1e-6) - tot(q - 1e-6)) / 2e-6, rtol=1e-6)


def test_affine_forms_of_latents():
    src = ('let a <- sample(addr!("a"), Normal(0.0, 1.0)); let b <- sample(addr!("b"), Normal(0.0, 1.0));'
           'let m = 0.5 * a - b / 4.0 + 1.5; let z <- sample(addr!("z"), Normal(-m, 1.0)); pure(z)')
    cm = dsl.compile(src, "")
    terms = cm.builder.ops[2]["dist"].params[0].terms
    assert terms == [(8, 0, 0, 0, -0.5), (8, 1, 0, 0, 0.25), (1, 0, 0, 0, -1.5)]


@pytest.mark.gpu
def test_grammar_doc_example_on_device():
    import oracle
    from fugue_b200 import api
    cm = dsl.compile(GRAMMAR_EXAMPLE, '{"y": [0.1, 0.2, 0.3, -0.4, 0.9]}')
    O, P = oracle.OracleProgram(cm.builder), api.Program(cm)
    q = np.linspace(0.05, 0.95, 19)[None, :]
    olp, oll, olf, og = O.logp(q, grad=True)
    lp, ll, lf = P.logp(q)
    for a, b in ((lp, olp), (ll, oll), (lf, olf)):
        assert np.all(np.abs(a - b) <= 1e-10 * np.maximum(np.abs(a), np.abs(b)))
    tot, g = P.logp_grad(q)
    assert np.all(np.abs(g - og) <= 1e-10 * np.maximum(np.abs(g), np.abs(og)) + 1e-13)
    draws, _, _, _ = api.adaptive_mcmc_chain(cm, n_samples=400, n_warmup=300, n_chains=256, seed=1, thin=4)
    p = draws[:, 0, :]
    assert ((p > 0) & (p < 1)).all()
    # posterior mean by quadrature of the same density
    grid = np.linspace(1e-4, 1 - 1e-4, 4001)[None, :]
    w = np.exp(sum(O.logp(grid)) - sum(O.logp(grid)).max())
    assert abs(p.mean() - (w * grid[0]).sum() / w.sum()) < 0.01
