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


def test_syntax_module_tokens_precedence_and_messages():
    """fugue_b200/dsl_syntax.py on its own: tokens of `0..n`, exponent numbers, associativity and binding
    powers, postfix forms, and the wording of the static errors (dsl.rs:1248-1269 behaviour)."""
    from fugue_b200 import dsl_syntax as sx
    toks = sx.tokenize("for i in 0..xs.len() { } // tail\n1.5e-3 2E4 7")
    assert [(t.kind, t.value) for t in toks[:9]] == [("ident", "for"), ("ident", "i"), ("ident", "in"), ("num", 0), ("sym", ".."),
                                                      ("ident", "xs"), ("sym", "."), ("ident", "len"), ("sym", "(")]
    assert [t.value for t in toks if t.kind == "num"] == [0, 1.5e-3, 2e4, 7] and isinstance(toks[-2].value, int)
    e = sx.Parser("a - b - c * -d[i + 1] / xs.len()").p_expr()
    assert e == ("bin", "-", ("bin", "-", ("var", "a"), ("var", "b")),
                 ("bin", "/", ("bin", "*", ("var", "c"), ("neg", ("index", ("var", "d"), ("bin", "+", ("var", "i"), ("int", 1))))), ("len", ("var", "xs"))))
    assert sx.Parser("pow(2.0, (x))").p_expr() == ("call", "pow", [("num", 2.0), ("var", "x")])
    stmts, ret = sx.parse_program('for i in 0..3 { let t = i; factor(t); }\nlet z <- sample(addr!("z", 2), Categorical(0.2, 0.3, 0.5)); pure(z);')
    assert stmts[0][0] == "for" and stmts[0][4][0] == ("let", "t", ("var", "i")) and stmts[1] == ("sample", "z", ("z", ("int", 2)), ("Categorical", [("num", 0.2), ("num", 0.3), ("num", 0.5)]))
    for text, msg in (('let x <- sample(addr!("x") Normal(0,1)); pure(x)', 'line 1: expected `,`, found `Normal`'),
                      ("let x <- draw(1); pure(x)", "line 1: expected `sample`, found `draw`"),
                      ("for i 0..3 { } pure(1)", "line 1: expected `in`, found `0`"),
                      ("for i in 0..3 { factor(1.0);", "line 1: expected `}`, found end of input"),
                      ('observe(addr!(x), Normal(0,1), 1.0); pure(1)', "line 1: expected a string literal address name, found `x`"),
                      ("pure(1) pure(2)", "line 1: expected end of input after `pure(..)`, found `pure`"),
                      ("\n\nwhile", "line 3: expected a statement (`let`, `observe`, `factor`, `for`, or `pure`), found `while`"),
                      ("let x = ;", "line 1: expected an expression, found `;`"),
                      ("let x = 1.0;", "model must end with `pure(<expr>)`"),
                      ('let s = "abc', "unterminated string at byte 8"),
                      ("let s = 1 # 2;", "unexpected character `#` at byte 10"),
                      ("pure(sqrt(1, 2))", "`sqrt` takes 1 argument(s)"),
                      ("pure(nope(1))", "unknown function `nope`"),
                      ('let x <- sample(addr!("x"), Normal::with(0, 1)); pure(x)', "unknown distribution constructor `Normal::with`"),
                      ('let x <- sample(addr!("x"), StudentT(1.0)); pure(x)', "`StudentT` takes 3 argument(s), got 1")):
        with pytest.raises(dsl.DslError) as err:
            sx.parse_program(text)
        assert str(err.value) == msg, (text, str(err.value))
    # scoping: a loop's variable and its lets end with the loop
    stmts, ret = sx.parse_program("for i in 0..2 { let t = i; } pure(t)")
    scope = []
    sx.check_stmts(stmts, scope)
    with pytest.raises(dsl.DslError, match="unknown variable `t`"):
        sx.check_expr(ret, scope)


def test_k_way_categorical_lowers_to_a_probability_column():
    """Categorical(p0, .., pK-1) with constant probabilities (distribution.rs:679-713) is K-way on the device
    (fc_op.probs_col); the latent-dependent form stays two-way; invalid vectors are ErrorCode 102."""
    cm = dsl.compile('let z <- sample(addr!("z"), Categorical(0.2, 0.3, 0.5)); observe(addr!("c"), Categorical(0.1, 0.2, 0.3, 0.4), 3); pure(z)', "")
    ops = cm.builder.ops
    assert [k for k, _, _ in _kinds(cm)] == [KIND_SAMPLE, KIND_OBSERVE]
    assert np.allclose(ops[0]["dist"].probs, [0.2, 0.3, 0.5]) and np.allclose(ops[1]["dist"].probs, [0.1, 0.2, 0.3, 0.4]) and ops[1]["value"] == 3.0
    bad = dsl.compile('let z <- sample(addr!("z"), Categorical(0.2, 0.3, 0.6)); pure(z)', "")  # soft error (dsl.rs:926-944)
    assert [op["value"] for op in bad.builder.ops if op["kind"] == KIND_FACTOR] == [-math.inf] and "sum to 1" in bad.take_warnings()[0]
    with pytest.raises(FugueError) as e:
        dsl.compile('let p <- sample(addr!("p"), Beta(1.0, 1.0)); let z <- sample(addr!("z"), Categorical(p, 0.5, 0.5)); pure(z)', "")
    assert e.value.code == 1000


@pytest.mark.gpu
def test_k_way_categorical_density_on_device():
    from fugue_b200 import api
    cm = dsl.compile('let z <- sample(addr!("z"), Categorical(0.2, 0.3, 0.5)); pure(z)', "")
    P = api.Program(cm)
    lp = P.logp(np.zeros((0, 4)), np.array([[0, 1, 2, 3]]))[0]
    assert np.allclose(lp[:3], np.log([0.2, 0.3, 0.5]), rtol=1e-14) and lp[3] == -math.inf
