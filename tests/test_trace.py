"""Trace materialisation and the Vec<Vec<Trace>>-shaped diagnostics front (fugue_b200/trace.py).
Reference: src/runtime/trace.rs:168-343 (typed getters, error codes), src/inference/diagnostics.rs
:331-451 (quantile rule, table format)."""
import math
import numpy as np
import pytest
from fugue_b200 import trace as T, model as M
from fugue_b200.model import ModelBuilder, Normal, Gamma, Bernoulli, Poisson, Categorical
from fugue_b200._ir import FugueError


def _model():
    m = ModelBuilder()
    mu = m.sample("mu", Normal(0.0, 2.0))
    m.sample("flag", Bernoulli(0.3))
    m.sample("k", Poisson(4.0))
    m.sample("z", Categorical(0.25))
    m.plate_sample("alpha", 12, Normal(mu, 1.0))
    m.plate_observe("y", Normal(mu, 1.0), m.data(np.array([1.5, 2.0, 2.5])))
    return m


def test_site_table_is_btreemap_order_with_types():
    cont, disc = T.site_table(_model())
    assert [a for a, _ in cont] == sorted([f"alpha#{i}" for i in range(12)]) + ["mu"]
    assert cont[2][0] == "alpha#10" and cont[3][0] == "alpha#11" and cont[4][0] == "alpha#2"  # string order
    assert disc == [("flag", "bool"), ("k", "u64"), ("z", "usize")]


def test_traces_from_draws_types_and_getters():
    m = _model()
    S, Cn = 3, 2
    rng = np.random.default_rng(0)
    draws = rng.normal(size=(S, 13, Cn))
    dd = np.zeros((S, 3, Cn), dtype=np.int64)
    dd[:, 0, :] = 1; dd[:, 1, :] = 7; dd[:, 2, :] = 1
    chains = T.traces_from_draws(m, draws, dd, score=False)
    assert len(chains) == Cn and len(chains[0]) == S
    t = chains[1][2]
    assert list(t.choices) == sorted(t.choices)
    assert t.get_f64("mu") == draws[2, 12, 1]
    assert t.get_f64("alpha#10") == draws[2, 2, 1]
    assert t.get_bool("flag") is True and t.get_u64("k") == 7 and t.get_usize("z") == 1
    assert t.get_f64("flag") is None and t.get_i64("k") is None and t.get_f64("nope") is None
    with pytest.raises(FugueError) as e:
        t.get_f64_result("nope")
    assert e.value.code == 500
    with pytest.raises(FugueError) as e:
        t.get_f64_result("flag")
    assert e.value.code == 600
    assert t.get_u64_result("k") == 7


def test_total_log_weight_order():
    t = T.Trace({}, 1e16, 1.0, -1e16)
    assert t.total_log_weight() == (1e16 + 1.0) + -1e16


def _chains_from(x):  # x [S][C] -> chains of traces with one site "p"
    return [[T.Trace({"p": T.Choice("p", float(v), "f64")}) for v in x[:, c]] for c in range(x.shape[1])]


def test_quantile_rule_is_round_of_len_minus_one_times_p():
    """diagnostics.rs:361-372: idx = round((len-1)*p), no interpolation."""
    x = np.arange(10.0).reshape(10, 1)
    per = [T.extract_f64_values(ch, "p") for ch in _chains_from(x)]
    assert per == [list(map(float, range(10)))]
    srt = sorted(per[0])
    for name, p, want in (("2.5%", 0.025, 0.0), ("25%", 0.25, 2.0), ("50%", 0.5, 5.0), ("75%", 0.75, 7.0), ("97.5%", 0.975, 9.0)):
        assert srt[T._rust_round(9 * p)] == want


def test_empty_parameter_summary():
    s = T.summarize_f64_parameter([[T.Trace()]], "missing")
    assert math.isnan(s["mean"]) and math.isnan(s["r_hat"]) and s["ess"] == 0.0 and s["quantiles"] == {}
    assert T.format_diagnostics([]) == "No chains or empty chains provided\n"


@pytest.mark.gpu
def test_summary_matches_oracle_and_table_format():
    import oracle
    rng = np.random.default_rng(5)
    x = rng.normal(1.0, 2.0, size=(200, 4))
    chains = _chains_from(x)
    s = T.summarize_f64_parameter(chains, "p")
    allv = x.T.reshape(-1)
    assert abs(s["mean"] - allv.mean()) < 1e-12
    assert abs(s["std"] - allv.std(ddof=1)) < 1e-12
    srt = np.sort(allv)
    assert s["quantiles"]["50%"] == srt[int(math.floor(799 * 0.5 + 0.5))]
    assert abs(s["r_hat"] - oracle.split_rhat(x)) < 1e-12
    assert abs(s["ess"] - oracle.ess_multichain(x)) < 1e-9 * max(1.0, s["ess"])
    txt = T.format_diagnostics(chains)
    lines = txt.splitlines()
    assert lines[0] == "MCMC Diagnostics:"
    assert lines[1] == "Parameter           Mean      Std     2.5%      50%    97.5%    R-hat      ESS"
    assert lines[2] == "-" * 80
    assert lines[3].startswith("p" + " " * 14 + " ") and len(lines[3]) == 15 + 7 * 9
    assert "Convergence Assessment:" in txt and "Average R-hat:" in txt


@pytest.mark.gpu
def test_materialised_traces_carry_device_scored_accumulators():
    from fugue_b200 import api
    m = ModelBuilder()
    mu = m.sample("mu", Normal(0.0, 2.0))
    m.plate_observe("y", Normal(mu, 1.0), m.data(np.array([1.5, 2.0, 2.5, 1.8, 2.2])))
    draws, lp, st = api.hmc_chain(m, n_samples=20, n_warmup=20, n_chains=4, seed=1)
    chains = T.traces_from_draws(m, draws)
    assert len(chains) == 4 and len(chains[0]) == 20
    for c in range(4):
        for s in range(20):
            t = chains[c][s]
            v = t.get_f64("mu")
            assert v == draws[s, 0, c]
            want_prior = -0.5 * (v / 2.0) ** 2 - math.log(2.0) - 0.5 * math.log(2 * math.pi)
            assert abs(t.log_prior - want_prior) <= 1e-10 * abs(want_prior)
            assert abs(t.total_log_weight() - lp[s, c]) <= 1e-10 * abs(lp[s, c])
            assert t.log_factors == 0.0
