"""C3 (BASELINE configs[2]: hierarchical normal model, HMC chains with R-hat / ESS) at a scaled shape, run
the way the full shape is: hmc_chain with adapt_mass, the draws left in HBM (fc_hmc_run_dev), diagnostics
on them in place (fc_diagnostics_dev; diagnostics.rs:240-304, mcmc_utils.rs:231-339). Checks: the
device-resident pair returns exactly what the host pair does; split-R-hat < 1.05 on every parameter;
posterior moments agree with chains driven by the ORACLE's hmc_transition (hmc.rs:419-473 restated)
within Monte-Carlo error. The full shape (10 000 groups, 1024 chains, 500 + 500) is
tools/bench_configs.py c3 -> profiles/."""
import numpy as np
import pytest
import oracle
from fugue_b200 import api, models

pytestmark = pytest.mark.gpu


def test_device_resident_run_and_diagnostics_equal_the_host_path():
    m = models.hierarchical(G=40, per_group=6)[0]
    P = api.Program(m)
    cfg = api.HMCConfig(n_leapfrog=8, adapt_mass=True)
    h_draws, h_lp, h_st = api.hmc_chain(P, n_samples=60, n_warmup=40, config=cfg, n_chains=96, seed=4)
    d_draws, d_lp, d_st = api.hmc_chain_dev(P, n_samples=60, n_warmup=40, config=cfg, n_chains=96, seed=4)
    assert d_draws.is_cuda and tuple(d_draws.shape) == h_draws.shape
    assert np.array_equal(d_draws.cpu().numpy(), h_draws) and np.array_equal(d_lp.cpu().numpy(), h_lp)
    assert np.array_equal(h_st["n_accept"], d_st["n_accept"]) and np.array_equal(h_st["step_size"], d_st["step_size"])
    a, b = api.diagnostics(h_draws), api.diagnostics_dev(d_draws)
    for k in ("mean", "std", "r_hat", "ess"):
        assert np.array_equal(a[k], b[k]), k
    # and against the oracle's diagnostics on a few parameters
    for k in (0, 3, 17):
        x = h_draws[:, k, :]
        assert abs(b["r_hat"][k] - oracle.split_rhat(x)) <= 1e-12 * b["r_hat"][k]
        assert abs(b["ess"][k] - oracle.ess_multichain(x)) <= 1e-9 * b["ess"][k]


def test_c3_scaled_converges_and_matches_oracle_chains():
    G, per = 1000, 15
    m, truth = models.hierarchical(G=G, per_group=per)
    P, O = api.Program(m), oracle.OracleProgram(m)
    d, Cn, W, S = P.d_cont, 256, 500, 500
    assert d == G + 4
    # starting points as the full-shape run takes them (tools/bench_configs.py c3): jittered group means relaxed
    # by an identity-mass pilot, so that the Welford window of adapt_mass (hmc.rs:868-893) sees the posterior
    pilot = api.HmcSession(P, n_chains=Cn, n_warmup=W, config=api.HMCConfig(n_leapfrog=16), seed=1, q0=models.hierarchical_inits(truth, P.sites, Cn, 0.5, 0))
    pilot.step(W)
    q0 = pilot.position()
    pilot.close()
    draws, lp, st = api.hmc_chain_dev(P, n_samples=S, n_warmup=W, config=api.HMCConfig(n_leapfrog=16, adapt_mass=True), n_chains=Cn, seed=11, q0=q0)
    dg = api.diagnostics_dev(draws)
    assert np.all(np.isfinite(dg["r_hat"])) and dg["r_hat"].max() < 1.05, (dg["r_hat"].max(), int(np.argmax(dg["r_hat"])))
    assert dg["ess"].min() > 0.01 * S * Cn and np.median(dg["ess"]) > 0.2 * S * Cn
    assert st["n_divergent"].sum() <= 0.02 * Cn * (W + S)
    # the sampler found the data-generating values (alpha within 5 posterior sd, globals close)
    names = P.sites
    mean = dict(zip(names, dg["mean"]))
    assert abs(mean["beta"] - 0.6) < 0.05 and abs(mean["sigma_y"] - 0.8) < 0.05 and abs(mean["sigma_alpha"] - 1.0) < 0.25
    # chains of the oracle's own transition started from device draws (stationary), under a fixed diagonal
    # metric (the pooled posterior variances: any fixed metric leaves the target invariant)
    x = draws[-1].cpu().numpy()      # [d][Cn]
    ljs = lp[-1].cpu().numpy()
    rng = np.random.default_rng(5)
    n_or, n_it, eps = 12, 60, 0.15
    m_inv = dg["std"] ** 2
    mass_sqrt = 1.0 / dg["std"]
    acc = []
    samples = np.zeros((n_or, n_it, d))
    for c in range(n_or):
        q, lj = x[:, c].copy(), float(ljs[c])
        for it in range(n_it):
            q, lj, info = O.hmc_transition(q, lj, rng.standard_normal(d), rng.random(), eps, 16, m_inv=m_inv, mass_sqrt=mass_sqrt)
            acc.append(info["accepted"])
            samples[c, it] = q
    assert np.mean(acc) > 0.5
    o_mean = samples.reshape(-1, d).mean(0)
    o_chain_means = samples.mean(1)                                   # [n_or][d]
    o_se = o_chain_means.std(0, ddof=1) / np.sqrt(n_or) + 1e-12       # between-chain standard error
    dev_se = dg["std"] / np.sqrt(np.maximum(dg["ess"], 1.0))
    z = (dg["mean"] - o_mean) / np.sqrt(o_se ** 2 + dev_se ** 2)
    # t-like with 11 degrees of freedom over 1004 parameters: allow 8, and the bulk within 3
    assert np.abs(z).max() < 8.0 and np.mean(np.abs(z) < 3.0) > 0.9, (np.abs(z).max(), np.mean(np.abs(z) < 3.0))
    o_std = samples.reshape(-1, d).std(0)
    assert np.median(np.abs(o_std / dg["std"] - 1.0)) < 0.15


@pytest.mark.parametrize("G,per,adapt", [(137, 7, False), (3, 1500, False), (64, 15, True)])
def test_fused_hierarchical_leapfrog_equals_generic_path(G, per, adapt):
    """fc_hier.cu (force of a group latent = prior term + its observations, kick and drift applied in the same
    kernel; two launches per leapfrog step) against the generic nine-launch path on the same seed: the same
    trajectories up to the order in which per-CTA partials of the GLOBAL rows are folded (1e-10 after several
    transitions; group rows of one step are bit-identical). Shapes: ragged against the 8-group batches, groups
    larger than the 1024-element shared-memory chunk, and a run through mass adaptation + the step-size search."""
    m, truth = models.hierarchical(G=G, per_group=per)
    out = []
    for hier in (1, 0):
        api.debug_set("hmc_hier", hier)
        try:
            P = api.Program(m)
            q0 = models.hierarchical_inits(truth, P.sites, 160, 0.1, 3)
            S = api.HmcSession(P, n_chains=160, n_warmup=8, config=api.HMCConfig(n_leapfrog=5, adapt_mass=adapt), seed=5, q0=q0)
            l0 = P.launch_count
            S.step(1)
            launches = P.launch_count - l0
            first = (S.position(), S.log_joint(), S.last_step_info()["accept_prob"], S.trace())
            S.step(11)   # through the end of warm-up (and the mass switch at transition 4 when adapt)
            out.append((first, S.position(), S.log_joint(), S.stats(), launches))
            S.close()
        finally:
            api.debug_set("hmc_hier", 1)
    (fa, qa, la, sa, na), (fb, qb, lb, sb, nb) = out
    assert 2 * na < nb, (na, nb)      # 2 launches per leapfrog step instead of 8 (+ 2 memsets)
    assert np.allclose(fa[2], fb[2], rtol=1e-9, atol=1e-12)
    same = (fa[0] == fb[0]).all(axis=0) | np.isclose(fa[2], fb[2], rtol=1e-9, atol=1e-12)
    assert np.allclose(fa[0], fb[0], rtol=1e-11, atol=1e-13) and np.allclose(fa[1], fb[1], rtol=1e-12)
    for x, y in zip(fa[3], fb[3]):
        assert np.allclose(x, y, rtol=1e-12, atol=1e-12)
    # several transitions later: the same chains unless an accept decision sat on the rounding noise
    agree = np.isclose(la, lb, rtol=1e-8)
    assert agree.mean() > 0.97, agree.mean()
    assert np.allclose(qa[:, agree], qb[:, agree], rtol=1e-7, atol=1e-9)
    assert np.array_equal(sa["n_accept"][agree], sb["n_accept"][agree])
    assert np.allclose(sa["step_size"][agree], sb["step_size"][agree], rtol=1e-7)
