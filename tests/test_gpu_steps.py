"""Step-level parity of the single-site MH step and of the tempered SMC move against the oracle's
restatement (oracle/fo_smc.hpp: mh.rs:128-295, 698-744; smc.rs:631-688) ON IDENTICAL RANDOMNESS: the
device draws from Philox4x32-10 keyed by (seed, chain); the test regenerates the same site pick, normal
draw and accept uniform (oracle.philox_uniforms) and feeds them to the oracle's step, chain by chain.
Also normalize_particles and the whole adaptive_smc ladder against the oracle's (distributional)."""
import math
import numpy as np
import pytest
import oracle
from fugue_b200 import api, models, model as M

pytestmark = pytest.mark.gpu


def _mh_model():
    rng = np.random.default_rng(3)
    m = M.ModelBuilder()
    mu = m.sample("mu", M.Normal(0.0, 3.0))          # Gaussian walk
    sg = m.sample("sigma", M.Gamma(2.0, 1.0))        # positive support -> log-space walk (mh.rs:351)
    m.sample("flag", M.Bernoulli(0.4))               # bool -> flip
    m.sample("count", M.Poisson(3.0))                # u64 -> discrete walk reflected about -1/2
    m.plate_observe("y", M.Normal(mu, sg), m.data(rng.normal(1.0, 1.5, 25)))
    return m


def _draws(seed, chain, tag):
    u = oracle.philox_uniforms(seed, chain, tag, 4)  # blocks (0, tag) and (1, tag)
    return u[0], u[1], u[2], u[3]


def test_single_site_mh_step_matches_oracle_on_identical_draws():
    m = _mh_model()
    P, O = api.Program(m), oracle.OracleProgram(m)
    Cn, seed = 384, 17
    S = api.MhSession(P, n_chains=Cn, n_warmup=0, seed=seed)
    n_sites = O.n_sites()
    assert n_sites == S.n_sites == 4
    kinds_seen = set()
    for it in range(6):
        g0 = S.get()
        S.step(1)
        g1 = S.get()
        n_acc = 0
        for c in range(Cn):
            u_site, u_acc, u1, u2 = _draws(seed, c, it)
            g = min(int(u_site * n_sites), n_sites - 1)
            z = math.sqrt(-2.0 * math.log(max(u1, 1e-10))) * math.cos(2.0 * math.pi * u2)
            kinds_seen.add((g, O.mh_kind(g0["q"][:, c], g0["qd"][:, c], g)))
            q, qd, acc, la, lw = O.mh_step(g0["q"][:, c], g0["qd"][:, c], g0["log_joint"][c], g, z, u_acc, scale=1.0)
            margin = abs(la - math.log(u_acc)) if la < 0 else abs(la)
            if margin < 1e-9:
                continue  # the decision sits on the rounding noise of the two scorings
            assert np.array_equal(qd, g1["qd"][:, c]), (it, c, g)
            assert np.allclose(q, g1["q"][:, c], rtol=1e-13, atol=0.0), (it, c, g, q, g1["q"][:, c])
            assert abs(lw - g1["log_joint"][c]) <= 1e-10 * max(1.0, abs(lw))
            n_acc += acc
        assert 0 < n_acc < Cn
    # sites in address order: count, flag, mu, sigma -> kinds -1 (discrete), -1, Gaussian 0, LogSpace 1
    assert {(2, 0), (3, 1)} <= kinds_seen and (0, -1) in kinds_seen and (1, -1) in kinds_seen


@pytest.mark.parametrize("generic", [0, 1])
def test_tempered_move_matches_oracle_on_identical_draws(generic):
    api.debug_set("smc_generic", generic)
    try:
        m = models.gaussian_mixture(n_obs=16)[0]
        P, O = api.Program(m), oracle.OracleProgram(m)
        n, seed, beta = 512, 23, 0.37
        S = api.SmcSession(P, n, api.SMCConfig(rejuvenation_steps=1), seed)
        assert S.counters()["fused"] == (not generic)
        eng_seed = seed ^ 0x5bd1e995
        d = P.d_cont
        for mv in range(3):
            r0 = S.result()
            S.rejuvenate_particles(beta, 1)
            r1 = S.result()
            ll1 = S.prior_weights().weights()[0]  # log_weight := particle_log_likelihood of the kept state
            n_acc = 0
            for c in range(n):
                u = oracle.philox_uniforms(eng_seed, c, mv, 4)
                u_site, u_acc = u[0], u[1]
                z = math.sqrt(-2.0 * math.log(1.0 - u[2])) * math.cos(2.0 * math.pi * u[3])  # Philox::normal2, first branch
                k = min(int(u_site * d), d - 1)
                scale = 1.0 if mv == 0 else None
                if scale is None:
                    break  # later sweeps use the adapted shared scale: covered by the accept-rate check below
                q, acc, la, lp3 = O.tempered_step(r0.values[:, c], r0.dvalues[:, c], k, beta, scale, z, u_acc)
                margin = abs(la - math.log(u_acc)) if la < 0 else abs(la)
                if not (margin > 1e-9):
                    continue
                assert np.allclose(q, r1.values[:, c], rtol=1e-13, atol=0.0), (c, k, q, r1.values[:, c])
                assert abs((lp3[1] + lp3[2]) - ll1[c]) <= 1e-10 * max(1.0, abs(ll1[c])) or (math.isinf(ll1[c]) and ll1[c] == lp3[1] + lp3[2])
                n_acc += acc
            if mv == 0:
                assert 0 < n_acc < n
            assert np.array_equal(r0.dvalues, r1.dvalues)  # discrete sites never move in this kernel (smc.rs:639-645)
        S.close()
    finally:
        api.debug_set("smc_generic", 0)


def test_normalize_particles_matches_oracle():
    """normalize_particles (smc.rs:719-753): exp(lw - lse); all -inf -> uniform."""
    m = M.ModelBuilder()
    s = m.sample("s", M.Normal(0.0, 1.0))
    m.observe("y", M.Normal(0.5, s), 0.3)  # sigma <= 0 for half of the prior draws: log-weight -inf
    S = api.smc_prior_particles(m, 5000, seed=2)
    lw, w = S.weights()
    ow = oracle.normalize_particles(lw)
    assert np.any(np.isinf(lw)) and np.allclose(w, ow, rtol=1e-12, atol=0.0) and abs(w.sum() - 1.0) < 1e-12
    ln, ess = S.normalize()
    assert abs(ln - oracle.log_sum_exp(lw)) <= 1e-12 * abs(ln) and abs(ess - oracle.ess(ow)) <= 1e-10 * ess
    S.close()
    m2 = M.ModelBuilder()
    m2.sample("s", M.Normal(0.0, 1.0))
    m2.observe("y", M.Uniform(5.0, 6.0), 0.0)  # outside the support for every particle
    S2 = api.smc_prior_particles(m2, 1000, seed=2)
    lw2, w2 = S2.weights()
    assert np.all(np.isinf(lw2)) and np.allclose(w2, 1e-3) and np.allclose(oracle.normalize_particles(lw2), 1e-3)
    S2.close()


def test_adaptive_smc_ladder_agrees_with_the_oracles():
    """The whole ladder (smc.rs:455-581), device vs the oracle's own adaptive_smc (different generators:
    agreement within Monte-Carlo error) on the conjugate model of tests/f_smc_smc.rs; same number of
    tempering steps, evidence and moments against the closed form."""
    Y = [1.5, 2.0, 2.5, 1.8, 2.2]
    m = M.ModelBuilder()
    mu = m.sample("mu", M.Normal(0.0, 2.0))
    m.plate_observe("y", M.Normal(mu, 1.0), m.data(np.array(Y)))
    O = oracle.OracleProgram(m)
    n = 20000
    o = O.adaptive_smc(n, rejuvenation_steps=2, seed=5)
    d = api.adaptive_smc(m, n, api.SMCConfig(rejuvenation_steps=2), seed=5)
    cov = np.eye(5) + 4.0 * np.ones((5, 5))
    y = np.array(Y)
    log_z = -0.5 * (y @ np.linalg.solve(cov, y)) - 0.5 * np.linalg.slogdet(cov)[1] - 2.5 * math.log(2 * math.pi)
    assert o["steps"] == d.n_temper_steps
    assert abs(o["log_evidence"] - log_z) < 0.04 and abs(d.log_evidence - log_z) < 0.04
    for w, x in ((o["weight"], o["values"][0]), (d.weight, d.values[0])):
        mean = (w * x).sum()
        var = (w * (x - mean) ** 2).sum()
        assert abs(mean - 1.90476) < 0.03 and abs(var - 0.190476) < 0.03
    # pure importance sampling (rejuvenation_steps == 0, smc.rs:484-493): one reweight, no resample
    o0 = O.adaptive_smc(n, rejuvenation_steps=0, seed=6)
    d0 = api.adaptive_smc(m, n, api.SMCConfig(rejuvenation_steps=0), seed=6)
    assert o0["steps"] == d0.n_temper_steps == 1 and abs(o0["log_evidence"] - d0.log_evidence) < 0.1


def test_mixture_sufficient_statistics_equal_term_by_term_scoring():
    """The fused SMC kernels score a mixture plate (z#i two-valued, x#i ~ Normal(mu[z_i], sigma[z_i])) from the
    per-particle, per-component sufficient statistics kept in the record (fc_smc.cu: mix_stats). A/B against the
    term-by-term loop (fc_debug_set("smc_stats", 0)) on the same seed: same population, accumulators to 1e-12
    relative, and the same moves -- through resampling (statistics are cloned with the particle) and 3 sweeps."""
    m = models.gaussian_mixture(n_obs=37)[0]   # not a multiple of 16: the vector loop's tail as well
    out = []
    # by-value program + 128-byte records (z as bits); interpreter on the same records; 192-byte; 144-byte term by term
    for stats, zbits, comp in ((1, 1, 1), (1, 1, 0), (1, 0, 1), (0, 0, 0)):
        api.debug_set("smc_stats", stats)
        api.debug_set("smc_zbits", zbits)
        api.debug_set("smc_compiled", comp)
        try:
            S = api.SmcSession(api.Program(m), 4096, api.SMCConfig(rejuvenation_steps=1), 31)
            assert S.counters()["fused"]
            r0 = S.result()
            lw0 = S.prior_weights().weights()[0]
            S.reweight(0.2)
            idx = S.resample_particles(api.ResamplingMethod.Systematic)
            for _ in range(3):
                S.rejuvenate_particles(0.2, 1)
            r1 = S.result()
            lw1 = S.prior_weights().weights()[0]
            out.append((r0.values, r0.dvalues, lw0, np.asarray(idx), r1.values, r1.dvalues, lw1))
            S.close()
        finally:
            api.debug_set("smc_stats", 1)
            api.debug_set("smc_zbits", 1)
            api.debug_set("smc_compiled", 1)
    b = out[3]
    for a in out[:3]:
        assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1])          # the same prior population
        assert set(np.unique(a[1])) == {0, 1}
        fin = np.isfinite(b[2])
        assert np.array_equal(np.isfinite(a[2]), fin) and np.allclose(a[2][fin], b[2][fin], rtol=1e-12, atol=0.0)
        assert np.array_equal(a[3], b[3])                                          # identical ancestors
        same = np.all(a[4] == b[4], axis=0)                                        # a decision on the rounding noise may differ
        assert same.mean() > 0.999 and np.array_equal(a[5], b[5])
        fin = np.isfinite(b[6]) & same
        assert np.allclose(a[6][fin], b[6][fin], rtol=1e-12, atol=0.0)
    for o in out[1:3]:   # compiled vs interpreted, bits vs bytes: the same arithmetic, bit for bit
        assert np.array_equal(out[0][4], o[4]) and np.array_equal(out[0][6], o[6])
