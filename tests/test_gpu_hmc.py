"""GPU HMC: transition parity with the CPU oracle on identical randomness, step-size search,
dual averaging / mass adaptation, and posterior moments within Monte-Carlo error (the
reference's own statistical templates: tests/f_hmc_posterior.rs, src/inference/hmc.rs:960-1020)."""
import numpy as np
import pytest
import oracle
from fugue_b200 import models
from fugue_b200.model import ModelBuilder, Normal, Gamma

pytestmark = pytest.mark.gpu


def _init_q(O, rng, Cn):
    q = rng.normal(0, 0.4, (O.d_cont, Cn))
    for s, nm in enumerate(O.sites()):
        if nm.startswith("sigma"):
            q[s] = rng.uniform(0.6, 1.6, Cn)
    return q


@pytest.mark.parametrize("name", ["linear_regression", "logistic", "hierarchical"])
def test_transition_matches_oracle_on_same_randomness(name):
    from fugue_b200 import api
    rng = np.random.default_rng(21)
    if name == "linear_regression":
        m = models.linear_regression()[0]
    elif name == "logistic":
        m = models.logistic_regression(N=600, D=5)[0]
    else:
        m = models.hierarchical(G=9, per_group=5)[0]
    O = oracle.OracleProgram(m)
    Cn, L, eps = 24, 7, 0.02
    q0 = _init_q(O, rng, Cn)
    S = api.HmcSession(m, n_chains=Cn, n_warmup=0, config=api.HMCConfig(n_leapfrog=L, init_step_size=eps), seed=5, q0=q0)
    lj0 = S.log_joint()
    olp = O.logp(q0)
    assert np.allclose(lj0, olp[0] + olp[1] + olp[2], rtol=1e-12)
    for it in range(3):
        qc = S.position()
        ljc = S.log_joint()
        z = rng.standard_normal((O.d_cont, Cn))
        u = rng.random(Cn)
        S.set_explicit_randomness(z, u)
        S.step(1)
        qn, info, ljn = S.position(), S.last_step_info(), S.log_joint()
        for c in range(Cn):
            oq, olj, oinfo = O.hmc_transition(qc[:, c], ljc[c], z[:, c], u[c], eps, L)
            assert oinfo["divergent"] == info["divergent"][c]
            assert abs(oinfo["accept_prob"] - info["accept_prob"][c]) <= 1e-8 * max(1.0, oinfo["accept_prob"])
            if abs(u[c] - oinfo["accept_prob"]) > 1e-6:  # not a knife-edge accept
                assert oinfo["accepted"] == info["accepted"][c]
                assert np.allclose(oq, qn[:, c], rtol=1e-9, atol=1e-11)
                assert abs(olj - ljn[c]) <= 1e-9 * abs(olj)
    assert S.grad_evals() == 3 * L  # cached force: L (not L+1) gradient passes per trajectory
    S.close()


def test_trajectory_leaving_the_support_is_divergent_as_in_the_reference():
    """A bounded site and a large step: interior leapfrog points leave the support. The reference's finite-difference
    force is NaN there, so the trajectory is divergent and rejected (hmc.rs:304-329, 384-386); the analytic force
    must say the same (it used to be a finite zero, letting the trajectory glide on and possibly re-enter)."""
    from fugue_b200 import api
    rng = np.random.default_rng(8)
    m = ModelBuilder()
    s = m.sample("sigma", Gamma(2.0, 2.0))
    mu = m.sample("mu", Normal(0.0, 1.0))
    m.plate_observe("y", Normal(mu, s), m.data(rng.normal(0.3, 0.5, 12)))
    O = oracle.OracleProgram(m)
    Cn, L, eps = 64, 9, 0.35
    q0 = np.stack([rng.normal(0, 0.4, Cn), rng.uniform(0.05, 0.5, Cn)])  # sorted sites: mu, sigma (close to 0)
    S = api.HmcSession(m, n_chains=Cn, n_warmup=0, config=api.HMCConfig(n_leapfrog=L, init_step_size=eps), seed=5, q0=q0)
    n_div = 0
    for it in range(3):
        qc, ljc = S.position(), S.log_joint()
        z = rng.standard_normal((O.d_cont, Cn))
        u = rng.random(Cn)
        S.set_explicit_randomness(z, u)
        S.step(1)
        qn, info = S.position(), S.last_step_info()
        for c in range(Cn):
            oq, olj, oinfo = O.hmc_transition(qc[:, c], ljc[c], z[:, c], u[c], eps, L)
            assert oinfo["divergent"] == info["divergent"][c], (it, c)
            if oinfo["divergent"]:
                assert not info["accepted"][c] and np.array_equal(qn[:, c], qc[:, c])
                n_div += 1
            elif abs(u[c] - oinfo["accept_prob"]) > 1e-6:
                assert oinfo["accepted"] == info["accepted"][c] and np.allclose(oq, qn[:, c], rtol=1e-9, atol=1e-11)
    assert n_div >= 5  # the case is really exercised
    assert S.stats()["n_divergent"].sum() == n_div
    S.close()


def test_discrete_only_model_yields_fresh_prior_draws():
    """d == 0: HmcSession::advance draws a fresh prior sample each step (hmc.rs:826-846)."""
    from fugue_b200 import api
    from fugue_b200.model import Bernoulli
    m = ModelBuilder()
    m.sample("b", Bernoulli(0.3))
    m.observe("y", Normal(0.0, 1.0), 0.2)
    S = api.HmcSession(m, n_chains=4096, n_warmup=0, seed=3)
    lj = []
    for _ in range(6):
        S.step(1)
        lj.append(S.log_joint().copy())
    lj = np.array(lj)
    p1 = np.isclose(lj, np.log(0.3) + (-0.5 * 0.04 - 0.5 * np.log(2 * np.pi))).mean(axis=1)
    assert np.all(np.abs(p1 - 0.3) < 0.04) and (lj[0] != lj[1]).any() and (lj[1] != lj[2]).any()
    S.close()


def test_find_reasonable_epsilon_matches_oracle():
    from fugue_b200 import api
    rng = np.random.default_rng(4)
    m = models.logistic_regression(N=500, D=4)[0]
    O = oracle.OracleProgram(m)
    Cn = 16
    q0 = rng.normal(0, 0.3, (O.d_cont, Cn))
    S = api.HmcSession(m, n_chains=Cn, n_warmup=10, config=api.HMCConfig(init_step_size=0.1), seed=1, q0=q0)
    z = rng.standard_normal((O.d_cont, Cn))
    S.set_explicit_randomness(z, np.zeros(Cn))
    e = S.find_reasonable_epsilon()
    lj = S.log_joint()
    for c in range(Cn):
        oe = O.find_reasonable_epsilon(q0[:, c], lj[c], z[:, c])
        assert e[c] == oe, (c, e[c], oe)  # powers of two: exact


def test_finite_difference_mode_matches_reference_force():
    """grad_mode=1 runs the reference's own force (hmc.rs:304-329) on device."""
    from fugue_b200 import api
    rng = np.random.default_rng(8)
    m = models.linear_regression()[0]
    O = oracle.OracleProgram(m)
    Cn, L, eps = 8, 4, 0.01
    q0 = _init_q(O, rng, Cn)
    S = api.HmcSession(m, n_chains=Cn, config=api.HMCConfig(n_leapfrog=L, init_step_size=eps, grad_mode=1), seed=2, q0=q0)
    qc, ljc = S.position(), S.log_joint()
    z, u = rng.standard_normal((3, Cn)), rng.random(Cn)
    S.set_explicit_randomness(z, u)
    S.step(1)
    qn, info = S.position(), S.last_step_info()
    for c in range(Cn):
        oq, olj, oinfo = O.hmc_transition(qc[:, c], ljc[c], z[:, c], u[c], eps, L, grad_mode=1)
        assert abs(oinfo["accept_prob"] - info["accept_prob"][c]) < 1e-6
        if abs(u[c] - oinfo["accept_prob"]) > 1e-4:
            assert np.allclose(oq, qn[:, c], rtol=1e-6, atol=1e-8)


def test_conjugate_normal_posterior_moments():
    """tests/f_hmc_posterior.rs: mu ~ N(0,2) [var 4], y_i ~ N(mu,1), ys = [1.5,2.0,2.5,1.8,2.2]:
    posterior mean 1.90476, variance 0.190476; mean within 3 SE, variance within 12%."""
    from fugue_b200 import api
    m = ModelBuilder()
    mu = m.sample("mu", Normal(0.0, 2.0))
    m.plate_observe("y", Normal(mu, 1.0), m.data(np.array([1.5, 2.0, 2.5, 1.8, 2.2])))
    draws, lp, st = api.hmc_chain(m, n_samples=400, n_warmup=300, n_chains=256, seed=3)
    x = draws[:, 0, :]
    post_var = 1.0 / (1.0 / 4.0 + 5.0)
    post_mean = post_var * 10.0
    assert abs(post_mean - 1.90476) < 1e-4
    se = np.sqrt(post_var / (x.size / 4.0))
    assert abs(x.mean() - post_mean) < 4 * se + 0.01
    assert abs(x.var() / post_var - 1.0) < 0.12
    assert 0.5 < (st["n_accept"] / 700.0).mean() <= 1.0
    # dual averaging steers acceptance towards the 0.8 target (hmc.rs:934-957 direction test)
    assert (st["n_divergent"] == 0).all()


def test_gamma_target_stays_in_support():
    """hmc.rs:960-1020 template: x ~ Gamma(3,1): all draws > 0, mean 3 +- 0.3."""
    from fugue_b200 import api
    m = ModelBuilder()
    m.sample("x", Gamma(3.0, 1.0))
    draws, lp, st = api.hmc_chain(m, n_samples=300, n_warmup=300, n_chains=128, seed=9)
    x = draws[:, 0, :]
    assert (x > 0).all()
    assert abs(x.mean() - 3.0) < 0.3
    assert abs(x.var() - 3.0) < 0.6


def test_mass_adaptation_recovers_scales():
    """adapt_mass: N(0,1) x N(0,10) -> M^-1 ~ marginal variances; both scales sampled correctly."""
    from fugue_b200 import api
    m = ModelBuilder()
    m.sample("a", Normal(0.0, 1.0))
    m.sample("b", Normal(0.0, 10.0))
    draws, lp, st = api.hmc_chain(m, n_samples=400, n_warmup=400, config=api.HMCConfig(adapt_mass=True), n_chains=128, seed=12)
    a, b = draws[:, 0, :], draws[:, 1, :]
    assert abs(a.std() - 1.0) < 0.1 and abs(b.std() - 10.0) < 1.0
    assert abs(a.mean()) < 0.1 and abs(b.mean()) < 1.0


def test_session_equals_batch_and_sharding_streams():
    """hmc.rs:1025-1052 (session == batch); and a sharded session reproduces the same chains."""
    from fugue_b200 import api
    m = models.linear_regression()[0]
    draws, lp, st = api.hmc_chain(m, n_samples=5, n_warmup=20, n_chains=8, seed=77)
    S = api.HmcSession(m, n_chains=8, n_warmup=20, seed=77)
    S.step(20)
    out = []
    for _ in range(5):
        S.step(1)
        out.append(S.position())
    assert np.array_equal(np.stack(out), draws)
    # chains 4..7 as their own shard
    S2 = api.HmcSession(m, n_chains=4, n_warmup=20, seed=77, first_chain=4)
    S2.step(25)
    assert np.array_equal(S2.position(), draws[-1][:, 4:])


def test_returned_trace_has_fresh_weights():
    """fg31 (hmc.rs:975-989): the session's accumulators equal a re-score of its position."""
    from fugue_b200 import api
    m = models.hierarchical(G=6, per_group=4)[0]
    S = api.HmcSession(m, n_chains=16, n_warmup=10, seed=3)
    S.step(15)
    q = S.position()
    lp, ll, lf = S.trace()
    P = S.program
    rlp, rll, rlf = P.logp(q)
    assert np.allclose(lp, rlp, rtol=1e-12) and np.allclose(ll, rll, rtol=1e-12)
    assert np.allclose(S.log_joint(), rlp + rll + rlf, rtol=1e-12)


def test_step_recorded_trajectory_matches_restated_leapfrog():
    """HmcSession::step_recorded (hmc.rs:811-817): LeapfrogPoint{q, h} at the start and after every
    step equals the leapfrog of hmc.rs:353-403 restated here with the oracle as scorer, on the same
    momentum; the recorded endpoint is the proposal, and a recorded step is a normal transition."""
    from fugue_b200 import api
    rng = np.random.default_rng(4)
    m = models.hierarchical(G=6, per_group=4)[0]
    O = oracle.OracleProgram(m)
    Cn, L, eps = 8, 5, 0.03
    q0 = _init_q(O, rng, Cn)
    S = api.HmcSession(m, n_chains=Cn, n_warmup=0, config=api.HMCConfig(n_leapfrog=L, init_step_size=eps), seed=5, q0=q0)
    z = rng.standard_normal((O.d_cont, Cn))
    u = rng.random(Cn)
    S.set_explicit_randomness(z, u)
    rq, rh = S.step_recorded()
    assert rq.shape == (L + 1, O.d_cont, Cn) and rh.shape == (L + 1, Cn)
    assert np.array_equal(rq[0], q0)

    def lj_grad(q):
        lp, ll, lf, g = O.logp(q, grad=True)
        return lp + ll + lf, g

    q, p = q0.copy(), z.copy()  # unit mass: p = z
    lj, g = lj_grad(q)
    hs, qs = [-lj + 0.5 * (p * p).sum(0)], [q.copy()]
    for _ in range(L):
        p = p + 0.5 * eps * g
        q = q + eps * p
        lj, g = lj_grad(q)
        p = p + 0.5 * eps * g
        hs.append(-lj + 0.5 * (p * p).sum(0))
        qs.append(q.copy())
    assert np.allclose(rq, np.array(qs), rtol=1e-10, atol=1e-12)
    assert np.allclose(rh, np.array(hs), rtol=1e-10, atol=1e-10)
    info, qn = S.last_step_info(), S.position()
    ap = np.minimum(1.0, np.exp(hs[0] - hs[-1]))
    assert np.allclose(info["accept_prob"], ap, rtol=1e-8)
    for c in range(Cn):
        want = rq[L, :, c] if info["accepted"][c] else q0[:, c]
        assert np.array_equal(qn[:, c], want)
    assert S.iterations()[0] == 1 if isinstance(S.iterations(), tuple) else True
    # the next, unrecorded transition is unaffected by the recorder
    S.step(1)
    assert S.grad_evals() == 2 * L
    S.close()
