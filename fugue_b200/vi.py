"""Mean-field VI as a BATCHED SCORING CLIENT of the device path: the model's
log-joint of every guide draw -- the part the reference spends its time in, one closure
interpretation per draw (`run(ScoreGivenTrace{base: guide_trace}, model_fn())`,
/root/reference/src/inference/vi.rs:639-675) -- is ONE `fc_logp` launch over all draws; with the
Jacobi finite-difference update of vi.rs:784-860 every coordinate's +/- pair of an iteration rides in
the same launch. The guide (a handful of 1-D densities) stays on the host: sampling and log q follow
vi.rs:210-404 (Normal / LogNormal / Beta families, `for_support`, `init_log_sigma`, the update clamps).
"""
import math
import numpy as np
from . import _ir
from ._ir import FugueError

LOG_SCALE_MIN, LOG_SCALE_MAX, MU_ABS_MAX = -20.0, 20.0, 1.0e6  # vi.rs:100-102
REAL, POSITIVE, UNIT = "Real", "Positive", "Unit"               # Support, vi.rs:119-133
LOCATION, SCALE = 0, 1                                          # ParamCoord, vi.rs:167-172
_LN_2PI = math.log(2.0 * math.pi)


def init_log_sigma(value):  # vi.rs:413-416
    scale = abs(value) if math.isfinite(value) else 1.0
    return math.log(max(0.1 * scale, 0.1))


class VariationalParam:
    """VariationalParam (vi.rs:210-404): kind in {"Normal", "LogNormal", "Beta"}, two unconstrained
    coordinates (mu, log_sigma) or (log_alpha, log_beta)."""

    def __init__(self, kind, a, b):
        self.kind, self.a, self.b = kind, float(a), float(b)

    @staticmethod
    def for_support(support, init_value):  # vi.rs:245-292
        if support == REAL:
            return VariationalParam("Normal", init_value, init_log_sigma(init_value))
        if support == POSITIVE:
            safe = init_value if (math.isfinite(init_value) and init_value > 0.0) else 1.0
            return VariationalParam("LogNormal", math.log(safe), math.log(0.5))
        m = min(max(init_value, 1e-3), 1.0 - 1e-3) if math.isfinite(init_value) else 0.5
        return VariationalParam("Beta", math.log(2.0 * m), math.log(2.0 * (1.0 - m)))

    def copy(self):
        return VariationalParam(self.kind, self.a, self.b)

    def shifted(self, coord, delta):  # vi.rs:418-462
        p = self.copy()
        if coord == LOCATION:
            p.a += delta
        else:
            p.b += delta
        return p

    def apply_update(self, coord, delta):  # vi.rs:464-496 (clamped)
        if self.kind == "Beta":
            if coord == LOCATION:
                self.a = min(max(self.a + delta, LOG_SCALE_MIN), LOG_SCALE_MAX)
            else:
                self.b = min(max(self.b + delta, LOG_SCALE_MIN), LOG_SCALE_MAX)
        elif coord == LOCATION:
            self.a = min(max(self.a + delta, -MU_ABS_MAX), MU_ABS_MAX)
        else:
            self.b = min(max(self.b + delta, LOG_SCALE_MIN), LOG_SCALE_MAX)

    def _valid(self):
        if self.kind == "Beta":
            al, be = math.exp(self.a), math.exp(self.b)
            return math.isfinite(al) and math.isfinite(be) and al > 0.0 and be > 0.0
        s = math.exp(self.b)
        return math.isfinite(self.a) and math.isfinite(s) and s > 0.0

    def sample(self, rng, n):
        """n draws (vi.rs:294-330); NaN when the parameters are invalid. Normal / LogNormal draws are
        mu + sigma * z of the generator's standard normals, so equal seeds give common random numbers."""
        if not self._valid():
            return np.full(n, math.nan)
        if self.kind == "Beta":
            return rng.beta(math.exp(self.a), math.exp(self.b), n)
        z = self.a + math.exp(self.b) * rng.standard_normal(n)
        return np.exp(z) if self.kind == "LogNormal" else z

    def log_prob(self, x):
        """log q(x), the reference distributions' log_prob (distribution.rs Normal / LogNormal / Beta)."""
        x = np.asarray(x, dtype=np.float64)
        with np.errstate(all="ignore"):
            if self.kind == "Beta":
                al, be = math.exp(self.a), math.exp(self.b)
                ln_b = math.lgamma(al) + math.lgamma(be) - math.lgamma(al + be)
                out = (al - 1.0) * np.log(x) + (be - 1.0) * np.log1p(-x) - ln_b
                return np.where((x > 0.0) & (x < 1.0), out, -math.inf)
            s = math.exp(self.b)
            if self.kind == "Normal":
                z = (x - self.a) / s
                return -0.5 * z * z - math.log(s) - 0.5 * _LN_2PI
            lx = np.log(x)
            z = (lx - self.a) / s
            return np.where(x > 0.0, -0.5 * z * z - math.log(s) - 0.5 * _LN_2PI - lx, -math.inf)


class MeanFieldGuide:
    """MeanFieldGuide (vi.rs:520-636): address -> VariationalParam."""

    def __init__(self):
        self.params = {}

    def add_latent(self, addr, support, init_value):  # vi.rs:559-562
        self.params[addr] = VariationalParam.for_support(support, init_value)

    @staticmethod
    def from_values(values, types=None):
        """from_trace (vi.rs:577-606): a Normal factor per continuous choice; a discrete choice is the
        typed GuideError::UnsupportedDiscreteLatent. `values` maps address -> value."""
        g = MeanFieldGuide()
        for a in sorted(values):
            if types and types.get(a, "f64") != "f64":
                raise FugueError(_ir.FC_ERR_UNSUPPORTED_OP, f"UnsupportedDiscreteLatent: `{a}` is {types[a]}; mean-field VI handles continuous latents only")
            g.params[a] = VariationalParam("Normal", values[a], init_log_sigma(values[a]))
        return g

    def copy(self):
        g = MeanFieldGuide()
        g.params = {a: p.copy() for a, p in self.params.items()}
        return g

    def sample(self, rng, n):
        """n joint draws in sorted-address order (sample_trace, vi.rs:609-635): (z [len(params)][n], log q [n])."""
        addrs = sorted(self.params)
        z = np.empty((len(addrs), n))
        lq = np.zeros(n)
        for i, a in enumerate(addrs):
            z[i] = self.params[a].sample(rng, n)
            lq += self.params[a].log_prob(z[i])
        return addrs, z, lq


def _rng(seed):
    return np.random.Generator(np.random.Philox(key=int(seed) & (2**64 - 1)))


def _program(model):
    from . import api
    return model if isinstance(model, api.Program) else api.Program(model)


def _elbo_terms(P, guides_and_seeds, n):
    """For each (guide, seed): n draws, scored by the model in ONE device launch. Returns the list of
    per-guide ELBO estimates mean(log p(x, z) - log q(z)) (vi.rs:639-675). Guide latents the model
    does not sample are ignored; model latents the guide lacks are a typed error (the reference's
    ScoreGivenTrace would fail on the missing address)."""
    sites = list(P.sites)
    if P.d_disc:
        raise FugueError(_ir.FC_ERR_UNSUPPORTED_OP, "mean-field VI handles continuous latents only (the model has discrete sites)")
    G = len(guides_and_seeds)
    q = np.empty((len(sites), G * n))
    lqs = []
    for gi, (guide, seed) in enumerate(guides_and_seeds):
        missing = [s for s in sites if s not in guide.params]
        if missing:
            raise FugueError(500, f"guide has no factor for `{missing[0]}` (TraceAddressNotFound)")
        addrs, z, _ = guide.sample(_rng(seed), n)
        row = {a: i for i, a in enumerate(addrs)}
        lq = np.zeros(n)
        for si, s in enumerate(sites):  # only the latents the model samples count (vi.rs:662-668)
            q[si, gi * n:(gi + 1) * n] = z[row[s]]
            lq += guide.params[s].log_prob(z[row[s]])
        lqs.append(lq)
    lp, ll, lf = P.logp(q)
    tot = (lp + ll) + lf
    out = []
    for gi in range(G):
        t = tot[gi * n:(gi + 1) * n] - lqs[gi]
        out.append(float(np.sum(t) / n) if n else math.nan)
    return out


def elbo_with_guide(model, guide, num_samples, seed=0):
    """elbo_with_guide (vi.rs:639-675): E_q[log p(x, z) - log q(z)] from num_samples guide draws."""
    return _elbo_terms(_program(model), [(guide, seed)], int(num_samples))[0]


def elbo_gradient_fd(seed, model, guide, addr, coord, eps, num_samples):
    """elbo_gradient_fd (vi.rs:687-727): central difference in one unconstrained coordinate with common
    random numbers (the same seed for the + and - guide)."""
    if addr not in guide.params:
        return 0.0
    gp, gm = guide.copy(), guide.copy()
    gp.params[addr] = guide.params[addr].shifted(coord, eps)
    gm.params[addr] = guide.params[addr].shifted(coord, -eps)
    ep, em = _elbo_terms(_program(model), [(gp, seed), (gm, seed)], int(num_samples))
    return (ep - em) / (2.0 * eps)


class VIConfig:  # vi.rs:729-753
    def __init__(self, n_iterations=1000, n_samples_per_iter=16, base_learning_rate=0.1, fd_eps=0.01, convergence_tol=1e-4,
                 convergence_window=20, step_decay_exponent=0.6):
        self.n_iterations, self.n_samples_per_iter, self.base_learning_rate = n_iterations, n_samples_per_iter, base_learning_rate
        self.fd_eps, self.convergence_tol, self.convergence_window, self.step_decay_exponent = fd_eps, convergence_tol, convergence_window, step_decay_exponent


class VIResult:  # vi.rs:763-772
    def __init__(self, guide, elbo_history, converged, iterations):
        self.guide, self.elbo_history, self.converged, self.iterations = guide, elbo_history, converged, iterations


def optimize_meanfield_vi_with_config(model, initial_guide, config=None, seed=0):
    """optimize_meanfield_vi_with_config (vi.rs:784-866): per iteration the monitor ELBO and the +/-
    pair of every (address, coordinate) -- 1 + 4 * |guide| guides x n_samples draws -- are scored in
    ONE device launch; Jacobi update from the snapshot, Robbins-Monro step, plateau stop."""
    cfg = config or VIConfig()
    P = _program(model)
    guide = initial_guide.copy()
    master = _rng(seed)
    hist, converged, iterations = [], False, 0
    for it in range(cfg.n_iterations):
        iterations = it + 1
        snapshot = guide.copy()
        addrs = sorted(snapshot.params)
        batch = [(snapshot, int(master.integers(0, 2**63)))]  # the monitor ELBO of this iteration
        for a in addrs:
            for coord in (LOCATION, SCALE):
                s = int(master.integers(0, 2**63))  # one seed per coordinate, shared by the +/- pair (CRN)
                gp, gm = snapshot.copy(), snapshot.copy()
                gp.params[a] = snapshot.params[a].shifted(coord, cfg.fd_eps)
                gm.params[a] = snapshot.params[a].shifted(coord, -cfg.fd_eps)
                batch += [(gp, s), (gm, s)]
        vals = _elbo_terms(P, batch, cfg.n_samples_per_iter)
        hist.append(vals[0])
        w = cfg.convergence_window
        if w > 0 and len(hist) >= 2 * w:
            recent, previous = sum(hist[-w:]) / w, sum(hist[-2 * w:-w]) / w
            if abs(recent - previous) / max(abs(previous), 1e-8) < cfg.convergence_tol:
                converged = True
                break
        step = cfg.base_learning_rate * float(it + 1) ** (-cfg.step_decay_exponent)
        k = 1
        for a in addrs:
            for coord in (LOCATION, SCALE):
                grad = (vals[k] - vals[k + 1]) / (2.0 * cfg.fd_eps)
                k += 2
                if math.isfinite(grad) and math.isfinite(step * grad):
                    guide.params[a].apply_update(coord, step * grad)
    return VIResult(guide, hist, converged, iterations)


def optimize_meanfield_vi(model, initial_guide, n_iterations, n_samples_per_iter, learning_rate, seed=0):
    """optimize_meanfield_vi (vi.rs:874-892)."""
    return optimize_meanfield_vi_with_config(model, initial_guide, VIConfig(n_iterations=n_iterations, n_samples_per_iter=n_samples_per_iter,
                                                                            base_learning_rate=learning_rate), seed).guide


def estimate_elbo(model, num_samples, seed=0):
    """estimate_elbo (vi.rs:905-922): ELBO with q = prior = E_prior[log-likelihood + factors]; the prior
    draws come from the device's prior sampler (a particle session), scored in the same launch."""
    from . import api
    S = api.SmcSession(model, int(num_samples), api.SMCConfig(), seed)
    r = S.prior_weights().result(values=False)  # log_weight = log-likelihood (+ factors) of each prior draw, normalised
    ln, _ = S.normalize()                        # log sum exp of the raw log weights
    S.close()
    # mean of ll_i = mean(normalised log weights) + lse
    return float(np.mean(r.log_weight) + ln)
