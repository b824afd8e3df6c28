"""Host-side `Trace` materialisation of device draws, and the `Vec<Vec<Trace>>`-shaped diagnostics
front (`extract_f64_values`, `summarize_f64_parameter`, `print_diagnostics`) so results drop into
reference call sites.

Reference: /root/reference/src/runtime/trace.rs:32-47 (ChoiceValue), :168-177 (Trace, three separate
accumulators), :198-200 (total_log_weight), :228-343 (typed getters and their errors);
/root/reference/src/inference/diagnostics.rs:200-228 (extract_f64_values), :321-391
(summarize_f64_parameter), :394-451 (print_diagnostics table format).

The samplers return arrays `[draw][site][chain]`; `traces_from_draws` turns them into
`chains[c][s] -> Trace` with byte-identical addresses and the reference's value types. The three
accumulators of every trace are re-scored on the DEVICE in one batched `fc_logp` call (not on the
host). `Choice.logp` (the per-site log-density the reference caches for its MH proposals) is not
materialised: the device keeps accumulators only; it is None here.
"""
import math
import numpy as np
from . import _ir, model as M
from ._ir import FugueError

FC_ERR_TRACE_ADDRESS_NOT_FOUND, FC_ERR_TYPE_MISMATCH = 500, 600
_TYPE_NAMES = {"f64": "f64", "bool": "bool", "u64": "u64", "usize": "usize", "i64": "i64"}


class Choice:
    """trace.rs:140-147: (addr, value, logp). `kind` is the ChoiceValue variant name."""
    __slots__ = ("addr", "value", "kind", "logp")

    def __init__(self, addr, value, kind, logp=None):
        self.addr, self.value, self.kind, self.logp = addr, value, kind, logp

    def __repr__(self):
        return f"Choice({self.addr!r}, {self.kind}({self.value!r}))"


class Trace:
    """trace.rs:168-177. `choices` is ordered by address (BTreeMap order)."""
    __slots__ = ("choices", "log_prior", "log_likelihood", "log_factors")

    def __init__(self, choices=None, log_prior=0.0, log_likelihood=0.0, log_factors=0.0):
        self.choices = choices if choices is not None else {}
        self.log_prior, self.log_likelihood, self.log_factors = log_prior, log_likelihood, log_factors

    def total_log_weight(self):
        """trace.rs:198-200: (prior + likelihood) + factors, in this order."""
        return self.log_prior + self.log_likelihood + self.log_factors

    def _get(self, addr, kind):
        c = self.choices.get(addr)
        return c.value if c is not None and c.kind == kind else None

    def get_f64(self, addr):
        return self._get(addr, "f64")

    def get_bool(self, addr):
        return self._get(addr, "bool")

    def get_u64(self, addr):
        return self._get(addr, "u64")

    def get_usize(self, addr):
        return self._get(addr, "usize")

    def get_i64(self, addr):
        return self._get(addr, "i64")

    def _get_result(self, addr, kind):
        """trace.rs:228-317: TraceAddressNotFound (500) / TypeMismatch (600)."""
        c = self.choices.get(addr)
        if c is None:
            raise FugueError(FC_ERR_TRACE_ADDRESS_NOT_FOUND, f"get_{kind}: Address not found in trace: {addr}")
        if c.kind != kind:
            raise FugueError(FC_ERR_TYPE_MISMATCH, f"Type mismatch at {addr}: expected {kind}, found {_TYPE_NAMES[c.kind]}")
        return c.value

    def get_f64_result(self, addr):
        return self._get_result(addr, "f64")

    def get_bool_result(self, addr):
        return self._get_result(addr, "bool")

    def get_u64_result(self, addr):
        return self._get_result(addr, "u64")

    def get_usize_result(self, addr):
        return self._get_result(addr, "usize")

    def get_i64_result(self, addr):
        return self._get_result(addr, "i64")

    def insert_choice(self, addr, value, kind="f64", logp=None):
        self.choices[addr] = Choice(addr, value, kind, logp)


def site_table(builder):
    """(continuous, discrete) site lists of a model, each sorted by address (the state row order,
    hmc.rs:238-248): entries are (address, value type name)."""
    builder = getattr(builder, "builder", builder)
    cont, disc = [], []
    for op in builder.ops:
        if op["kind"] not in (_ir.FC_OP_SAMPLE, _ir.FC_OP_PLATE_SAMPLE):
            continue
        vt = op["dist"].value_type
        names = [M.addr(op["name"])] if op["kind"] == _ir.FC_OP_SAMPLE else [M.addr(op["name"], i) for i in range(op["plate_len"])]
        (cont if vt == "f64" else disc).extend((a, vt) for a in names)
    return sorted(cont), sorted(disc)


def _convert(kind, v):
    if kind == "f64":
        return float(v)
    if kind == "bool":
        return bool(v)
    return int(v)


def traces_from_draws(model, draws, ddraws=None, program=None, score=True):
    """chains[c][s] -> Trace from draws [S][d_cont][C] (+ ddraws [S][d_disc][C]).
    With score=True the accumulators come from one batched device scoring of all S*C points."""
    cont, disc = site_table(model)
    draws = np.asarray(draws, dtype=np.float64)
    S, d, Cn = draws.shape if draws.ndim == 3 else (ddraws.shape[0], 0, ddraws.shape[2])
    if d != len(cont):
        raise FugueError(302, f"draws have {d} continuous rows, the model has {len(cont)} sites")
    if disc and (ddraws is None or ddraws.shape[1] != len(disc)):
        raise FugueError(302, f"the model has {len(disc)} discrete sites but ddraws is missing or mis-shaped")
    lp = ll = lf = None
    if score:
        from . import api
        P = program or api.Program(model)
        q = np.ascontiguousarray(draws.transpose(1, 0, 2).reshape(d, S * Cn)) if d else np.zeros((0, S * Cn))
        qd = np.ascontiguousarray(np.asarray(ddraws).transpose(1, 0, 2).reshape(len(disc), S * Cn)) if disc else None
        lp, ll, lf = (a.reshape(S, Cn) for a in P.logp(q, qd))
    chains = []
    for c in range(Cn):
        chain = []
        for s in range(S):
            ch = {}
            for r, (a, vt) in enumerate(cont):
                ch[a] = Choice(a, float(draws[s, r, c]), "f64")
            for r, (a, vt) in enumerate(disc):
                ch[a] = Choice(a, _convert(vt, ddraws[s, r, c]), vt)
            ch = dict(sorted(ch.items()))
            t = Trace(ch)
            if score:
                t.log_prior, t.log_likelihood, t.log_factors = float(lp[s, c]), float(ll[s, c]), float(lf[s, c])
            chain.append(t)
        chains.append(chain)
    return chains


def extract_f64_values(chain, addr):
    """diagnostics.rs:200-207: the f64 values at `addr`, skipping traces that lack it."""
    out = []
    for t in chain:
        v = t.get_f64(addr)
        if v is not None:
            out.append(v)
    return out


def _rust_round(v):
    return int(math.floor(v + 0.5)) if v >= 0 else -int(math.floor(-v + 0.5))


def summarize_f64_parameter(chains, addr):
    """diagnostics.rs:331-391 on `chains[c][s] -> Trace`. Pooled mean / std / quantiles as the
    reference computes them; split-R-hat and the multichain ESS run on the device
    (fc_split_rhat / fc_ess_multichain) when the chains are rectangular."""
    per_chain = [extract_f64_values(ch, addr) for ch in chains]
    allv = [v for ch in per_chain for v in ch]
    if not allv:
        return dict(mean=math.nan, std=math.nan, quantiles={}, r_hat=math.nan, ess=0.0)
    n = len(allv)
    mean = sum(allv) / n  # left-to-right sum, as iter().sum()
    var = sum((x - mean) ** 2 for x in allv) / (n - 1) if n > 1 else math.nan
    srt = sorted(allv)
    qs = {name: srt[_rust_round((n - 1) * p)] for name, p in (("2.5%", 0.025), ("25%", 0.25), ("50%", 0.5), ("75%", 0.75), ("97.5%", 0.975))}
    lens = {len(c) for c in per_chain}
    if len(per_chain) < 1 or 0 in lens:
        r_hat, ess = math.nan, 0.0
    elif len(lens) == 1:
        from . import api
        x = np.ascontiguousarray(np.array(per_chain, dtype=np.float64).T)  # [S][C]
        r_hat, ess = api.r_hat_f64(x), api.effective_sample_size_multichain(x)
    else:
        raise FugueError(_ir.FC_ERR_UNSUPPORTED_OP, "ragged chains: the device diagnostics need equal-length chains")
    return dict(mean=mean, std=math.sqrt(var) if var == var else math.nan, quantiles=qs, r_hat=r_hat, ess=ess)


def format_diagnostics(chains):
    """The text print_diagnostics writes (diagnostics.rs:394-451), as a string. Parameters are
    listed in address order (the reference iterates a HashSet, i.e. in unspecified order)."""
    if not chains or not chains[0]:
        return "No chains or empty chains provided\n"
    addrs = sorted({a for ch in chains for t in ch for a in t.choices})
    lines = ["MCMC Diagnostics:",
             "{:<15} {:>8} {:>8} {:>8} {:>8} {:>8} {:>8} {:>8}".format("Parameter", "Mean", "Std", "2.5%", "50%", "97.5%", "R-hat", "ESS"),
             "-" * 80]
    rhats = []
    for a in addrs:
        s = summarize_f64_parameter(chains, a)
        q = s["quantiles"]
        lines.append("{:<15} {:>8.3f} {:>8.3f} {:>8.3f} {:>8.3f} {:>8.3f} {:>8.3f} {:>8.0f}".format(
            a, s["mean"], s["std"], q.get("2.5%", math.nan), q.get("50%", math.nan), q.get("97.5%", math.nan), s["r_hat"], s["ess"]).replace("nan", "NaN"))
        if math.isfinite(s["r_hat"]):
            rhats.append(s["r_hat"])
    if rhats:
        mx, avg = max(rhats), sum(rhats) / len(rhats)
        lines.append("")
        lines.append("Convergence Assessment:")
        if mx < 1.01:
            lines.append("✓ Excellent convergence (max R-hat = {:.3f})".format(mx))
        elif mx < 1.1:
            lines.append("⚠ Good convergence (max R-hat = {:.3f})".format(mx))
        else:
            lines.append("✗ Poor convergence (max R-hat = {:.3f}) - consider more samples".format(mx))
        lines.append("  Average R-hat: {:.3f}".format(avg))
    return "\n".join(lines) + "\n"


def print_diagnostics(chains):
    print(format_diagnostics(chains), end="")
