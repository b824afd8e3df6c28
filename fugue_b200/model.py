"""Model surface: distributions + the op-list builder (lowering front-end).

Mirrors the reference's user-facing names: the 17 distribution constructors with their
validation (src/core/distribution.rs `new` fns -> ErrorCode 100..106, src/error.rs:40-49),
sample/observe/factor (src/core/model.rs:144,258,397) and plate! (src/macros/mod.rs:82-86).
A parameter may be a float or an Expr over latents and data columns (see _ir.py); that is
the information the reference hides inside closures and the device needs.
"""
import math
import numpy as np
from . import _ir
from ._ir import (Column, Expr, Site, Plate, as_expr, FugueError, dense, select2, exp, sigmoid_clamped, clamp)

_VT = {"f64": _ir.FC_T_F64, "i64": _ir.FC_T_I64, "u64": _ir.FC_T_U64, "usize": _ir.FC_T_USIZE, "bool": _ir.FC_T_BOOL}


class Distribution:
    name, value_type, codes = "", "f64", ()

    def __init__(self, *params):
        self.params = [p if isinstance(p, Expr) else float(p) for p in params]
        self.validate()

    def _const(self, i):
        p = self.params[i]
        return None if isinstance(p, Expr) else p

    def validate(self):
        pass

    def _need(self, ok, code, msg):
        if not ok:
            raise FugueError(code, f"{self.name}: {msg}")

    def _pos(self, i, code, what):
        v = self._const(i)
        if v is not None:
            self._need(v > 0.0 and math.isfinite(v), code, f"{what} must be positive and finite, got {v}")

    def _finite(self, i, code, what):
        v = self._const(i)
        if v is not None:
            self._need(math.isfinite(v), code, f"{what} must be finite, got {v}")


class Normal(Distribution):
    name = "Normal"

    def validate(self):  # distribution.rs:133-153
        self._finite(0, 100, "mu"); self._pos(1, 101, "sigma")


class Uniform(Distribution):
    name = "Uniform"

    def validate(self):
        self._finite(0, 103, "low"); self._finite(1, 103, "high")
        lo, hi = self._const(0), self._const(1)
        if lo is not None and hi is not None:
            self._need(lo < hi, 103, "low must be < high")


class LogNormal(Distribution):
    name = "LogNormal"

    def validate(self):
        self._finite(0, 100, "mu"); self._pos(1, 101, "sigma")


class Exponential(Distribution):
    name = "Exponential"

    def validate(self):
        self._pos(0, 105, "rate")


class _Prob(Distribution):
    def _prob(self, i):
        v = self._const(i)
        if v is not None:
            self._need(math.isfinite(v) and 0.0 <= v <= 1.0, 102, f"probability must be in [0,1], got {v}")


class Bernoulli(_Prob):
    name, value_type = "Bernoulli", "bool"

    def validate(self):
        self._prob(0)


class Categorical(_Prob):
    """Categorical(p): the two-way form [p, 1-p] of examples/mixture_models.rs:94 (p may depend on
    latents); Categorical([p0, .., pK-1]): K-way with constant probabilities (Categorical::new,
    distribution.rs:679-713: non-empty, finite, non-negative, sum within 1e-6 of 1 -> ErrorCode 102)."""
    name, value_type = "Categorical", "usize"

    def __init__(self, p):
        if isinstance(p, (list, tuple, np.ndarray)):
            self.probs = np.ascontiguousarray(p, dtype=np.float64).reshape(-1)
            self.params = []
            self._need(self.probs.size > 0, 102, "Probability vector cannot be empty")
            self._need(abs(float(self.probs.sum()) - 1.0) <= 1e-6, 102, "Probabilities must sum to 1.0")
            self._need(bool(np.all(np.isfinite(self.probs) & (self.probs >= 0.0))), 102, "All probabilities must be non-negative and finite")
        else:
            self.probs = None
            super().__init__(p)

    def validate(self):
        self._prob(0)


class Beta(Distribution):
    name = "Beta"

    def validate(self):
        self._pos(0, 104, "alpha"); self._pos(1, 104, "beta")


class Gamma(Distribution):
    name = "Gamma"

    def validate(self):
        self._pos(0, 104, "shape"); self._pos(1, 105, "rate")


class Binomial(_Prob):
    name, value_type = "Binomial", "u64"

    def validate(self):
        self._prob(1)


class Poisson(Distribution):
    name, value_type = "Poisson", "u64"

    def validate(self):
        self._pos(0, 105, "lambda")


class StudentT(Distribution):
    name = "StudentT"

    def validate(self):
        self._pos(0, 104, "df"); self._finite(1, 100, "loc"); self._pos(2, 101, "scale")


class Cauchy(Distribution):
    name = "Cauchy"

    def validate(self):
        self._finite(0, 100, "loc"); self._pos(1, 101, "scale")


class Laplace(Distribution):
    name = "Laplace"

    def validate(self):
        self._finite(0, 100, "loc"); self._pos(1, 101, "scale")


class Weibull(Distribution):
    name = "Weibull"

    def validate(self):
        self._pos(0, 104, "shape"); self._pos(1, 101, "scale")


class ChiSquared(Distribution):
    name = "ChiSquared"

    def validate(self):
        self._pos(0, 104, "k")


class InverseGamma(Distribution):
    name = "InverseGamma"

    def validate(self):
        self._pos(0, 104, "shape"); self._pos(1, 105, "rate")


class DiscreteUniform(Distribution):
    name, value_type = "DiscreteUniform", "i64"

    def validate(self):
        lo, hi = self._const(0), self._const(1)
        if lo is not None and hi is not None:
            self._need(lo <= hi, 103, "low must be <= high")


def addr(name, i=None):
    """addr!(name) / addr!(name, i) with the reference's escaping (address.rs:175-223)."""
    esc = lambda s: str(s).replace("\\", "\\\\").replace("#", "\\#")
    return esc(name) if i is None else f"{esc(name)}#{esc(i)}"


class ModelBuilder:
    """Records ops in program order; `lower()` yields ctypes arrays for fc_program_create."""

    def __init__(self):
        self.ops, self.cols, self._keep = [], [], []
        self.n_cont = self.n_disc = 0

    # ---- data ----
    def data(self, arr):
        a = np.ascontiguousarray(arr)
        if a.dtype == np.bool_:
            a = a.astype(np.uint8)
        if a.dtype not in (np.float64, np.uint8, np.uint32, np.int64):
            a = a.astype(np.int64 if np.issubdtype(a.dtype, np.integer) else np.float64)
        col = Column(len(self.cols), a)
        self.cols.append(col)
        return col

    # ---- ops ----
    def _op(self, kind, dist, name, plate_len=0, value_col=-1, value=0.0):
        d = dict(kind=kind, dist=dist, name=name, plate_len=plate_len, value_col=value_col, value=value)
        if getattr(dist, "probs", None) is not None:  # K-way Categorical: the probabilities travel as a data column
            d["probs_col"] = self.data(dist.probs).cid
        self.ops.append(d)
        return len(self.ops) - 1

    def sample(self, name, dist):
        disc = dist.value_type != "f64"
        self._op(_ir.FC_OP_SAMPLE, dist, name)
        if disc:
            self.n_disc += 1
            return Site(self.n_disc - 1, True)
        self.n_cont += 1
        return Site(self.n_cont - 1)

    def plate_sample(self, name, n, dist):
        disc = dist.value_type != "f64"
        oi = self._op(_ir.FC_OP_PLATE_SAMPLE, dist, name, plate_len=int(n))
        if disc:
            self.n_disc += n
            return Plate(self.n_disc - n, n, True, oi)
        self.n_cont += n
        return Plate(self.n_cont - n, n, False, oi)

    def observe(self, name, dist, value):
        self._op(_ir.FC_OP_OBSERVE, dist, name, value=float(value))

    def plate_observe(self, name, dist, values):
        assert isinstance(values, Column)
        self._op(_ir.FC_OP_PLATE_OBSERVE, dist, name, plate_len=len(values), value_col=values.cid)

    def factor(self, logw, nan_is_neg_inf=False):
        """factor(log_weight) (src/core/model.rs:97-104). A number is a constant factor; a `Sym`
        (or a bare latent) is evaluated per chain on the device with its analytic gradient."""
        if isinstance(logw, (_ir.Sym, _ir.Expr)):
            self._op(_ir.FC_OP_FACTOR_EXPR, None, "", value=1.0 if nan_is_neg_inf else 0.0)
            self.ops[-1]["expr"] = _ir.Sym.lift(logw)
            return
        self._op(_ir.FC_OP_FACTOR, None, "", value=float(logw))

    # ---- lowering ----
    def lower(self):
        n = len(self.ops)
        node_list = []
        ops = (_ir.fc_op * max(n, 1))()
        for i, d in enumerate(self.ops):
            o = ops[i]
            o.kind, o.plate_len, o.value_col, o.value = d["kind"], d["plate_len"], d["value_col"], d["value"]
            o.name = d["name"].encode()
            dist = d["dist"]
            if d["kind"] == _ir.FC_OP_FACTOR_EXPR:
                o.expr_first, o.expr_len = len(node_list), len(d["expr"].nodes)
                node_list.extend(d["expr"].nodes)
            if dist is None:
                continue
            o.probs_col = d.get("probs_col", 0)
            o.dist = _ir.DIST_IDS[dist.name]
            o.value_type = _VT[dist.value_type]
            o.n_params = len(dist.params)
            for j, p in enumerate(dist.params):
                e = as_expr(p)
                if len(e.terms) > 3:
                    raise FugueError(_ir.FC_ERR_UNSUPPORTED_OP, "parameter expression has more than 3 terms")
                pr = o.param[j]
                pr.link, pr.lo, pr.hi = e.link, e.lo, e.hi
                for t, (k, slot, slot_b, col, c) in enumerate(e.terms):
                    tm = pr.term[t]
                    tm.kind, tm.slot, tm.slot_b, tm.col, tm.c = k, slot, slot_b, col, c
        ncol = len(self.cols)
        cols = (_ir.fc_column * max(ncol, 1))()
        dt = {np.dtype(np.float64): 0, np.dtype(np.uint8): 1, np.dtype(np.uint32): 2, np.dtype(np.int64): 3}
        for i, c in enumerate(self.cols):
            cols[i].dtype = dt[c.arr.dtype]
            cols[i].width = 1 if c.arr.ndim == 1 else c.arr.shape[1]
            cols[i].rows = c.arr.shape[0]
            cols[i].data = c.arr.ctypes.data
        nodes = (_ir.fc_expr_node * max(len(node_list), 1))()
        for i, (o, a, b, sl, c) in enumerate(node_list):
            nodes[i].op, nodes[i].a, nodes[i].b, nodes[i].slot, nodes[i].c = o, a, b, sl, c
        self.nodes_lowered = (nodes, len(node_list))  # the FACTOR_EXPR table (fc_program_create_ex)
        return ops, n, cols, ncol


# ---- functions of factor expressions (the playground's set, crates/fugue-wasm/src/dsl.rs:775-786) ----
def _fn(op):
    return lambda *a: _ir.Sym.apply(op, *a)


ln, sqrt, abs_, floor, sin, cos, tanh = (_fn(o) for o in (_ir.FC_X_LN, _ir.FC_X_SQRT, _ir.FC_X_ABS, _ir.FC_X_FLOOR,
                                                             _ir.FC_X_SIN, _ir.FC_X_COS, _ir.FC_X_TANH))
_link_exp = exp


def exp(e):
    """exp as a LINK on a parameter expression (hierarchical / GLM parameters) or, on a general
    expression, as a function node of a factor expression."""
    return _ir.Sym.apply(_ir.FC_X_EXP, e) if isinstance(e, _ir.Sym) else _link_exp(e)


pow_, min_, max_ = _fn(_ir.FC_X_POW), _fn(_ir.FC_X_MIN), _fn(_ir.FC_X_MAX)
