"""Text front-end: the `prob!`-subset model language of the reference's playground, lowered to the
flat op list of include/fugue_cuda.h instead of being interpreted into closures.

The language is the one /root/reference/crates/fugue-wasm/src/dsl.rs documents (grammar :10-20,
distribution table :807-818, evaluation policy of constant expressions :714-803, soft errors :926-944
and :1002-1007, data binding :1055-1107); its syntax lives in dsl_syntax.py (tokens, productions as
data, precedence climbing), this file is the lowering:

    let p <- sample(addr!("p"), Beta(2.0, 2.0));
    for i in 0..data.len() {
        observe(addr!("flip", i), Bernoulli(p), data[i]);
    }
    pure(p)

What differs is the back end. The reference folds the AST into `sample`/`observe`/`factor`/`bind`
closures that are re-run for every density evaluation; here the program is run ONCE, symbolically:

  * a `for` body whose statements address `addr!(name, i)` with the loop variable and read data as
    `arr[i]` becomes ONE plate op (PLATE_OBSERVE / PLATE_SAMPLE) over a device-resident column,
    with byte-identical addresses `name#i` (src/core/address.rs:208-223);
  * parameters must be of the form link(term + term + term) with terms const | latent |
    latent * arr[i] (see _ir.py); anything else -- a structure-varying or non-affine model --
    is rejected with the typed error FC_ERR_UNSUPPORTED_OP (1000): there is no CPU fallback;
  * loops that are not plates (other index expressions, lower bound != 0) are unrolled into
    scalar ops when every parameter is still expressible.

Static errors (syntax, unknown variables / distributions / functions, wrong arities) raise
DslError with the reference's messages ("line N: expected ..., found ..."). Soft errors
(invalid constant parameters, index out of bounds) degrade to `factor(-inf)` plus a warning, as
the reference does (dsl.rs:926-944, 1002-1007).
"""
import json
import math
import numpy as np
from . import _ir, model as M
from ._ir import FugueError


from .dsl_syntax import DslError, MATH_FNS, DIST_ARITY, _fmt_num  # noqa: F401 (re-exported)
from .dsl_syntax import parse_program as _parse_program, check_stmts as _check_stmts, check_expr as _check_expr


# ---------------------------------------------------------------------------------------------
# Symbolic values of the single lowering run
#   constants: bool | int | float (the reference's Value::{Bool,Int,F64}), np.ndarray (Value::Arr)
#   _LoopVar : the index of the enclosing plate-candidate loop
#   _Elem    : arr[i] for the loop index i -- a whole derived column; arithmetic with constants and
#              other _Elems is folded element-wise (the same f64 operations the reference performs
#              per iteration)
#   _ir.Expr : link(term + term + term) over latents (the fc_param form)
# ---------------------------------------------------------------------------------------------
class _LoopVar:
    pass


class _Elem:
    def __init__(self, arr):
        self.arr = np.asarray(arr, dtype=np.float64)


class _Opaque:
    """A discrete plate element: addressable, but not usable in parameter expressions."""


class _NotPlate(Exception):
    """The loop body does not have the plate shape; the caller unrolls instead."""


def _unsupported(msg):
    return FugueError(_ir.FC_ERR_UNSUPPORTED_OP, msg + " (not expressible as link(term+term+term) over latents; no CPU fallback)")


def _is_const(v):
    return isinstance(v, (bool, int, float))


def _as_f64(v):  # Value::as_f64 (dsl.rs:55-69)
    if isinstance(v, bool):
        return 1.0 if v else 0.0
    if isinstance(v, (int, float)):
        return float(v)
    return float("nan")


def _as_index(v):  # Value::as_index (dsl.rs:71-77)
    if isinstance(v, bool):
        return None
    if isinstance(v, int):
        return v
    if isinstance(v, float) and math.isfinite(v) and v == math.floor(v):
        return int(v)
    return None


def _math(name, a):
    try:
        if name == "exp": return math.exp(a[0])
        if name in ("ln", "log"): return math.log(a[0]) if a[0] > 0 else (float("-inf") if a[0] == 0 else float("nan"))
        if name == "sqrt": return math.sqrt(a[0]) if a[0] >= 0 else float("nan")
        if name == "abs": return abs(a[0])
        if name == "floor": return float(math.floor(a[0])) if math.isfinite(a[0]) else a[0]
        if name == "sin": return math.sin(a[0])
        if name == "cos": return math.cos(a[0])
        if name == "tanh": return math.tanh(a[0])
        if name == "pow": return math.pow(a[0], a[1])
        if name == "min": return a[1] if a[0] != a[0] else (a[0] if a[1] != a[1] else min(a[0], a[1]))
        if name == "max": return a[1] if a[0] != a[0] else (a[0] if a[1] != a[1] else max(a[0], a[1]))
    except OverflowError:
        return float("inf")
    except ValueError:
        return float("nan")
    return float("nan")


def _bin_const(op, x, y):  # dsl.rs:755-772
    if isinstance(x, int) and isinstance(y, int) and not isinstance(x, bool) and not isinstance(y, bool) and op in "+-*":
        return x + y if op == "+" else (x - y if op == "-" else x * y)
    a, b = _as_f64(x), _as_f64(y)
    if op == "+": return a + b
    if op == "-": return a - b
    if op == "*": return a * b
    if b == 0.0:
        return float("nan") if (a == 0.0 or a != a) else math.copysign(float("inf"), a) * (math.copysign(1.0, b))
    return a / b


def _bare_latent(e):
    return isinstance(e, _ir.Expr) and e.link == _ir.FC_LINK_IDENTITY and len(e.terms) == 1 and e.terms[0][0] == _ir.FC_TERM_LATENT


class _Lowering:
    def __init__(self, builder, warnings):
        self.m, self.warnings = builder, warnings
        self.plate_n = None     # length of the plate-candidate loop being lowered, else None
        self.col_cache = {}

    def warn(self, msg):
        if len(self.warnings) < 64:
            self.warnings.append(msg)

    def column(self, arr):
        key = (arr.dtype.str, arr.tobytes())
        if key not in self.col_cache:
            self.col_cache[key] = self.m.data(arr)
        return self.col_cache[key]

    # ---- expressions ----
    def ev(self, e, env):
        t = e[0]
        if t in ("num", "int", "bool"):
            return e[1]
        if t == "var":
            if isinstance(env[e[1]], _ir.Sym):
                raise _unsupported(f"`{e[1]}` is a non-affine latent expression: usable in factor(..) and pure(..) only")
            return env[e[1]]
        if t == "len":
            a = self.ev(e[1], env)
            if isinstance(a, np.ndarray):
                return int(a.shape[0])
            self.warn("`.len()` on a non-array value")
            return 0
        if t == "index":
            a, i = self.ev(e[1], env), self.ev(e[2], env)
            if isinstance(a, np.ndarray) and isinstance(i, _LoopVar):
                if a.shape[0] < self.plate_n:
                    raise _NotPlate()  # the tail would be out of bounds: unroll to get the per-element warnings
                return _Elem(a[:self.plate_n])
            if isinstance(i, (_LoopVar, _Elem, _ir.Expr, _Opaque)):
                raise _NotPlate() if self.plate_n is not None else _unsupported("array index depends on a latent")
            k = _as_index(i)
            if isinstance(a, np.ndarray) and k is not None:
                if 0 <= k < a.shape[0]:
                    return float(a[k])
                self.warn(f"index {k} out of bounds (len {a.shape[0]})")
                return float("nan")
            self.warn("indexing a non-array value")
            return float("nan")
        if t == "neg":
            a = self.ev(e[1], env)
            if isinstance(a, int) and not isinstance(a, bool):
                return -a
            if _is_const(a):
                return -_as_f64(a)
            if isinstance(a, _Elem):
                return _Elem(-a.arr)
            if isinstance(a, _ir.Expr) and a.is_affine():
                return a.scale(-1.0)
            raise _unsupported("negated latent expression")
        if t == "bin":
            sig = self.sigmoid_pattern(e, env)
            if sig is not None:
                return sig
            return self.bin(e[1], self.ev(e[2], env), self.ev(e[3], env))
        if t == "call":
            return self.call(e[1], [self.ev(a, env) for a in e[2]])
        raise AssertionError(t)

    _SYM_FN = {"exp": _ir.FC_X_EXP, "ln": _ir.FC_X_LN, "log": _ir.FC_X_LN, "sqrt": _ir.FC_X_SQRT, "abs": _ir.FC_X_ABS, "floor": _ir.FC_X_FLOOR,
               "sin": _ir.FC_X_SIN, "cos": _ir.FC_X_COS, "tanh": _ir.FC_X_TANH, "pow": _ir.FC_X_POW, "min": _ir.FC_X_MIN, "max": _ir.FC_X_MAX}
    _SYM_BIN = {"+": _ir.FC_X_ADD, "-": _ir.FC_X_SUB, "*": _ir.FC_X_MUL, "/": _ir.FC_X_DIV}

    def sym(self, e, env):
        """The expression as a general function of the continuous latents (fc_expr_node list), with
        constant sub-expressions folded by the same rules as ev(). Returns a number or an _ir.Sym."""
        t = e[0]
        if t in ("num", "int", "bool"):
            return e[1]
        if t == "var":
            v = env[e[1]]
            if _is_const(v) or isinstance(v, _ir.Sym):
                return v
            if isinstance(v, _ir.Expr):
                return _ir.Sym.lift(v)  # const + coef * latent + ..., link applied (typed error for discrete / per-element terms)
            raise _unsupported("this value cannot enter a factor expression")
        if t in ("len", "index"):
            v = self.ev(e, env)
            if not _is_const(v):
                raise _unsupported("array element indexed by a latent")
            return v
        if t == "neg":
            a = self.sym(e[1], env)
            if _is_const(a):
                return -a if (isinstance(a, int) and not isinstance(a, bool)) else -_as_f64(a)
            return _ir.Sym.apply(_ir.FC_X_NEG, a)
        if t == "bin":
            x, y = self.sym(e[2], env), self.sym(e[3], env)
            if _is_const(x) and _is_const(y):
                return _bin_const(e[1], x, y)
            return _ir.Sym.apply(self._SYM_BIN[e[1]], _as_f64(x) if _is_const(x) else x, _as_f64(y) if _is_const(y) else y)
        if t == "call":
            args = [self.sym(a, env) for a in e[2]]
            if all(_is_const(a) for a in args):
                return _math(e[1], [_as_f64(a) for a in args])
            return _ir.Sym.apply(self._SYM_FN[e[1]], *[_as_f64(a) if _is_const(a) else a for a in args])
        raise AssertionError(t)

    def sigmoid_pattern(self, e, env):
        """1.0 / (1.0 + exp(-eta)) (examples/classification.rs:124) -> the SIGMOID link."""
        def is_one(x):
            return x[0] in ("num", "int") and float(x[1]) == 1.0
        if e[1] != "/" or not is_one(e[2]) or e[3][0] != "bin" or e[3][1] != "+":
            return None
        l, r = e[3][2], e[3][3]
        ex = r if is_one(l) else (l if is_one(r) else None)
        if ex is None or ex[0] != "call" or ex[1] != "exp" or ex[2][0][0] != "neg":
            return None
        eta = self.ev(ex[2][0][1], env)
        if not isinstance(eta, _ir.Expr) or eta.link != _ir.FC_LINK_IDENTITY:
            return None
        return _ir.sigmoid_clamped(eta, 0.0, 1.0)

    def bin(self, op, x, y):
        if _is_const(x) and _is_const(y):
            return _bin_const(op, x, y)
        for v in (x, y):
            if isinstance(v, (np.ndarray, _LoopVar, _Opaque)):
                if isinstance(v, _LoopVar):
                    raise _NotPlate()
                raise _unsupported("arithmetic on an array / discrete plate element")
        if isinstance(x, _Elem) or isinstance(y, _Elem):
            if isinstance(x, _ir.Expr) or isinstance(y, _ir.Expr):
                lat, el = (x, y) if isinstance(x, _ir.Expr) else (y, x)
                if op == "*" and _bare_latent(lat):
                    return _ir.Expr([(_ir.FC_TERM_LATENT_X, lat.terms[0][1], 0, self.column(el.arr).cid, 0.0)])
                raise _unsupported(f"`{op}` between a latent expression and a data element")
            xa = x.arr if isinstance(x, _Elem) else _as_f64(x)
            ya = y.arr if isinstance(y, _Elem) else _as_f64(y)
            with np.errstate(all="ignore"):
                return _Elem({"+": np.add, "-": np.subtract, "*": np.multiply, "/": np.divide}[op](xa, ya))
        # at least one latent expression, the other a constant or a latent expression: affine forms
        # const + coef * latent + ... stay parameter expressions (`let mu = 2.0 * p - 1.0`, dsl.rs:12-13)
        xe, ye = isinstance(x, _ir.Expr), isinstance(y, _ir.Expr)
        for v in (x, y):
            if isinstance(v, _ir.Expr) and (v.link != _ir.FC_LINK_IDENTITY or getattr(v, "discrete", False)):
                raise _unsupported("arithmetic after a link function / on a discrete latent")
        if op in "+-":
            a = x if xe else _ir.Expr([(_ir.FC_TERM_CONST, 0, 0, 0, _as_f64(x))])
            b = y if ye else _ir.Expr([(_ir.FC_TERM_CONST, 0, 0, 0, _as_f64(y))])
            if op == "-":
                if not b.is_affine():
                    raise _unsupported("subtracting a per-element latent term")
                b = b.scale(-1.0)
            out = _ir.Expr(a.terms + b.terms)
            # fold the constants into one term (keeps `a*x - b + c` within three terms)
            consts = [t for t in out.terms if t[0] == _ir.FC_TERM_CONST]
            if len(consts) > 1 and len(out.terms) > 3:
                first = out.terms.index(consts[0])
                rest = [t for t in out.terms if t[0] != _ir.FC_TERM_CONST]
                rest.insert(min(first, len(rest)), (_ir.FC_TERM_CONST, 0, 0, 0, sum(t[4] for t in consts)))
                out = _ir.Expr(rest)
            if len(out.terms) > 3:
                raise _unsupported("more than three affine terms in one expression")
            return out
        if op == "*":
            lat, c = (x, y) if xe else (y, x)
            if _is_const(c) and lat.is_affine():
                return lat if _as_f64(c) == 1.0 else lat.scale(_as_f64(c))
            raise _unsupported("product of latent expressions")
        if xe and _is_const(y) and _as_f64(y) != 0.0 and x.is_affine():
            return x.scale(1.0 / _as_f64(y))
        raise _unsupported("division by a latent expression")

    def call(self, name, args):
        if all(_is_const(a) for a in args):
            return _math(name, [_as_f64(a) for a in args])
        if any(isinstance(a, _LoopVar) for a in args):
            raise _NotPlate()
        if all(_is_const(a) or isinstance(a, _Elem) for a in args):
            n = next(a.arr.shape[0] for a in args if isinstance(a, _Elem))
            cols = [a.arr if isinstance(a, _Elem) else np.full(n, _as_f64(a)) for a in args]
            return _Elem(np.array([_math(name, [float(c[k]) for c in cols]) for k in range(n)]))
        if name == "exp" and isinstance(args[0], _ir.Expr) and args[0].link == _ir.FC_LINK_IDENTITY:
            return _ir.exp(args[0])
        if name in ("min", "max"):
            ex, c = (args[0], args[1]) if isinstance(args[0], _ir.Expr) else (args[1], args[0])
            if isinstance(ex, _ir.Expr) and _is_const(c):
                c = _as_f64(c)
                if ex.link == _ir.FC_LINK_IDENTITY:
                    lo, hi, link = float("-inf"), float("inf"), _ir.FC_LINK_CLAMP
                elif ex.link in (_ir.FC_LINK_CLAMP, _ir.FC_LINK_SIGMOID_CLAMPED):
                    lo, hi, link = ex.lo, ex.hi, ex.link
                else:
                    raise _unsupported("min/max after exp()")
                lo, hi = (max(lo, c), hi) if name == "max" else (lo, min(hi, c))
                if lo > hi:
                    raise _unsupported("clamp with lower bound above upper bound")
                return _ir.Expr(ex.terms, link, lo, hi)
        raise _unsupported(f"`{name}` of a latent expression")

    # ---- distributions (dsl.rs:828-871) ----
    def dist(self, spec, env):
        name, arg_asts = spec
        args = [self.ev(a, env) for a in arg_asts]
        for a in args:
            if isinstance(a, _LoopVar):
                raise _NotPlate()
            if not (_is_const(a) or isinstance(a, _ir.Expr)):
                raise _unsupported(f"`{name}` parameter is a data element or an array")
        p = [a if isinstance(a, _ir.Expr) else _as_f64(a) for a in args]
        if name == "Binomial":
            if isinstance(p[0], _ir.Expr):
                raise _unsupported("Binomial n depends on a latent")
            if p[0] < 0.0 or p[0] != math.floor(p[0]):
                raise FugueError(106, f"Binomial n must be a non-negative integer, got {_fmt_num(p[0])}")
        if name == "DiscreteUniform":
            if any(isinstance(v, _ir.Expr) for v in p):
                raise _unsupported("DiscreteUniform bounds depend on a latent")
            p = [float(int(v)) if math.isfinite(v) else v for v in p]  # `a[0] as i64`
        if name == "Categorical":
            # constant probabilities: K-way (fc_op.probs_col); latent ones: the two-way [p, 1-p] form
            latent = [isinstance(v, _ir.Expr) for v in p]
            if any(latent):
                if len(p) != 2:
                    raise _unsupported(f"Categorical with {len(p)} latent-dependent categories (the latent form is two-way)")
                a0, a1 = arg_asts
                comp = a1[0] == "bin" and a1[1] == "-" and a1[2][0] in ("num", "int") and float(a1[2][1]) == 1.0 and a1[3] == a0
                if not comp:
                    raise _unsupported("Categorical(p, q) with latent probabilities needs q written as `1.0 - p`")
                return M.Categorical(p[0])
            if not (all(math.isfinite(v) and v >= 0.0 for v in p) and abs(sum(p) - 1.0) <= 1e-6):  # distribution.rs:679-713
                raise FugueError(102, f"Categorical probabilities must be non-negative and sum to 1, got {p}")
            return M.Categorical(p[0]) if len(p) == 2 else M.Categorical([float(v) for v in p])
        return getattr(M, name)(*p)

    @staticmethod
    def coerce_value(vt, v):
        """observe value coercions (dsl.rs:985-1000) for a constant or an array of constants."""
        arr = np.asarray(v, dtype=np.float64)
        if vt == "f64":
            return arr
        if vt == "bool":
            return (arr != 0.0).astype(np.uint8)
        idx = np.where(np.isfinite(arr) & (arr == np.floor(arr)), arr, 0.0)  # as_index().unwrap_or(0)
        if vt in ("u64", "usize"):
            idx = np.maximum(idx, 0.0)
        return idx.astype(np.int64)

    def address(self, spec, env):
        """(name, kind): kind None = scalar address (already escaped), "plate" = addr!(name, loop index)."""
        name, ix = spec
        if ix is None:
            return M.addr(name), None
        v = self.ev(ix, env)
        if isinstance(v, _LoopVar):
            return M.addr(name), "plate"
        if not _is_const(v):
            raise _NotPlate() if self.plate_n is not None else _unsupported("address index depends on a latent")
        k = _as_index(v)
        return M.addr(name, k if k is not None else _rust_f64(_as_f64(v))), None

    # ---- statements ----
    def stmts(self, stmts, env):
        for s in stmts:
            kind = s[0]
            if kind == "let":
                try:
                    env[s[1]] = self.ev(s[2], env)
                except FugueError as ex:
                    if ex.code != _ir.FC_ERR_UNSUPPORTED_OP or self.plate_n is not None:
                        raise
                    env[s[1]] = self.sym(s[2], env)  # usable by factor(..) and pure(..) only
            elif kind == "factor":
                try:
                    lw = self.ev(s[1], env)
                except FugueError as ex:
                    if ex.code != _ir.FC_ERR_UNSUPPORTED_OP:
                        raise
                    lw = None  # not affine: a general expression over the latents
                if lw is None or isinstance(lw, (_ir.Expr, _ir.Sym)):
                    if self.plate_n is not None:
                        raise _NotPlate()  # a latent-dependent factor inside a loop: unroll into scalar factors
                    sy = self.sym(s[1], env)
                    if isinstance(sy, _ir.Sym):
                        # the playground maps a NaN log-weight to -inf (dsl.rs:922-926)
                        self.m.factor(sy, nan_is_neg_inf=True)
                        continue
                    lw = sy
                if not _is_const(lw):
                    if isinstance(lw, (_LoopVar, _Elem)):
                        raise _NotPlate()
                    raise _unsupported("factor() of this expression")
                lw = _as_f64(lw)
                reps = self.plate_n if self.plate_n is not None else 1
                if reps > 4096:
                    raise _unsupported("factor() inside a loop of more than 4096 iterations")
                for _ in range(reps):
                    self.m.factor(float("-inf") if lw != lw else lw)
            elif kind == "sample":
                _, var, aspec, dspec = s
                a, ak = self.address(aspec, env)
                if self.plate_n is not None and ak != "plate":
                    raise _NotPlate()
                try:
                    d = self.dist(dspec, env)
                except FugueError as e:
                    if e.code >= 1000:
                        raise
                    # impossible parameter region: placeholder draw + factor(-inf) (dsl.rs:926-944)
                    self.warn(f"sample `{var}`: {e}")
                    if ak == "plate":
                        site = self.m.plate_sample(a, self.plate_n, M.Normal(0.0, 1.0)).elem()
                    else:
                        site = self.m.sample(a, M.Normal(0.0, 1.0))
                    self.m.factor(float("-inf"))
                    env[var] = site
                    continue
                if ak == "plate":
                    pl = self.m.plate_sample(a, self.plate_n, d)
                    env[var] = _Opaque() if pl.discrete else pl.elem()
                else:
                    site = self.m.sample(a, d)
                    env[var] = _Opaque() if site.discrete else site
            elif kind == "observe":
                _, aspec, dspec, vast = s
                a, ak = self.address(aspec, env)
                if self.plate_n is not None and ak != "plate":
                    raise _NotPlate()
                v = self.ev(vast, env)
                if isinstance(v, _LoopVar):
                    raise _NotPlate()
                if not (_is_const(v) or isinstance(v, _Elem)):
                    raise _unsupported("observed value depends on a latent")
                try:
                    d = self.dist(dspec, env)
                except FugueError as e:
                    if e.code >= 1000:
                        raise
                    self.warn(f"observe at `{a}`: {e}")  # dsl.rs:1002-1007
                    self.m.factor(float("-inf"))
                    continue
                if ak == "plate":
                    col = v.arr if isinstance(v, _Elem) else np.full(self.plate_n, _as_f64(v))
                    self.m.plate_observe(a, d, self.column(self.coerce_value(d.value_type, col)))
                else:
                    if isinstance(v, _Elem):
                        raise _NotPlate()
                    cv = self.coerce_value(d.value_type, _as_f64(v) if not isinstance(v, bool) else float(v))
                    self.m.observe(a, d, float(cv))
            else:
                self.loop(s, env)

    def loop(self, s, env):
        _, var, lo_e, hi_e, body = s
        lo, hi = self.ev(lo_e, env), self.ev(hi_e, env)
        if not (_is_const(lo) and _is_const(hi)):
            raise _NotPlate() if self.plate_n is not None else _unsupported("loop bound depends on a latent")
        lo_v, hi_v = _as_index(lo), _as_index(hi)
        if lo_v is None:
            self.warn("`for` lower bound is not an integer"); lo_v = 0
        if hi_v is None:
            self.warn("`for` upper bound is not an integer"); hi_v = 0
        if hi_v <= lo_v:
            return
        if self.plate_n is not None:
            raise _NotPlate()  # nested loops: unroll the outer one
        # 1) as a plate: index 0..n, every address addr!(name, i), data read as arr[i]
        if lo_v == 0:
            snap = (len(self.m.ops), len(self.m.cols), self.m.n_cont, self.m.n_disc, dict(self.col_cache), len(self.warnings))
            env2 = dict(env)
            env2[var] = _LoopVar()
            self.plate_n = hi_v
            try:
                self.stmts(body, env2)
                return
            except _NotPlate:
                del self.m.ops[snap[0]:]
                del self.m.cols[snap[1]:]
                self.m.n_cont, self.m.n_disc, self.col_cache = snap[2], snap[3], snap[4]
                del self.warnings[snap[5]:]
            finally:
                self.plate_n = None
        # 2) unrolled into scalar ops
        if hi_v - lo_v > 4096:
            raise _unsupported(f"loop of {hi_v - lo_v} iterations is not a plate (addresses must be addr!(name, {var}) and data must be read as arr[{var}])")
        for i in range(lo_v, hi_v):
            env2 = dict(env)
            env2[var] = i
            self.stmts(body, env2)


def _rust_f64(x):
    """Display of an f64 as Rust prints it in make_indexed (address.rs:208-223)."""
    if x != x:
        return "NaN"
    if math.isinf(x):
        return "inf" if x > 0 else "-inf"
    return str(int(x)) if x == math.floor(x) and abs(x) < 1e16 else repr(x)


# ---------------------------------------------------------------------------------------------
# Public handle (dsl.rs:1047-1139)
# ---------------------------------------------------------------------------------------------
def _bind_data(data):
    """JSON text (object of number arrays, or a bare array bound to `data`), or the same as Python
    objects (dict of array-likes / one array-like). Booleans are coerced to 0/1 (dsl.rs:1062-1094)."""
    if data is None:
        return {}
    if isinstance(data, str):
        trimmed = data.strip()
        if trimmed == "" or trimmed == "null":
            return {}
        try:
            data = json.loads(trimmed)
        except json.JSONDecodeError as e:
            raise DslError(f"data is not valid JSON: {e}")

    def to_arr(v):
        if isinstance(v, np.ndarray):
            if v.ndim != 1:
                raise DslError("data arrays must be JSON arrays")
            return v.astype(np.float64)
        if not isinstance(v, (list, tuple)):
            raise DslError("data arrays must be JSON arrays")
        for x in v:
            if not isinstance(x, (bool, int, float)):
                raise DslError("data arrays must hold numbers or booleans")
        return np.array([float(x) for x in v], dtype=np.float64)

    if isinstance(data, (list, tuple, np.ndarray)):
        return {"data": to_arr(data)}
    if isinstance(data, dict):
        return {k: to_arr(v) for k, v in data.items()}
    raise DslError("data must be a JSON array or object of arrays")


class CompiledModel:
    """A parsed, validated and lowered model. `builder` is the op list (a ModelBuilder accepted by
    every entry point of fugue_b200.api); `warnings` holds the soft errors that were mapped to
    factor(-inf); `return_value(draws)` evaluates the `pure(..)` expression on returned draws."""

    def __init__(self, stmts, ret, data, builder, warnings):
        self.stmts, self.ret, self.data, self.builder, self._warnings = stmts, ret, data, builder, warnings

    def build(self):
        return self.builder

    def take_warnings(self):
        w, self._warnings[:] = list(self._warnings), []
        return w

    def data_summary(self):
        return "".join(f"{k}[{len(v)}] " for k, v in self.data.items())

    def return_value(self, draws):
        """`pure(expr)` for every draw: `draws` maps a top-level site address to an array of values
        (any shape; broadcast together). Only top-level `let` bindings are in scope at `pure`."""
        env = dict(self.data)
        for s in self.stmts:
            if s[0] == "sample":
                name, ix = s[2]
                if ix is None or ix[0] in ("num", "int"):
                    a = M.addr(name) if ix is None else M.addr(name, _as_index(ix[1]) if _as_index(ix[1]) is not None else ix[1])
                    if a in draws:
                        env[s[1]] = np.asarray(draws[a], dtype=np.float64)
            elif s[0] == "let":
                try:
                    env[s[1]] = _num_eval(s[2], env)
                except KeyError:
                    pass
        return _num_eval(self.ret, env)


def _num_eval(e, env):
    t = e[0]
    if t in ("num", "int", "bool"):
        return float(e[1])
    if t == "var":
        return env[e[1]]
    if t == "len":
        return float(len(_num_eval(e[1], env)))
    if t == "index":
        a, i = _num_eval(e[1], env), _num_eval(e[2], env)
        k = int(i)
        return float(a[k]) if 0 <= k < len(a) else float("nan")
    if t == "neg":
        return -_num_eval(e[1], env)
    with np.errstate(all="ignore"):
        if t == "bin":
            x, y = _num_eval(e[2], env), _num_eval(e[3], env)
            return {"+": np.add, "-": np.subtract, "*": np.multiply, "/": np.divide}[e[1]](x, y)
        a = [_num_eval(x, env) for x in e[2]]
        f = {"exp": np.exp, "ln": np.log, "log": np.log, "sqrt": np.sqrt, "abs": np.abs, "floor": np.floor, "sin": np.sin, "cos": np.cos,
             "tanh": np.tanh, "pow": np.power, "min": np.fmin, "max": np.fmax}[e[1]]
        return f(*a)


def compile(source, data=None):  # noqa: A001 (the reference's name: CompiledModel::compile)
    """Parse `source`, bind `data`, validate names statically and lower to the op list."""
    stmts, ret = _parse_program(source)
    env = _bind_data(data)
    scope = list(env.keys())
    _check_stmts(stmts, scope)
    _check_expr(ret, scope)
    builder, warnings = M.ModelBuilder(), []
    low = _Lowering(builder, warnings)
    try:
        low.stmts(stmts, dict(env))
    except _NotPlate:
        raise _unsupported("loop index used outside addr!(name, i) / arr[i]")
    return CompiledModel(stmts, ret, env, builder, warnings)
