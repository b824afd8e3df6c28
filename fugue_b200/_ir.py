"""ctypes mirror of include/fugue_cuda.h plus the flat op-list builder.

The builder is the Python-side stand-in for the lowering handler north_star describes
(run the model once, record its static sample/observe structure). Surface names follow the
reference: sample / observe / factor (src/core/model.rs:144,258,397), plate
(src/macros/mod.rs:82-86), addresses "name" / "name#i" (src/core/address.rs:208-223).
"""
import ctypes as C
import numpy as np

# ---- enums (include/fugue_cuda.h) ----
(FC_NORMAL, FC_UNIFORM, FC_LOGNORMAL, FC_EXPONENTIAL, FC_BERNOULLI, FC_CATEGORICAL, FC_BETA, FC_GAMMA,
 FC_BINOMIAL, FC_POISSON, FC_STUDENT_T, FC_CAUCHY, FC_LAPLACE, FC_WEIBULL, FC_CHI_SQUARED,
 FC_INVERSE_GAMMA, FC_DISCRETE_UNIFORM) = range(17)
DIST_IDS = {"Normal": 0, "Uniform": 1, "LogNormal": 2, "Exponential": 3, "Bernoulli": 4, "Categorical": 5,
            "Beta": 6, "Gamma": 7, "Binomial": 8, "Poisson": 9, "StudentT": 10, "Cauchy": 11, "Laplace": 12,
            "Weibull": 13, "ChiSquared": 14, "InverseGamma": 15, "DiscreteUniform": 16}
FC_OP_SAMPLE, FC_OP_OBSERVE, FC_OP_FACTOR, FC_OP_PLATE_SAMPLE, FC_OP_PLATE_OBSERVE, FC_OP_FACTOR_EXPR = range(6)
(FC_X_CONST, FC_X_LATENT, FC_X_ADD, FC_X_SUB, FC_X_MUL, FC_X_DIV, FC_X_NEG, FC_X_EXP, FC_X_LN, FC_X_SQRT, FC_X_ABS, FC_X_FLOOR,
 FC_X_SIN, FC_X_COS, FC_X_TANH, FC_X_POW, FC_X_MIN, FC_X_MAX) = range(18)
FC_EXPR_MAX_NODES = 64
FC_T_F64, FC_T_I64, FC_T_U64, FC_T_USIZE, FC_T_BOOL = range(5)
(FC_TERM_NONE, FC_TERM_CONST, FC_TERM_LATENT, FC_TERM_LATENT_X, FC_TERM_DENSE, FC_TERM_GATHER,
 FC_TERM_SELECT2, FC_TERM_PLATE_LATENT, FC_TERM_SCALED) = range(9)
FC_LINK_IDENTITY, FC_LINK_EXP, FC_LINK_SIGMOID_CLAMPED, FC_LINK_CLAMP = range(4)
FC_COL_F64, FC_COL_U8, FC_COL_U32, FC_COL_I64 = range(4)
FC_RESAMPLE_MULTINOMIAL, FC_RESAMPLE_SYSTEMATIC, FC_RESAMPLE_STRATIFIED = range(3)

FC_OK = 0
FC_ERR_ADDRESS_CONFLICT = 301
FC_ERR_UNEXPECTED_MODEL_STRUCTURE = 302
FC_ERR_UNSUPPORTED_OP = 1000
FC_ERR_CUDA = 1001
FC_ERR_INVALID_ARGUMENT = 1003


class fc_term(C.Structure):
    _fields_ = [("kind", C.c_int32), ("slot", C.c_int32), ("slot_b", C.c_int32), ("col", C.c_int32),
                ("c", C.c_double)]


class fc_param(C.Structure):
    _fields_ = [("term", fc_term * 3), ("link", C.c_int32), ("_pad", C.c_int32), ("lo", C.c_double),
                ("hi", C.c_double)]


class fc_op(C.Structure):
    _fields_ = [("kind", C.c_int32), ("dist", C.c_int32), ("value_type", C.c_int32), ("n_params", C.c_int32),
                ("name", C.c_char_p), ("plate_len", C.c_int64), ("value_col", C.c_int32), ("probs_col", C.c_int32),
                ("value", C.c_double), ("param", fc_param * 3), ("expr_first", C.c_int32), ("expr_len", C.c_int32)]


class fc_expr_node(C.Structure):
    _fields_ = [("op", C.c_int32), ("a", C.c_int32), ("b", C.c_int32), ("slot", C.c_int32), ("c", C.c_double)]


class fc_column(C.Structure):
    _fields_ = [("dtype", C.c_int32), ("width", C.c_int32), ("rows", C.c_int64), ("data", C.c_void_p)]


class fc_hmc_config(C.Structure):
    _fields_ = [("n_leapfrog", C.c_uint64), ("target_accept", C.c_double), ("init_step_size", C.c_double),
                ("finite_diff_eps", C.c_double), ("adapt_mass", C.c_int32), ("grad_mode", C.c_int32)]


class fc_hmc_stats(C.Structure):
    _fields_ = [("n_accept", C.c_uint64), ("n_divergent", C.c_uint64), ("step_size", C.c_double)]


class fc_mh_config(C.Structure):
    _fields_ = [("target_accept", C.c_double), ("gamma", C.c_double)]


class fc_mh_stats(C.Structure):
    _fields_ = [("n_accept", C.c_uint64), ("n_steps", C.c_uint64)]


class fc_smc_config(C.Structure):
    _fields_ = [("resampling_method", C.c_int32), ("_pad", C.c_int32), ("ess_threshold", C.c_double),
                ("rejuvenation_steps", C.c_uint64)]


class FugueError(Exception):
    """Typed error mirroring FugueError/ErrorCode (src/error.rs:40-58,161-190)."""

    def __init__(self, code, msg=""):
        super().__init__(f"[{code}] {msg}")
        self.code = code


# ---------------------------------------------------------------------------------------------
# Expressions
# ---------------------------------------------------------------------------------------------
class Column:
    def __init__(self, cid, arr):
        self.cid, self.arr = cid, arr

    def __len__(self):
        return self.arr.shape[0]


class Expr:
    """link(term0 + term1 + term2), evaluated left to right (fc_param)."""

    def __init__(self, terms, link=FC_LINK_IDENTITY, lo=0.0, hi=0.0):
        self.terms, self.link, self.lo, self.hi = list(terms), link, lo, hi

    def __add__(self, o):
        if isinstance(o, Sym):
            return Sym.apply(FC_X_ADD, self, o)
        o = as_expr(o)
        if self.link != FC_LINK_IDENTITY or o.link != FC_LINK_IDENTITY:
            return Sym.apply(FC_X_ADD, self, o)  # arithmetic after a link: a general (factor) expression
        return Expr(self.terms + o.terms)

    __radd__ = lambda self, o: as_expr(o).__add__(self)

    def is_affine(self):
        """const + coef * latent + ... with the identity link (what scale() / negation keep affine)."""
        return self.link == FC_LINK_IDENTITY and not getattr(self, "discrete", False) and all(k in (FC_TERM_CONST, FC_TERM_LATENT, FC_TERM_SCALED) for k, *_ in self.terms)

    def scale(self, c):
        """c * (const + coef * latent + ...): CONST terms are multiplied, latents become SCALED terms."""
        c = float(c)
        out = []
        for k, slot, sb, col, v in self.terms:
            if k == FC_TERM_CONST:
                out.append((FC_TERM_CONST, 0, 0, 0, c * v))
            elif k == FC_TERM_LATENT:
                out.append((FC_TERM_SCALED, slot, 0, 0, c))
            else:
                out.append((FC_TERM_SCALED, slot, 0, 0, c * v))
        return Expr(out)

    def __mul__(self, col):
        if isinstance(col, Column):
            assert len(self.terms) == 1 and self.terms[0][0] == FC_TERM_LATENT
            return Expr([(FC_TERM_LATENT_X, self.terms[0][1], 0, col.cid, 0.0)])
        if isinstance(col, (int, float)) and self.is_affine():
            return self.scale(col)
        return Sym.apply(FC_X_MUL, self, col)  # latent * latent, ...: a general (factor) expression

    def __rmul__(self, o):
        return self.scale(o) if (isinstance(o, (int, float)) and self.is_affine()) else Sym.apply(FC_X_MUL, o, self)

    def __neg__(self):
        return self.scale(-1.0) if self.is_affine() else Sym.apply(FC_X_NEG, self)

    def __sub__(self, o):
        if isinstance(o, (int, float)) and self.is_affine():
            return Expr(self.terms + [(FC_TERM_CONST, 0, 0, 0, -float(o))])
        if isinstance(o, Expr) and not isinstance(o, Sym) and self.is_affine() and o.is_affine() and len(self.terms) + len(o.terms) <= 3:
            return Expr(self.terms + o.scale(-1.0).terms)
        return Sym.apply(FC_X_SUB, self, o)

    def __rsub__(self, o):
        if isinstance(o, (int, float)) and self.is_affine():
            return Expr([(FC_TERM_CONST, 0, 0, 0, float(o))] + self.scale(-1.0).terms)
        return Sym.apply(FC_X_SUB, o, self)

    def __truediv__(self, o):
        if isinstance(o, (int, float)) and o != 0 and self.is_affine():
            return self.scale(1.0 / float(o))
        return Sym.apply(FC_X_DIV, self, o)

    # everything that is not link(term + term + term) becomes a general expression (factor(..) only)
    __rtruediv__ = lambda self, o: Sym.apply(FC_X_DIV, o, self)
    __pow__ = lambda self, o: Sym.apply(FC_X_POW, self, o)


class Sym:
    """A general scalar expression over continuous latents (for factor(..)): SSA nodes
    (op, a, b, slot, c) in evaluation order -- fc_expr_node. Built with + - * / unary -, ** and the
    functions exp ln sqrt abs floor sin cos tanh pow min max of `fugue_b200.model`."""

    def __init__(self, nodes):
        self.nodes = nodes

    @staticmethod
    def lift(x):
        if isinstance(x, Sym):
            return x
        if isinstance(x, Expr):  # a scalar parameter expression: link(const + latent + ...)
            if getattr(x, "discrete", False) or any(k not in (FC_TERM_CONST, FC_TERM_LATENT, FC_TERM_SCALED) for k, *_ in x.terms):
                raise FugueError(FC_ERR_UNSUPPORTED_OP, "only continuous scalar latents and constants can enter a factor expression")
            acc = None
            for k, slot, _, _, c in x.terms:
                t = Sym([(FC_X_LATENT, 0, 0, slot, 0.0)]) if k != FC_TERM_CONST else Sym([(FC_X_CONST, 0, 0, 0, float(c))])
                if k == FC_TERM_SCALED:
                    t = Sym.apply(FC_X_MUL, float(c), t)
                acc = t if acc is None else Sym.apply(FC_X_ADD, acc, t)
            if x.link == FC_LINK_EXP:
                acc = Sym.apply(FC_X_EXP, acc)
            elif x.link == FC_LINK_SIGMOID_CLAMPED:
                acc = Sym.apply(FC_X_DIV, 1.0, Sym.apply(FC_X_ADD, 1.0, Sym.apply(FC_X_EXP, Sym.apply(FC_X_NEG, acc))))
            if x.link in (FC_LINK_SIGMOID_CLAMPED, FC_LINK_CLAMP):
                acc = Sym.apply(FC_X_MAX, float(x.lo), Sym.apply(FC_X_MIN, acc, float(x.hi)))
            return acc
        return Sym([(FC_X_CONST, 0, 0, 0, float(x))])

    @staticmethod
    def apply(op, *args):
        nodes, roots = [], []
        for a in args:
            a = Sym.lift(a)
            off = len(nodes)
            for (o, x, y, s, c) in a.nodes:
                nodes.append((o, x + off if o >= FC_X_ADD else 0, y + off if ((FC_X_ADD <= o <= FC_X_DIV) or o >= FC_X_POW) else 0, s, c))
            roots.append(len(nodes) - 1)
        nodes.append((op, roots[0], roots[1] if len(roots) > 1 else 0, 0, 0.0))
        if len(nodes) > FC_EXPR_MAX_NODES:
            raise FugueError(FC_ERR_UNSUPPORTED_OP, f"factor expression has more than {FC_EXPR_MAX_NODES} nodes")
        return Sym(nodes)

    __add__ = lambda a, b: Sym.apply(FC_X_ADD, a, b)
    __radd__ = lambda a, b: Sym.apply(FC_X_ADD, b, a)
    __sub__ = lambda a, b: Sym.apply(FC_X_SUB, a, b)
    __rsub__ = lambda a, b: Sym.apply(FC_X_SUB, b, a)
    __mul__ = lambda a, b: Sym.apply(FC_X_MUL, a, b)
    __rmul__ = lambda a, b: Sym.apply(FC_X_MUL, b, a)
    __truediv__ = lambda a, b: Sym.apply(FC_X_DIV, a, b)
    __rtruediv__ = lambda a, b: Sym.apply(FC_X_DIV, b, a)
    __neg__ = lambda a: Sym.apply(FC_X_NEG, a)
    __pow__ = lambda a, b: Sym.apply(FC_X_POW, a, b)


class Site(Expr):
    """A scalar latent (result of sample())."""

    def __init__(self, slot, discrete=False):
        super().__init__([(FC_TERM_LATENT, slot, 0, 0, 0.0)])
        self.slot, self.discrete = slot, discrete


class Plate:
    """Result of a plate sample: `len` consecutive slots starting at slot0."""

    def __init__(self, slot0, n, discrete, op_index):
        self.slot0, self.n, self.discrete, self.op_index = slot0, n, discrete, op_index

    def __getitem__(self, idx):  # alphas[group_ids[i]] -> gather
        assert isinstance(idx, Column) and not self.discrete
        return Expr([(FC_TERM_GATHER, self.slot0, 0, idx.cid, 0.0)])

    def elem(self):  # element i of a continuous plate, aligned with the enclosing plate
        return Expr([(FC_TERM_PLATE_LATENT, self.slot0, 0, 0, 0.0)])


def as_expr(v):
    if isinstance(v, Expr):
        return v
    return Expr([(FC_TERM_CONST, 0, 0, 0, float(v))])


def dense(X, plate):
    """sum_k plate[k] * X[i][k], accumulated left to right from 0.0 (classification.rs:119-122)."""
    assert X.arr.ndim == 2 and X.arr.shape[1] == plate.n and X.arr.dtype == np.float64
    return Expr([(FC_TERM_DENSE, plate.slot0, 0, X.cid, 0.0)])


def select2(z_plate, a, b):
    """if z[i] == 0 { a } else { b } (mixture_models.rs:99-102)."""
    assert z_plate.discrete and isinstance(a, Site) and isinstance(b, Site)
    return Expr([(FC_TERM_SELECT2, a.slot, b.slot, z_plate.op_index, 0.0)])


def _link(e, link, lo=0.0, hi=0.0):
    e = as_expr(e)
    assert e.link == FC_LINK_IDENTITY
    return Expr(e.terms, link, lo, hi)


def exp(e):
    return _link(e, FC_LINK_EXP)


def sigmoid_clamped(e, lo, hi):
    """(1/(1+exp(-e))).clamp(lo, hi) (classification.rs:124-127)."""
    return _link(e, FC_LINK_SIGMOID_CLAMPED, lo, hi)


def clamp(e, lo, hi):
    return _link(e, FC_LINK_CLAMP, lo, hi)
