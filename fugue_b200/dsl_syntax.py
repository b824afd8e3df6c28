"""Syntax of the model text: tokens, grammar, name resolution.

Written from the GRAMMAR the reference documents for its playground language
(crates/fugue-wasm/src/dsl.rs:10-20) and from its user-visible behaviour -- which texts are accepted, the
arities of the distribution / function tables (dsl.rs:807-818, 755-786) and the wording of its error
messages ("line N: expected ..., found ...") -- not from its lexer / parser code:

    program := stmt* 'pure' '(' expr ')' [';']
    stmt    := 'let' IDENT '<-' 'sample' '(' addr ',' dist ')' ';'
             | 'let' IDENT '=' expr ';'
             | 'observe' '(' addr ',' dist ',' expr ')' ';'
             | 'factor' '(' expr ')' ';'
             | 'for' IDENT 'in' expr '..' expr '{' stmt* '}'
    addr    := 'addr' '!' '(' STRING [',' expr] ')'
    dist    := IDENT ['::' 'new'] '(' [expr {',' expr}] ')' ['.' 'unwrap' '(' ')']
    expr    := infix over + - * /, prefix -, postfix [expr] and .len(), calls f(expr, ..), literals, names

Implementation: one master regular expression produces the token list; statements are PRODUCTIONS written
as data (tuples of literals, keywords and nonterminals) executed by one small interpreter; expressions go
through a precedence-climbing (Pratt) loop driven by a binding-power table; name resolution is a table of
"which fields of a statement are expressions, which name does it bind".

AST (consumed by the lowering in dsl.py): tuples
   expr: ("num", v) ("int", i) ("bool", b) ("var", n) ("index", a, i) ("len", a) ("neg", a)
         ("bin", op, a, b) ("call", f, [args])
   stmt: ("sample", name, addr, dist) ("let", name, expr) ("observe", addr, dist, value)
         ("factor", expr) ("for", var, lo, hi, [body])
   addr: (name, index_expr | None); dist: (name, [args])
"""
import re
from collections import namedtuple


class DslError(ValueError):
    """A static (compile-time) error of the model text."""


# function name -> number of arguments; distribution name -> (min, max) number of arguments
MATH_FNS = dict(exp=1, ln=1, log=1, sqrt=1, abs=1, floor=1, sin=1, cos=1, tanh=1, pow=2, min=2, max=2)
DIST_ARITY = dict(Normal=(2, 2), Uniform=(2, 2), LogNormal=(2, 2), Beta=(2, 2), Gamma=(2, 2), InverseGamma=(2, 2), Cauchy=(2, 2),
                  Laplace=(2, 2), Weibull=(2, 2), Binomial=(2, 2), DiscreteUniform=(2, 2), Exponential=(1, 1), Poisson=(1, 1),
                  Bernoulli=(1, 1), ChiSquared=(1, 1), StudentT=(3, 3), Categorical=(1, 64))


def _fmt_num(v):
    return str(int(v)) if float(v).is_integer() and abs(v) < 1e16 else repr(v)


# ---------------------------------------------------------------------------------------------
# tokens
# ---------------------------------------------------------------------------------------------
Tok = namedtuple("Tok", "kind text value pos")  # kind: ident | num | str | sym | eof

_SCAN = re.compile(r"""
    (?P<skip>  \s+ | //[^\n]* )
  | (?P<num>   [0-9]+ (?:\.[0-9]+)? (?:[eE](?:[+-][0-9]*|[0-9]+))? )     # `0..n` is 0 then `..`
  | (?P<ident> [A-Za-z_][A-Za-z0-9_]* )
  | (?P<str>   "[^"]*" )
  | (?P<sym>   <- | :: | \.\. | [(){}\[\],;+\-*/=.!] )
""", re.X)


def tokenize(src):
    out, pos = [], 0
    while pos < len(src):
        m = _SCAN.match(src, pos)
        if m is None:
            if src[pos] == '"':
                raise DslError(f"unterminated string at byte {pos}")
            raise DslError(f"unexpected character `{src[pos]}` at byte {pos}")
        kind, text = m.lastgroup, m.group()
        if kind == "num":
            try:
                v = float(text)
            except ValueError:
                raise DslError(f"bad number `{text}` at byte {pos}")
            is_float = any(ch in text for ch in ".eE")
            out.append(Tok("num", text, v if is_float else int(text), pos))
        elif kind == "str":
            out.append(Tok("str", text, text[1:-1], pos))
        elif kind != "skip":
            out.append(Tok(kind, text, text, pos))
        pos = m.end()
    out.append(Tok("eof", "", None, len(src)))
    return out


# ---------------------------------------------------------------------------------------------
# productions as data
# ---------------------------------------------------------------------------------------------
class K(str):
    """a keyword: an identifier token with exactly this text"""


class N(str):
    """a nonterminal: the parser method of this name; its result is captured"""


IDENT, STRING = N("ident"), N("string")
PRODUCTIONS = {
    "let_sample": (K("let"), IDENT, "<-", K("sample"), "(", N("addr"), ",", N("dist"), ")", ";"),
    "let_value":  (K("let"), IDENT, "=", N("expr"), ";"),
    "observe":    (K("observe"), "(", N("addr"), ",", N("dist"), ",", N("expr"), ")", ";"),
    "factor":     (K("factor"), "(", N("expr"), ")", ";"),
    "for":        (K("for"), IDENT, K("in"), N("expr"), "..", N("expr"), "{", N("block"), "}"),
    "pure":       (K("pure"), "(", N("expr"), ")"),
}
# left binding power of the infix operators; prefix minus binds tighter, postfix tightest
INFIX_BP = {"+": 10, "-": 10, "*": 20, "/": 20}


class Parser:
    def __init__(self, src):
        self.src, self.toks, self.i = src, tokenize(src), 0

    # ---- token access ----
    @property
    def cur(self):
        return self.toks[self.i]

    def ahead(self, k):
        return self.toks[min(self.i + k, len(self.toks) - 1)]

    def is_sym(self, text, k=0):
        t = self.ahead(k)
        return t.kind == "sym" and t.text == text

    def is_word(self, text, k=0):
        t = self.ahead(k)
        return t.kind == "ident" and t.text == text

    def advance(self):
        t = self.cur
        if t.kind != "eof":
            self.i += 1
        return t

    def fail(self, expected):
        t = self.cur
        if t.kind == "eof":
            shown = "end of input"
        elif t.kind == "str":
            shown = f'"{t.value}"'
        else:
            shown = f"`{_fmt_num(t.value) if t.kind == 'num' else t.text}`"
        line = self.src.count("\n", 0, t.pos) + 1
        return DslError(f"line {line}: expected {expected}, found {shown}")

    # ---- the production interpreter ----
    def run(self, name):
        got = []
        for item in PRODUCTIONS[name]:
            if isinstance(item, N):
                got.append(getattr(self, "p_" + item)())
            elif isinstance(item, K):
                if not self.is_word(item):
                    raise self.fail(f"`{item}`")
                self.advance()
            else:
                if not self.is_sym(item):
                    raise self.fail(f"`{item}`")
                self.advance()
        return got

    # ---- nonterminals ----
    def p_ident(self):
        if self.cur.kind != "ident":
            raise self.fail("an identifier")
        return self.advance().text

    def p_string(self):
        if self.cur.kind != "str":
            raise self.fail("a string literal address name")
        return self.advance().value

    def p_block(self):
        body = []
        while not self.is_sym("}"):
            if self.cur.kind == "eof":
                raise self.fail("`}`")
            body.append(self.p_stmt())
        return body

    def p_stmt(self):
        t = self.cur
        word = t.text if t.kind == "ident" else None
        if word == "let":
            name = "let_sample" if self.is_sym("<-", 2) else "let_value"
            got = self.run(name)
            return ("sample", got[0], got[1], got[2]) if name == "let_sample" else ("let", got[0], got[1])
        if word == "observe":
            addr, dist, value = self.run("observe")
            return ("observe", addr, dist, value)
        if word == "factor":
            return ("factor", self.run("factor")[0])
        if word == "for":
            var, lo, hi, body = self.run("for")
            return ("for", var, lo, hi, body)
        raise self.fail("a statement (`let`, `observe`, `factor`, `for`, or `pure`)")

    def p_program(self):
        stmts = []
        while not self.is_word("pure"):
            if self.cur.kind == "eof":
                raise DslError("model must end with `pure(<expr>)`")
            stmts.append(self.p_stmt())
        ret = self.run("pure")[0]
        if self.is_sym(";"):
            self.advance()
        if self.cur.kind != "eof":
            raise self.fail("end of input after `pure(..)`")
        return stmts, ret

    def p_addr(self):
        if not self.is_word("addr"):
            raise self.fail("`addr!(..)`")
        self.advance()
        for s in "!(":
            if not self.is_sym(s):
                raise self.fail(f"`{s}`")
            self.advance()
        name = self.p_string()
        index = None
        if self.is_sym(","):
            self.advance()
            index = self.p_expr()
        if not self.is_sym(")"):
            raise self.fail("`)`")
        self.advance()
        return (name, index)

    def args_until_paren(self):
        """'(' [expr {',' expr}] ')'"""
        if not self.is_sym("("):
            raise self.fail("`(`")
        self.advance()
        args = []
        while not self.is_sym(")"):
            args.append(self.p_expr())
            if not self.is_sym(","):
                break
            self.advance()
        if not self.is_sym(")"):
            raise self.fail("`)`")
        self.advance()
        return args

    def p_dist(self):
        name = self.p_ident()
        if self.is_sym("::"):
            self.advance()
            ctor = self.p_ident()
            if ctor != "new":
                raise DslError(f"unknown distribution constructor `{name}::{ctor}`")
        args = self.args_until_paren()
        if self.is_sym(".") and self.is_word("unwrap", 1):  # Rust-ism, accepted and ignored
            self.advance(); self.advance()
            if self.args_until_paren():
                raise DslError("`unwrap` takes no arguments")
        if name not in DIST_ARITY:
            raise DslError(f"unknown distribution `{name}`")
        lo, hi = DIST_ARITY[name]
        if not lo <= len(args) <= hi:
            raise DslError(f"`{name}` takes {lo if lo == hi else f'{lo}..={hi}'} argument(s), got {len(args)}")
        return (name, args)

    # ---- expressions: precedence climbing ----
    def p_expr(self, min_bp=0):
        left = self.operand()
        while self.cur.kind == "sym" and INFIX_BP.get(self.cur.text, -1) > min_bp:
            op = self.advance().text
            left = ("bin", op, left, self.p_expr(INFIX_BP[op]))  # left-associative: the right side stops at equal power
        return left

    def operand(self):
        if self.is_sym("-"):
            self.advance()
            return ("neg", self.operand())  # binds tighter than every infix operator
        e = self.atom()
        while True:  # postfix
            if self.is_sym("["):
                self.advance()
                idx = self.p_expr()
                if not self.is_sym("]"):
                    raise self.fail("`]`")
                self.advance()
                e = ("index", e, idx)
            elif self.is_sym(".") and self.is_word("len", 1):
                self.advance(); self.advance()
                if self.args_until_paren():
                    raise DslError("`len` takes no arguments")
                e = ("len", e)
            else:
                return e

    def atom(self):
        t = self.cur
        if t.kind == "num":
            self.advance()
            return ("int", t.value) if isinstance(t.value, int) else ("num", t.value)
        if t.kind == "sym" and t.text == "(":
            self.advance()
            e = self.p_expr()
            if not self.is_sym(")"):
                raise self.fail("`)`")
            self.advance()
            return e
        if t.kind == "ident":
            self.advance()
            if t.text in ("true", "false"):
                return ("bool", t.text == "true")
            if self.is_sym("("):
                if t.text not in MATH_FNS:
                    raise DslError(f"unknown function `{t.text}`")
                args = self.args_until_paren()
                if len(args) != MATH_FNS[t.text]:
                    raise DslError(f"`{t.text}` takes {MATH_FNS[t.text]} argument(s)")
                return ("call", t.text, args)
            return ("var", t.text)
        raise self.fail("an expression")


def parse_program(src):
    return Parser(src).p_program()


# ---------------------------------------------------------------------------------------------
# name resolution: every variable must be bound by data, an earlier `let`, or an enclosing `for`
# ---------------------------------------------------------------------------------------------
def _subexprs(e):
    tag = e[0]
    if tag in ("index", "bin"):
        return e[-2:]
    if tag in ("len", "neg"):
        return e[1:2]
    if tag == "call":
        return e[2]
    return ()


def check_expr(e, scope):
    todo = [e]
    while todo:
        x = todo.pop()
        if x[0] == "var" and x[1] not in scope:
            raise DslError(f"unknown variable `{x[1]}`")
        todo.extend(reversed(_subexprs(x)))


def _addr_dist_exprs(addr, dist):
    return ([addr[1]] if addr[1] is not None else []) + list(dist[1])


# statement tag -> (expressions read, in order; name bound afterwards or None)
_READS = {
    "sample":  lambda s: (_addr_dist_exprs(s[2], s[3]), s[1]),
    "let":     lambda s: ([s[2]], s[1]),
    "observe": lambda s: (_addr_dist_exprs(s[1], s[2]) + [s[3]], None),
    "factor":  lambda s: ([s[1]], None),
}


def check_stmts(stmts, scope):
    for s in stmts:
        if s[0] == "for":
            _, var, lo, hi, body = s
            check_expr(lo, scope)
            check_expr(hi, scope)
            check_stmts(body, scope + [var])  # the loop variable and the body's lets end with the loop
            continue
        reads, binds = _READS[s[0]](s)
        for e in reads:
            check_expr(e, scope)
        if binds is not None:
            scope.append(binds)
