#!/usr/bin/env python3
"""Extracts the reference's known-answer log_prob / numerics tests into JSON fixtures.

Run HERE (the container that has /root/reference); the fixtures are committed because
/root/reference does not exist on the GPU box. Sources scanned:
  src/core/distribution.rs  (#[cfg(test)] module, lines ~2073-2593)
  tests/f_dist_distributions.rs, tests/f_dist_numerical.rs, src/core/numerical.rs (tests)
Only the *test expectations* (distribution, constructor args, point, expected value) are
extracted -- no reference source code is copied.
"""
import json, math, os, re, sys

REF = "/root/reference"
OUT = os.path.dirname(os.path.abspath(__file__))

DISTS = ["Normal", "Uniform", "LogNormal", "Exponential", "Bernoulli", "Categorical", "Beta", "Gamma",
         "Binomial", "Poisson", "StudentT", "Cauchy", "Laplace", "Weibull", "ChiSquared", "InverseGamma",
         "DiscreteUniform"]


def rust_num(s):
    """Evaluate a small Rust numeric expression."""
    s = s.strip().rstrip(",").strip()
    s = s.replace("std::f64::consts::LN_2", "math.log(2.0)").replace("f64::NEG_INFINITY", "float('-inf')")
    s = s.replace("f64::INFINITY", "float('inf')").replace("f64::MIN_POSITIVE", "2.2250738585072014e-308")
    s = s.replace("f64::NAN", "float('nan')").replace("f64::MAX", "1.7976931348623157e308")
    s = s.replace("i64::MIN", "(-2**63)").replace("i64::MAX", "(2**63-1)").replace("u64::MAX", "(2**64-1)")
    s = re.sub(r"\(?(-?[\d.]+)_f64\)?\.ln\(\)", r"math.log(\1)", s)
    s = re.sub(r"(\d)_(\d)", r"\1\2", s)
    s = re.sub(r"(\d)(_?[fiu](64|32|size))", r"\1", s)
    s = s.replace("true", "1").replace("false", "0")
    if "vec!" in s:
        s = s.replace("vec!", "")
    return eval(s, {"math": math, "__builtins__": {}, "float": float})


def split_args(s):
    out, depth, cur = [], 0, ""
    for ch in s:
        if ch in "([": depth += 1
        if ch in ")]": depth -= 1
        if ch == "," and depth == 0:
            out.append(cur); cur = ""
        else:
            cur += ch
    if cur.strip(): out.append(cur)
    return out


def scan(path, lo=0, hi=10**9):
    src = open(os.path.join(REF, path)).read().split("\n")
    text = "\n".join(src[lo:hi])
    text = re.sub(r"//[^\n]*", "", text)
    flat = re.sub(r"\s+", " ", text)
    cases, skipped = [], 0
    # let-bound distributions, scoped per fn
    for fn in re.split(r"\bfn\s+", flat):
        fname = fn.split("(")[0]
        binds = {}
        for m in re.finditer(r"let (\w+) = (?:fugue::)?(\w+)::new\(((?:[^()]|\([^()]*\))*)\)\s*\.unwrap\(\);", fn):
            if m.group(2) in DISTS: binds[m.group(1)] = (m.group(2), m.group(3))
        pat = r"(close|assert_eq!)\(\s*(?:(?:fugue::)?(\w+)::new\(((?:[^()]|\([^()]*\))*)\)\s*\.unwrap\(\)|(\w+))\s*\.log_prob\(&((?:[^()]|\([^()]*\))*?)\),\s*([^;]*?)\s*,?\s*\);"
        for m in re.finditer(pat, fn):
            kind, dist, args, var, x, exp = m.groups()
            if var:
                if var not in binds: skipped += 1; continue
                dist, args = binds[var]
            if dist not in DISTS: skipped += 1; continue
            try:
                a = [rust_num(t) for t in split_args(args)]
                xv = rust_num(x)
                ev = rust_num(exp)
            except Exception:
                skipped += 1; continue
            cases.append({"dist": dist, "args": a, "x": xv, "expected": ev, "exact": kind == "assert_eq!",
                          "src": f"{path}::{fname}"})
    return cases, skipped


def enc(o):
    if isinstance(o, float):
        if math.isinf(o): return "inf" if o > 0 else "-inf"
        if math.isnan(o): return "nan"
    if isinstance(o, list): return [enc(v) for v in o]
    if isinstance(o, dict): return {k: enc(v) for k, v in o.items()}
    return o


def main():
    all_cases, sk = [], 0
    for path, lo, hi in [("src/core/distribution.rs", 2040, 2600), ("tests/f_dist_distributions.rs", 0, 10**9)]:
        c, s = scan(path, lo, hi)
        all_cases += c; sk += s
    json.dump(enc(all_cases), open(os.path.join(OUT, "dist_logprob.json"), "w"), indent=1)
    print(f"dist_logprob.json: {len(all_cases)} cases ({sk} expressions skipped)")
    by = {}
    for c in all_cases: by[c["dist"]] = by.get(c["dist"], 0) + 1
    print(by)


if __name__ == "__main__":
    main()
