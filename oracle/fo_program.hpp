// oracle/fo_program.hpp -- TEST INFRASTRUCTURE ONLY.
// Sequential interpreter of the flat op list: the CPU restatement of
//   run(ScoreGivenTrace{base}, model)   src/runtime/interpreters.rs:76-84,138-163
//   Trace accumulators                   src/runtime/trace.rs:168-200
//   the trampoline's program order       src/runtime/handler.rs:124-209
// One op at a time, one plate element at a time, left-to-right f64 accumulation.
// Model link forms follow examples/classification.rs:118-130, hierarchical_models.rs:26-28,
// mixture_models.rs:93-103, linear_regression.rs:41.
#pragma once
#include <algorithm>
#include <cstring>
#include <string>
#include "fo_grad.hpp"

namespace fo {

struct Column { int dtype, width; int64_t rows; const void* data; };

inline double col_f64(const Column& c, int64_t i, int k = 0) {
  int64_t j = i * c.width + k;
  switch (c.dtype) {
    case FC_COL_F64: return ((const double*)c.data)[j];
    case FC_COL_U8: return (double)((const uint8_t*)c.data)[j];
    case FC_COL_U32: return (double)((const uint32_t*)c.data)[j];
    default: return (double)((const int64_t*)c.data)[j];
  }
}
inline int64_t col_i64(const Column& c, int64_t i) {
  switch (c.dtype) {
    case FC_COL_F64: return (int64_t)((const double*)c.data)[i];
    case FC_COL_U8: return ((const uint8_t*)c.data)[i];
    case FC_COL_U32: return ((const uint32_t*)c.data)[i];
    default: return ((const int64_t*)c.data)[i];
  }
}

// src/core/address.rs:175-223
inline std::string escape_seg(const std::string& s) {
  std::string o;
  for (char ch : s) { if (ch == '\\') o += "\\\\"; else if (ch == '#') o += "\\#"; else o += ch; }
  return o;
}

struct Program {
  std::vector<fc_op> ops;
  std::vector<std::string> names;
  std::vector<Column> cols;
  std::vector<int64_t> cslot0, dslot0;  // per op: first continuous / discrete program slot (-1)
  int64_t d_cont = 0, d_disc = 0;
  std::vector<int64_t> perm_c, perm_d;  // sorted index -> program slot
  std::vector<std::string> site_c, site_d;  // program-order addresses
  std::vector<fc_expr_node> nodes;          // FACTOR_EXPR table (latent slots in PROGRAM order)

  int build(const fc_op* o, size_t n, const fc_column* c, size_t nc, const fc_expr_node* nd = nullptr, size_t nn = 0) {
    nodes.assign(nd, nd + (nd ? nn : 0));
    ops.assign(o, o + n);
    names.resize(n);
    for (size_t i = 0; i < n; i++) { names[i] = o[i].name ? o[i].name : ""; ops[i].name = nullptr; }
    cols.resize(nc);
    for (size_t i = 0; i < nc; i++) cols[i] = Column{c[i].dtype, c[i].width, c[i].rows, c[i].data};
    cslot0.assign(n, -1); dslot0.assign(n, -1);
    for (size_t i = 0; i < n; i++) {
      const fc_op& op = ops[i];
      if (op.kind != FC_OP_SAMPLE && op.kind != FC_OP_PLATE_SAMPLE) continue;
      int64_t len = op.kind == FC_OP_SAMPLE ? 1 : op.plate_len;
      bool cont = op.value_type == FC_T_F64;
      (cont ? cslot0 : dslot0)[i] = cont ? d_cont : d_disc;
      for (int64_t e = 0; e < len; e++) {
        std::string a = escape_seg(names[i]);
        if (op.kind == FC_OP_PLATE_SAMPLE) a += "#" + std::to_string(e);
        (cont ? site_c : site_d).push_back(a);
      }
      (cont ? d_cont : d_disc) += len;
    }
    auto mk = [](const std::vector<std::string>& s, std::vector<int64_t>& p) {
      p.resize(s.size());
      for (size_t i = 0; i < s.size(); i++) p[i] = (int64_t)i;
      std::sort(p.begin(), p.end(), [&](int64_t a, int64_t b) { return s[a] < s[b]; });
    };
    mk(site_c, perm_c); mk(site_d, perm_d);
    for (size_t i = 1; i < perm_c.size(); i++) if (site_c[perm_c[i]] == site_c[perm_c[i - 1]]) return FC_ERR_ADDRESS_CONFLICT;
    for (size_t i = 1; i < perm_d.size(); i++) if (site_d[perm_d[i]] == site_d[perm_d[i - 1]]) return FC_ERR_ADDRESS_CONFLICT;
    return FC_OK;
  }
};

struct Score { double log_prior = 0.0, log_lik = 0.0, log_factors = 0.0; double total() const { return log_prior + log_lik + log_factors; } };

// Evaluates one parameter expression at plate element i; q/qd in PROGRAM order.
// If dq != nullptr, adds  w * d(param)/dq  into dq (program order).
struct ParamEval {
  const Program& P; const double* q; const int64_t* qd;
  double value(const fc_param& pr, int64_t i, double* dlink) const {
    double s = 0.0; bool first = true;
    for (int t = 0; t < 3; t++) {
      const fc_term& tm = pr.term[t];
      if (tm.kind == FC_TERM_NONE) continue;
      double v = 0.0;
      switch (tm.kind) {
        case FC_TERM_CONST: v = tm.c; break;
        case FC_TERM_LATENT: v = q[tm.slot]; break;
        case FC_TERM_SCALED: v = tm.c * q[tm.slot]; break;
        case FC_TERM_LATENT_X: v = q[tm.slot] * col_f64(P.cols[tm.col], i); break;
        case FC_TERM_DENSE: {  // classification.rs:119-122: linear_pred = 0.0; += coef*x
          const Column& X = P.cols[tm.col];
          double a = 0.0;
          for (int k = 0; k < X.width; k++) a += q[tm.slot + k] * col_f64(X, i, k);
          v = a;
        } break;
        case FC_TERM_GATHER: v = q[tm.slot + col_i64(P.cols[tm.col], i)]; break;
        case FC_TERM_SELECT2: v = (qd[P.dslot0[tm.col] + i] == 0) ? q[tm.slot] : q[tm.slot_b]; break;
        case FC_TERM_PLATE_LATENT: v = q[tm.slot + i]; break;
      }
      if (first) { s = v; first = false; } else s = s + v;
    }
    switch (pr.link) {
      case FC_LINK_EXP: { double e = std::exp(s); *dlink = e; return e; }
      case FC_LINK_SIGMOID_CLAMPED: {  // classification.rs:124-127
        double p = 1.0 / (1.0 + std::exp(-s));
        bool inside = (p > pr.lo && p < pr.hi);
        *dlink = inside ? p * (1.0 - p) : 0.0;
        return p < pr.lo ? pr.lo : (p > pr.hi ? pr.hi : p);
      }
      case FC_LINK_CLAMP: {  // mixture_models.rs:93
        bool inside = (s > pr.lo && s < pr.hi);
        *dlink = inside ? 1.0 : 0.0;
        return s < pr.lo ? pr.lo : (s > pr.hi ? pr.hi : s);
      }
      default: *dlink = 1.0; return s;
    }
  }
  void backprop(const fc_param& pr, int64_t i, double w, double* dq) const {
    for (int t = 0; t < 3; t++) {
      const fc_term& tm = pr.term[t];
      switch (tm.kind) {
        case FC_TERM_LATENT: dq[tm.slot] += w; break;
        case FC_TERM_SCALED: dq[tm.slot] += w * tm.c; break;
        case FC_TERM_LATENT_X: dq[tm.slot] += w * col_f64(P.cols[tm.col], i); break;
        case FC_TERM_DENSE: { const Column& X = P.cols[tm.col]; for (int k = 0; k < X.width; k++) dq[tm.slot + k] += w * col_f64(X, i, k); } break;
        case FC_TERM_GATHER: dq[tm.slot + col_i64(P.cols[tm.col], i)] += w; break;
        case FC_TERM_SELECT2: dq[(qd[P.dslot0[tm.col] + i] == 0) ? tm.slot : tm.slot_b] += w; break;
        case FC_TERM_PLATE_LATENT: dq[tm.slot + i] += w; break;
        default: break;
      }
    }
  }
};

// factor(f(latents)) -- the value the playground's evaluator computes for the expression
// (crates/fugue-wasm/src/dsl.rs:755-786: f64 arithmetic, exp ln sqrt abs floor sin cos tanh, powf,
// f64::min / f64::max) and its gradient by FORWARD-mode differentiation, one latent at a time
// (independent of the device's reverse sweep). NaN -> -inf when the op asks for the playground's
// policy (dsl.rs:922-926).
struct Dual { double v, d; };
inline Dual expr_forward(const fc_expr_node* nd, int n, const double* q, int wrt_slot) {
  Dual t[FC_EXPR_MAX_NODES];
  for (int k = 0; k < n; k++) {
    const fc_expr_node& e = nd[k];
    Dual a = e.op >= FC_X_ADD ? t[e.a] : Dual{0, 0};
    Dual b = ((e.op >= FC_X_ADD && e.op <= FC_X_DIV) || e.op >= FC_X_POW) ? t[e.b] : Dual{0, 0};
    Dual r{0, 0};
    switch (e.op) {
      case FC_X_CONST: r = {e.c, 0.0}; break;
      case FC_X_LATENT: r = {q[e.slot], e.slot == wrt_slot ? 1.0 : 0.0}; break;
      case FC_X_ADD: r = {a.v + b.v, a.d + b.d}; break;
      case FC_X_SUB: r = {a.v - b.v, a.d - b.d}; break;
      case FC_X_MUL: r = {a.v * b.v, a.d * b.v + a.v * b.d}; break;
      case FC_X_DIV: r = {a.v / b.v, a.d / b.v - a.v * b.d / (b.v * b.v)}; break;
      case FC_X_NEG: r = {-a.v, -a.d}; break;
      case FC_X_EXP: { double x = std::exp(a.v); r = {x, x * a.d}; } break;
      case FC_X_LN: r = {std::log(a.v), a.d / a.v}; break;
      case FC_X_SQRT: { double x = std::sqrt(a.v); r = {x, 0.5 * a.d / x}; } break;
      case FC_X_ABS: r = {std::fabs(a.v), a.v < 0.0 ? -a.d : a.d}; break;
      case FC_X_FLOOR: r = {std::floor(a.v), 0.0}; break;
      case FC_X_SIN: r = {std::sin(a.v), std::cos(a.v) * a.d}; break;
      case FC_X_COS: r = {std::cos(a.v), -std::sin(a.v) * a.d}; break;
      case FC_X_TANH: { double x = std::tanh(a.v); r = {x, (1.0 - x * x) * a.d}; } break;
      case FC_X_POW: {
        double x = std::pow(a.v, b.v);
        double d = b.v * std::pow(a.v, b.v - 1.0) * a.d;
        if (nd[e.b].op != FC_X_CONST) d += x * std::log(a.v) * b.d;
        r = {x, d};
      } break;
      case FC_X_MIN: r = (a.v <= b.v || b.v != b.v) ? a : b; if (a.v != a.v) r = b; break;  // f64::min ignores a NaN operand
      default: r = (a.v >= b.v || b.v != b.v) ? a : b; if (a.v != a.v) r = b; break;
    }
    t[k] = r;
  }
  return t[n - 1];
}

// q, qd in PROGRAM order. dq (program order, may be null) receives d total / dq.
inline Score score_program_order(const Program& P, const double* q, const int64_t* qd, double* dq) {
  Score sc;
  if (dq) std::fill(dq, dq + P.d_cont, 0.0);
  ParamEval pe{P, q, qd};
  for (size_t oi = 0; oi < P.ops.size(); oi++) {
    const fc_op& op = P.ops[oi];
    if (op.kind == FC_OP_FACTOR) { sc.log_factors += op.value; continue; }
    if (op.kind == FC_OP_FACTOR_EXPR) {
      const fc_expr_node* nd = P.nodes.data() + op.expr_first;
      double v = expr_forward(nd, op.expr_len, q, -1).v;
      if (op.value != 0.0 && v != v) v = -INFINITY;
      sc.log_factors += v;
      if (dq && std::isfinite(v))
        for (int k = 0; k < op.expr_len; k++)
          if (nd[k].op == FC_X_LATENT) {
            bool seen = false;  // one forward sweep per distinct latent
            for (int k2 = 0; k2 < k; k2++) seen |= (nd[k2].op == FC_X_LATENT && nd[k2].slot == nd[k].slot);
            if (!seen) dq[nd[k].slot] += expr_forward(nd, op.expr_len, q, nd[k].slot).d;
          }
      continue;
    }
    bool is_sample = (op.kind == FC_OP_SAMPLE || op.kind == FC_OP_PLATE_SAMPLE);
    bool plate = (op.kind == FC_OP_PLATE_SAMPLE || op.kind == FC_OP_PLATE_OBSERVE);
    int64_t len = plate ? op.plate_len : 1;
    bool cont = op.value_type == FC_T_F64;
    for (int64_t i = 0; i < len; i++) {
      double pv[3] = {0, 0, 0}, dl[3] = {0, 0, 0};
      for (int j = 0; j < op.n_params; j++) pv[j] = pe.value(op.param[j], i, &dl[j]);
      double x = 0.0; int64_t xi = 0;
      if (is_sample) {
        if (cont) x = q[P.cslot0[oi] + i]; else xi = qd[P.dslot0[oi] + i];
      } else if (plate) {
        if (cont) x = col_f64(P.cols[op.value_col], i); else xi = col_i64(P.cols[op.value_col], i);
      } else {
        if (cont) x = op.value; else xi = (int64_t)op.value;
      }
      LpGrad g;
      if (op.dist == FC_CATEGORICAL && op.n_params == 0) {  // K-way, constant probabilities (distribution.rs:785-791)
        const Column& pc = P.cols[op.probs_col];
        g = LpGrad{(xi >= 0) ? categorical_lp((const double*)pc.data, (size_t)pc.rows, (uint64_t)xi) : NEG_INF, 0.0, {0.0, 0.0, 0.0}};
      } else g = dist_eval(op.dist, pv, x, xi);
      if (is_sample) sc.log_prior += g.lp; else sc.log_lik += g.lp;
      if (dq && g.lp == NEG_INF) {  // outside the support there is no gradient: the reference's FD force is NaN (hmc.rs:304-329)
        for (int64_t k = 0; k < P.d_cont; k++) dq[k] = std::nan("");
      }
      if (dq) {
        if (is_sample && cont) dq[P.cslot0[oi] + i] += g.dx;
        for (int j = 0; j < op.n_params; j++) {
          double w = g.dp[j] * dl[j];
          if (w != 0.0) pe.backprop(op.param[j], i, w, dq);
        }
      }
    }
  }
  return sc;
}

// q, qd, grad in SORTED order, strided (element s at q[s*stride]).
inline Score score_sorted(const Program& P, const double* q, const int64_t* qd, int64_t stride, double* grad) {
  std::vector<double> qp(P.d_cont), gp(grad ? P.d_cont : 0);
  std::vector<int64_t> dp(P.d_disc);
  for (int64_t s = 0; s < P.d_cont; s++) qp[P.perm_c[s]] = q[s * stride];
  for (int64_t s = 0; s < P.d_disc; s++) dp[P.perm_d[s]] = qd[s * stride];
  Score sc = score_program_order(P, qp.data(), dp.data(), grad ? gp.data() : nullptr);
  if (grad) for (int64_t s = 0; s < P.d_cont; s++) grad[s * stride] = gp[P.perm_c[s]];
  return sc;
}

// Reference-faithful central finite-difference force: src/inference/hmc.rs:304-329.
// Returns ok=false if any component is non-finite.
inline bool grad_fd_sorted(const Program& P, const double* q, const int64_t* qd, int64_t stride, double h, double* grad) {
  std::vector<double> w(P.d_cont);
  for (int64_t s = 0; s < P.d_cont; s++) w[s] = q[s * stride];
  std::vector<int64_t> dd(P.d_disc);
  for (int64_t s = 0; s < P.d_disc; s++) dd[s] = qd[s * stride];
  bool ok = true;
  for (int64_t s = 0; s < P.d_cont; s++) {
    double orig = w[s];
    w[s] = orig + h; double lp = score_sorted(P, w.data(), dd.data(), 1, nullptr).total();
    w[s] = orig - h; double lm = score_sorted(P, w.data(), dd.data(), 1, nullptr).total();
    w[s] = orig;
    double g = (lp - lm) / (2.0 * h);
    if (!fin(g)) ok = false;
    grad[s * stride] = g;
  }
  return ok;
}
}  // namespace fo
