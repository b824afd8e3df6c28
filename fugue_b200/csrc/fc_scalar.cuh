// fc_scalar.cuh -- the scalar-op interpreter (sample / observe / factor ops outside plates) as a device
// function: shared by k_scalar (fc_score.cu) and the fused SMC kernels (fc_smc.cu).
// Replaces the per-site virtual calls of ScoreGivenTrace (src/runtime/interpreters.rs:138-163).
#pragma once
#include "fc_plate.cuh"

namespace fc {

// factor(f(latents)): forward sweep over the SSA nodes, then (GRAD) a reverse sweep that pushes the
// adjoints into the gradient rows. Functions as the playground evaluates them (dsl.rs:755-786);
// abs / floor / min / max use the derivative of the active branch (0 for floor).
template <bool GRAD>
__device__ __noinline__ double factor_expr_eval(const fc_expr_node* __restrict__ nd, int n, const double* __restrict__ q, int64_t C, int64_t c, double* __restrict__ grad, bool nan_is_neg_inf) {
  double v[FC_EXPR_MAX_NODES];
  for (int k = 0; k < n; k++) {
    const fc_expr_node e = nd[k];
    const double a = e.op >= FC_X_ADD ? v[e.a] : 0.0;
    const double b = ((e.op >= FC_X_ADD && e.op <= FC_X_DIV) || e.op >= FC_X_POW) ? v[e.b] : 0.0;
    double r;
    switch (e.op) {
      case FC_X_CONST: r = e.c; break;
      case FC_X_LATENT: r = q[(int64_t)e.slot * C + c]; break;
      case FC_X_ADD: r = a + b; break;
      case FC_X_SUB: r = a - b; break;
      case FC_X_MUL: r = a * b; break;
      case FC_X_DIV: r = a / b; break;
      case FC_X_NEG: r = -a; break;
      case FC_X_EXP: r = exp(a); break;
      case FC_X_LN: r = log(a); break;
      case FC_X_SQRT: r = sqrt(a); break;
      case FC_X_ABS: r = fabs(a); break;
      case FC_X_FLOOR: r = floor(a); break;
      case FC_X_SIN: r = sin(a); break;
      case FC_X_COS: r = cos(a); break;
      case FC_X_TANH: r = tanh(a); break;
      case FC_X_POW: r = pow(a, b); break;
      case FC_X_MIN: r = fmin(a, b); break;  // f64::min: the non-NaN operand
      default: r = fmax(a, b); break;
    }
    v[k] = r;
  }
  double out = v[n - 1];
  if (nan_is_neg_inf && out != out) return d_neg_inf();
  if (GRAD && d_fin(out)) {
    double adj[FC_EXPR_MAX_NODES];
    for (int k = 0; k < n; k++) adj[k] = 0.0;
    adj[n - 1] = 1.0;
    for (int k = n - 1; k >= 0; k--) {
      const double g = adj[k];
      if (g == 0.0) continue;
      const fc_expr_node e = nd[k];
      const double a = e.op >= FC_X_ADD ? v[e.a] : 0.0;
      const double b = ((e.op >= FC_X_ADD && e.op <= FC_X_DIV) || e.op >= FC_X_POW) ? v[e.b] : 0.0;
      switch (e.op) {
        case FC_X_LATENT: grad[(int64_t)e.slot * C + c] += g; break;
        case FC_X_ADD: adj[e.a] += g; adj[e.b] += g; break;
        case FC_X_SUB: adj[e.a] += g; adj[e.b] -= g; break;
        case FC_X_MUL: adj[e.a] += g * b; adj[e.b] += g * a; break;
        case FC_X_DIV: adj[e.a] += g / b; adj[e.b] -= g * a / (b * b); break;
        case FC_X_NEG: adj[e.a] -= g; break;
        case FC_X_EXP: adj[e.a] += g * v[k]; break;
        case FC_X_LN: adj[e.a] += g / a; break;
        case FC_X_SQRT: adj[e.a] += g * 0.5 / v[k]; break;
        case FC_X_ABS: adj[e.a] += a < 0.0 ? -g : g; break;
        case FC_X_SIN: adj[e.a] += g * cos(a); break;
        case FC_X_COS: adj[e.a] -= g * sin(a); break;
        case FC_X_TANH: adj[e.a] += g * (1.0 - v[k] * v[k]); break;
        case FC_X_POW:
          adj[e.a] += g * b * pow(a, b - 1.0);
          if (nd[e.b].op != FC_X_CONST) adj[e.b] += g * v[k] * log(a);  // constant exponents (the common case) take no adjoint
          break;
        case FC_X_MIN: if (a <= b || b != b) adj[e.a] += g; else adj[e.b] += g; break;
        case FC_X_MAX: if (a >= b || b != b) adj[e.a] += g; else adj[e.b] += g; break;
        default: break;  // CONST, FLOOR
      }
    }
  }
  return out;
}

// One chain's walk over the scalar ops in program order. `za(dslot)` returns the value of the discrete
// program slot `dslot` for this chain (the caller knows where discrete values live).
template <bool GRAD, class ZA>
__device__ __forceinline__ void scalar_ops_eval(const DOp* __restrict__ ops, int n_ops, const int32_t* __restrict__ crow, const double* __restrict__ q, int64_t C, int64_t c,
                                                const ZA& za, double* __restrict__ grad, const fc_expr_node* __restrict__ nodes, double& a_prior, double& a_lik, double& a_fac) {
  for (int oi = 0; oi < n_ops; oi++) {
    const DOp& op = ops[oi];
    if (op.kind == FC_OP_FACTOR) { a_fac += op.value; continue; }
    if (op.kind == FC_OP_FACTOR_EXPR) { a_fac += factor_expr_eval<GRAD>(nodes + op.expr_first, op.expr_len, q, C, c, grad, op.value != 0.0); continue; }
    double pv[3] = {0.0, 0.0, 0.0}, dl[3] = {0.0, 0.0, 0.0};
    for (int j = 0; j < op.n_params; j++) {
      const DParam& pr = op.p[j];
      double s = 0.0;
      bool first = true;
      for (int t = 0; t < 3; t++) {
        const DTerm& tm = pr.t[t];
        if (tm.kind == FC_TERM_NONE) continue;
        double v = tm.kind == FC_TERM_CONST ? tm.c : tm.c * q[(int64_t)tm.row * C + c];  // LATENT: coef * q
        s = first ? v : s + v;
        first = false;
      }
      pv[j] = apply_link(pr.link, s, pr.lo, pr.hi, dl[j]);
    }
    double x = 0.0;
    long long xi = 0;
    if (op.is_sample) {
      if (op.cont) x = q[(int64_t)crow[op.cslot0] * C + c];
      else xi = za((int)op.dslot0);
    } else {
      if (op.cont) x = op.value;
      else xi = (long long)op.value;
    }
    DistOut g = op_dist_eval<GRAD>(op, pv[0], pv[1], pv[2], x, xi, __longlong_as_double(0x7ff8000000000000LL));
    if (op.is_sample) a_prior += g.lp;
    else a_lik += g.lp;
    if (GRAD) {
      if (op.is_sample && op.cont) grad[(int64_t)crow[op.cslot0] * C + c] += g.dx;
      double dpj[3] = {g.dp0, g.dp1, g.dp2};
      for (int j = 0; j < op.n_params; j++) {
        double w = dpj[j] * dl[j];
        if (w != 0.0)
          for (int t = 0; t < 3; t++)
            if (op.p[j].t[t].kind == FC_TERM_LATENT) grad[(int64_t)op.p[j].t[t].row * C + c] += w * op.p[j].t[t].c;
      }
    }
  }
}

}  // namespace fc
