// oracle/fo_grad.hpp -- TEST INFRASTRUCTURE ONLY.
// Value (via fo_dist.hpp, reference-faithful) + closed-form derivatives of the 17 log-densities
// w.r.t. the value and each parameter. The reference has NO analytic gradient
// (src/inference/hmc.rs:304-329 is central finite differences); these closed forms are new and
// are validated in tests/test_oracle_grad.py against central differences of fo_dist.hpp.
#pragma once
#include "../include/fugue_cuda.h"
#include "fo_dist.hpp"

namespace fo {

// digamma: recurrence up to x>=6 then asymptotic series (abs err < 1e-15 for x > 0)
inline double digamma(double x) {
  double r = 0.0;
  while (x < 10.0) { r -= 1.0 / x; x += 1.0; }
  double f = 1.0 / (x * x);
  double t = f * (-1.0 / 12.0 + f * (1.0 / 120.0 + f * (-1.0 / 252.0 + f * (1.0 / 240.0 +
             f * (-1.0 / 132.0 + f * (691.0 / 32760.0 + f * (-1.0 / 12.0)))))));
  return r + std::log(x) - 0.5 / x + t;
}

struct LpGrad { double lp, dx, dp[3]; };

// P: parameters in constructor order; x: continuous value; xi: discrete value (bool as 0/1).
inline LpGrad dist_eval(int dist, const double* P, double x, int64_t xi) {
  LpGrad g{NEG_INF, 0.0, {0.0, 0.0, 0.0}};
  switch (dist) {
    case FC_NORMAL: {
      g.lp = normal_lp(P[0], P[1], x);
      if (fin(g.lp)) { double z = (x - P[0]) / P[1]; g.dx = -z / P[1]; g.dp[0] = z / P[1]; g.dp[1] = (z * z - 1.0) / P[1]; }
    } break;
    case FC_UNIFORM: {
      g.lp = uniform_lp(P[0], P[1], x);
      if (fin(g.lp)) { double w = P[1] - P[0]; g.dp[0] = 1.0 / w; g.dp[1] = -1.0 / w; }
    } break;
    case FC_LOGNORMAL: {
      g.lp = lognormal_lp(P[0], P[1], x);
      if (fin(g.lp)) { double z = (std::log(x) - P[0]) / P[1]; g.dx = (-z / P[1] - 1.0) / x; g.dp[0] = z / P[1]; g.dp[1] = (z * z - 1.0) / P[1]; }
    } break;
    case FC_EXPONENTIAL: {
      g.lp = exponential_lp(P[0], x);
      if (fin(g.lp)) { g.dx = -P[0]; g.dp[0] = 1.0 / P[0] - x; }
    } break;
    case FC_BERNOULLI: {
      g.lp = bernoulli_lp(P[0], xi != 0);
      if (fin(g.lp)) g.dp[0] = xi ? 1.0 / P[0] : -1.0 / (1.0 - P[0]);
    } break;
    case FC_CATEGORICAL: {  // 2-way: probs = [p, 1-p] (examples/mixture_models.rs:94)
      double probs[2] = {P[0], 1.0 - P[0]};
      g.lp = (xi >= 0) ? categorical_lp(probs, 2, (uint64_t)xi) : NEG_INF;
      if (fin(g.lp)) g.dp[0] = (xi == 0) ? 1.0 / probs[0] : -1.0 / probs[1];
    } break;
    case FC_BETA: {
      g.lp = beta_lp(P[0], P[1], x);
      if (fin(g.lp) && x > 0.0 && x < 1.0) {
        double pab = digamma(P[0] + P[1]);
        g.dx = (P[0] - 1.0) / x - (P[1] - 1.0) / (1.0 - x);
        g.dp[0] = std::log(x) - (digamma(P[0]) - pab);
        g.dp[1] = std::log(1.0 - x) - (digamma(P[1]) - pab);
      }
    } break;
    case FC_GAMMA: {
      g.lp = gamma_lp(P[0], P[1], x);
      if (fin(g.lp)) { g.dx = (P[0] - 1.0) / x - P[1]; g.dp[0] = std::log(P[1]) + std::log(x) - digamma(P[0]); g.dp[1] = P[0] / P[1] - x; }
    } break;
    case FC_BINOMIAL: {
      uint64_t n = (uint64_t)P[0];
      g.lp = (xi >= 0) ? binomial_lp(n, P[1], (uint64_t)xi) : NEG_INF;
      if (fin(g.lp) && P[1] > 0.0 && P[1] < 1.0) g.dp[1] = (double)xi / P[1] - (double)(n - (uint64_t)xi) / (1.0 - P[1]);
    } break;
    case FC_POISSON: {
      g.lp = (xi >= 0) ? poisson_lp(P[0], (uint64_t)xi) : NEG_INF;
      if (fin(g.lp)) g.dp[0] = (double)xi / P[0] - 1.0;
    } break;
    case FC_STUDENT_T: {
      g.lp = student_t_lp(P[0], P[1], P[2], x);
      if (fin(g.lp)) {
        double df = P[0], s = P[2], z = (x - P[1]) / s, w = z * z / df, o = 1.0 + w;
        g.dx = -(df + 1.0) * z / (s * df * o);
        g.dp[1] = -g.dx;
        g.dp[2] = -1.0 / s + (df + 1.0) * z * z / (df * s * o);
        g.dp[0] = 0.5 * digamma((df + 1.0) / 2.0) - 0.5 * digamma(df / 2.0) - 0.5 / df -
                  0.5 * std::log1p(w) + 0.5 * (df + 1.0) * (w / df) / o;
      }
    } break;
    case FC_CAUCHY: {
      g.lp = cauchy_lp(P[0], P[1], x);
      if (fin(g.lp)) { double s = P[1], z = (x - P[0]) / s, o = 1.0 + z * z; g.dx = -2.0 * z / (s * o); g.dp[0] = -g.dx; g.dp[1] = -1.0 / s + 2.0 * z * z / (s * o); }
    } break;
    case FC_LAPLACE: {
      g.lp = laplace_lp(P[0], P[1], x);
      if (fin(g.lp)) { double b = P[1], r = x - P[0], sg = (r > 0.0) - (r < 0.0); g.dx = -sg / b; g.dp[0] = sg / b; g.dp[1] = -1.0 / b + std::fabs(r) / (b * b); }
    } break;
    case FC_WEIBULL: {
      g.lp = weibull_lp(P[0], P[1], x);
      if (fin(g.lp) && x > 0.0) {
        double k = P[0], l = P[1], t = std::pow(x / l, k);
        g.dx = (k - 1.0) / x - k * t / x;
        g.dp[0] = 1.0 / k - std::log(l) + std::log(x) - t * std::log(x / l);
        g.dp[1] = -k / l + k * t / l;
      }
    } break;
    case FC_CHI_SQUARED: {
      g.lp = chi_squared_lp(P[0], x);
      if (fin(g.lp)) { double h = P[0] / 2.0; g.dx = (h - 1.0) / x - 0.5; g.dp[0] = 0.5 * (-LN_2 - digamma(h) + std::log(x)); }
    } break;
    case FC_INVERSE_GAMMA: {
      g.lp = inverse_gamma_lp(P[0], P[1], x);
      if (fin(g.lp)) { g.dx = -(P[0] + 1.0) / x + P[1] / (x * x); g.dp[0] = std::log(P[1]) - digamma(P[0]) - std::log(x); g.dp[1] = P[0] / P[1] - 1.0 / x; }
    } break;
    case FC_DISCRETE_UNIFORM: {
      g.lp = discrete_uniform_lp((int64_t)P[0], (int64_t)P[1], xi);
    } break;
    default: break;
  }
  return g;
}
}  // namespace fo
