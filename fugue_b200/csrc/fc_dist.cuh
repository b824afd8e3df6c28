// fc_dist.cuh -- device log-densities of fugue's 17 distributions with analytic derivatives.
// Value formulas follow /root/reference/src/core/distribution.rs (line ranges beside each
// case); every guard returns -inf exactly where the reference does. The derivatives
// (d/dx and d/dparam_j) are new -- the reference differentiates numerically
// (src/inference/hmc.rs:304-329) -- and are checked against the CPU oracle's closed forms.
#pragma once
#include "fc_common.cuh"

namespace fc {

#define FC_LN_2PI 1.8378770664093456  /* distribution.rs:206 */
#define FC_LN_PI 1.1447298858494002   /* distribution.rs:1372 */
#define FC_LN_2 0.6931471805599453

struct DistOut {
  double lp, dx, dp0, dp1, dp2;
};

// digamma for x > 0: upward recurrence to x >= 10, then the asymptotic series.
__device__ __forceinline__ double d_digamma(double x) {
  double r = 0.0;
  while (x < 10.0) {
    r -= 1.0 / x;
    x += 1.0;
  }
  double f = 1.0 / (x * x);
  double t = f * (-1.0 / 12.0 + f * (1.0 / 120.0 + f * (-1.0 / 252.0 + f * (1.0 / 240.0 + f * (-1.0 / 132.0 + f * (691.0 / 32760.0 + f * (-1.0 / 12.0)))))));
  return r + log(x) - 0.5 / x + t;
}

// lgc, lgc2: optional functions of CONSTANT parameters evaluated once at lowering (DOp::lgc / lgc2; NaN =>
// compute here): Normal ln sigma; Gamma lgamma(k), ln rate; Beta ln B(a, b).
template <bool GRAD>
__device__ __forceinline__ DistOut dist_eval(int dist, double P0, double P1, double P2, double x, long long xi, double lgam_kp1, double lgc, double lgc2);
template <bool GRAD>
__device__ __forceinline__ DistOut dist_eval(int dist, double P0, double P1, double P2, double x, long long xi, double lgam_kp1) {
  const double qn = __longlong_as_double(0x7ff8000000000000LL);
  return dist_eval<GRAD>(dist, P0, P1, P2, x, xi, lgam_kp1, qn, qn);
}
template <bool GRAD>
__device__ __forceinline__ DistOut dist_eval(int dist, double P0, double P1, double P2, double x, long long xi, double lgam_kp1, double lgc) {
  return dist_eval<GRAD>(dist, P0, P1, P2, x, xi, lgam_kp1, lgc, __longlong_as_double(0x7ff8000000000000LL));
}
// lgam_kp1: optional precomputed lgamma(k+1) for count-valued data (NaN => compute here).
// A term whose coefficient is exactly zero ((k - 1) ln x at k = 1, ...) is not evaluated in value-only
// mode: 0 * ln x is +-0 for every x the guards let through, so the sum is unchanged.
template <bool GRAD>
__device__ __forceinline__ DistOut dist_eval(int dist, double P0, double P1, double P2, double x, long long xi, double lgam_kp1, double lgc, double lgc2) {
  DistOut g;
  g.lp = d_neg_inf();
  g.dx = g.dp0 = g.dp1 = g.dp2 = 0.0;
  switch (dist) {
    case FC_NORMAL: {  // distribution.rs:189-208
      if (P1 <= 0.0 || !d_fin(P1) || !d_fin(P0) || !d_fin(x)) break;
      double z = (x - P0) / P1;
      g.lp = -0.5 * z * z - (isnan(lgc) ? log(P1) : lgc) - 0.5 * FC_LN_2PI;
      if (GRAD && d_fin(g.lp)) {
        g.dp0 = z / P1;
        g.dx = -g.dp0;
        g.dp1 = (z * z - 1.0) / P1;
      }
    } break;
    case FC_UNIFORM: {  // distribution.rs:309-330
      if (P0 >= P1 || !d_fin(P0) || !d_fin(P1) || !d_fin(x)) break;
      if (x < P0 || x >= P1) break;
      double w = P1 - P0;
      if (w <= 0.0) break;
      g.lp = -log(w);
      if (GRAD && d_fin(g.lp)) {
        g.dp0 = 1.0 / w;
        g.dp1 = -1.0 / w;
      }
    } break;
    case FC_LOGNORMAL: {  // distribution.rs:413-434
      if (P1 <= 0.0 || !d_fin(P1) || !d_fin(P0)) break;
      if (x <= 0.0 || !d_fin(x)) break;
      double lx = log(x);
      double z = (lx - P0) / P1;
      g.lp = -0.5 * z * z - lx - log(P1) - 0.5 * FC_LN_2PI;
      if (GRAD && d_fin(g.lp)) {
        g.dx = (-z / P1 - 1.0) / x;
        g.dp0 = z / P1;
        g.dp1 = (z * z - 1.0) / P1;
      }
    } break;
    case FC_EXPONENTIAL: {  // distribution.rs:503-518
      if (P0 <= 0.0 || !d_fin(P0) || !d_fin(x)) break;
      if (x < 0.0) break;
      g.lp = log(P0) - P0 * x;
      if (GRAD && d_fin(g.lp)) {
        g.dx = -P0;
        g.dp0 = 1.0 / P0 - x;
      }
    } break;
    case FC_BERNOULLI: {  // distribution.rs:598-619
      if (P0 < 0.0 || P0 > 1.0 || !d_fin(P0)) break;
      if (xi != 0) {
        if (P0 <= 0.0) break;
        g.lp = log(P0);
        if (GRAD && d_fin(g.lp)) g.dp0 = 1.0 / P0;
      } else {
        if (P0 >= 1.0) break;
        g.lp = log(1.0 - P0);
        if (GRAD && d_fin(g.lp)) g.dp0 = -1.0 / (1.0 - P0);
      }
    } break;
    case FC_CATEGORICAL: {  // distribution.rs:785-791, two-way probs = [p, 1-p]
      double pr = (xi == 0) ? P0 : 1.0 - P0;
      if (xi < 0 || xi > 1 || !(pr > 0.0)) break;
      g.lp = log(pr);
      if (GRAD && d_fin(g.lp)) g.dp0 = (xi == 0) ? 1.0 / pr : -1.0 / pr;
    } break;
    case FC_BETA: {  // distribution.rs:897-958
      if (P0 <= 0.0 || P1 <= 0.0 || !d_fin(P0) || !d_fin(P1) || !d_fin(x)) break;
      if (!(x >= 0.0 && x <= 1.0)) break;
      double log_beta_fn = isnan(lgc) ? lgamma(P0) + lgamma(P1) - lgamma(P0 + P1) : lgc;
      if (x == 0.0) {
        g.lp = P0 > 1.0 ? d_neg_inf() : (P0 < 1.0 ? d_pos_inf() : -log_beta_fn);
        break;
      }
      if (x == 1.0) {
        g.lp = P1 > 1.0 ? d_neg_inf() : (P1 < 1.0 ? d_pos_inf() : -log_beta_fn);
        break;
      }
      const double ln_x = (GRAD || P0 != 1.0) ? log(x) : 0.0, ln_1mx = (GRAD || P1 != 1.0) ? log(1.0 - x) : 0.0;
      g.lp = (P0 - 1.0) * ln_x + (P1 - 1.0) * ln_1mx - log_beta_fn;
      if (GRAD && d_fin(g.lp)) {
        double pab = d_digamma(P0 + P1);
        g.dx = (P0 - 1.0) / x - (P1 - 1.0) / (1.0 - x);
        g.dp0 = ln_x - (d_digamma(P0) - pab);
        g.dp1 = ln_1mx - (d_digamma(P1) - pab);
      }
    } break;
    case FC_GAMMA: {  // distribution.rs:1039-1068 (shape, rate)
      if (P0 <= 0.0 || P1 <= 0.0 || !d_fin(P0) || !d_fin(P1) || !d_fin(x)) break;
      if (x <= 0.0) break;
      const double log_rate = isnan(lgc2) ? log(P1) : lgc2, log_x = (GRAD || P0 != 1.0) ? log(x) : 0.0;
      g.lp = P0 * log_rate + (P0 - 1.0) * log_x - P1 * x - (isnan(lgc) ? lgamma(P0) : lgc);
      if (GRAD && d_fin(g.lp)) {
        g.dx = (P0 - 1.0) / x - P1;
        g.dp0 = log_rate + log_x - d_digamma(P0);
        g.dp1 = P0 / P1 - x;
      }
    } break;
    case FC_BINOMIAL: {  // distribution.rs:1138-1165; P0 = n (as f64), P1 = p
      if (!d_fin(P1) || !(P1 >= 0.0 && P1 <= 1.0)) break;
      if (xi < 0) break;
      unsigned long long n = (unsigned long long)P0, k = (unsigned long long)xi;
      if (k > n) break;
      if (P1 == 0.0) { g.lp = (k == 0) ? 0.0 : d_neg_inf(); break; }
      if (P1 == 1.0) { g.lp = (k == n) ? 0.0 : d_neg_inf(); break; }
      double lk = isnan(lgam_kp1) ? lgamma((double)k + 1.0) : lgam_kp1;
      double c = lgamma((double)n + 1.0) - lk - lgamma((double)(n - k) + 1.0);
      g.lp = c + (double)k * log(P1) + (double)(n - k) * log(1.0 - P1);
      if (GRAD && d_fin(g.lp)) g.dp1 = (double)k / P1 - (double)(n - k) / (1.0 - P1);
    } break;
    case FC_POISSON: {  // distribution.rs:1237-1258
      if (P0 <= 0.0 || !d_fin(P0)) break;
      if (xi < 0) break;
      double kf = (double)(unsigned long long)xi;
      if (P0 > 700.0 && xi == 0) {
        g.lp = -P0;
      } else {
        double lf = isnan(lgam_kp1) ? lgamma(kf + 1.0) : lgam_kp1;
        g.lp = kf * log(P0) - P0 - lf;
      }
      if (GRAD && d_fin(g.lp)) g.dp0 = kf / P0 - 1.0;
    } break;
    case FC_STUDENT_T: {  // distribution.rs:1362-1381 (df, loc, scale)
      if (P0 <= 0.0 || P2 <= 0.0 || !d_fin(P0) || !d_fin(P2) || !d_fin(P1) || !d_fin(x)) break;
      double z = (x - P1) / P2;
      double w = z * z / P0;
      g.lp = lgamma((P0 + 1.0) / 2.0) - lgamma(P0 / 2.0) - 0.5 * (log(P0) + FC_LN_PI) - log(P2) - 0.5 * (P0 + 1.0) * log1p(w);
      if (GRAD && d_fin(g.lp)) {
        double o = 1.0 + w;
        g.dx = -(P0 + 1.0) * z / (P2 * P0 * o);
        g.dp1 = -g.dx;
        g.dp2 = -1.0 / P2 + (P0 + 1.0) * z * z / (P0 * P2 * o);
        g.dp0 = 0.5 * d_digamma((P0 + 1.0) / 2.0) - 0.5 * d_digamma(P0 / 2.0) - 0.5 / P0 - 0.5 * log1p(w) + 0.5 * (P0 + 1.0) * (w / P0) / o;
      }
    } break;
    case FC_CAUCHY: {  // distribution.rs:1451-1459
      if (P1 <= 0.0 || !d_fin(P1) || !d_fin(P0) || !d_fin(x)) break;
      double z = (x - P0) / P1;
      g.lp = -FC_LN_PI - log(P1) - log1p(z * z);
      if (GRAD && d_fin(g.lp)) {
        double o = 1.0 + z * z;
        g.dx = -2.0 * z / (P1 * o);
        g.dp0 = -g.dx;
        g.dp1 = -1.0 / P1 + 2.0 * z * z / (P1 * o);
      }
    } break;
    case FC_LAPLACE: {  // distribution.rs:1535-1541
      if (P1 <= 0.0 || !d_fin(P1) || !d_fin(P0) || !d_fin(x)) break;
      double r = x - P0;
      g.lp = -log(2.0 * P1) - fabs(r) / P1;
      if (GRAD && d_fin(g.lp)) {
        double sg = (r > 0.0) ? 1.0 : ((r < 0.0) ? -1.0 : 0.0);
        g.dx = -sg / P1;
        g.dp0 = sg / P1;
        g.dp1 = -1.0 / P1 + fabs(r) / (P1 * P1);
      }
    } break;
    case FC_WEIBULL: {  // distribution.rs:1618-1644 (shape k, scale lambda)
      if (P0 <= 0.0 || P1 <= 0.0 || !d_fin(P0) || !d_fin(P1) || !d_fin(x)) break;
      if (x < 0.0) break;
      if (x == 0.0) {
        g.lp = P0 > 1.0 ? d_neg_inf() : (P0 < 1.0 ? d_pos_inf() : -log(P1));
        break;
      }
      double t = pow(x / P1, P0);
      g.lp = log(P0) - P0 * log(P1) + (P0 - 1.0) * log(x) - t;
      if (GRAD && d_fin(g.lp)) {
        g.dx = (P0 - 1.0) / x - P0 * t / x;
        g.dp0 = 1.0 / P0 - log(P1) + log(x) - t * log(x / P1);
        g.dp1 = -P0 / P1 + P0 * t / P1;
      }
    } break;
    case FC_CHI_SQUARED: {  // distribution.rs:1699-1709
      if (P0 <= 0.0 || !d_fin(P0) || !d_fin(x)) break;
      if (x <= 0.0) break;
      double h = P0 / 2.0;
      g.lp = -h * FC_LN_2 - lgamma(h) + (h - 1.0) * log(x) - x / 2.0;
      if (GRAD && d_fin(g.lp)) {
        g.dx = (h - 1.0) / x - 0.5;
        g.dp0 = 0.5 * (-FC_LN_2 - d_digamma(h) + log(x));
      }
    } break;
    case FC_INVERSE_GAMMA: {  // distribution.rs:1789-1806 (shape, rate)
      if (P0 <= 0.0 || P1 <= 0.0 || !d_fin(P0) || !d_fin(P1) || !d_fin(x)) break;
      if (x <= 0.0) break;
      g.lp = P0 * log(P1) - lgamma(P0) - (P0 + 1.0) * log(x) - P1 / x;
      if (GRAD && d_fin(g.lp)) {
        g.dx = -(P0 + 1.0) / x + P1 / (x * x);
        g.dp0 = log(P1) - d_digamma(P0) - log(x);
        g.dp1 = P0 / P1 - 1.0 / x;
      }
    } break;
    case FC_DISCRETE_UNIFORM: {  // distribution.rs:1917-1932; P0 = lo, P1 = hi (as f64)
      long long lo = (long long)P0, hi = (long long)P1;
      if (hi < lo || xi < lo || xi > hi) break;
      unsigned long long span = (unsigned long long)hi - (unsigned long long)lo;  // count - 1
      g.lp = (span == 0xffffffffffffffffULL) ? -(64.0 * FC_LN_2) : -log((double)span + 1.0);
    } break;
    default: break;
  }
  // A point outside the support (or an invalid parameter) has no gradient: the reference's finite-difference force
  // is NaN there, which marks the trajectory divergent (hmc.rs:304-329, 384-386). A zero force would let it glide on.
  if (GRAD && g.lp == d_neg_inf()) g.dx = g.dp0 = g.dp1 = g.dp2 = __longlong_as_double(0x7ff8000000000000LL);
  return g;
}

}  // namespace fc
