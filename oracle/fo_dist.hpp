// oracle/fo_dist.hpp -- TEST INFRASTRUCTURE ONLY (never linked into the product).
// CPU restatement of fugue's 17 log-densities and numerics, line-for-line in arithmetic order.
// Parity pinning: checked against the reference's known-answer tests (tests/golden/*.json,
// extracted from /root/reference/src/core/distribution.rs:2073-2355 and
// /root/reference/tests/f_dist_distributions.rs, f_dist_numerical.rs).
// Third-party arithmetic: the reference calls libm 0.2.15 (Cargo.lock:1859) lgamma; here glibc
// lgamma_r (same FreeBSD msun lineage). Build with -O2 -ffp-contract=off (rustc never fuses).
#pragma once
#include <cmath>
#include <cstdint>
#include <limits>
#include <vector>

namespace fo {
static const double NEG_INF = -std::numeric_limits<double>::infinity();
static const double POS_INF = std::numeric_limits<double>::infinity();
static const double LN_2PI = 1.8378770664093456;  // distribution.rs:206
static const double LN_PI = 1.1447298858494002;   // distribution.rs:1372
static const double LN_2 = 0.6931471805599453;    // std::f64::consts::LN_2

inline double lgam(double x) { int s; return lgamma_r(x, &s); }
inline bool fin(double x) { return std::isfinite(x); }

// distribution.rs:189-208
inline double normal_lp(double mu, double sigma, double x) {
  if (sigma <= 0.0 || !fin(sigma) || !fin(mu) || !fin(x)) return NEG_INF;
  double z = (x - mu) / sigma;
  return -0.5 * z * z - std::log(sigma) - 0.5 * LN_2PI;
}
// distribution.rs:309-330
inline double uniform_lp(double lo, double hi, double x) {
  if (lo >= hi || !fin(lo) || !fin(hi) || !fin(x)) return NEG_INF;
  if (x < lo || x >= hi) return NEG_INF;
  double width = hi - lo;
  if (width <= 0.0) return NEG_INF;
  return -std::log(width);
}
// distribution.rs:413-434
inline double lognormal_lp(double mu, double sigma, double x) {
  if (sigma <= 0.0 || !fin(sigma) || !fin(mu)) return NEG_INF;
  if (x <= 0.0 || !fin(x)) return NEG_INF;
  double lx = std::log(x);
  double z = (lx - mu) / sigma;
  return -0.5 * z * z - lx - std::log(sigma) - 0.5 * LN_2PI;
}
// distribution.rs:503-518
inline double exponential_lp(double rate, double x) {
  if (rate <= 0.0 || !fin(rate) || !fin(x)) return NEG_INF;
  if (x < 0.0) return NEG_INF;
  return std::log(rate) - rate * x;
}
// distribution.rs:598-619
inline double bernoulli_lp(double p, bool x) {
  if (p < 0.0 || p > 1.0 || !fin(p)) return NEG_INF;
  if (x) return p <= 0.0 ? NEG_INF : std::log(p);
  return p >= 1.0 ? NEG_INF : std::log(1.0 - p);
}
// distribution.rs:785-791
inline double categorical_lp(const double* probs, size_t k, uint64_t x) {
  if (x < k && probs[x] > 0.0) return std::log(probs[x]);
  return NEG_INF;
}
// distribution.rs:897-958
inline double beta_lp(double a, double b, double x) {
  if (a <= 0.0 || b <= 0.0 || !fin(a) || !fin(b) || !fin(x)) return NEG_INF;
  if (!(x >= 0.0 && x <= 1.0)) return NEG_INF;
  double log_beta_fn = lgam(a) + lgam(b) - lgam(a + b);
  if (x == 0.0) return a > 1.0 ? NEG_INF : (a < 1.0 ? POS_INF : -log_beta_fn);
  if (x == 1.0) return b > 1.0 ? NEG_INF : (b < 1.0 ? POS_INF : -log_beta_fn);
  double ln_x = std::log(x);
  double ln_1mx = std::log(1.0 - x);
  return (a - 1.0) * ln_x + (b - 1.0) * ln_1mx - log_beta_fn;
}
// distribution.rs:1039-1068 (rate parameterisation)
inline double gamma_lp(double shape, double rate, double x) {
  if (shape <= 0.0 || rate <= 0.0 || !fin(shape) || !fin(rate) || !fin(x)) return NEG_INF;
  if (x <= 0.0) return NEG_INF;
  double log_rate = std::log(rate);
  double log_x = std::log(x);
  double lg = lgam(shape);
  return shape * log_rate + (shape - 1.0) * log_x - rate * x - lg;
}
// distribution.rs:1138-1165
inline double binomial_lp(uint64_t n, double p, uint64_t k) {
  if (!fin(p) || !(p >= 0.0 && p <= 1.0)) return NEG_INF;
  if (k > n) return NEG_INF;
  if (p == 0.0) return k == 0 ? 0.0 : NEG_INF;
  if (p == 1.0) return k == n ? 0.0 : NEG_INF;
  double c = lgam((double)n + 1.0) - lgam((double)k + 1.0) - lgam((double)(n - k) + 1.0);
  return c + (double)k * std::log(p) + (double)(n - k) * std::log(1.0 - p);
}
// distribution.rs:1237-1258
inline double poisson_lp(double lambda, uint64_t k) {
  if (lambda <= 0.0 || !fin(lambda)) return NEG_INF;
  if (lambda > 700.0 && k == 0) return -lambda;
  double kf = (double)k;
  double log_lambda = std::log(lambda);
  double log_fact = lgam(kf + 1.0);
  return kf * log_lambda - lambda - log_fact;
}
// distribution.rs:1362-1381
inline double student_t_lp(double df, double loc, double scale, double x) {
  if (df <= 0.0 || scale <= 0.0 || !fin(df) || !fin(scale) || !fin(loc) || !fin(x)) return NEG_INF;
  double z = (x - loc) / scale;
  return lgam((df + 1.0) / 2.0) - lgam(df / 2.0) - 0.5 * (std::log(df) + LN_PI) - std::log(scale) -
         0.5 * (df + 1.0) * std::log1p(z * z / df);
}
// distribution.rs:1451-1459
inline double cauchy_lp(double loc, double scale, double x) {
  if (scale <= 0.0 || !fin(scale) || !fin(loc) || !fin(x)) return NEG_INF;
  double z = (x - loc) / scale;
  return -LN_PI - std::log(scale) - std::log1p(z * z);
}
// distribution.rs:1535-1541
inline double laplace_lp(double loc, double scale, double x) {
  if (scale <= 0.0 || !fin(scale) || !fin(loc) || !fin(x)) return NEG_INF;
  return -std::log(2.0 * scale) - std::fabs(x - loc) / scale;
}
// distribution.rs:1618-1644
inline double weibull_lp(double shape, double scale, double x) {
  if (shape <= 0.0 || scale <= 0.0 || !fin(shape) || !fin(scale) || !fin(x)) return NEG_INF;
  if (x < 0.0) return NEG_INF;
  if (x == 0.0) return shape > 1.0 ? NEG_INF : (shape < 1.0 ? POS_INF : -std::log(scale));
  return std::log(shape) - shape * std::log(scale) + (shape - 1.0) * std::log(x) -
         std::pow(x / scale, shape);
}
// distribution.rs:1699-1709
inline double chi_squared_lp(double k, double x) {
  if (k <= 0.0 || !fin(k) || !fin(x)) return NEG_INF;
  if (x <= 0.0) return NEG_INF;
  double h = k / 2.0;
  return -h * LN_2 - lgam(h) + (h - 1.0) * std::log(x) - x / 2.0;
}
// distribution.rs:1789-1806
inline double inverse_gamma_lp(double shape, double rate, double x) {
  if (shape <= 0.0 || rate <= 0.0 || !fin(shape) || !fin(rate) || !fin(x)) return NEG_INF;
  if (x <= 0.0) return NEG_INF;
  return shape * std::log(rate) - lgam(shape) - (shape + 1.0) * std::log(x) - rate / x;
}
// distribution.rs:1917-1932 (count in u128; full i64 range -> -64 ln 2)
inline double discrete_uniform_lp(int64_t lo, int64_t hi, int64_t x) {
  if (hi < lo) return NEG_INF;
  if (x < lo || x > hi) return NEG_INF;
  if (lo == INT64_MIN && hi == INT64_MAX) return -(64.0 * LN_2);
  unsigned __int128 count = (unsigned __int128)((__int128)hi - (__int128)lo) + 1;
  return -std::log((double)count);
}

// ---- numerics: src/core/numerical.rs ----
// numerical.rs:15-38
inline double log_sum_exp(const double* x, size_t n) {
  double mx = NEG_INF;
  for (size_t i = 0; i < n; i++) mx = std::fmax(mx, x[i]);  // fold(NEG_INFINITY, f64::max)
  if (n == 0) return NEG_INF;
  if (mx == NEG_INF) return NEG_INF;
  double s = 0.0;
  for (size_t i = 0; i < n; i++) s += std::exp(x[i] - mx);
  if (s == 0.0) return NEG_INF;
  return mx + std::log(s);
}
// numerical.rs:87-90
inline void normalize_log_probs(const double* x, size_t n, double* out) {
  double lse = log_sum_exp(x, n);
  for (size_t i = 0; i < n; i++) out[i] = std::exp(x[i] - lse);
}
// numerical.rs:101-112
// numerical.rs:123-129
inline double safe_ln(double x) { return (x <= 0.0 || !fin(x)) ? NEG_INF : std::log(x); }
inline double log1p_exp(double x) {
  if (x > 33.3) return x;
  if (x > -37.0) return std::log1p(std::exp(x));
  return std::exp(x);
}
}  // namespace fo
