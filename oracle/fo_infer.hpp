// oracle/fo_infer.hpp -- TEST INFRASTRUCTURE ONLY.
// CPU restatement of the sampler-side arithmetic. Randomness is passed in explicitly
// (normals z, uniforms u) so device and oracle can be compared on identical draws.
#pragma once
#include "fo_program.hpp"

namespace fo {

// ---- src/inference/hmc.rs:141-184 ----
struct DualAveraging {
  double mu, log_eps_bar = 0.0, h_bar = 0.0; uint64_t m = 0;
  double gamma = 0.05, t0 = 10.0, kappa = 0.75, target;
  DualAveraging(double eps0, double tgt) : mu(std::log(10.0 * eps0)), target(tgt) {}
  double update(double alpha) {
    m += 1; double mf = (double)m;
    double a = alpha < 0.0 ? 0.0 : (alpha > 1.0 ? 1.0 : alpha);
    double frac = 1.0 / (mf + t0);
    h_bar = (1.0 - frac) * h_bar + frac * (target - a);
    double log_eps = mu - (std::sqrt(mf) / gamma) * h_bar;
    double w = std::pow(mf, -kappa);
    log_eps_bar = w * log_eps + (1.0 - w) * log_eps_bar;
    return std::exp(log_eps);
  }
  double frozen_step() const { return std::exp(log_eps_bar); }
};

// ---- src/inference/hmc.rs:187-233 ----
struct Welford {
  uint64_t n = 0; std::vector<double> mean, m2;
  explicit Welford(size_t d) : mean(d, 0.0), m2(d, 0.0) {}
  void push(const double* x) {
    n += 1; double nf = (double)n;
    for (size_t i = 0; i < mean.size(); i++) {
      double delta = x[i] - mean[i]; mean[i] += delta / nf; double d2 = x[i] - mean[i]; m2[i] += delta * d2;
    }
  }
  std::vector<double> variances() const {
    std::vector<double> v(mean.size(), 1.0);
    if (n < 2) return v;
    double denom = (double)(n - 1);
    for (size_t i = 0; i < v.size(); i++) { double s = m2[i] / denom; v[i] = (fin(s) && s > 1e-8) ? s : 1.0; }
    return v;
  }
};

// grad_mode 0 = analytic closed form, 1 = reference central differences (hmc.rs:304-329)
struct HmcTarget {
  const Program& P; const int64_t* qd; int grad_mode; double h;
  double logp(const double* q) const { return score_sorted(P, q, qd, 1, nullptr).total(); }
  bool grad(const double* q, double* g) const {
    if (grad_mode == 1) return grad_fd_sorted(P, q, qd, 1, h, g);
    score_sorted(P, q, qd, 1, g);
    bool ok = true;
    for (int64_t i = 0; i < P.d_cont; i++) if (!fin(g[i])) ok = false;
    return ok;
  }
};

// ---- src/inference/hmc.rs:353-407; returns divergent ----
inline bool leapfrog(const HmcTarget& T, std::vector<double>& q, std::vector<double>& p, double eps, size_t l, const double* m_inv) {
  size_t d = q.size(); std::vector<double> g(d);
  if (!T.grad(q.data(), g.data())) return true;
  for (size_t s = 0; s < l; s++) {
    for (size_t i = 0; i < d; i++) p[i] += 0.5 * eps * g[i];
    for (size_t i = 0; i < d; i++) q[i] += eps * m_inv[i] * p[i];
    if (!T.grad(q.data(), g.data())) return true;
    for (size_t i = 0; i < d; i++) p[i] += 0.5 * eps * g[i];
  }
  return false;
}
inline double kinetic(const std::vector<double>& p, const double* m_inv) {
  double s = 0.0; for (size_t i = 0; i < p.size(); i++) s += p[i] * p[i] * m_inv[i]; return 0.5 * s;
}

struct TransitionOut { bool accepted, divergent; double accept_prob, lj_new; };
// ---- src/inference/hmc.rs:419-473; z = d StandardNormal draws, u = the accept uniform ----
inline TransitionOut hmc_transition(const HmcTarget& T, std::vector<double>& q, double& lj_cur, const double* z, double u,
                                    double eps, size_t l, const double* m_inv, const double* mass_sqrt) {
  size_t d = q.size(); std::vector<double> p(d), qn(q);
  for (size_t i = 0; i < d; i++) p[i] = z[i] * mass_sqrt[i];
  double h0 = -lj_cur + kinetic(p, m_inv);
  if (leapfrog(T, qn, p, eps, l, m_inv)) return {false, true, 0.0, lj_cur};
  double lj_new = T.logp(qn.data());
  if (!fin(lj_new)) return {false, true, 0.0, lj_cur};
  double h_new = -lj_new + kinetic(p, m_inv);
  double ap = std::fmin(std::exp(h0 - h_new), 1.0);
  bool acc = u < ap;
  if (acc) { q = qn; lj_cur = lj_new; }
  return {acc, false, ap, lj_new};
}

// ---- src/inference/hmc.rs:479-535 ----
inline double find_reasonable_epsilon(const HmcTarget& T, const std::vector<double>& q, double lj_q, const double* z,
                                      const double* m_inv, const double* mass_sqrt) {
  size_t d = q.size(); std::vector<double> p0(d);
  for (size_t i = 0; i < d; i++) p0[i] = z[i] * mass_sqrt[i];
  double h0 = -lj_q + kinetic(p0, m_inv);
  auto lr_at = [&](double eps) -> double {
    std::vector<double> q1(q), p1(p0);
    if (leapfrog(T, q1, p1, eps, 1, m_inv)) return NEG_INF;
    double lj1 = T.logp(q1.data());
    if (!fin(lj1)) return NEG_INF;
    return h0 - (-lj1 + kinetic(p1, m_inv));
  };
  double eps = 1.0, lr = lr_at(eps);
  double ln_half = std::log(0.5), ln2 = std::log(2.0);
  double a = lr > ln_half ? 1.0 : -1.0;
  unsigned iters = 0;
  while (a * lr > -a * ln2) {
    eps *= std::pow(2.0, a);
    lr = lr_at(eps);
    iters++;
    if (iters > 100 || !(eps >= 1e-12 && eps <= 1e12)) break;
    if (a > 0.0 && lr == NEG_INF) { eps /= 2.0; break; }
  }
  return eps < 1e-6 ? 1e-6 : (eps > 1e3 ? 1e3 : eps);
}

// ---- src/inference/mcmc_utils.rs:51-150 (one address) ----
struct DimAdapt {
  double scale = 1.0, log_scale = 0.0; uint64_t acc = 0, tot = 0; double target, gamma;
  DimAdapt(double t = 0.44, double g = 0.7) : target(t), gamma(g) {}
  void update(bool accepted) {
    tot += 1; if (accepted) acc += 1;
    if (tot < 10) return;
    double rate = (double)acc / (double)tot;
    double step = 1.0 / std::pow((double)tot, gamma);
    log_scale += step * (rate - target);
    double ns = std::exp(log_scale);
    scale = (fin(ns) && ns > 0.0) ? (ns < 0.001 ? 0.001 : (ns > 100.0 ? 100.0 : ns)) : 1.0;
    log_scale = (scale == 1.0) ? 0.0 : std::log(scale);
  }
};

// ---- src/inference/smc.rs:230-233 ----
inline double ess(const double* w, size_t n) { double s = 0.0; for (size_t i = 0; i < n; i++) s += w[i] * w[i]; return 1.0 / s; }

// ---- src/inference/smc.rs:255-272. u_raw = rng.gen::<f64>() BEFORE the division by n.
// (i - 1) with i == 0 wraps in release builds -> min(n-1) gives n-1. ----
inline void systematic_indices(const double* w, size_t n, double u_raw, uint64_t* idx) {
  double u = u_raw / (double)n; double cum = 0.0; size_t i = 0;
  for (size_t j = 0; j < n; j++) {
    double thr = u + (double)j / (double)n;
    while (cum < thr && i < n) { cum += w[i]; i++; }
    size_t k = i - 1; idx[j] = (uint64_t)(k < n - 1 ? k : n - 1);
  }
}
// ---- src/inference/smc.rs:275-292 ----
inline void stratified_indices(const double* w, size_t n, const double* us, uint64_t* idx) {
  double cum = 0.0; size_t i = 0;
  for (size_t j = 0; j < n; j++) {
    double thr = ((double)j + us[j]) / (double)n;
    while (cum < thr && i < n) { cum += w[i]; i++; }
    size_t k = i - 1; idx[j] = (uint64_t)(k < n - 1 ? k : n - 1);
  }
}
// ---- src/inference/smc.rs:295-314 ----
inline void multinomial_indices(const double* w, size_t n, const double* us, uint64_t* idx) {
  for (size_t j = 0; j < n; j++) {
    double cum = 0.0; size_t sel = n - 1;
    for (size_t i = 0; i < n; i++) { cum += w[i]; if (us[j] <= cum) { sel = i; break; } }
    idx[j] = sel;
  }
}

// ---- src/inference/smc.rs:588-622 ----
inline double next_beta(double beta, const double* log_w, const double* ll, size_t n, double target_ess) {
  std::vector<double> lv(n), lv2(n);
  auto ess_at = [&](double b) -> double {
    for (size_t i = 0; i < n; i++) { lv[i] = log_w[i] + (b - beta) * ll[i]; }
    double l1 = log_sum_exp(lv.data(), n);
    for (size_t i = 0; i < n; i++) lv2[i] = 2.0 * lv[i];
    double l2 = log_sum_exp(lv2.data(), n);
    if (!fin(l1) || !fin(l2)) return (double)n;
    return std::exp(2.0 * l1 - l2);
  };
  if (ess_at(1.0) >= target_ess) return 1.0;
  double lo = beta, hi = 1.0;
  for (int it = 0; it < 64; it++) { double mid = 0.5 * (lo + hi); if (ess_at(mid) < target_ess) hi = mid; else lo = mid; }
  return std::fmin(std::fmax(hi, beta + 1e-9), 1.0);
}

// ---- src/inference/diagnostics.rs:262-304; chains[j] has len n ----
inline double r_hat(const std::vector<std::vector<double>>& ch) {
  if (ch.size() < 2) return 1.0;
  for (auto& v : ch) if (v.empty()) return std::nan("");
  double m = (double)ch.size(), n = (double)ch[0].size();
  std::vector<double> means(ch.size());
  for (size_t j = 0; j < ch.size(); j++) { double s = 0.0; for (double x : ch[j]) s += x; means[j] = s / (double)ch[j].size(); }
  double overall = 0.0; for (double v : means) overall += v; overall /= m;
  double sb = 0.0; for (double v : means) sb += (v - overall) * (v - overall);
  double b = n / (m - 1.0) * sb;
  double sw = 0.0;
  for (size_t j = 0; j < ch.size(); j++) { double s = 0.0; for (double x : ch[j]) s += (x - means[j]) * (x - means[j]); sw += s / (n - 1.0); }
  double w = sw / m;
  double var_plus = ((n - 1.0) / n) * w + (1.0 / n) * b;
  return std::sqrt(var_plus / w);
}
// ---- src/inference/diagnostics.rs:240-260 ----
inline double split_r_hat(const std::vector<std::vector<double>>& ch) {
  std::vector<std::vector<double>> sp;
  for (auto& c : ch) {
    size_t half = c.size() / 2;
    if (half == 0) { sp.push_back(c); continue; }
    sp.emplace_back(c.begin(), c.begin() + half);
    sp.emplace_back(c.begin() + half, c.begin() + 2 * half);
  }
  return r_hat(sp);
}
// ---- src/inference/mcmc_utils.rs:231-339 ----
inline double ess_from_chains(const std::vector<std::vector<double>>& ch) {
  size_t m = ch.size(); if (m == 0) return 0.0;
  size_t n = ch[0].size();
  bool ragged = false; size_t total = 0;
  for (auto& c : ch) { if (c.size() != n) ragged = true; total += c.size(); }
  if (n < 4 || ragged) return (double)(total > 1 ? total : 1);
  size_t max_lag = std::min(n - 1, (size_t)2048);
  double nf = (double)n, mf = (double)m;
  std::vector<std::vector<double>> acov(m, std::vector<double>(max_lag + 1));
  std::vector<double> means(m), cen(n);
  for (size_t j = 0; j < m; j++) {
    double s = 0.0; for (double x : ch[j]) s += x; double mean = s / nf; means[j] = mean;
    for (size_t i = 0; i < n; i++) cen[i] = ch[j][i] - mean;
    for (size_t lag = 0; lag <= max_lag; lag++) { double a = 0.0; for (size_t i = 0; i < n - lag; i++) a += cen[i] * cen[i + lag]; acov[j][lag] = a / nf; }
  }
  double mean_var = 0.0; for (size_t j = 0; j < m; j++) mean_var += acov[j][0] * nf / (nf - 1.0); mean_var /= mf;
  if (mean_var <= 0.0) return (double)(m * n);
  double var_plus = mean_var * (nf - 1.0) / nf;
  if (m > 1) {
    double overall = 0.0; for (double v : means) overall += v; overall /= mf;
    double b = 0.0; for (double v : means) b += (v - overall) * (v - overall);
    var_plus += b / (mf - 1.0);
  }
  auto rho = [&](size_t t) { double a = 0.0; for (size_t j = 0; j < m; j++) a += acov[j][t]; a /= mf; return 1.0 - (mean_var - a) / var_plus; };
  std::vector<double> rh(max_lag + 1, 0.0); rh[0] = 1.0; if (max_lag >= 1) rh[1] = rho(1);
  size_t t = 1, max_t = std::min((size_t)1, max_lag);
  while (t + 2 <= max_lag) {
    double re = rho(t + 1), ro = rho(t + 2);
    if (re + ro < 0.0) break;
    rh[t + 1] = re; rh[t + 2] = ro; max_t = t + 2; t += 2;
  }
  size_t k = 1;
  while (k + 2 <= max_t) {
    double prev = rh[k - 1] + rh[k], cur = rh[k + 1] + rh[k + 2];
    if (cur > prev) { double avg = prev / 2.0; rh[k + 1] = avg; rh[k + 2] = avg; }
    k += 2;
  }
  double sum_rho = 0.0; for (size_t i = 0; i <= max_t; i++) sum_rho += rh[i];
  double tau = std::fmax(-1.0 + 2.0 * sum_rho, 1.0);
  return (double)(m * n) / tau;
}
}  // namespace fo
