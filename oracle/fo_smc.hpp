// oracle/fo_smc.hpp -- TEST INFRASTRUCTURE ONLY.
// CPU restatement of the single-site MH step and of the tempered SMC driver, for fixed-structure
// programs. Follows /root/reference/src/inference/mh.rs:128-295 (proposals), :339-358 (support probe),
// :698-744 (single_site_mh_step); src/inference/smc.rs:455-581 (adaptive_smc), :631-688
// (tempered_single_site_mh), :719-753 (normalize_particles), :764-790 (smc_prior_particles).
// Step functions take their randomness explicitly (site pick, normal draw, accept uniform) so the device
// and the oracle can be compared on identical draws; the whole-ladder driver owns a small generator
// (the reference's ChaCha12 stream cannot be reproduced: parity there is distributional).
#pragma once
#include "fo_infer.hpp"

namespace fo {

// one sample site of the program in the global address order adaptive_mcmc_chain uses (mh.rs:965)
struct SiteRef { bool cont; int64_t slot; size_t op; int64_t elem; std::string name; };
inline std::vector<SiteRef> all_sites(const Program& P) {
  std::vector<SiteRef> v;
  for (size_t oi = 0; oi < P.ops.size(); oi++) {
    const fc_op& op = P.ops[oi];
    if (op.kind != FC_OP_SAMPLE && op.kind != FC_OP_PLATE_SAMPLE) continue;
    bool cont = op.value_type == FC_T_F64;
    int64_t len = op.kind == FC_OP_SAMPLE ? 1 : op.plate_len;
    for (int64_t e = 0; e < len; e++) {
      int64_t slot = (cont ? P.cslot0[oi] : P.dslot0[oi]) + e;
      v.push_back(SiteRef{cont, slot, oi, e, cont ? P.site_c[slot] : P.site_d[slot]});
    }
  }
  std::sort(v.begin(), v.end(), [](const SiteRef& a, const SiteRef& b) { return a.name < b.name; });
  return v;
}
// the f64 sites in BTreeMap order (what tempered_single_site_mh collects, smc.rs:639-645)
inline std::vector<SiteRef> f64_sites(const Program& P) {
  std::vector<SiteRef> v;
  for (auto& s : all_sites(P)) if (s.cont) v.push_back(s);
  return v;
}

// state in PROGRAM order
struct State { std::vector<double> q; std::vector<int64_t> qd; };
inline Score score_state(const Program& P, const State& s) { return score_program_order(P, s.q.data(), s.qd.data(), nullptr); }
inline double loglik_of(const Score& s) { return s.log_lik + s.log_factors; }  // particle_log_likelihood, smc.rs:381-383

// parameters of the site's distribution at the current state
inline void site_params(const Program& P, const State& s, const SiteRef& site, double* pv) {
  const fc_op& op = P.ops[site.op];
  ParamEval pe{P, s.q.data(), s.qd.data()};
  double dl;
  for (int j = 0; j < 3; j++) pv[j] = j < op.n_params ? pe.value(op.param[j], site.elem, &dl) : 0.0;
}

// ---- mh.rs:128-138 ----
inline double gaussian_z(double u1, double u2) { u1 = u1 > 1e-10 ? u1 : 1e-10; return std::sqrt(-2.0 * std::log(u1)) * std::cos(2.0 * M_PI * u2); }
inline double normal_logpdf(double x, double mean, double sd) { double z = (x - mean) / sd; return -0.5 * z * z - std::log(sd) - 0.5 * std::log(2.0 * M_PI); }

enum { PK_GAUSSIAN = 0, PK_LOGSPACE = 1, PK_REFLECT = 2, PK_PRIOR = 3 };
// the support probe that picks the f64 proposal kind (mh.rs:339-358)
inline int f64_kind(const Program& P, const State& s, const SiteRef& site) {
  double pv[3]; site_params(P, s, site, pv);
  double cur = s.q[site.slot];
  double lp = dist_eval(P.ops[site.op].dist, pv, -1.0, 0).lp;
  return (cur > 0.0 && !fin(lp)) ? PK_LOGSPACE : PK_GAUSSIAN;
}

struct MhStepOut { bool accept; double log_alpha, lw_new; };
// single_site_mh_step (mh.rs:698-744) on site `site` with the standard-normal draw z (gaussian_z of the two
// uniforms), the accept uniform u and, for prior-resample proposals, the value drawn from the prior.
// kind < 0: automatic (support probe for f64; Flip / DiscreteWalk / i64 walk / prior resample by value type).
inline MhStepOut single_site_mh_step(const Program& P, State& s, double cur_lw, const SiteRef& site, int kind, double scale, double z, double u,
                                     double lower = 0.0, double upper = 0.0, double prior_x = 0.0, int64_t prior_xi = 0) {
  State prop = s;
  double lqf = 0.0, lqr = 0.0;
  const fc_op& op = P.ops[site.op];
  double pv[3]; site_params(P, s, site, pv);
  if (site.cont) {
    double cur = s.q[site.slot], p = cur;
    if (kind < 0) kind = f64_kind(P, s, site);
    switch (kind) {
      case PK_LOGSPACE: {  // mh.rs:199-224
        if (cur <= 0.0) p = 2.2250738585072014e-308;
        else { double pr = std::exp(std::log(cur) + scale * z); p = fin(pr) ? (pr > 2.2250738585072014e-308 ? pr : 2.2250738585072014e-308) : 1.7976931348623157e308; }
        if (cur > 0.0 && p > 0.0) { lqf = normal_logpdf(std::log(p), std::log(cur), scale) - std::log(p); lqr = normal_logpdf(std::log(cur), std::log(p), scale) - std::log(cur); }
      } break;
      case PK_REFLECT: {  // mh.rs:230-258
        p = cur + scale * z;
        if (upper - lower <= 0.0) p = cur;
        else { while (p < lower || p > upper) { if (p < lower) p = 2.0 * lower - p; if (p > upper) p = 2.0 * upper - p; } p = p < lower ? lower : (p > upper ? upper : p); }
      } break;
      case PK_PRIOR: p = prior_x; lqf = dist_eval(op.dist, pv, p, 0).lp; lqr = dist_eval(op.dist, pv, cur, 0).lp; break;  // mh.rs:400-403
      default: p = cur + scale * z; break;  // mh.rs:181-187
    }
    prop.q[site.slot] = p;
  } else {
    int64_t cur = s.qd[site.slot], p = cur;
    switch (op.value_type) {
      case FC_T_BOOL: p = cur ? 0 : 1; break;                                                       // mh.rs:261-270
      case FC_T_U64: { int64_t k = cur + (int64_t)std::llround(scale * z); p = k >= 0 ? k : -k - 1; } break;  // mh.rs:283-295
      case FC_T_I64: p = cur + (int64_t)std::llround(scale * z); break;                              // mh.rs:557-567
      default: p = prior_xi; lqf = dist_eval(op.dist, pv, 0.0, p).lp; lqr = dist_eval(op.dist, pv, 0.0, cur).lp; break;  // usize: mh.rs:516-530
    }
    prop.qd[site.slot] = p;
  }
  double prop_lw = score_state(P, prop).total();
  double log_alpha = prop_lw - cur_lw + (lqr - lqf) + 0.0;  // dim_term = 0 for a fixed structure (mh.rs:723-731)
  bool accept = log_alpha >= 0.0 || u < std::exp(log_alpha);
  if (accept) s = prop;
  return MhStepOut{accept, log_alpha, accept ? prop_lw : cur_lw};
}

// tempered_single_site_mh (smc.rs:631-688): Gaussian walk on f64 site `site` (an entry of f64_sites), both
// the current and the proposed state are scored (two model runs, as the reference does)
struct TemperedOut { bool accept; double log_alpha; Score score; };
inline TemperedOut tempered_single_site_mh(const Program& P, State& s, const SiteRef& site, double beta, double scale, double z, double u) {
  State prop = s;
  prop.q[site.slot] = s.q[site.slot] + scale * z;
  Score cs = score_state(P, s), ps = score_state(P, prop);
  double log_alpha = (ps.log_prior - cs.log_prior) + beta * (loglik_of(ps) - loglik_of(cs));
  bool accept = log_alpha >= 0.0 || u < std::exp(log_alpha);
  if (accept) s = prop;
  return TemperedOut{accept, log_alpha, accept ? ps : cs};
}

// normalize_particles (smc.rs:719-753): w = exp(lw - lse), then divided by their sum; all -inf -> uniform
inline void normalize_particles(const double* log_w, size_t n, double* w) {
  if (n == 0) return;
  double log_norm = log_sum_exp(log_w, n);
  if (std::isinf(log_norm) && log_norm < 0.0) { for (size_t i = 0; i < n; i++) w[i] = 1.0 / (double)n; return; }
  for (size_t i = 0; i < n; i++) w[i] = std::exp(log_w[i] - log_norm);
  double sum = 0.0;
  for (size_t i = 0; i < n; i++) sum += w[i];
  if (sum > 0.0) for (size_t i = 0; i < n; i++) w[i] /= sum;
}

// ---- a small generator for the whole-ladder driver (xoshiro256++; streams seeded by splitmix64) ----
struct Rng {
  uint64_t s[4];
  explicit Rng(uint64_t seed) {
    for (int i = 0; i < 4; i++) { seed += 0x9E3779B97F4A7C15ULL; uint64_t z = seed; z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ULL; z = (z ^ (z >> 27)) * 0x94D049BB133111EBULL; s[i] = z ^ (z >> 31); }
  }
  static uint64_t rotl(uint64_t x, int k) { return (x << k) | (x >> (64 - k)); }
  uint64_t next() {
    uint64_t r = rotl(s[0] + s[3], 23) + s[0], t = s[1] << 17;
    s[2] ^= s[0]; s[3] ^= s[1]; s[1] ^= s[2]; s[0] ^= s[3]; s[2] ^= t; s[3] = rotl(s[3], 45);
    return r;
  }
  double uniform() { return (double)(next() >> 11) * 0x1.0p-53; }  // rng.gen::<f64>(): a multiple of 2^-53 in [0,1)
  double normal() { double u1 = 1.0 - uniform(), u2 = uniform(); return std::sqrt(-2.0 * std::log(u1)) * std::cos(2.0 * M_PI * u2); }
  double gamma(double k) {  // Marsaglia & Tsang, unit rate
    double boost = 1.0;
    if (k < 1.0) { boost = std::pow(1.0 - uniform(), 1.0 / k); k += 1.0; }
    double d = k - 1.0 / 3.0, c = 1.0 / std::sqrt(9.0 * d);
    for (;;) {
      double x = normal(), v = 1.0 + c * x;
      if (v <= 0.0) continue;
      v = v * v * v;
      double u = 1.0 - uniform();
      if (u < 1.0 - 0.0331 * x * x * x * x || std::log(u) < 0.5 * x * x + d * (1.0 - v + std::log(v))) return d * v * boost;
    }
  }
};
// one prior draw of a site (the families the BASELINE models sample from; others: not restated -> false)
inline bool prior_draw(Rng& r, int dist, const double* pv, double& x, int64_t& xi) {
  x = 0.0; xi = 0;
  switch (dist) {
    case FC_NORMAL: x = pv[0] + pv[1] * r.normal(); return true;
    case FC_UNIFORM: x = pv[0] + (pv[1] - pv[0]) * r.uniform(); return true;
    case FC_EXPONENTIAL: x = -std::log(1.0 - r.uniform()) / pv[0]; return true;
    case FC_LOGNORMAL: x = std::exp(pv[0] + pv[1] * r.normal()); return true;
    case FC_GAMMA: x = r.gamma(pv[0]) / pv[1]; return true;
    case FC_BETA: { double a = r.gamma(pv[0]), b = r.gamma(pv[1]); x = a / (a + b); } return true;
    case FC_BERNOULLI: xi = r.uniform() < pv[0] ? 1 : 0; return true;
    case FC_CATEGORICAL: xi = r.uniform() < pv[0] ? 0 : 1; return true;  // CDF search over [p, 1-p] (distribution.rs:771-784)
    default: return false;
  }
}
// run(PriorHandler, model) (interpreters.rs:88-104): every sample site from its state-dependent prior, program order
inline bool prior_state(const Program& P, Rng& r, State& s) {
  s.q.assign(P.d_cont, 0.0); s.qd.assign(P.d_disc, 0);
  for (size_t oi = 0; oi < P.ops.size(); oi++) {
    const fc_op& op = P.ops[oi];
    if (op.kind != FC_OP_SAMPLE && op.kind != FC_OP_PLATE_SAMPLE) continue;
    bool cont = op.value_type == FC_T_F64;
    int64_t len = op.kind == FC_OP_SAMPLE ? 1 : op.plate_len;
    ParamEval pe{P, s.q.data(), s.qd.data()};
    for (int64_t e = 0; e < len; e++) {
      double pv[3] = {0, 0, 0}, dl;
      for (int j = 0; j < op.n_params; j++) pv[j] = pe.value(op.param[j], e, &dl);
      double x; int64_t xi;
      if (!prior_draw(r, op.dist, pv, x, xi)) return false;
      if (cont) s.q[P.cslot0[oi] + e] = x; else s.qd[P.dslot0[oi] + e] = xi;
    }
  }
  return true;
}

// ---- adaptive_smc (smc.rs:455-581) ----
struct SmcOut { std::vector<State> particles; std::vector<double> log_w, weight; double log_evidence = 0.0; uint64_t steps = 0; };
template <class ParFor>
inline bool adaptive_smc(const Program& P, size_t n, int method, double ess_threshold, uint64_t rejuvenation_steps, uint64_t seed, int nthreads, ParFor par_for, SmcOut& out) {
  out = SmcOut();
  if (n == 0) return true;
  // smc_prior_particles (smc.rs:764-790); with nthreads == 1 one generator is threaded through everything, as
  // the reference's `rng: &mut R` is; with more threads every particle owns a stream (seed, index)
  Rng master(seed);
  std::vector<State> parts(n);
  std::vector<Score> scores(n);
  std::vector<double> ll(n);
  bool ok = true;
  if (nthreads <= 1) {
    for (size_t i = 0; i < n; i++) { ok &= prior_state(P, master, parts[i]); scores[i] = score_state(P, parts[i]); ll[i] = loglik_of(scores[i]); }
  } else {
    par_for(n, nthreads, [&](size_t i) { Rng r(seed ^ (0x9E3779B97F4A7C15ULL * (i + 1))); if (!prior_state(P, r, parts[i])) ok = false; scores[i] = score_state(P, parts[i]); ll[i] = loglik_of(scores[i]); });
  }
  if (!ok) return false;
  const double nf = (double)n;
  std::vector<double> log_w(n, -std::log(nf));
  double beta = 0.0, log_evidence = 0.0;
  double target_ess = ess_threshold * nf; target_ess = target_ess < 1.0 ? 1.0 : (target_ess > nf ? nf : target_ess);
  const std::vector<SiteRef> fsites = f64_sites(P);
  std::vector<DimAdapt> adapt(fsites.size(), DimAdapt(0.44, 0.7));
  uint64_t steps = 0;
  if (rejuvenation_steps == 0) {  // smc.rs:484-493
    for (size_t i = 0; i < n; i++) log_w[i] = -std::log(nf) + ll[i];
    log_evidence = log_sum_exp(log_w.data(), n);
    beta = 1.0; steps = 1;
  } else {
    std::vector<double> comb(n), w(n), us;
    std::vector<uint64_t> idx(n);
    while (beta < 1.0) {
      steps += 1;
      double beta_new = next_beta(beta, log_w.data(), ll.data(), n, target_ess);
      if (steps >= 10000) beta_new = 1.0;
      double d_beta = beta_new - beta;
      for (size_t i = 0; i < n; i++) comb[i] = log_w[i] + d_beta * ll[i];
      double log_norm = log_sum_exp(comb.data(), n);
      log_evidence += log_norm;
      if (fin(log_norm)) for (size_t i = 0; i < n; i++) log_w[i] = comb[i] - log_norm;
      else for (size_t i = 0; i < n; i++) log_w[i] = -std::log(nf);
      beta = beta_new;
      if (beta < 1.0) {
        for (size_t i = 0; i < n; i++) w[i] = std::exp(log_w[i]);
        us.assign(method == FC_RESAMPLE_SYSTEMATIC ? 1 : n, 0.0);
        for (auto& u : us) u = master.uniform();
        if (method == FC_RESAMPLE_SYSTEMATIC) systematic_indices(w.data(), n, us[0], idx.data());
        else if (method == FC_RESAMPLE_STRATIFIED) stratified_indices(w.data(), n, us.data(), idx.data());
        else multinomial_indices(w.data(), n, us.data(), idx.data());
        std::vector<State> np(n); std::vector<Score> ns(n);
        for (size_t j = 0; j < n; j++) { np[j] = parts[idx[j]]; ns[j] = scores[idx[j]]; }
        parts.swap(np); scores.swap(ns);
        for (size_t i = 0; i < n; i++) log_w[i] = -std::log(nf);
        if (!fsites.empty()) {
          if (nthreads <= 1) {  // one generator, one adaptation object shared by all particles (smc.rs:482,551)
            for (size_t i = 0; i < n; i++)
              for (uint64_t r = 0; r < rejuvenation_steps; r++) {
                size_t si = (size_t)(master.uniform() * (double)fsites.size()); if (si >= fsites.size()) si = fsites.size() - 1;
                double z = master.normal();
                // the accept uniform is only drawn when log alpha < 0 (short-circuit `||`, smc.rs:681); drawing it always
                // changes the stream, not the law
                TemperedOut t = tempered_single_site_mh(P, parts[i], fsites[si], beta, adapt[si].scale, z, master.uniform());
                adapt[si].update(t.accept);
                scores[i] = t.score;
              }
          } else {  // particles in parallel: per-particle streams, the scales frozen during a sweep
            for (uint64_t r = 0; r < rejuvenation_steps; r++) {
              std::vector<uint8_t> acc(n), site(n);
              par_for(n, nthreads, [&](size_t i) {
                Rng rr(seed ^ (0xD1B54A32D192ED03ULL * (i + 1)) ^ (steps * 1315423911ULL + r));
                size_t si = (size_t)(rr.uniform() * (double)fsites.size()); if (si >= fsites.size()) si = fsites.size() - 1;
                TemperedOut t = tempered_single_site_mh(P, parts[i], fsites[si], beta, adapt[si].scale, rr.normal(), rr.uniform());
                scores[i] = t.score; acc[i] = t.accept; site[i] = (uint8_t)si;
              });
              for (size_t i = 0; i < n; i++) adapt[site[i]].update(acc[i] != 0);
            }
          }
        }
        for (size_t i = 0; i < n; i++) ll[i] = loglik_of(scores[i]);
      }
    }
  }
  // final normalisation (smc.rs:565-575)
  double log_norm = log_sum_exp(log_w.data(), n);
  out.log_w.resize(n); out.weight.resize(n);
  for (size_t i = 0; i < n; i++) {
    if (fin(log_norm)) { out.log_w[i] = log_w[i] - log_norm; out.weight[i] = std::exp(out.log_w[i]); }
    else { out.log_w[i] = -std::log(nf); out.weight[i] = 1.0 / nf; }
  }
  out.particles.swap(parts);
  out.log_evidence = log_evidence;
  out.steps = steps;
  return true;
}

}  // namespace fo
