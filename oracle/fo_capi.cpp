// oracle/fo_capi.cpp -- TEST INFRASTRUCTURE ONLY: C entry points so tests/ and bench.py's
// cpu_baseline leg can drive the CPU restatement through ctypes. Nothing under fugue_b200/
// may load this library.
#include <atomic>
#include <thread>
#include "fo_smc.hpp"

using namespace fo;

// dynamic-schedule parallel for over chains (no OpenMP runtime in this image)
template <class F> static void par_for(size_t n, int nthreads, F f) {
  if (nthreads <= 1 || n <= 1) { for (size_t i = 0; i < n; i++) f(i); return; }
  std::atomic<size_t> next{0};
  std::vector<std::thread> th;
  for (int t = 0; t < nthreads; t++) th.emplace_back([&] { for (;;) { size_t i = next.fetch_add(1); if (i >= n) break; f(i); } });
  for (auto& t : th) t.join();
}

extern "C" {

double fo_dist_logp(int dist, const double* P, double x, int64_t xi) { return dist_eval(dist, P, x, xi).lp; }
void fo_dist_eval(int dist, const double* P, double x, int64_t xi, double* out5) {
  LpGrad g = dist_eval(dist, P, x, xi);
  out5[0] = g.lp; out5[1] = g.dx; out5[2] = g.dp[0]; out5[3] = g.dp[1]; out5[4] = g.dp[2];
}
double fo_categorical_lp(const double* probs, size_t k, uint64_t x) { return categorical_lp(probs, k, x); }
double fo_binomial_lp(uint64_t n, double p, uint64_t k) { return binomial_lp(n, p, k); }
double fo_discrete_uniform_lp(int64_t lo, int64_t hi, int64_t x) { return discrete_uniform_lp(lo, hi, x); }
double fo_digamma(double x) { return digamma(x); }
double fo_log_sum_exp(const double* x, size_t n) { return log_sum_exp(x, n); }
void fo_normalize_log_probs(const double* x, size_t n, double* out) { normalize_log_probs(x, n, out); }
double fo_log1p_exp(double x) { return log1p_exp(x); }
double fo_safe_ln(double x) { return safe_ln(x); }

int fo_program_create(const fc_op* ops, size_t n, const fc_column* cols, size_t nc, void** out) {
  Program* p = new Program();
  int rc = p->build(ops, n, cols, nc);
  if (rc != FC_OK) { delete p; *out = nullptr; return rc; }
  *out = p; return FC_OK;
}
int fo_program_create_ex(const fc_op* ops, size_t n, const fc_column* cols, size_t nc, const fc_expr_node* nodes, size_t nn, void** out) {
  Program* p = new Program();
  int rc = p->build(ops, n, cols, nc, nodes, nn);
  if (rc != FC_OK) { delete p; *out = nullptr; return rc; }
  *out = p; return FC_OK;
}
void fo_program_destroy(void* p) { delete (Program*)p; }
void fo_program_dims(void* p, int64_t* dc, int64_t* dd) { *dc = ((Program*)p)->d_cont; *dd = ((Program*)p)->d_disc; }
const char* fo_program_site(void* p, int64_t sorted_idx, int discrete) {
  Program* P = (Program*)p;
  return discrete ? P->site_d[P->perm_d[sorted_idx]].c_str() : P->site_c[P->perm_c[sorted_idx]].c_str();
}
int64_t fo_program_perm(void* p, int64_t sorted_idx, int discrete) {
  Program* P = (Program*)p; return discrete ? P->perm_d[sorted_idx] : P->perm_c[sorted_idx];
}

// q [d_cont][C], qd [d_disc][C], chain-minor. grad may be null. nthreads<=1 => serial.
void fo_logp(void* p, const double* q, const int64_t* qd, size_t C, double* lp, double* ll, double* lf, double* grad, int nthreads) {
  Program* P = (Program*)p;
  par_for(C, nthreads, [&](size_t c) {
    Score s = score_sorted(*P, q + c, qd ? qd + c : nullptr, (int64_t)C, grad ? grad + c : nullptr);
    if (lp) lp[c] = s.log_prior; if (ll) ll[c] = s.log_lik; if (lf) lf[c] = s.log_factors;
  });
}
// Reference-faithful finite-difference force (hmc.rs:304-329) per chain; ok[c] = all finite.
void fo_grad_fd(void* p, const double* q, const int64_t* qd, size_t C, double h, double* grad, int32_t* ok, int nthreads) {
  Program* P = (Program*)p;
  par_for(C, nthreads, [&](size_t c) {
    bool o = grad_fd_sorted(*P, q + c, qd ? qd + c : nullptr, (int64_t)C, h, grad + c);
    if (ok) ok[c] = o;
  });
}

// One hmc_transition (hmc.rs:419-473) for one chain. q[d] in/out, lj in/out.
// out4 = {accepted, divergent, accept_prob, lj_new}
void fo_hmc_transition(void* p, const int64_t* qd, double* q, double* lj, const double* z, double u, double eps, uint64_t l,
                       const double* m_inv, const double* mass_sqrt, int grad_mode, double h, double* out4) {
  Program* P = (Program*)p; HmcTarget T{*P, qd, grad_mode, h};
  std::vector<double> qv(q, q + P->d_cont);
  TransitionOut t = hmc_transition(T, qv, *lj, z, u, eps, l, m_inv, mass_sqrt);
  std::copy(qv.begin(), qv.end(), q);
  out4[0] = t.accepted; out4[1] = t.divergent; out4[2] = t.accept_prob; out4[3] = t.lj_new;
}
double fo_find_reasonable_epsilon(void* p, const int64_t* qd, const double* q, double lj, const double* z, const double* m_inv,
                                  const double* mass_sqrt, int grad_mode, double h) {
  Program* P = (Program*)p; HmcTarget T{*P, qd, grad_mode, h};
  std::vector<double> qv(q, q + P->d_cont);
  return find_reasonable_epsilon(T, qv, lj, z, m_inv, mass_sqrt);
}
// eps_out[i] = DualAveraging::update(alphas[i]); returns frozen_step()
double fo_dual_averaging(double eps0, double target, const double* alphas, size_t n, double* eps_out) {
  DualAveraging da(eps0, target);
  for (size_t i = 0; i < n; i++) { double e = da.update(alphas[i]); if (eps_out) eps_out[i] = e; }
  return da.frozen_step();
}
void fo_welford(const double* x, size_t n, size_t d, double* var_out) {
  Welford w(d); for (size_t i = 0; i < n; i++) w.push(x + i * d);
  auto v = w.variances(); std::copy(v.begin(), v.end(), var_out);
}
void fo_dim_adapt(const uint8_t* accepts, size_t n, double target, double gamma, double* scales_out) {
  DimAdapt a(target, gamma); for (size_t i = 0; i < n; i++) { a.update(accepts[i] != 0); scales_out[i] = a.scale; }
}

void fo_resample(int method, const double* w, size_t n, const double* us, uint64_t* idx) {
  if (method == FC_RESAMPLE_SYSTEMATIC) systematic_indices(w, n, us[0], idx);
  else if (method == FC_RESAMPLE_STRATIFIED) stratified_indices(w, n, us, idx);
  else multinomial_indices(w, n, us, idx);
}
double fo_ess(const double* w, size_t n) { return ess(w, n); }
double fo_next_beta(double beta, const double* log_w, const double* ll, size_t n, double target_ess) { return next_beta(beta, log_w, ll, n, target_ess); }

static std::vector<std::vector<double>> to_chains(const double* draws, size_t S, size_t C) {
  std::vector<std::vector<double>> ch(C, std::vector<double>(S));
  for (size_t s = 0; s < S; s++) for (size_t c = 0; c < C; c++) ch[c][s] = draws[s * C + c];
  return ch;
}
double fo_split_rhat(const double* draws, size_t S, size_t C) { return split_r_hat(to_chains(draws, S, C)); }
double fo_classic_rhat(const double* draws, size_t S, size_t C) { return r_hat(to_chains(draws, S, C)); }
// effective_sample_size_multichain (mcmc_utils.rs:214-224)
double fo_ess_multichain(const double* draws, size_t S, size_t C) {
  if (C == 0) return 0.0;
  return ess_from_chains(to_chains(draws, S, C));
}
// ---- single-site MH / tempered SMC (fo_smc.hpp). State arrays are in SORTED site order (the C ABI's order). ----
static State state_from_sorted(const Program& P, const double* q, const int64_t* qd) {
  State s; s.q.resize(P.d_cont); s.qd.resize(P.d_disc);
  for (int64_t k = 0; k < P.d_cont; k++) s.q[P.perm_c[k]] = q[k];
  for (int64_t k = 0; k < P.d_disc; k++) s.qd[P.perm_d[k]] = qd[k];
  return s;
}
static void state_to_sorted(const Program& P, const State& s, double* q, int64_t* qd) {
  for (int64_t k = 0; k < P.d_cont; k++) q[k] = s.q[P.perm_c[k]];
  for (int64_t k = 0; k < P.d_disc; k++) if (qd) qd[k] = s.qd[P.perm_d[k]];
}
int64_t fo_n_sites(void* p) { return (int64_t)all_sites(*(Program*)p).size(); }
// single_site_mh_step on global site g (address order over ALL sample sites); out3 = {accept, log_alpha, lw_new}
void fo_mh_step(void* p, double* q, int64_t* qd, double cur_lw, int64_t g, int kind, double scale, double z, double u, double lower, double upper,
                double prior_x, int64_t prior_xi, double* out3) {
  Program* P = (Program*)p;
  State s = state_from_sorted(*P, q, qd);
  auto sites = all_sites(*P);
  MhStepOut o = single_site_mh_step(*P, s, cur_lw, sites[(size_t)g], kind, scale, z, u, lower, upper, prior_x, prior_xi);
  state_to_sorted(*P, s, q, qd);
  out3[0] = o.accept; out3[1] = o.log_alpha; out3[2] = o.lw_new;
}
// the automatic f64 proposal kind of global site g at this state (mh.rs:339-358); -1 for a discrete site
int fo_mh_kind(void* p, const double* q, const int64_t* qd, int64_t g) {
  Program* P = (Program*)p;
  State s = state_from_sorted(*P, q, qd);
  auto sites = all_sites(*P);
  return sites[(size_t)g].cont ? f64_kind(*P, s, sites[(size_t)g]) : -1;
}
// tempered_single_site_mh on the k-th f64 site (sorted); out5 = {accept, log_alpha, log_prior, log_lik, log_factors} of the kept state
void fo_tempered_step(void* p, double* q, int64_t* qd, int64_t k, double beta, double scale, double z, double u, double* out5) {
  Program* P = (Program*)p;
  State s = state_from_sorted(*P, q, qd);
  auto fs = f64_sites(*P);
  TemperedOut t = tempered_single_site_mh(*P, s, fs[(size_t)k], beta, scale, z, u);
  state_to_sorted(*P, s, q, qd);
  out5[0] = t.accept; out5[1] = t.log_alpha; out5[2] = t.score.log_prior; out5[3] = t.score.log_lik; out5[4] = t.score.log_factors;
}
void fo_normalize_particles(const double* log_w, size_t n, double* w) { normalize_particles(log_w, n, w); }
// adaptive_smc (smc.rs:455-581). values [d_cont][n] / dvalues [d_disc][n] sorted rows, particle-minor; outputs may be null.
// nthreads <= 1: one generator and one adaptation object threaded through all particles, as the reference does.
int fo_adaptive_smc(void* p, size_t n, int method, double ess_threshold, uint64_t rejuvenation_steps, uint64_t seed, int nthreads, double* values,
                    int64_t* dvalues, double* weight, double* log_weight, double* log_evidence, uint64_t* steps) {
  Program* P = (Program*)p;
  SmcOut o;
  auto pf = [](size_t cnt, int nt, auto f) { par_for(cnt, nt, f); };
  if (!adaptive_smc(*P, n, method, ess_threshold, rejuvenation_steps, seed, nthreads, pf, o)) return FC_ERR_UNSUPPORTED_OP;
  for (size_t i = 0; i < n; i++) {
    if (values) for (int64_t k = 0; k < P->d_cont; k++) values[(size_t)k * n + i] = o.particles[i].q[P->perm_c[k]];
    if (dvalues) for (int64_t k = 0; k < P->d_disc; k++) dvalues[(size_t)k * n + i] = o.particles[i].qd[P->perm_d[k]];
    if (weight) weight[i] = o.weight[i];
    if (log_weight) log_weight[i] = o.log_w[i];
  }
  if (log_evidence) *log_evidence = o.log_evidence;
  if (steps) *steps = o.steps;
  return FC_OK;
}
int fo_max_threads() { unsigned n = std::thread::hardware_concurrency(); return n ? (int)n : 1; }
}
