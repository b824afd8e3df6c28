// fc_smc.cu -- adaptive (likelihood-tempered) SMC with all particles resident on one B200.
//
// Device restatement of src/inference/smc.rs: adaptive_smc (:455-581), next_beta (:588-622),
// tempered_single_site_mh (:631-688, via the move engine of fc_mh.cu), resample + clone
// (:326-349, 534-541) and smc_prior_particles (:764-790). Particle state is SoA
// [site][particle]; the cached per-particle accumulators (log_prior, log_lik, log_factors)
// travel with the particle through resampling, so a rejuvenation move scores the proposal only
// (the reference re-scores the current state as well: 2 model runs per move, smc.rs:659-676).
//
// Documented deviation: the reference threads ONE DiminishingAdaptation through all particles
// sequentially (smc.rs:482,551) -- a cross-particle serial dependency. Here every particle of a
// sweep uses the same per-site scale and the scale is updated once per sweep from the sweep's
// aggregate acceptance rate with gain 2 * sweep^-gamma. Any positive scale leaves pi_beta
// invariant, so only efficiency is affected, not the target.
#include "fc_mh.cuh"

namespace fc {
int32_t dev_reduce(int op, const double* d_x, int64_t n, double m, double* d_scratch, double* d_out, cudaStream_t st);
int32_t dev_log_sum_exp(const double* d_x, int64_t n, double* d_scratch, double* h_out, cudaStream_t st);
struct ResampleWork;
ResampleWork* resample_work_new();
void resample_work_free(ResampleWork*);
int32_t resample_dev(int method, const double* d_w, int64_t n, const double* d_us, uint64_t* d_idx, ResampleWork& W, cudaStream_t st, uint64_t* launches);

#define RDB 1024  // must match RD_BLOCKS of fc_resample.cu

__global__ void k_loglik(const double* __restrict__ lp3, double* __restrict__ ll, int64_t P) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < P) ll[i] = lp3[P + i] + lp3[2 * P + i];  // particle_log_likelihood (smc.rs:381-383)
}
__global__ void k_fill2(double* p, double v, int64_t n) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) p[i] = v;
}
// combined = log_w + d_beta * ll   (smc.rs:512-517)
__global__ void k_combine(const double* __restrict__ log_w, const double* __restrict__ ll, double d_beta, double* __restrict__ out, int64_t P) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < P) out[i] = log_w[i] + d_beta * ll[i];
}
__global__ void k_sub_exp(const double* __restrict__ x, double sub, double* __restrict__ out_log, double* __restrict__ out_exp, int64_t P) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= P) return;
  double v = x[i] - sub;
  if (out_log) out_log[i] = v;
  if (out_exp) out_exp[i] = exp(v);
}

// ---- next_beta (smc.rs:588-622): bisection state lives on the device; 65 ESS evaluations, each
// two fixed-order reductions over (log_w, ll) ----
struct Bisect {
  double beta, lo, hi, mid, target, n;
  double m1;       // max of lv at `mid`
  int done;
};
__global__ void k_nb_max(const double* __restrict__ log_w, const double* __restrict__ ll, const Bisect* __restrict__ B, double* __restrict__ part, int64_t P) {
  double d = B->mid - B->beta;
  double acc = d_neg_inf();
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < P; i += (int64_t)gridDim.x * blockDim.x) acc = fmax(acc, log_w[i] + d * ll[i]);
  __shared__ double sm[256];
  sm[threadIdx.x] = acc;
  __syncthreads();
  for (int o = 128; o > 0; o >>= 1) { if (threadIdx.x < o) sm[threadIdx.x] = fmax(sm[threadIdx.x], sm[threadIdx.x + o]); __syncthreads(); }
  if (threadIdx.x == 0) part[blockIdx.x] = sm[0];
}
__global__ void k_nb_max_final(const double* __restrict__ part, int nb, Bisect* __restrict__ B) {
  __shared__ double sm[256];
  double acc = d_neg_inf();
  for (int i = threadIdx.x; i < nb; i += 256) acc = fmax(acc, part[i]);
  sm[threadIdx.x] = acc;
  __syncthreads();
  for (int o = 128; o > 0; o >>= 1) { if (threadIdx.x < o) sm[threadIdx.x] = fmax(sm[threadIdx.x], sm[threadIdx.x + o]); __syncthreads(); }
  if (threadIdx.x == 0) B->m1 = sm[0];
}
__global__ void k_nb_sums(const double* __restrict__ log_w, const double* __restrict__ ll, const Bisect* __restrict__ B, double* __restrict__ part, int64_t P) {
  double d = B->mid - B->beta, m1 = B->m1;
  double s1 = 0.0, s2 = 0.0;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < P; i += (int64_t)gridDim.x * blockDim.x) {
    double lv = log_w[i] + d * ll[i];
    s1 += exp(lv - m1);
    s2 += exp(2.0 * lv - 2.0 * m1);
  }
  __shared__ double sa[256], sb[256];
  sa[threadIdx.x] = s1; sb[threadIdx.x] = s2;
  __syncthreads();
  for (int o = 128; o > 0; o >>= 1) { if (threadIdx.x < o) { sa[threadIdx.x] += sa[threadIdx.x + o]; sb[threadIdx.x] += sb[threadIdx.x + o]; } __syncthreads(); }
  if (threadIdx.x == 0) { part[blockIdx.x] = sa[0]; part[RDB + blockIdx.x] = sb[0]; }
}
// first = 1: this was the ess_at(1.0) probe (smc.rs:605-607)
__global__ void k_nb_update(const double* __restrict__ part, int nb, Bisect* __restrict__ B, int first) {
  __shared__ double sa[256], sb[256];
  double a = 0.0, b = 0.0;
  for (int i = threadIdx.x; i < nb; i += 256) { a += part[i]; b += part[RDB + i]; }
  sa[threadIdx.x] = a; sb[threadIdx.x] = b;
  __syncthreads();
  for (int o = 128; o > 0; o >>= 1) { if (threadIdx.x < o) { sa[threadIdx.x] += sa[threadIdx.x + o]; sb[threadIdx.x] += sb[threadIdx.x + o]; } __syncthreads(); }
  if (threadIdx.x != 0) return;
  if (B->done) return;
  double m1 = B->m1;
  double lse1 = (m1 == d_neg_inf()) ? d_neg_inf() : (sa[0] == 0.0 ? d_neg_inf() : m1 + log(sa[0]));
  double lse2 = (m1 == d_neg_inf()) ? d_neg_inf() : (sb[0] == 0.0 ? d_neg_inf() : 2.0 * m1 + log(sb[0]));
  double ess = (!isfinite(lse1) || !isfinite(lse2)) ? B->n : exp(2.0 * lse1 - lse2);
  if (first) {
    if (ess >= B->target) { B->hi = 1.0; B->done = 1; return; }
    B->lo = B->beta; B->hi = 1.0;
  } else {
    if (ess < B->target) B->hi = B->mid; else B->lo = B->mid;
  }
  B->mid = 0.5 * (B->lo + B->hi);
}

// gather particle planes by ancestor index (resample_particles, smc.rs:326-349)
template <class T>
__global__ void k_gather(const T* __restrict__ src, T* __restrict__ dst, const uint64_t* __restrict__ idx, int64_t rows, int64_t P) {
  int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= P) return;
  int64_t a = (int64_t)idx[j];
  for (int64_t r = blockIdx.y; r < rows; r += gridDim.y) dst[r * P + j] = src[r * P + a];
}
__global__ void k_uniform_one(double* out, uint64_t seed, uint64_t tag, int64_t n) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i * 2 >= n) return;
  Philox ph(seed, 0x8000000000000000ULL);
  double a, b;
  ph.uniform2((uint64_t)i, tag, a, b);
  out[2 * i] = a;
  if (2 * i + 1 < n) out[2 * i + 1] = b;
}

}  // namespace fc

struct fc_smc_session {
  fc_program* P = nullptr;
  fc_smc_config cfg;
  uint64_t seed = 0;
  int64_t n = 0;
  fc::MoveEngine eng;
  fc::DevBuf<double> q, q2, lp3, lp3b, total, ll, log_w, comb, w, scratch, us;
  fc::DevBuf<fc::disc_t> qd, qd2;
  fc::DevBuf<uint64_t> idx, n_accept;
  fc::DevBuf<fc::Bisect> bis;
  fc::ResampleWork* rw = nullptr;
  double beta = 0.0, log_evidence = 0.0;
  uint64_t steps = 0, moves = 0;
  bool finished = false;
};

namespace fc {

static int32_t smc_next_beta(fc_smc_session* S, double target_ess, double* out) {
  cudaStream_t st = S->P->stream;
  int64_t n = S->n;
  int nb = (int)std::max<int64_t>(1, std::min<int64_t>(RDB, cdiv(n, 256)));
  Bisect hb;
  hb.beta = S->beta; hb.lo = S->beta; hb.hi = 1.0; hb.mid = 1.0; hb.target = target_ess; hb.n = (double)n; hb.m1 = 0.0; hb.done = 0;
  FC_CUDA(cudaMemcpyAsync(S->bis.p, &hb, sizeof(hb), cudaMemcpyHostToDevice, st));
  for (int it = 0; it < 65; it++) {
    k_nb_max<<<nb, 256, 0, st>>>(S->log_w.p, S->ll.p, S->bis.p, S->scratch.p, n);
    k_nb_max_final<<<1, 256, 0, st>>>(S->scratch.p, nb, S->bis.p);
    k_nb_sums<<<nb, 256, 0, st>>>(S->log_w.p, S->ll.p, S->bis.p, S->scratch.p, n);
    k_nb_update<<<1, 256, 0, st>>>(S->scratch.p, nb, S->bis.p, it == 0 ? 1 : 0);
    S->P->n_kernel_launches += 4;
    if (it == 0) {  // early exit when a full jump to beta = 1 keeps the ESS above target
      FC_CUDA(cudaMemcpyAsync(&hb, S->bis.p, sizeof(hb), cudaMemcpyDeviceToHost, st));
      FC_CUDA(cudaStreamSynchronize(st));
      if (hb.done) { *out = 1.0; return FC_OK; }
    }
  }
  FC_KERNEL_CHECK();
  FC_CUDA(cudaMemcpyAsync(&hb, S->bis.p, sizeof(hb), cudaMemcpyDeviceToHost, st));
  FC_CUDA(cudaStreamSynchronize(st));
  *out = std::fmin(std::fmax(hb.hi, S->beta + 1e-9), 1.0);  // smc.rs:621
  return FC_OK;
}

// one tempering step (the body of the while loop, smc.rs:502-559); returns beta
static int32_t smc_temper_step(fc_smc_session* S) {
  fc_program* P = S->P;
  cudaStream_t st = P->stream;
  int64_t n = S->n;
  unsigned nbk = (unsigned)cdiv(n, 256);
  const double neg_ln_n = -std::log((double)n);
  double target_ess = std::fmin(std::fmax(S->cfg.ess_threshold * (double)n, 1.0), (double)n);
  S->steps += 1;
  double beta_new;
  FC_TRY(smc_next_beta(S, target_ess, &beta_new));
  if (S->steps >= 10000) beta_new = 1.0;
  double d_beta = beta_new - S->beta;
  k_combine<<<nbk, 256, 0, st>>>(S->log_w.p, S->ll.p, d_beta, S->comb.p, n);
  P->n_kernel_launches += 1;
  double log_norm;
  FC_TRY(dev_log_sum_exp(S->comb.p, n, S->scratch.p, &log_norm, st));
  P->n_kernel_launches += 4;
  S->log_evidence += log_norm;
  if (std::isfinite(log_norm)) k_sub_exp<<<nbk, 256, 0, st>>>(S->comb.p, log_norm, S->log_w.p, S->w.p, n);
  else { k_fill2<<<nbk, 256, 0, st>>>(S->log_w.p, neg_ln_n, n); k_fill2<<<nbk, 256, 0, st>>>(S->w.p, 1.0 / (double)n, n); }
  P->n_kernel_launches += 1;
  S->beta = beta_new;
  if (S->beta < 1.0) {
    // resample (smc.rs:534-541)
    int64_t nu = S->cfg.resampling_method == FC_RESAMPLE_SYSTEMATIC ? 1 : n;
    FC_CUDA(S->us.reserve(nu + 1));
    k_uniform_one<<<(unsigned)cdiv(cdiv(nu, 2), 256), 256, 0, st>>>(S->us.p, S->seed, 0x1000000000000000ULL | S->steps, nu);
    FC_TRY(resample_dev(S->cfg.resampling_method, S->w.p, n, S->us.p, S->idx.p, *S->rw, st, &P->n_kernel_launches));
    dim3 gg(nbk, 1);
    if (P->d_cont) { gg.y = (unsigned)std::min<int64_t>(P->d_cont, 64); k_gather<double><<<gg, 256, 0, st>>>(S->q.p, S->q2.p, S->idx.p, P->d_cont, n); std::swap(S->q.p, S->q2.p); std::swap(S->q.n, S->q2.n); }
    if (P->d_disc) { gg.y = (unsigned)std::min<int64_t>(P->d_disc, 64); k_gather<disc_t><<<gg, 256, 0, st>>>(S->qd.p, S->qd2.p, S->idx.p, P->d_disc, n); std::swap(S->qd.p, S->qd2.p); std::swap(S->qd.n, S->qd2.n); }
    gg.y = 3;
    k_gather<double><<<gg, 256, 0, st>>>(S->lp3.p, S->lp3b.p, S->idx.p, 3, n);
    std::swap(S->lp3.p, S->lp3b.p); std::swap(S->lp3.n, S->lp3b.n);
    k_fill2<<<nbk, 256, 0, st>>>(S->log_w.p, neg_ln_n, n);
    FC_KERNEL_CHECK();
    P->n_kernel_launches += 5;
    // pi_beta-invariant rejuvenation (smc.rs:544-554)
    for (uint64_t r = 0; r < S->cfg.rejuvenation_steps; r++) {
      FC_TRY(S->eng.step(S->q.p, S->qd.p, S->lp3.p, S->total.p, S->n_accept.p, S->moves, S->beta, true, 0.44, 0.7));
      FC_TRY(S->eng.shared_adapt(0.44, 0.7));
      S->moves += 1;
    }
    k_loglik<<<nbk, 256, 0, st>>>(S->lp3.p, S->ll.p, n);
    FC_KERNEL_CHECK();
    P->n_kernel_launches += 1;
  }
  return FC_OK;
}

}  // namespace fc

extern "C" {

void fc_smc_config_default(fc_smc_config* cfg) {  // SMCConfig::default (smc.rs:181-189)
  cfg->resampling_method = FC_RESAMPLE_SYSTEMATIC;
  cfg->_pad = 0;
  cfg->ess_threshold = 0.5;
  cfg->rejuvenation_steps = 0;
}

int32_t fc_smc_destroy(fc_smc_session* s) {
  if (!s) return FC_OK;
  if (s->P && s->P->stream) cudaStreamSynchronize(s->P->stream);
  if (s->rw) fc::resample_work_free(s->rw);
  delete s;
  return FC_OK;
}

// smc_prior_particles (smc.rs:764-790) + the state adaptive_smc keeps (smc.rs:470-482)
int32_t fc_smc_create(fc_program* p, const fc_smc_config* cfg, uint64_t seed, size_t n, uint64_t first_particle, fc_smc_session** out) {
  using namespace fc;
  if (!p || !cfg || !out) { set_error("NULL argument"); return FC_ERR_INVALID_ARGUMENT; }
  *out = nullptr;
  if (n == 0) { set_error("num_particles must be > 0"); return FC_ERR_INVALID_ARGUMENT; }
  FC_CUDA(cudaSetDevice(p->device));
  std::unique_ptr<fc_smc_session> S(new fc_smc_session());
  S->P = p; S->cfg = *cfg; S->seed = seed; S->n = (int64_t)n;
  int64_t N = S->n;
  cudaStream_t st = p->stream;
  int64_t nq = std::max<int64_t>(1, p->d_cont * N), nd = std::max<int64_t>(1, p->d_disc * N);
  FC_CUDA(S->q.reserve(nq)); FC_CUDA(S->q2.reserve(nq)); FC_CUDA(S->qd.reserve(nd)); FC_CUDA(S->qd2.reserve(nd));
  FC_CUDA(S->lp3.reserve(3 * N)); FC_CUDA(S->lp3b.reserve(3 * N));
  for (auto* b : {&S->total, &S->ll, &S->log_w, &S->comb, &S->w}) FC_CUDA(b->reserve(N));
  FC_CUDA(S->scratch.reserve(2 * RDB + 16));
  FC_CUDA(S->idx.reserve(N)); FC_CUDA(S->n_accept.reserve(N)); FC_CUDA(S->bis.reserve(1));
  FC_CUDA(cudaMemsetAsync(S->n_accept.p, 0, N * 8, st));
  S->rw = resample_work_new();
  FC_TRY(S->eng.init(p, N, true, seed ^ 0x5bd1e995u, first_particle));
  FC_TRY(prior_sample_dev(p, S->q.p, S->qd.p, N, seed, first_particle, 0x7000000000000000ULL));
  FC_TRY(score_dev(p, S->q.p, S->qd.p, N, S->lp3.p, S->total.p, nullptr));
  unsigned nbk = (unsigned)cdiv(N, 256);
  k_loglik<<<nbk, 256, 0, st>>>(S->lp3.p, S->ll.p, N);
  k_fill2<<<nbk, 256, 0, st>>>(S->log_w.p, -std::log((double)N), N);
  FC_KERNEL_CHECK();
  p->n_kernel_launches += 2;
  FC_CUDA(cudaStreamSynchronize(st));
  *out = S.release();
  return FC_OK;
}

// One step of the tempering ladder (no-op once beta == 1). With rejuvenation_steps == 0 the single
// "step" is the pure importance-sampling reweight (smc.rs:484-493).
int32_t fc_smc_step(fc_smc_session* s, double* beta_out) {
  using namespace fc;
  if (!s) { set_error("session is NULL"); return FC_ERR_INVALID_ARGUMENT; }
  FC_CUDA(cudaSetDevice(s->P->device));
  if (!s->finished) {
    if (s->cfg.rejuvenation_steps == 0) {
      cudaStream_t st = s->P->stream;
      unsigned nbk = (unsigned)cdiv(s->n, 256);
      k_combine<<<nbk, 256, 0, st>>>(s->log_w.p, s->ll.p, 1.0, s->comb.p, s->n);
      FC_KERNEL_CHECK();
      double le;
      FC_TRY(dev_log_sum_exp(s->comb.p, s->n, s->scratch.p, &le, st));
      s->P->n_kernel_launches += 5;
      s->log_evidence = le;
      s->beta = 1.0;
      std::swap(s->log_w.p, s->comb.p); std::swap(s->log_w.n, s->comb.n);
      s->steps = 1;
    } else {
      FC_TRY(smc_temper_step(s));
    }
    if (s->beta >= 1.0) s->finished = true;
  }
  if (beta_out) *beta_out = s->beta;
  return FC_OK;
}

// final normalisation (smc.rs:565-575) and copy-out. Any pointer may be NULL.
int32_t fc_smc_get(fc_smc_session* s, double* h_values, int64_t* h_dvalues, double* h_weight, double* h_log_weight, double* log_evidence, uint64_t* n_temper_steps) {
  using namespace fc;
  if (!s) { set_error("session is NULL"); return FC_ERR_INVALID_ARGUMENT; }
  fc_program* p = s->P;
  cudaStream_t st = p->stream;
  int64_t n = s->n;
  unsigned nbk = (unsigned)cdiv(n, 256);
  if (h_weight || h_log_weight) {
    double log_norm;
    FC_TRY(dev_log_sum_exp(s->log_w.p, n, s->scratch.p, &log_norm, st));
    if (std::isfinite(log_norm)) k_sub_exp<<<nbk, 256, 0, st>>>(s->log_w.p, log_norm, s->comb.p, s->w.p, n);
    else { k_fill2<<<nbk, 256, 0, st>>>(s->comb.p, -std::log((double)n), n); k_fill2<<<nbk, 256, 0, st>>>(s->w.p, 1.0 / (double)n, n); }
    FC_KERNEL_CHECK();
    p->n_kernel_launches += 5;
    if (h_weight) FC_CUDA(cudaMemcpyAsync(h_weight, s->w.p, n * 8, cudaMemcpyDeviceToHost, st));
    if (h_log_weight) FC_CUDA(cudaMemcpyAsync(h_log_weight, s->comb.p, n * 8, cudaMemcpyDeviceToHost, st));
  }
  if (h_values && p->d_cont) FC_CUDA(cudaMemcpyAsync(h_values, s->q.p, p->d_cont * n * 8, cudaMemcpyDeviceToHost, st));
  if (h_dvalues && p->d_disc) FC_TRY(disc_to_host(p, s->qd.p, p->d_disc * n, h_dvalues));
  if (log_evidence) *log_evidence = s->log_evidence;
  if (n_temper_steps) *n_temper_steps = s->steps;
  FC_CUDA(cudaStreamSynchronize(st));
  return FC_OK;
}

int32_t fc_smc_device_state(fc_smc_session* s, double** d_values, int32_t** d_dvalues, double** d_log_w, double** d_loglik) {
  if (d_values) *d_values = s->q.p;
  if (d_dvalues) *d_dvalues = s->qd.p;
  if (d_log_w) *d_log_w = s->log_w.p;
  if (d_loglik) *d_loglik = s->ll.p;
  return FC_OK;
}

// adaptive_smc (smc.rs:455-581). h_values [d_cont][n], h_dvalues [d_disc][n]; outputs may be NULL.
int32_t fc_smc_run(fc_program* p, const fc_smc_config* cfg, uint64_t seed, size_t n, double* h_values, int64_t* h_dvalues, double* h_weight, double* h_log_weight,
                   double* log_evidence, uint64_t* n_temper_steps) {
  if (n == 0) {  // smc.rs:462-467
    if (log_evidence) *log_evidence = 0.0;
    if (n_temper_steps) *n_temper_steps = 0;
    return FC_OK;
  }
  fc_smc_session* s = nullptr;
  FC_TRY(fc_smc_create(p, cfg, seed, n, 0, &s));
  int32_t rc = FC_OK;
  double beta = 0.0;
  while (rc == FC_OK && beta < 1.0) rc = fc_smc_step(s, &beta);
  if (rc == FC_OK) rc = fc_smc_get(s, h_values, h_dvalues, h_weight, h_log_weight, log_evidence, n_temper_steps);
  fc_smc_destroy(s);
  return rc;
}
}
