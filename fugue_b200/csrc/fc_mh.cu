// fc_mh.cu -- single-site Metropolis-Hastings for C independent chains, and the tempered
// single-site move SMC rejuvenation uses.
//
// Device restatement of src/inference/mh.rs (fixed-structure subset): proposals
// GaussianWalk / LogSpaceWalk / ReflectionWalk / PriorResample / Flip / DiscreteWalk / i64 walk
// (:128-295, 400-567), the support probe that picks LogSpace for positive sites (:339-358),
// single_site_mh_step (:698-744), adaptive_mcmc_chain[_with_overrides] (:921-1014) and
// DiminishingAdaptation (src/inference/mcmc_utils.rs:29-150); tempered_single_site_mh is
// src/inference/smc.rs:631-688. Reversible-jump bookkeeping (birth/death terms) is identically
// zero for fixed-structure programs and is not built (structure-varying models are rejected at
// lowering with ErrorCode 302).
//
// One step = every chain proposes a move of ONE of its own sites (chains pick different sites),
// the whole batch is re-scored by the plate kernels (value only), and each chain accepts or
// restores its single changed value. The current state is scored once and cached, so a step is
// one model evaluation, as in the reference (mh.rs:637-683).
#include "fc_mh.cuh"
#include "fc_prior.cuh"

namespace fc {

__device__ __forceinline__ double gaussian_z(double u1, double u2) {  // mh.rs:128-132
  u1 = fmax(u1, 1e-10);
  return sqrt(-2.0 * log(u1)) * cos(2.0 * 3.141592653589793 * u2);
}
__device__ __forceinline__ double normal_logpdf(double x, double mean, double sd) {  // mh.rs:135-138
  double z = (x - mean) / sd;
  return -0.5 * z * z - log(sd) - 0.5 * log(2.0 * 3.141592653589793);
}

struct MoveArgs {
  const DOp* ops;
  const SiteInfo* sites;   // global order (MH) ; for SMC the continuous rows are used directly
  const int32_t* crow;
  const int32_t* drow;
  const SiteOverride* overrides;  // per global site (or null)
  double* q;
  disc_t* qd;
  int32_t* target;   // [C]
  double* old_x;     // [C]
  double* lqdiff;    // [C] log q(reverse) - log q(forward)
  double* scale;     // [n_sites][C] (per-chain) or [n_sites] (shared)
  int8_t* kind;      // [n_sites][C] cached f64 proposal kind (-1 = not probed yet)
  int64_t C, n_sites, d_cont;
  uint64_t seed, chain0, step_tag;
  int32_t smc_mode;      // 1: pick among continuous sites, Gaussian walk, shared scale, StandardNormal draw
};

__global__ void k_move_propose(const __grid_constant__ MoveArgs A) {
  int64_t c = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= A.C) return;
  const int64_t C = A.C;
  Philox ph(A.seed, A.chain0 + (uint64_t)c);
  double u_site, u_acc, u1, u2;
  ph.uniform2(0, A.step_tag, u_site, u_acc);
  ph.uniform2(1, A.step_tag, u1, u2);
  A.lqdiff[c] = 0.0;
  if (A.smc_mode) {
    if (A.d_cont == 0) { A.target[c] = -1; return; }
    int64_t s = (int64_t)(u_site * (double)A.d_cont);
    if (s >= A.d_cont) s = A.d_cont - 1;
    double z0, z1;
    ph.normal2(1, A.step_tag, z0, z1);  // Normal::new(0,1).sample (smc.rs:656)
    double cur = A.q[s * C + c];
    A.target[c] = (int32_t)s;
    A.old_x[c] = cur;
    A.q[s * C + c] = cur + A.scale[s] * z0;
    return;
  }
  if (A.n_sites == 0) { A.target[c] = -1; return; }
  int64_t g = (int64_t)(u_site * (double)A.n_sites);  // rng.gen_range(0..sites.len()) (mh.rs:716)
  if (g >= A.n_sites) g = A.n_sites - 1;
  const SiteInfo si = A.sites[g];
  const DOp& op = A.ops[si.op];
  const double scale = A.scale[g * C + c];
  A.target[c] = (int32_t)g;
  const double z = gaussian_z(u1, u2);
  if (si.is_cont) {
    const int64_t idx = (int64_t)si.row * C + c;
    const double cur = A.q[idx];
    int kind = A.overrides ? A.overrides[g].kind : -1;
    double pv[3] = {0.0, 0.0, 0.0};
    bool have_params = false;
    if (kind < 0) {
      kind = A.kind[g * C + c];
      if (kind < 0) {  // support probe, cached per address (mh.rs:339-358)
        for (int j = 0; j < op.n_params; j++) pv[j] = eval_param_value(op, j, si.elem, A.q, A.qd, c, C, A.crow, A.drow);
        have_params = true;
        DistOut o = dist_eval<false>(op.dist, pv[0], pv[1], pv[2], -1.0, 0, __longlong_as_double(0x7ff8000000000000LL));
        kind = (cur > 0.0 && !isfinite(o.lp)) ? FC_PROPOSAL_LOGSPACE : FC_PROPOSAL_GAUSSIAN;
        A.kind[g * C + c] = (int8_t)kind;
      }
    }
    double prop = cur, lqf = 0.0, lqr = 0.0;
    switch (kind) {
      case FC_PROPOSAL_LOGSPACE: {  // mh.rs:199-224
        if (cur <= 0.0) prop = 2.2250738585072014e-308;
        else {
          double pr = exp(log(cur) + scale * z);
          prop = isfinite(pr) ? fmax(pr, 2.2250738585072014e-308) : 1.7976931348623157e308;
        }
        if (cur > 0.0 && prop > 0.0) {
          lqf = normal_logpdf(log(prop), log(cur), scale) - log(prop);
          lqr = normal_logpdf(log(cur), log(prop), scale) - log(cur);
        }
      } break;
      case FC_PROPOSAL_REFLECT: {  // mh.rs:230-258
        double lo = A.overrides[g].lower, hi = A.overrides[g].upper;
        prop = cur + scale * z;
        if (hi - lo <= 0.0) prop = cur;
        else {
          for (int it = 0; it < 100000 && (prop < lo || prop > hi); it++) {
            if (prop < lo) prop = 2.0 * lo - prop;
            if (prop > hi) prop = 2.0 * hi - prop;
          }
          prop = fmin(fmax(prop, lo), hi);
        }
      } break;
      case FC_PROPOSAL_PRIOR_RESAMPLE: {  // mh.rs:400-403
        if (!have_params) for (int j = 0; j < op.n_params; j++) pv[j] = eval_param_value(op, j, si.elem, A.q, A.qd, c, C, A.crow, A.drow);
        DevRng r(A.seed, A.chain0 + (uint64_t)c, 0x2000000000000000ULL | A.step_tag);
        long long xi;
        op_dist_sample(r, op, pv[0], pv[1], pv[2], prop, xi);
        const double qn = __longlong_as_double(0x7ff8000000000000LL);
        lqf = op_dist_eval<false>(op, pv[0], pv[1], pv[2], prop, 0, qn).lp;
        lqr = op_dist_eval<false>(op, pv[0], pv[1], pv[2], cur, 0, qn).lp;
      } break;
      default: prop = cur + scale * z; break;  // Gaussian walk (mh.rs:181-187)
    }
    A.old_x[c] = cur;
    A.q[idx] = prop;
    A.lqdiff[c] = lqr - lqf;
  } else {
    const int64_t idx = (int64_t)si.row * C + c;
    const long long cur = A.qd[idx];
    long long prop = cur;
    double lqf = 0.0, lqr = 0.0;
    switch (si.value_type) {
      case FC_T_BOOL: prop = cur ? 0 : 1; break;  // FlipProposal (mh.rs:261-270)
      case FC_T_U64: {                              // DiscreteWalkProposal (mh.rs:283-295)
        long long k = cur + (long long)rint(scale * z);  // f64::round is half-away-from-zero; ties have probability 0
        prop = k >= 0 ? k : -k - 1;
      } break;
      case FC_T_I64: prop = cur + (long long)rint(scale * z); break;  // mh.rs:557-567
      default: {  // usize: resample from the prior (mh.rs:516-530)
        double pv[3] = {0.0, 0.0, 0.0};
        for (int j = 0; j < op.n_params; j++) pv[j] = eval_param_value(op, j, si.elem, A.q, A.qd, c, C, A.crow, A.drow);
        DevRng r(A.seed, A.chain0 + (uint64_t)c, 0x2000000000000000ULL | A.step_tag);
        double xd;
        op_dist_sample(r, op, pv[0], pv[1], pv[2], xd, prop);
        const double qn = __longlong_as_double(0x7ff8000000000000LL);
        lqf = op_dist_eval<false>(op, pv[0], pv[1], pv[2], 0.0, prop, qn).lp;
        lqr = op_dist_eval<false>(op, pv[0], pv[1], pv[2], 0.0, cur, qn).lp;
      } break;
    }
    A.old_x[c] = (double)cur;
    A.qd[idx] = (disc_t)prop;
    A.lqdiff[c] = lqr - lqf;
  }
}

struct AcceptArgs {
  const SiteInfo* sites;
  double* q;
  disc_t* qd;
  const int32_t* target;
  const double* old_x;
  const double* lqdiff;
  double* lp3_cur;        // [3][C]
  const double* lp3_new;  // [3][C]
  double* total_cur;      // [C]
  const double* total_new;
  double* scale;          // per-chain [n_sites][C]
  double* log_scale;
  uint32_t* acc_cnt;      // per-chain [n_sites][C]  (MH)  |  shared [n_sites] (SMC)
  uint32_t* tot_cnt;
  uint64_t* n_accept;     // [C]
  int64_t C, n_sites;
  uint64_t seed, chain0, step_tag;
  double beta, target_rate, gamma;
  int32_t smc_mode, adapt;
};

__global__ void k_move_accept(const __grid_constant__ AcceptArgs A) {
  int64_t c = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t C = A.C;
  int g = c < C ? A.target[c] : -1;
  const bool valid = g >= 0;
  bool accept = false;
  if (valid) {
    double log_alpha;
    if (A.smc_mode) {  // smc.rs:679-680
      double dprior = A.lp3_new[c] - A.lp3_cur[c];
      double dll = (A.lp3_new[C + c] + A.lp3_new[2 * C + c]) - (A.lp3_cur[C + c] + A.lp3_cur[2 * C + c]);
      log_alpha = dprior + A.beta * dll;
    } else {  // mh.rs:731-732 (dim_term = 0 for fixed structure)
      log_alpha = (A.total_new[c] - A.total_cur[c]) + A.lqdiff[c];
    }
    Philox ph(A.seed, A.chain0 + (uint64_t)c);
    double u_site, u_acc;
    ph.uniform2(0, A.step_tag, u_site, u_acc);
    accept = log_alpha >= 0.0 || u_acc < exp(log_alpha);  // mh.rs:733
    if (accept) {
      for (int a = 0; a < 3; a++) A.lp3_cur[a * C + c] = A.lp3_new[a * C + c];
      A.total_cur[c] = A.total_new[c];
      if (A.n_accept) A.n_accept[c] += 1;
    } else {
      if (A.smc_mode) A.q[(int64_t)g * C + c] = A.old_x[c];
      else {
        const SiteInfo si = A.sites[g];
        if (si.is_cont) A.q[(int64_t)si.row * C + c] = A.old_x[c];
        else A.qd[(int64_t)si.row * C + c] = (disc_t)A.old_x[c];
      }
    }
  }
  if (A.smc_mode) {
    if (!A.adapt) return;
    // shared per-site counters: aggregate in shared memory, one global atomic per site per CTA
    __shared__ unsigned s_tot[64], s_acc[64];
    if (A.n_sites <= 64) {
      for (int i = threadIdx.x; i < 64; i += blockDim.x) { s_tot[i] = 0; s_acc[i] = 0; }
      __syncthreads();
      if (valid) { atomicAdd(&s_tot[g], 1u); if (accept) atomicAdd(&s_acc[g], 1u); }
      __syncthreads();
      for (int i = threadIdx.x; i < A.n_sites; i += blockDim.x) {
        if (s_tot[i]) atomicAdd(&A.tot_cnt[i], s_tot[i]);
        if (s_acc[i]) atomicAdd(&A.acc_cnt[i], s_acc[i]);
      }
    } else if (valid) { atomicAdd(&A.tot_cnt[g], 1u); if (accept) atomicAdd(&A.acc_cnt[g], 1u); }
  } else if (A.adapt && valid) {  // DiminishingAdaptation::update (mcmc_utils.rs:88-150), per chain and address
    int64_t i = (int64_t)g * C + c;
    uint32_t tot = A.tot_cnt[i] + 1, acc = A.acc_cnt[i] + (accept ? 1u : 0u);
    A.tot_cnt[i] = tot;
    A.acc_cnt[i] = acc;
    if (tot >= 10) {
      double rate = (double)acc / (double)tot;
      double step = 1.0 / pow((double)tot, A.gamma);
      double ls = A.log_scale[i] + step * (rate - A.target_rate);
      double ns = exp(ls);
      double sc = (isfinite(ns) && ns > 0.0) ? fmin(fmax(ns, 0.001), 100.0) : 1.0;
      A.scale[i] = sc;
      A.log_scale[i] = (sc == 1.0) ? 0.0 : log(sc);
    }
  }
}

// shared-scale update after one SMC rejuvenation sweep (see fc_smc.cu for the documented deviation)
__global__ void k_shared_adapt(double* scale, double* log_scale, uint32_t* acc, uint32_t* tot, double* sweeps, int64_t n_sites, double target, double gamma) {
  int64_t s = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= n_sites) return;
  uint32_t t = tot[s], a = acc[s];
  tot[s] = 0; acc[s] = 0;
  if (t < 10) return;
  double k = sweeps[s] + 1.0;
  sweeps[s] = k;
  double rate = (double)a / (double)t;
  double ls = log_scale[s] + 2.0 * pow(k, -gamma) * (rate - target);
  double ns = exp(ls);
  double sc = (isfinite(ns) && ns > 0.0) ? fmin(fmax(ns, 0.001), 100.0) : 1.0;
  scale[s] = sc;
  log_scale[s] = (sc == 1.0) ? 0.0 : log(sc);
}

__global__ void k_fill_d(double* p, double v, int64_t n) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) p[i] = v;
}

int32_t MoveEngine::init(fc_program* P_, int64_t C_, bool smc_mode_, uint64_t seed_, uint64_t chain0_) {
  P = P_; C = C_; smc_mode = smc_mode_; seed = seed_; chain0 = chain0_;
  n_sites = smc_mode ? P->d_cont : (int64_t)P->sites_all.size();
  cudaStream_t st = P->stream;
  int64_t ns = std::max<int64_t>(1, n_sites);
  int64_t per = smc_mode ? ns : ns * C;
  FC_CUDA(target.reserve(C)); FC_CUDA(old_x.reserve(C)); FC_CUDA(lqdiff.reserve(C));
  FC_CUDA(lp3_new.reserve(3 * C)); FC_CUDA(total_new.reserve(C));
  FC_CUDA(scale.reserve(per)); FC_CUDA(log_scale.reserve(per)); FC_CUDA(acc_cnt.reserve(per)); FC_CUDA(tot_cnt.reserve(per));
  FC_CUDA(sweeps.reserve(ns));
  if (!smc_mode) { FC_CUDA(kind.reserve(per)); FC_CUDA(cudaMemsetAsync(kind.p, 0xff, per, st)); }
  k_fill_d<<<(unsigned)cdiv(per, 256), 256, 0, st>>>(scale.p, 1.0, per);  // get_scale default (mcmc_utils.rs:70-77)
  FC_CUDA(cudaMemsetAsync(log_scale.p, 0, per * 8, st));
  FC_CUDA(cudaMemsetAsync(acc_cnt.p, 0, per * 4, st));
  FC_CUDA(cudaMemsetAsync(tot_cnt.p, 0, per * 4, st));
  FC_CUDA(cudaMemsetAsync(sweeps.p, 0, ns * 8, st));
  FC_KERNEL_CHECK();
  P->n_kernel_launches += 1;
  return FC_OK;
}

int32_t MoveEngine::set_overrides(const std::vector<SiteOverride>& ov) {
  FC_CUDA(overrides.reserve(ov.size()));
  FC_CUDA(cudaMemcpyAsync(overrides.p, ov.data(), ov.size() * sizeof(SiteOverride), cudaMemcpyHostToDevice, P->stream));
  FC_CUDA(cudaStreamSynchronize(P->stream));
  has_overrides = true;
  return FC_OK;
}

int32_t MoveEngine::step(double* q, disc_t* qd, double* lp3_cur, double* total_cur, uint64_t* n_accept, uint64_t step_tag, double beta, bool adapt, double target_rate, double gamma) {
  cudaStream_t st = P->stream;
  unsigned nb = (unsigned)cdiv(C, 128);
  MoveArgs M;
  M.ops = P->d_all_ops.p; M.sites = P->d_sites_all.p; M.crow = P->d_crow.p; M.drow = P->d_drow.p;
  M.overrides = has_overrides ? overrides.p : nullptr;
  M.q = q; M.qd = qd; M.target = target.p; M.old_x = old_x.p; M.lqdiff = lqdiff.p; M.scale = scale.p; M.kind = kind.p;
  M.C = C; M.n_sites = n_sites; M.d_cont = P->d_cont; M.seed = seed; M.chain0 = chain0; M.step_tag = step_tag; M.smc_mode = smc_mode ? 1 : 0;
  k_move_propose<<<nb, 128, 0, st>>>(M);
  FC_KERNEL_CHECK();
  FC_TRY(score_dev(P, q, qd, C, lp3_new.p, total_new.p, nullptr));
  AcceptArgs A;
  A.sites = P->d_sites_all.p; A.q = q; A.qd = qd; A.target = target.p; A.old_x = old_x.p; A.lqdiff = lqdiff.p;
  A.lp3_cur = lp3_cur; A.lp3_new = lp3_new.p; A.total_cur = total_cur; A.total_new = total_new.p;
  A.scale = scale.p; A.log_scale = log_scale.p; A.acc_cnt = acc_cnt.p; A.tot_cnt = tot_cnt.p; A.n_accept = n_accept;
  A.C = C; A.n_sites = n_sites; A.seed = seed; A.chain0 = chain0; A.step_tag = step_tag;
  A.beta = beta; A.target_rate = target_rate; A.gamma = gamma; A.smc_mode = smc_mode ? 1 : 0; A.adapt = adapt ? 1 : 0;
  k_move_accept<<<(unsigned)cdiv(C, 256), 256, 0, st>>>(A);
  FC_KERNEL_CHECK();
  P->n_kernel_launches += 2;
  return FC_OK;
}

int32_t MoveEngine::shared_adapt(double target_rate, double gamma) {
  k_shared_adapt<<<(unsigned)cdiv(std::max<int64_t>(1, n_sites), 128), 128, 0, P->stream>>>(scale.p, log_scale.p, acc_cnt.p, tot_cnt.p, sweeps.p, n_sites, target_rate, gamma);
  FC_KERNEL_CHECK();
  P->n_kernel_launches += 1;
  return FC_OK;
}

__global__ void k_disc_to_i64(const disc_t* __restrict__ in, long long* __restrict__ out, int64_t n) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) out[i] = in[i];
}
int32_t disc_to_host(fc_program* P, const disc_t* d_qd, int64_t n, int64_t* h_out) {
  if (n == 0 || !h_out) return FC_OK;
  FC_CUDA(P->ws.qd64.reserve(n + 1));
  k_disc_to_i64<<<(unsigned)cdiv(n, 256), 256, 0, P->stream>>>(d_qd, (long long*)P->ws.qd64.p, n);
  FC_KERNEL_CHECK();
  P->n_kernel_launches += 1;
  FC_CUDA(cudaMemcpyAsync(h_out, P->ws.qd64.p, n * 8, cudaMemcpyDeviceToHost, P->stream));
  return FC_OK;
}

}  // namespace fc

struct fc_mh_session {
  fc_program* P = nullptr;
  fc_mh_config cfg;
  int64_t C = 0;
  uint64_t n_warmup = 0, iter = 0;
  fc::MoveEngine eng;
  fc::DevBuf<double> q, lp3, total;
  fc::DevBuf<fc::disc_t> qd;
  fc::DevBuf<uint64_t> n_accept;
};

extern "C" {

void fc_mh_config_default(fc_mh_config* cfg) {  // DiminishingAdaptation::new(0.44, 0.7) (mh.rs:946)
  cfg->target_accept = 0.44;
  cfg->gamma = 0.7;
}

int32_t fc_mh_destroy(fc_mh_session* s) {
  if (!s) return FC_OK;
  if (s->P && s->P->stream) cudaStreamSynchronize(s->P->stream);
  delete s;
  return FC_OK;
}

int32_t fc_mh_create(fc_program* p, const fc_mh_config* cfg, uint64_t seed, size_t C, size_t n_warmup, const fc_site_proposal* overrides, size_t n_overrides,
                     uint64_t first_chain, fc_mh_session** out) {
  using namespace fc;
  if (!p || !cfg || !out || C == 0) { set_error("bad argument"); return FC_ERR_INVALID_ARGUMENT; }
  if (!(cfg->target_accept > 0.0 && cfg->target_accept < 1.0) || !(cfg->gamma > 0.5 && cfg->gamma <= 1.0)) {
    set_error("DiminishingAdaptation needs 0 < target < 1 and 0.5 < gamma <= 1 (mcmc_utils.rs:41-42)");
    return FC_ERR_INVALID_ARGUMENT;
  }
  *out = nullptr;
  FC_CUDA(cudaSetDevice(p->device));
  std::unique_ptr<fc_mh_session> S(new fc_mh_session());
  S->P = p; S->cfg = *cfg; S->C = (int64_t)C; S->n_warmup = n_warmup;
  int64_t Cn = S->C;
  FC_CUDA(S->q.reserve(std::max<int64_t>(1, p->d_cont * Cn)));
  FC_CUDA(S->qd.reserve(std::max<int64_t>(1, p->d_disc * Cn)));
  FC_CUDA(S->lp3.reserve(3 * Cn)); FC_CUDA(S->total.reserve(Cn)); FC_CUDA(S->n_accept.reserve(Cn));
  FC_CUDA(cudaMemsetAsync(S->n_accept.p, 0, Cn * 8, p->stream));
  FC_TRY(S->eng.init(p, Cn, false, seed, first_chain));
  if (n_overrides > 0) {
    std::vector<SiteOverride> ov(p->sites_all.size(), SiteOverride{-1, 0, 0.0, 0.0});
    for (size_t i = 0; i < n_overrides; i++) {
      std::string a = overrides[i].address ? overrides[i].address : "";
      bool found = false;
      for (size_t g = 0; g < p->sites_all_names.size(); g++)
        if (p->sites_all_names[g] == a) {
          if (!p->sites_all[g].is_cont) { set_error("proposal override for non-f64 site '%s'", a.c_str()); return FC_ERR_TYPE_MISMATCH; }
          ov[g] = SiteOverride{overrides[i].kind, 0, overrides[i].lower, overrides[i].upper};
          found = true;
        }
      if (!found) { set_error("proposal override for unknown address '%s'", a.c_str()); return FC_ERR_TRACE_ADDRESS_NOT_FOUND; }
    }
    FC_TRY(S->eng.set_overrides(ov));
  }
  FC_TRY(prior_sample_dev(p, S->q.p, S->qd.p, Cn, seed, first_chain, 0x7000000000000000ULL));  // mh.rs:950-957
  FC_TRY(score_dev(p, S->q.p, S->qd.p, Cn, S->lp3.p, S->total.p, nullptr));
  FC_CUDA(cudaStreamSynchronize(p->stream));
  *out = S.release();
  return FC_OK;
}

int32_t fc_mh_step(fc_mh_session* s, size_t n) {
  if (!s) { fc::set_error("session is NULL"); return FC_ERR_INVALID_ARGUMENT; }
  FC_CUDA(cudaSetDevice(s->P->device));
  for (size_t i = 0; i < n; i++) {
    bool adapt = s->iter < s->n_warmup;  // warmup adapts, sampling is frozen (mh.rs:968-1011)
    FC_TRY(s->eng.step(s->q.p, s->qd.p, s->lp3.p, s->total.p, s->n_accept.p, s->iter, 1.0, adapt, s->cfg.target_accept, s->cfg.gamma));
    s->iter++;
  }
  return FC_OK;
}

int32_t fc_mh_get(fc_mh_session* s, double* h_q, int64_t* h_qd, double* h_lp, fc_mh_stats* h_stats, double* h_scales) {
  if (!s) { fc::set_error("session is NULL"); return FC_ERR_INVALID_ARGUMENT; }
  fc_program* p = s->P;
  cudaStream_t st = p->stream;
  int64_t C = s->C;
  if (h_q && p->d_cont) FC_CUDA(cudaMemcpyAsync(h_q, s->q.p, p->d_cont * C * 8, cudaMemcpyDeviceToHost, st));
  if (h_qd && p->d_disc) FC_TRY(fc::disc_to_host(p, s->qd.p, p->d_disc * C, h_qd));
  if (h_lp) FC_CUDA(cudaMemcpyAsync(h_lp, s->total.p, C * 8, cudaMemcpyDeviceToHost, st));
  if (h_scales && s->eng.n_sites) FC_CUDA(cudaMemcpyAsync(h_scales, s->eng.scale.p, s->eng.n_sites * C * 8, cudaMemcpyDeviceToHost, st));
  if (h_stats) {
    std::vector<uint64_t> na(C);
    FC_CUDA(cudaMemcpyAsync(na.data(), s->n_accept.p, C * 8, cudaMemcpyDeviceToHost, st));
    FC_CUDA(cudaStreamSynchronize(st));
    for (int64_t c = 0; c < C; c++) { h_stats[c].n_accept = na[c]; h_stats[c].n_steps = s->iter; }
  }
  FC_CUDA(cudaStreamSynchronize(st));
  return FC_OK;
}

int32_t fc_mh_n_sites(const fc_mh_session* s, int64_t* n) { *n = s->eng.n_sites; return FC_OK; }

// adaptive_mcmc_chain (mh.rs:921-1014): n_warmup adapting steps, then n_samples frozen steps; every
// `thin`-th sampling step is stored. h_draws [n_kept][d_cont][C], h_ddraws [n_kept][d_disc][C],
// h_lp [n_kept][C], n_kept = floor(n_samples / thin) (every thin-th sampling step). Any output may be NULL.
int32_t fc_mh_run(fc_program* p, const fc_mh_config* cfg, uint64_t seed, size_t C, size_t n_warmup, size_t n_samples, size_t thin, double* h_draws,
                  int64_t* h_ddraws, double* h_lp, fc_mh_stats* h_stats) {
  fc_mh_session* s = nullptr;
  if (thin == 0) thin = 1;
  FC_TRY(fc_mh_create(p, cfg, seed, C, n_warmup, nullptr, 0, 0, &s));
  int32_t rc = fc_mh_step(s, n_warmup);
  size_t kept = 0;
  for (size_t i = 0; i < n_samples && rc == FC_OK; i++) {
    rc = fc_mh_step(s, 1);
    if (rc == FC_OK && (i % thin) == thin - 1) {
      rc = fc_mh_get(s, h_draws ? h_draws + kept * p->d_cont * C : nullptr, h_ddraws ? h_ddraws + kept * p->d_disc * C : nullptr, h_lp ? h_lp + kept * C : nullptr, nullptr, nullptr);
      kept++;
    }
  }
  if (rc == FC_OK && h_stats) rc = fc_mh_get(s, nullptr, nullptr, nullptr, h_stats, nullptr);
  fc_mh_destroy(s);
  return rc;
}
}
