// fc_hmc.cu -- C independent HMC chains resident on one B200.
//
// Device restatement of src/inference/hmc.rs: hmc_transition (:419-473), leapfrog (:353-407),
// find_reasonable_epsilon (:479-535), DualAveraging (:141-184), Welford (:187-233) and the
// HmcSession phase logic (:667-919). Each chain is one independent hmc_chain() call of the
// reference: own step size, own dual-averaging state, own diagonal mass matrix. All chains
// take the same number of leapfrog steps, so the trajectory is one kernel sequence for the
// whole batch; per-chain control flow (divergence, accept, eps search) is data, not branches
// on the host.
//
// Differences from the reference, by design:
//  * force = analytic gradient from the scoring kernels (grad_mode 0); the reference's central
//    differences (hmc.rs:304-329) are available as grad_mode 1 for small d (2d scoring passes).
//  * the gradient at the current state is cached across transitions, so a trajectory of L steps
//    costs L gradient evaluations instead of L+1, and the endpoint's log-joint comes from the
//    last gradient pass instead of an extra score_full (hmc.rs:451).
//  * momentum/accept randomness: Philox streams keyed by (seed, chain).
// State layout: [d][C] doubles, chain-minor (coalesced across chains).
#include "fc_internal.cuh"

struct fc_hmc_session {
  fc_program* P = nullptr;
  fc_hmc_config cfg;
  uint64_t seed = 0;
  int64_t C = 0, d = 0;
  uint64_t n_warmup = 0, iter = 0, L = 16;
  uint64_t chain0 = 0;  // global index of the first chain (multi-GPU sharding keeps streams distinct)
  bool frozen = false;  // post-warmup step size has been fixed (hmc.rs:819-838)
  bool pinned = false;  // set_step_size() was called
  uint64_t welford_n = 0;
  int64_t mass_adapt_at = -1;
  uint64_t grad_evals = 0;  // per chain
  fc::DevBuf<double> q, q_new, p, g_cur, g_new, m_inv, mass_sqrt, wf_mean, wf_m2;
  fc::DevBuf<double> lp3, lj, lj_new, lp3_cur, h0, kpart, eps, eps_trial, ap;
  fc::DevBuf<double> da;  // [4][C]: mu, log_eps_bar, h_bar, m
  fc::DevBuf<fc::disc_t> qd;
  fc::DevBuf<int32_t> divergent, accepted, fre_state;
  fc::DevBuf<uint64_t> n_accept, n_divergent;
  fc::DevBuf<double> fd_q, fd_lp;  // finite-difference scratch
  fc::DevBuf<double> p0;           // momentum of the step-size search
  int64_t ksplits = 1;
  fc::DevBuf<double> ext_z, ext_u;  // caller-supplied randomness for the next transition (parity tests)
  bool use_ext = false;
  bool use_hier = false;             // leapfrog steps run as the fused hierarchical kernels (fc_hier.cu)
  fc::HierPlan hier;
  int64_t kin_splits = 1;            // rows of kpart the last trajectory's kinetic energy is spread over
  bool recording = false;            // the next transition records its leapfrog points (step_recorded)
  fc::DevBuf<double> rec_q, rec_h;  // [L+1][d][C], [L+1][C]
};

namespace fc {

#define EW_TX 32
#define EW_TY 8

// momentum refresh (hmc.rs:436-441): p = z * mass_sqrt, q_new = q; kinetic partials.
__global__ void k_momentum(const double* __restrict__ q, double* __restrict__ q_new, double* __restrict__ p, const double* __restrict__ mass_sqrt,
                           const double* __restrict__ m_inv, double* __restrict__ kpart, int64_t d, int64_t C, int64_t rows_per_split,
                           uint64_t seed, uint64_t chain0, uint64_t iter_tag, const double* __restrict__ z_ext) {
  int64_t c = (int64_t)blockIdx.x * EW_TX + threadIdx.x;
  int64_t s0 = (int64_t)blockIdx.y * rows_per_split, s1 = min(d, s0 + rows_per_split);
  double k = 0.0;
  if (c < C) {
    Philox ph(seed, chain0 + (uint64_t)c);
    // rows are drawn in pairs (2 normals per Philox block): pair index = s / 2
    for (int64_t sp = s0 / 2 + threadIdx.y; sp * 2 < s1; sp += EW_TY) {
      double z0, z1;
      ph.normal2((uint64_t)sp, iter_tag, z0, z1);
#pragma unroll
      for (int h = 0; h < 2; h++) {
        int64_t s = sp * 2 + h;
        if (s >= s0 && s < s1) {
          int64_t idx = s * C + c;
          double zz = z_ext ? z_ext[idx] : (h ? z1 : z0);
          double pv = zz * mass_sqrt[idx];
          p[idx] = pv;
          q_new[idx] = q[idx];
          k += pv * pv * m_inv[idx];
        }
      }
    }
  }
  __shared__ double sm[EW_TY][EW_TX + 1];
  sm[threadIdx.y][threadIdx.x] = k;
  __syncthreads();
  if (threadIdx.y == 0 && c < C) {
    double s = 0.0;
#pragma unroll
    for (int y = 0; y < EW_TY; y++) s += sm[y][threadIdx.x];
    kpart[(int64_t)blockIdx.y * C + c] = s;
  }
}

// one fused leapfrog update (hmc.rs:388-401):
//   p += kick * eps * g            (kick = 0.5 for the first / last half step, 1.0 in between)
//   q += eps * m_inv * p           (if drift)
//   kinetic partials of the updated p (if want_k)
// A non-finite force marks the chain divergent (hmc.rs:384-386).
__global__ void k_leap(double* __restrict__ qn, double* __restrict__ p, const double* __restrict__ g, const double* __restrict__ m_inv,
                       const double* __restrict__ eps, int32_t* __restrict__ divergent, double* __restrict__ kpart, int64_t d, int64_t C,
                       int64_t rows_per_split, double kick, int drift, int want_k) {
  int64_t c = (int64_t)blockIdx.x * EW_TX + threadIdx.x;
  int64_t s0 = (int64_t)blockIdx.y * rows_per_split, s1 = min(d, s0 + rows_per_split);
  double k = 0.0;
  bool bad = false;
  if (c < C) {
    double e = eps[c];
    for (int64_t s = s0 + threadIdx.y; s < s1; s += EW_TY) {
      int64_t idx = s * C + c;
      double gv = g[idx];
      if (!isfinite(gv)) bad = true;
      double pv = p[idx] + kick * e * gv;
      p[idx] = pv;
      double mi = m_inv[idx];
      if (drift) qn[idx] += e * mi * pv;
      k += pv * pv * mi;
    }
    if (bad) atomicOr(&divergent[c], 1);
  }
  if (want_k) {
    __shared__ double sm[EW_TY][EW_TX + 1];
    sm[threadIdx.y][threadIdx.x] = k;
    __syncthreads();
    if (threadIdx.y == 0 && c < C) {
      double s = 0.0;
#pragma unroll
      for (int y = 0; y < EW_TY; y++) s += sm[y][threadIdx.x];
      kpart[(int64_t)blockIdx.y * C + c] = s;
    }
  }
}

// Hamiltonian of a recorded leapfrog point (hmc.rs:372-379): -log_joint + 0.5 * sum p^2 m_inv. A
// chain that has left the support records NaN from that point on (hmc.rs:335-337: the reference
// truncates the recording there).
__global__ void k_record_h(const double* __restrict__ lj, const double* __restrict__ kpart, int64_t ksplits, const int32_t* __restrict__ divergent, double* __restrict__ h_out, int64_t C) {
  int64_t c = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= C) return;
  double k = 0.0;
  for (int64_t s = 0; s < ksplits; s++) k += kpart[s * C + c];
  h_out[c] = divergent[c] ? __longlong_as_double(0x7ff8000000000000LL) : -lj[c] + 0.5 * k;
}

__global__ void k_h0(const double* __restrict__ lj, const double* __restrict__ kpart, int64_t ksplits, double* __restrict__ h0, int32_t* __restrict__ divergent, int64_t C) {
  int64_t c = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= C) return;
  double k = 0.0;
  for (int64_t s = 0; s < ksplits; s++) k += kpart[s * C + c];
  h0[c] = -lj[c] + 0.5 * k;  // hmc.rs:441-442
  divergent[c] = 0;
}

// Metropolis accept (hmc.rs:444-472) + dual averaging (hmc.rs:168-178), one thread per chain.
__global__ void k_accept(const double* __restrict__ lj_new, const double* __restrict__ kpart, int64_t ksplits, const double* __restrict__ h0,
                         double* __restrict__ lj, const int32_t* __restrict__ divergent, int32_t* __restrict__ accepted, double* __restrict__ ap_out,
                         uint64_t* __restrict__ n_accept, uint64_t* __restrict__ n_divergent, double* __restrict__ eps, double* __restrict__ da,
                         int64_t C, uint64_t seed, uint64_t chain0, uint64_t iter_tag, uint64_t u_slot, int warming, double target, const double* __restrict__ u_ext) {
  int64_t c = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= C) return;
  double k = 0.0;
  for (int64_t s = 0; s < ksplits; s++) k += kpart[s * C + c];
  double ljn = lj_new[c];
  bool div = divergent[c] != 0 || !isfinite(ljn);
  double ap = 0.0;
  int acc = 0;
  if (!div) {
    double h_new = -ljn + 0.5 * k;
    ap = fmin(exp(h0[c] - h_new), 1.0);  // f64::min ignores NaN, as fmin does
    Philox ph(seed, chain0 + (uint64_t)c);
    double u, u1;
    ph.uniform2(u_slot, iter_tag, u, u1);
    if (u_ext) u = u_ext[c];
    acc = (u < ap) ? 1 : 0;
  } else {
    n_divergent[c] += 1;
  }
  if (acc) { lj[c] = ljn; n_accept[c] += 1; }
  accepted[c] = acc;
  ap_out[c] = ap;
  if (warming) {
    double mu = da[0 * C + c], log_eps_bar = da[1 * C + c], h_bar = da[2 * C + c], m = da[3 * C + c];
    m += 1.0;
    double a = ap < 0.0 ? 0.0 : (ap > 1.0 ? 1.0 : ap);
    double frac = 1.0 / (m + 10.0);                      // t0 = 10
    h_bar = (1.0 - frac) * h_bar + frac * (target - a);
    double log_eps = mu - (sqrt(m) / 0.05) * h_bar;      // gamma = 0.05
    double w = pow(m, -0.75);                            // kappa = 0.75
    log_eps_bar = w * log_eps + (1.0 - w) * log_eps_bar;
    da[1 * C + c] = log_eps_bar; da[2 * C + c] = h_bar; da[3 * C + c] = m;
    eps[c] = exp(log_eps);
  }
}

// commit accepted proposals + Welford push of the post-transition state (hmc.rs:868-884)
__global__ void k_commit(double* __restrict__ q, const double* __restrict__ q_new, double* __restrict__ g_cur, const double* __restrict__ g_new,
                         double* __restrict__ lp3_cur, const double* __restrict__ lp3_new, const int32_t* __restrict__ accepted,
                         double* __restrict__ wf_mean, double* __restrict__ wf_m2, double wf_n, int do_welford, int64_t d, int64_t C) {
  int64_t c = (int64_t)blockIdx.x * EW_TX + threadIdx.x;
  if (c >= C) return;
  int acc = accepted[c];
  for (int64_t s = (int64_t)blockIdx.y * EW_TY + threadIdx.y; s < d; s += (int64_t)gridDim.y * EW_TY) {
    int64_t idx = s * C + c;
    double x = q[idx];
    if (acc) { x = q_new[idx]; q[idx] = x; g_cur[idx] = g_new[idx]; }
    if (do_welford) {  // hmc.rs:207-217
      double mean = wf_mean[idx];
      double delta = x - mean;
      mean += delta / wf_n;
      wf_mean[idx] = mean;
      wf_m2[idx] += delta * (x - mean);
    }
  }
  if (blockIdx.y == 0 && threadIdx.y == 0 && acc && lp3_cur)
    for (int a = 0; a < 3; a++) lp3_cur[a * C + c] = lp3_new[a * C + c];
}

// M^-1 := Welford variances (hmc.rs:219-233, 886-893)
__global__ void k_set_mass(const double* __restrict__ wf_m2, double wf_n, double* __restrict__ m_inv, double* __restrict__ mass_sqrt, int64_t n) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  double v = 1.0;
  if (wf_n >= 2.0) {
    double s = wf_m2[i] / (wf_n - 1.0);
    v = (isfinite(s) && s > 1e-8) ? s : 1.0;
  }
  m_inv[i] = v;
  mass_sqrt[i] = sqrt(1.0 / v);
}

__global__ void k_fill(double* p, double v, int64_t n) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) p[i] = v;
}

// DualAveraging::new (hmc.rs:152-166) from per-chain eps
__global__ void k_da_init(const double* __restrict__ eps, double* __restrict__ da, int64_t C) {
  int64_t c = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= C) return;
  da[0 * C + c] = log(10.0 * eps[c]);
  da[1 * C + c] = 0.0;
  da[2 * C + c] = 0.0;
  da[3 * C + c] = 0.0;
}
__global__ void k_freeze(const double* __restrict__ da, double* __restrict__ eps, int64_t C) {
  int64_t c = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (c < C) eps[c] = exp(da[1 * C + c]);  // frozen_step() (hmc.rs:181-183)
}

// find_reasonable_epsilon state machine (hmc.rs:517-534), one thread per chain.
// state: 0 = first probe pending, 1 = searching, 2 = done.  da rows reused: [0]=a, [1]=iters.
__global__ void k_fre_update(const double* __restrict__ lj1, const double* __restrict__ kpart, int64_t ksplits, const double* __restrict__ h0,
                             const int32_t* __restrict__ divergent, double* __restrict__ eps_trial, double* __restrict__ aux, int32_t* __restrict__ state,
                             int32_t* __restrict__ n_active, int64_t C) {
  int64_t c = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= C) return;
  int st = state[c];
  if (st == 2) return;
  double k = 0.0;
  for (int64_t s = 0; s < ksplits; s++) k += kpart[s * C + c];
  double l1 = lj1[c];
  double lr = (divergent[c] != 0 || !isfinite(l1)) ? d_neg_inf() : h0[c] - (-l1 + 0.5 * k);
  const double ln_half = -0.6931471805599453, ln2 = 0.6931471805599453;
  double e = eps_trial[c];
  double a, iters;
  if (st == 0) { a = lr > ln_half ? 1.0 : -1.0; iters = 0.0; aux[0 * C + c] = a; }
  else {
    a = aux[0 * C + c];
    iters = aux[1 * C + c] + 1.0;
    bool stop = iters > 100.0 || !(e >= 1e-12 && e <= 1e12);
    if (!stop && a > 0.0 && lr == d_neg_inf()) { e *= 0.5; stop = true; }
    if (stop) { eps_trial[c] = e; state[c] = 2; return; }
  }
  aux[1 * C + c] = iters;
  if (a * lr > -a * ln2) {
    eps_trial[c] = e * (a > 0.0 ? 2.0 : 0.5);
    state[c] = 1;
    atomicAdd(n_active, 1);
  } else {
    state[c] = 2;
  }
}
__global__ void k_fre_finish(const double* __restrict__ eps_trial, double* __restrict__ eps, int64_t C) {
  int64_t c = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (c < C) { double e = eps_trial[c]; eps[c] = e < 1e-6 ? 1e-6 : (e > 1e3 ? 1e3 : e); }
}
__global__ void k_copy_where(double* __restrict__ dst, const double* __restrict__ src, const int32_t* __restrict__ state, int64_t d, int64_t C) {
  // dst[s][c] = src[s][c] for chains still searching (state != 2)
  int64_t c = (int64_t)blockIdx.x * EW_TX + threadIdx.x;
  if (c >= C || state[c] == 2) return;
  for (int64_t s = (int64_t)blockIdx.y * EW_TY + threadIdx.y; s < d; s += (int64_t)gridDim.y * EW_TY) dst[s * C + c] = src[s * C + c];
}

// reference-faithful central finite differences (hmc.rs:315-327): g[s] = (lp(q+h e_s) - lp(q-h e_s)) / 2h
__global__ void k_fd_perturb(double* __restrict__ qw, const double* __restrict__ q, int64_t s, double delta, int64_t d, int64_t C) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= d * C) return;
  double v = q[i];
  if (i / C == s) v += delta;
  qw[i] = v;
}
__global__ void k_fd_store(double* __restrict__ g, const double* __restrict__ lpp, const double* __restrict__ lpm, int64_t s, double h, int64_t C) {
  int64_t c = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (c < C) g[s * C + c] = (lpp[c] - lpm[c]) / (2.0 * h);
}

struct Hmc {
  fc_hmc_session* S;
  fc_program* P;
  cudaStream_t st;
  int64_t C, d;
  dim3 ew_grid, ew_block;
  int64_t rows_per_split;
  unsigned cb;

  explicit Hmc(fc_hmc_session* s) : S(s), P(s->P), st(s->P->stream), C(s->C), d(s->d) {
    ew_block = dim3(EW_TX, EW_TY);
    rows_per_split = cdiv(std::max<int64_t>(d, 1), S->ksplits);
    if (rows_per_split & 1) rows_per_split++;  // momentum pairs must not straddle splits
    ew_grid = dim3((unsigned)cdiv(C, EW_TX), (unsigned)cdiv(std::max<int64_t>(d, 1), rows_per_split));
    cb = (unsigned)cdiv(C, 128);
  }
  int64_t nsplit() const { return ew_grid.y; }

  // value + force at qv -> (lp3, lj_out, g_out)
  // need_value=false: interior leapfrog step, only the force is consumed (lp3 / lj_out are scratch)
  int32_t force(const double* qv, double* lp3, double* lj_out, double* g_out, bool need_value = true) {
    if (S->cfg.grad_mode == 0) {
      FC_TRY(score_dev(P, qv, S->qd.p, C, lp3, lj_out, g_out, need_value));
    } else {
      double h = S->cfg.finite_diff_eps;
      FC_CUDA(S->fd_q.reserve(d * C));
      FC_CUDA(S->fd_lp.reserve(5 * C));
      unsigned nb = (unsigned)cdiv(d * C, 256);
      for (int64_t s = 0; s < d; s++) {
        k_fd_perturb<<<nb, 256, 0, st>>>(S->fd_q.p, qv, s, h, d, C);
        FC_TRY(score_dev(P, S->fd_q.p, S->qd.p, C, S->fd_lp.p + 2 * C, S->fd_lp.p, nullptr));
        k_fd_perturb<<<nb, 256, 0, st>>>(S->fd_q.p, qv, s, -h, d, C);
        FC_TRY(score_dev(P, S->fd_q.p, S->qd.p, C, S->fd_lp.p + 2 * C, S->fd_lp.p + C, nullptr));
        k_fd_store<<<cb, 128, 0, st>>>(g_out, S->fd_lp.p, S->fd_lp.p + C, s, h, C);
        P->n_kernel_launches += 3;
      }
      FC_KERNEL_CHECK();
      if (need_value) FC_TRY(score_dev(P, qv, S->qd.p, C, lp3, lj_out, nullptr));
    }
    S->grad_evals += 1;
    return FC_OK;
  }

  int32_t momentum(uint64_t tag) {
    k_momentum<<<ew_grid, ew_block, 0, st>>>(S->q.p, S->q_new.p, S->p.p, S->mass_sqrt.p, S->m_inv.p, S->kpart.p, d, C, rows_per_split, S->seed, S->chain0, tag, S->use_ext ? S->ext_z.p : nullptr);
    FC_KERNEL_CHECK();
    k_h0<<<cb, 128, 0, st>>>(S->lj.p, S->kpart.p, nsplit(), S->h0.p, S->divergent.p, C);
    FC_KERNEL_CHECK();
    P->n_kernel_launches += 2;
    return FC_OK;
  }
  int32_t leap(const double* g, const double* eps, double kick, int drift, int want_k) {
    k_leap<<<ew_grid, ew_block, 0, st>>>(S->q_new.p, S->p.p, g, S->m_inv.p, eps, S->divergent.p, S->kpart.p, d, C, rows_per_split, kick, drift, want_k);
    FC_KERNEL_CHECK();
    P->n_kernel_launches += 1;
    return FC_OK;
  }
  // L leapfrog steps from (q, p) using the cached force at q; leaves q_new, p, g_new, lj_new, kpart(K_new)
  int32_t trajectory(const double* eps, uint64_t L) {
    FC_TRY(leap(S->g_cur.p, eps, 0.5, 1, 0));
    if (S->use_hier) {
      // force + kick (+ drift) of one leapfrog step in two launches; the last step leaves the force in g_new,
      // the accumulators / log-joint of the end point and the whole kinetic sum in kpart[0]
      for (uint64_t l = 1; l <= L; l++) {
        const bool last = l == L;
        FC_TRY(hier_step(P, S->hier, S->q_new.p, S->p.p, S->m_inv.p, eps, S->g_new.p, S->divergent.p, S->kpart.p, S->lp3.p, S->lj_new.p, C,
                         last ? 0.5 : 1.0, !last, last, last));
        S->grad_evals += 1;
      }
      S->kin_splits = 1;
      return FC_OK;
    }
    S->kin_splits = nsplit();
    for (uint64_t l = 1; l <= L; l++) {
      FC_TRY(force(S->q_new.p, S->lp3.p, S->lj_new.p, S->g_new.p, l == L));
      if (l < L) FC_TRY(leap(S->g_new.p, eps, 1.0, 1, 0));
      else FC_TRY(leap(S->g_new.p, eps, 0.5, 0, 1));
    }
    return FC_OK;
  }

  // The same trajectory with every integration point recorded (leapfrog(.., record), hmc.rs:353-403):
  // the start, then (q, H) after the trailing half kick of each step. The kicks are issued as two
  // half kicks (as the reference does) so the momentum exists at each recorded point.
  int32_t trajectory_recorded(const double* eps, uint64_t L) {
    FC_CUDA(S->rec_q.reserve((int64_t)(L + 1) * d * C));
    FC_CUDA(S->rec_h.reserve((int64_t)(L + 1) * C));
    FC_CUDA(cudaMemcpyAsync(S->rec_q.p, S->q.p, d * C * 8, cudaMemcpyDeviceToDevice, st));
    FC_CUDA(cudaMemcpyAsync(S->rec_h.p, S->h0.p, C * 8, cudaMemcpyDeviceToDevice, st));
    S->kin_splits = nsplit();
    FC_TRY(leap(S->g_cur.p, eps, 0.5, 1, 0));
    for (uint64_t l = 1; l <= L; l++) {
      FC_TRY(force(S->q_new.p, S->lp3.p, S->lj_new.p, S->g_new.p, true));
      FC_TRY(leap(S->g_new.p, eps, 0.5, 0, 1));
      k_record_h<<<cb, 128, 0, st>>>(S->lj_new.p, S->kpart.p, nsplit(), S->divergent.p, S->rec_h.p + (int64_t)l * C, C);
      FC_KERNEL_CHECK();
      P->n_kernel_launches += 1;
      FC_CUDA(cudaMemcpyAsync(S->rec_q.p + (int64_t)l * d * C, S->q_new.p, d * C * 8, cudaMemcpyDeviceToDevice, st));
      if (l < L) FC_TRY(leap(S->g_new.p, eps, 0.5, 1, 0));
    }
    return FC_OK;
  }

  // hmc.rs:479-535 for all chains at once
  int32_t find_reasonable_epsilon() {
    uint64_t tag = 0x4000000000000000ULL | S->iter;  // momentum stream distinct from transitions
    FC_TRY(momentum(tag));
    FC_CUDA(S->eps_trial.reserve(C));
    FC_CUDA(S->fre_state.reserve(C + 1));
    k_fill<<<cb, 128, 0, st>>>(S->eps_trial.p, 1.0, C);
    FC_CUDA(cudaMemsetAsync(S->fre_state.p, 0, (C + 1) * 4, st));
    FC_CUDA(S->p0.reserve(d * C));
    FC_CUDA(cudaMemcpyAsync(S->p0.p, S->p.p, d * C * 8, cudaMemcpyDeviceToDevice, st));
    double* aux = S->ap.p;  // [2][C] scratch (ap is [2][C])
    for (int it = 0; it < 102; it++) {
      // probe: one leapfrog step of size eps_trial from (q, p0)
      if (it > 0) {
        FC_CUDA(cudaMemcpyAsync(S->p.p, S->p0.p, d * C * 8, cudaMemcpyDeviceToDevice, st));
        FC_CUDA(cudaMemcpyAsync(S->q_new.p, S->q.p, d * C * 8, cudaMemcpyDeviceToDevice, st));
        FC_CUDA(cudaMemsetAsync(S->divergent.p, 0, C * 4, st));
      }
      FC_TRY(trajectory(S->eps_trial.p, 1));
      int32_t* n_active = S->fre_state.p + C;
      FC_CUDA(cudaMemsetAsync(n_active, 0, 4, st));
      k_fre_update<<<cb, 128, 0, st>>>(S->lj_new.p, S->kpart.p, S->kin_splits, S->h0.p, S->divergent.p, S->eps_trial.p, aux, S->fre_state.p, n_active, C);
      FC_KERNEL_CHECK();
      P->n_kernel_launches += 1;
      int32_t h_active = 0;
      FC_CUDA(cudaMemcpyAsync(&h_active, n_active, 4, cudaMemcpyDeviceToHost, st));
      FC_CUDA(cudaStreamSynchronize(st));
      if (h_active == 0) break;
    }
    k_fre_finish<<<cb, 128, 0, st>>>(S->eps_trial.p, S->eps.p, C);
    FC_KERNEL_CHECK();
    P->n_kernel_launches += 2;
    return FC_OK;
  }

  // HmcSession::advance (hmc.rs:819-919)
  int32_t step() {
    bool warming = S->iter < S->n_warmup;
    if (!warming && !S->frozen) {
      if (S->n_warmup > 0 && !S->pinned) {
        k_freeze<<<cb, 128, 0, st>>>(S->da.p, S->eps.p, C);
        FC_KERNEL_CHECK();
        P->n_kernel_launches += 1;
      }
      S->frozen = true;
    }
    if (d == 0) {
      // no continuous sites: the reference's advance() draws a fresh prior sample each step and rescoring
      // updates the log-joint and the trace (hmc.rs:826-846)
      FC_TRY(prior_sample_dev(P, S->q.p, S->qd.p, C, S->seed, S->chain0, 0x7100000000000000ULL | (S->iter + 1)));
      FC_TRY(score_dev(P, S->q.p, S->qd.p, C, S->lp3_cur.p, S->lj.p, nullptr));
      S->iter++;
      return FC_OK;
    }
    uint64_t tag = S->iter;
    FC_TRY(momentum(tag));
    if (S->recording) FC_TRY(trajectory_recorded(S->eps.p, S->L));
    else FC_TRY(trajectory(S->eps.p, S->L));
    S->recording = false;
    k_accept<<<cb, 128, 0, st>>>(S->lj_new.p, S->kpart.p, S->kin_splits, S->h0.p, S->lj.p, S->divergent.p, S->accepted.p, S->ap.p, S->n_accept.p, S->n_divergent.p,
                                 S->eps.p, S->da.p, C, S->seed, S->chain0, tag, (uint64_t)((d + 1) / 2 + 1), warming ? 1 : 0, S->cfg.target_accept, S->use_ext ? S->ext_u.p : nullptr);
    FC_KERNEL_CHECK();
    S->use_ext = false;
    bool welford = warming && S->mass_adapt_at >= 0;
    if (welford) S->welford_n += 1;
    dim3 cg((unsigned)cdiv(C, EW_TX), (unsigned)std::min<int64_t>(std::max<int64_t>(1, cdiv(d, EW_TY * 4)), 1024));
    k_commit<<<cg, ew_block, 0, st>>>(S->q.p, S->q_new.p, S->g_cur.p, S->g_new.p, S->lp3_cur.p, S->lp3.p, S->accepted.p, S->wf_mean.p, S->wf_m2.p, (double)S->welford_n, welford ? 1 : 0, d, C);
    FC_KERNEL_CHECK();
    P->n_kernel_launches += 2;
    if (warming && (int64_t)(S->iter + 1) == S->mass_adapt_at) {
      k_set_mass<<<(unsigned)cdiv(d * C, 256), 256, 0, st>>>(S->wf_m2.p, (double)S->welford_n, S->m_inv.p, S->mass_sqrt.p, d * C);
      FC_KERNEL_CHECK();
      P->n_kernel_launches += 1;
      S->iter++;  // the eps search below draws a fresh momentum keyed by the iteration
      FC_TRY(find_reasonable_epsilon());
      k_da_init<<<cb, 128, 0, st>>>(S->eps.p, S->da.p, C);
      FC_KERNEL_CHECK();
      return FC_OK;
    }
    S->iter++;
    return FC_OK;
  }
};

}  // namespace fc

extern "C" {

void fc_hmc_config_default(fc_hmc_config* cfg) {  // HMCConfig::default (hmc.rs:125-135)
  cfg->n_leapfrog = 16;
  cfg->target_accept = 0.8;
  cfg->init_step_size = __builtin_nan("");
  cfg->finite_diff_eps = 1e-5;
  cfg->adapt_mass = 0;
  cfg->grad_mode = 0;
}

int32_t fc_hmc_destroy(fc_hmc_session* s) {
  if (!s) return FC_OK;
  if (s->P && s->P->stream) cudaStreamSynchronize(s->P->stream);
  delete s;
  return FC_OK;
}

int32_t fc_hmc_create_sharded(fc_program* p, const fc_hmc_config* cfg, uint64_t seed, size_t C, size_t n_warmup, const double* h_q0,
                              const int64_t* h_qd0, uint64_t first_chain, fc_hmc_session** out) {
  using namespace fc;
  if (!p || !cfg || !out) { set_error("NULL argument"); return FC_ERR_INVALID_ARGUMENT; }
  *out = nullptr;
  if (C == 0) { set_error("C must be > 0"); return FC_ERR_INVALID_ARGUMENT; }
  FC_CUDA(cudaSetDevice(p->device));
  std::unique_ptr<fc_hmc_session> S(new fc_hmc_session());
  S->P = p; S->cfg = *cfg; S->seed = seed; S->C = (int64_t)C; S->d = p->d_cont; S->n_warmup = n_warmup;
  S->L = cfg->n_leapfrog < 1 ? 1 : cfg->n_leapfrog;  // hmc.rs:690
  S->chain0 = first_chain;
  S->mass_adapt_at = (cfg->adapt_mass && n_warmup >= 4) ? (int64_t)(n_warmup / 2) : -1;  // hmc.rs:704-708
  int64_t d = S->d, Cn = S->C, n = std::max<int64_t>(1, d * Cn);
  S->ksplits = std::max<int64_t>(1, std::min<int64_t>(64, cdiv(d, EW_TY * 8)));
  S->use_hier = cfg->grad_mode == 0 && hier_detect(p, &S->hier);
  cudaStream_t st = p->stream;
  for (auto* b : {&S->q, &S->q_new, &S->p, &S->g_cur, &S->g_new, &S->m_inv, &S->mass_sqrt, &S->wf_mean, &S->wf_m2}) FC_CUDA(b->reserve(n));
  FC_CUDA(S->lp3.reserve(3 * Cn)); FC_CUDA(S->lp3_cur.reserve(3 * Cn));
  for (auto* b : {&S->lj, &S->lj_new, &S->h0, &S->eps}) FC_CUDA(b->reserve(Cn));
  FC_CUDA(S->ap.reserve(2 * Cn));
  FC_CUDA(S->kpart.reserve((S->ksplits + 2) * Cn));
  FC_CUDA(S->da.reserve(4 * Cn));
  FC_CUDA(S->qd.reserve(std::max<int64_t>(1, p->d_disc * Cn)));
  FC_CUDA(S->divergent.reserve(Cn)); FC_CUDA(S->accepted.reserve(Cn));
  FC_CUDA(S->n_accept.reserve(Cn)); FC_CUDA(S->n_divergent.reserve(Cn));
  FC_CUDA(cudaMemsetAsync(S->n_accept.p, 0, Cn * 8, st));
  FC_CUDA(cudaMemsetAsync(S->n_divergent.p, 0, Cn * 8, st));
  FC_CUDA(cudaMemsetAsync(S->wf_mean.p, 0, n * 8, st));
  FC_CUDA(cudaMemsetAsync(S->wf_m2.p, 0, n * 8, st));
  FC_CUDA(cudaMemsetAsync(S->divergent.p, 0, Cn * 4, st));
  k_fill<<<(unsigned)cdiv(n, 256), 256, 0, st>>>(S->m_inv.p, 1.0, n);
  k_fill<<<(unsigned)cdiv(n, 256), 256, 0, st>>>(S->mass_sqrt.p, 1.0, n);
  // initial state: prior draw (hmc.rs:673) unless the caller supplies one
  // (the prior pass walks a chain's sites serially: 12 ms at the C3 shape; skipped when the caller's q0 replaces all of it)
  const bool q0_covers_all = h_q0 && d > 0 && (p->d_disc == 0 || h_qd0);
  if (!q0_covers_all) FC_TRY(prior_sample_dev(p, S->q.p, S->qd.p, Cn, seed, first_chain, 0x7000000000000000ULL));
  if (h_q0 && d > 0) FC_CUDA(cudaMemcpyAsync(S->q.p, h_q0, d * Cn * 8, cudaMemcpyHostToDevice, st));
  if (h_qd0 && p->d_disc > 0) {
    const disc_t* conv = nullptr;
    FC_TRY(convert_qd(p, h_qd0, Cn, &conv));
    FC_CUDA(cudaMemcpyAsync(S->qd.p, conv, p->d_disc * Cn * sizeof(disc_t), cudaMemcpyDeviceToDevice, st));
  }
  Hmc H(S.get());
  FC_TRY(H.force(S->q.p, S->lp3_cur.p, S->lj.p, S->g_cur.p));
  S->grad_evals = 0;
  if (d > 0) {
    if (cfg->init_step_size == cfg->init_step_size) {  // Some(e)
      k_fill<<<(unsigned)cdiv(Cn, 256), 256, 0, st>>>(S->eps.p, cfg->init_step_size, Cn);
    } else {
      FC_TRY(H.find_reasonable_epsilon());
    }
  } else {
    k_fill<<<(unsigned)cdiv(Cn, 256), 256, 0, st>>>(S->eps.p, 1.0, Cn);
  }
  k_da_init<<<(unsigned)cdiv(Cn, 128), 128, 0, st>>>(S->eps.p, S->da.p, Cn);
  FC_KERNEL_CHECK();
  p->n_kernel_launches += 4;
  FC_CUDA(cudaStreamSynchronize(st));
  *out = S.release();
  return FC_OK;
}

int32_t fc_hmc_create(fc_program* p, const fc_hmc_config* cfg, uint64_t seed, size_t C, size_t n_warmup, const double* h_q0, const int64_t* h_qd0, fc_hmc_session** out) {
  return fc_hmc_create_sharded(p, cfg, seed, C, n_warmup, h_q0, h_qd0, 0, out);
}

int32_t fc_hmc_step(fc_hmc_session* s, size_t n) {
  if (!s) { fc::set_error("session is NULL"); return FC_ERR_INVALID_ARGUMENT; }
  FC_CUDA(cudaSetDevice(s->P->device));
  fc::Hmc H(s);
  for (size_t i = 0; i < n; i++) FC_TRY(H.step());
  return FC_OK;
}

// HmcSession::step_recorded (hmc.rs:811-817): one transition whose leapfrog points are recorded.
// h_q [L+1][d][C], h_h [L+1][C] (either may be NULL); point 0 is the start of the trajectory.
int32_t fc_hmc_step_recorded(fc_hmc_session* s, double* h_q, double* h_h, uint64_t* n_points) {
  if (!s) { fc::set_error("session is NULL"); return FC_ERR_INVALID_ARGUMENT; }
  FC_CUDA(cudaSetDevice(s->P->device));
  fc::Hmc H(s);
  const uint64_t L = s->L;
  s->recording = s->d > 0;
  FC_TRY(H.step());
  s->recording = false;
  cudaStream_t st = s->P->stream;
  if (n_points) *n_points = s->d > 0 ? L + 1 : 0;
  if (s->d > 0) {
    if (h_q) FC_CUDA(cudaMemcpyAsync(h_q, s->rec_q.p, (L + 1) * s->d * s->C * 8, cudaMemcpyDeviceToHost, st));
    if (h_h) FC_CUDA(cudaMemcpyAsync(h_h, s->rec_h.p, (L + 1) * s->C * 8, cudaMemcpyDeviceToHost, st));
  }
  FC_CUDA(cudaStreamSynchronize(st));
  return FC_OK;
}

// Parity hook: the NEXT momentum refresh uses h_z [d][C] as its standard normals and the next
// accept test uses h_u [C] as its uniform (the reference's rng.sample(StandardNormal) /
// rng.gen::<f64>() draws, hmc.rs:436-441,461), so a transition can be compared with the CPU
// oracle on identical randomness.
int32_t fc_hmc_set_explicit_randomness(fc_hmc_session* s, const double* h_z, const double* h_u) {
  if (!s || !h_z || !h_u) { fc::set_error("NULL argument"); return FC_ERR_INVALID_ARGUMENT; }
  cudaStream_t st = s->P->stream;
  FC_CUDA(s->ext_z.reserve(std::max<int64_t>(1, s->d * s->C)));
  FC_CUDA(s->ext_u.reserve(s->C));
  if (s->d > 0) FC_CUDA(cudaMemcpyAsync(s->ext_z.p, h_z, s->d * s->C * 8, cudaMemcpyHostToDevice, st));
  FC_CUDA(cudaMemcpyAsync(s->ext_u.p, h_u, s->C * 8, cudaMemcpyHostToDevice, st));
  FC_CUDA(cudaStreamSynchronize(st));
  s->use_ext = true;
  return FC_OK;
}

// find_reasonable_epsilon (hmc.rs:479-535) at the current state; the result becomes the
// session's step size and is returned per chain.
int32_t fc_hmc_find_reasonable_epsilon(fc_hmc_session* s, double* h_eps) {
  if (!s) { fc::set_error("session is NULL"); return FC_ERR_INVALID_ARGUMENT; }
  fc::Hmc H(s);
  FC_TRY(H.find_reasonable_epsilon());
  s->use_ext = false;
  if (h_eps) FC_CUDA(cudaMemcpyAsync(h_eps, s->eps.p, s->C * 8, cudaMemcpyDeviceToHost, s->P->stream));
  FC_CUDA(cudaStreamSynchronize(s->P->stream));
  return FC_OK;
}

int32_t fc_hmc_sync(fc_hmc_session* s) {
  FC_CUDA(cudaStreamSynchronize(s->P->stream));
  return FC_OK;
}

int32_t fc_hmc_get(fc_hmc_session* s, double* h_q, double* h_lp, fc_hmc_stats* h_stats) {
  if (!s) { fc::set_error("session is NULL"); return FC_ERR_INVALID_ARGUMENT; }
  cudaStream_t st = s->P->stream;
  int64_t C = s->C, d = s->d;
  if (h_q && d > 0) FC_CUDA(cudaMemcpyAsync(h_q, s->q.p, d * C * 8, cudaMemcpyDeviceToHost, st));
  if (h_lp) FC_CUDA(cudaMemcpyAsync(h_lp, s->lj.p, C * 8, cudaMemcpyDeviceToHost, st));
  if (h_stats) {
    std::vector<uint64_t> na(C), nd(C);
    std::vector<double> e(C), da(4 * C);
    FC_CUDA(cudaMemcpyAsync(na.data(), s->n_accept.p, C * 8, cudaMemcpyDeviceToHost, st));
    FC_CUDA(cudaMemcpyAsync(nd.data(), s->n_divergent.p, C * 8, cudaMemcpyDeviceToHost, st));
    FC_CUDA(cudaMemcpyAsync(e.data(), s->eps.p, C * 8, cudaMemcpyDeviceToHost, st));
    FC_CUDA(cudaMemcpyAsync(da.data(), s->da.p, 4 * C * 8, cudaMemcpyDeviceToHost, st));
    FC_CUDA(cudaStreamSynchronize(st));
    bool use_frozen = s->iter >= s->n_warmup && !s->frozen && s->n_warmup > 0 && !s->pinned;  // step_size() (hmc.rs:771-777)
    for (int64_t c = 0; c < C; c++) {
      h_stats[c].n_accept = na[c];
      h_stats[c].n_divergent = nd[c];
      h_stats[c].step_size = use_frozen ? std::exp(da[C + c]) : e[c];
    }
  }
  FC_CUDA(cudaStreamSynchronize(st));
  return FC_OK;
}

// Device pointers of the resident state (q [d][C], lj [C]) for zero-copy consumers.
int32_t fc_hmc_device_state(fc_hmc_session* s, double** d_q, double** d_lj, double** d_lp3) {
  if (d_q) *d_q = s->q.p;
  if (d_lj) *d_lj = s->lj.p;
  if (d_lp3) *d_lp3 = s->lp3_cur.p;
  return FC_OK;
}

int32_t fc_hmc_trace(fc_hmc_session* s, double* h_log_prior, double* h_log_lik, double* h_log_factors) {
  cudaStream_t st = s->P->stream;
  int64_t C = s->C;
  double* o[3] = {h_log_prior, h_log_lik, h_log_factors};
  for (int a = 0; a < 3; a++)
    if (o[a]) FC_CUDA(cudaMemcpyAsync(o[a], s->lp3_cur.p + a * C, C * 8, cudaMemcpyDeviceToHost, st));
  FC_CUDA(cudaStreamSynchronize(st));
  return FC_OK;
}

int32_t fc_hmc_set_position(fc_hmc_session* s, const double* h_q) {
  if (!s || !h_q) { fc::set_error("NULL argument"); return FC_ERR_INVALID_ARGUMENT; }
  cudaStream_t st = s->P->stream;
  if (s->d > 0) FC_CUDA(cudaMemcpyAsync(s->q.p, h_q, s->d * s->C * 8, cudaMemcpyHostToDevice, st));
  fc::Hmc H(s);
  uint64_t ge = s->grad_evals;
  FC_TRY(H.force(s->q.p, s->lp3_cur.p, s->lj.p, s->g_cur.p));
  s->grad_evals = ge;
  FC_CUDA(cudaStreamSynchronize(st));
  return FC_OK;
}

int32_t fc_hmc_set_step_size(fc_hmc_session* s, double eps) {  // hmc.rs:738-744
  if (!s) { fc::set_error("session is NULL"); return FC_ERR_INVALID_ARGUMENT; }
  eps = eps < 1e-12 ? 1e-12 : eps;
  fc::k_fill<<<(unsigned)fc::cdiv(s->C, 256), 256, 0, s->P->stream>>>(s->eps.p, eps, s->C);
  FC_KERNEL_CHECK();
  s->pinned = true;
  s->frozen = true;
  if (s->n_warmup > s->iter) s->n_warmup = s->iter;
  return FC_OK;
}

int32_t fc_hmc_set_n_leapfrog(fc_hmc_session* s, uint64_t l) {  // hmc.rs:748-750
  if (!s) { fc::set_error("session is NULL"); return FC_ERR_INVALID_ARGUMENT; }
  s->L = l < 1 ? 1 : l;
  return FC_OK;
}

int32_t fc_hmc_grad_evals(const fc_hmc_session* s, uint64_t* per_chain) {
  *per_chain = s->grad_evals;
  return FC_OK;
}

int32_t fc_hmc_iterations(const fc_hmc_session* s, uint64_t* iters, int32_t* warming) {
  if (iters) *iters = s->iter;
  if (warming) *warming = s->iter < s->n_warmup;
  return FC_OK;
}

int32_t fc_hmc_last_accept_prob(fc_hmc_session* s, double* h_ap, int32_t* h_accepted, int32_t* h_divergent) {
  cudaStream_t st = s->P->stream;
  int64_t C = s->C;
  if (h_ap) FC_CUDA(cudaMemcpyAsync(h_ap, s->ap.p, C * 8, cudaMemcpyDeviceToHost, st));
  if (h_accepted) FC_CUDA(cudaMemcpyAsync(h_accepted, s->accepted.p, C * 4, cudaMemcpyDeviceToHost, st));
  if (h_divergent) FC_CUDA(cudaMemcpyAsync(h_divergent, s->divergent.p, C * 4, cudaMemcpyDeviceToHost, st));
  FC_CUDA(cudaStreamSynchronize(st));
  return FC_OK;
}

// n further transitions with every post-transition state kept: out_draws [n][d][C], out_lp [n][C] (either
// may be NULL). `kind` says where they live: host pointers go through one async copy per draw, device
// pointers (the draws stay in HBM for fc_diagnostics_dev; 41 GB at the C3 shape) through a device copy.
static int32_t hmc_sample_into(fc_hmc_session* s, size_t n, double* out_draws, double* out_lp, cudaMemcpyKind kind) {
  cudaStream_t st = s->P->stream;
  const size_t d = (size_t)s->d, C = (size_t)s->C;
  for (size_t i = 0; i < n; i++) {
    FC_TRY(fc_hmc_step(s, 1));
    if (out_draws && d > 0) FC_CUDA(cudaMemcpyAsync(out_draws + i * d * C, s->q.p, d * C * 8, kind, st));
    if (out_lp) FC_CUDA(cudaMemcpyAsync(out_lp + i * C, s->lj.p, C * 8, kind, st));
  }
  return FC_OK;
}
int32_t fc_hmc_sample_dev(fc_hmc_session* s, size_t n, double* d_draws, double* d_lp) {
  if (!s) { fc::set_error("session is NULL"); return FC_ERR_INVALID_ARGUMENT; }
  return hmc_sample_into(s, n, d_draws, d_lp, cudaMemcpyDeviceToDevice);
}
static int32_t hmc_run_impl(fc_program* p, const fc_hmc_config* cfg, uint64_t seed, size_t C, size_t n_warmup, size_t n_samples, double* draws, double* lp,
                            fc_hmc_stats* h_stats, cudaMemcpyKind kind) {
  fc_hmc_session* s = nullptr;
  FC_TRY(fc_hmc_create(p, cfg, seed, C, n_warmup, nullptr, nullptr, &s));
  int32_t rc = fc_hmc_step(s, n_warmup);
  if (rc == FC_OK) rc = hmc_sample_into(s, n_samples, draws, lp, kind);
  if (rc == FC_OK && h_stats) rc = fc_hmc_get(s, nullptr, nullptr, h_stats);
  if (cudaStreamSynchronize(p->stream) != cudaSuccess && rc == FC_OK) rc = FC_ERR_CUDA;
  fc_hmc_destroy(s);
  return rc;
}
// hmc_chain (hmc.rs:566-583): warmup, then n_samples draws. h_draws [n_samples][d][C], h_lp [n_samples][C].
int32_t fc_hmc_run(fc_program* p, const fc_hmc_config* cfg, uint64_t seed, size_t C, size_t n_warmup, size_t n_samples, double* h_draws, double* h_lp, fc_hmc_stats* h_stats) {
  return hmc_run_impl(p, cfg, seed, C, n_warmup, n_samples, h_draws, h_lp, h_stats, cudaMemcpyDeviceToHost);
}
// The same run with the draws left on the device: d_draws / d_lp are DEVICE pointers of the same shapes.
int32_t fc_hmc_run_dev(fc_program* p, const fc_hmc_config* cfg, uint64_t seed, size_t C, size_t n_warmup, size_t n_samples, double* d_draws, double* d_lp, fc_hmc_stats* h_stats) {
  return hmc_run_impl(p, cfg, seed, C, n_warmup, n_samples, d_draws, d_lp, h_stats, cudaMemcpyDeviceToDevice);
}
}
