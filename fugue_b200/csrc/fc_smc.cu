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
#include "fc_fastmath.cuh"

namespace fc {
int32_t dev_reduce(int op, const double* d_x, int64_t n, double m, double* d_scratch, double* d_out, cudaStream_t st);
int32_t dev_log_sum_exp(const double* d_x, int64_t n, double* d_scratch, double* h_out, cudaStream_t st);
struct ResampleWork;
ResampleWork* resample_work_new();
void resample_work_free(ResampleWork*);
int32_t resample_dev(int method, const double* d_w, int64_t n, const double* d_us, uint64_t* d_idx, ResampleWork& W, cudaStream_t st, uint64_t* launches);
int32_t shard_scan(const double* d_w, int64_t n, double carry, ResampleWork& W, cudaStream_t st, uint64_t* launches, double* end_value);
int32_t shard_ancestors(ResampleWork& W, int64_t n, double n_total, double u_raw, int64_t jlo, int64_t cnt, uint64_t* d_idx, cudaStream_t st);

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

// ---- next_beta (smc.rs:588-622) and the reweight log-sum-exp (smc.rs:512-524) ----
// ess_at(b) = exp(2 lse(lv) - lse(2 lv)), lv_i = log_w_i + (b - beta) ll_i. One launch per
// evaluation: a single streaming pass keeps the running (max, sum e, sum e^2) triple with
// e = exp(lv - max) -- sum exp(2 lv - 2 max) is sum e^2, so one exp per particle serves both
// log-sum-exps -- the last block to finish folds the per-block triples in fixed order and
// advances the bisection state on the device. The bisection stops issuing work once the
// interval has collapsed (mid == lo or mid == hi is a fixed point of the reference's loop).
struct Bisect {
  double beta, lo, hi, mid, target, n;
  double lse1;     // lse(lv) of the last evaluation (the reweight normaliser when mid == beta_new)
  int done;        // 1: ess_at(1) >= target; 2: interval collapsed
  unsigned counter;
  double max_ll;   // pre-pass (uniform log_w only): max of ll
  int bad;         // pre-pass: some ll is NaN or +inf => every lse is non-finite => ESS := n (smc.rs:595-601)
  int _pad;
};
struct Tri { double m, s1, s2; };
__device__ __forceinline__ void tri_merge(Tri& a, const Tri& b) {
  if (b.m == d_neg_inf() && b.s1 == 0.0) return;
  if (a.m == d_neg_inf() && a.s1 == 0.0) { a = b; return; }
  double m = fmax(a.m, b.m);
  double fa = a.m == m ? 1.0 : exp(a.m - m), fb = b.m == m ? 1.0 : exp(b.m - m);
  a.s1 = a.s1 * fa + b.s1 * fb;
  a.s2 = a.s2 * (fa * fa) + b.s2 * (fb * fb);
  a.m = m;
}
#define NB_THREADS 256
__device__ __forceinline__ void tri_push8(Tri& acc, const double (&lv)[8]) {
  double cm = lv[0];
#pragma unroll
  for (int k = 1; k < 8; k++) cm = fmax(cm, lv[k]);  // NaN-ignoring, as f64::max (numerical.rs:19)
  {  // branch-free rescale to the new running max (f == 1 when the max is unchanged)
    const double mn = fmax(acc.m, cm);
    const double f = fast_exp(acc.m - mn);  // acc.m == -inf: the sums are still 0
    acc.s1 *= f; acc.s2 *= f * f; acc.m = mn;
  }
  if (acc.m == d_neg_inf()) {  // only -inf (or NaN) so far: exp(x - max) is NaN for NaN, else the lse is -inf
#pragma unroll
    for (int k = 0; k < 8; k++) if (lv[k] != lv[k]) { acc.s1 += lv[k]; acc.s2 += lv[k]; }
    return;
  }
#pragma unroll
  for (int k = 0; k < 8; k++) {
    double x = lv[k] - acc.m;
    double e = (x == x) ? fast_exp(x) : x;  // NaN propagates into the sums as exp(NaN) does
    if (x < -708.0) e = 0.0;
    acc.s1 += e;
    acc.s2 += e * e;
  }
}
// bisection update for one ESS evaluation (smc.rs:605-619)
__device__ __forceinline__ void nb_update(Bisect* B, double ess, int first) {
  if (first) {
    if (ess >= B->target) { B->hi = 1.0; B->done = 1; return; }
    B->lo = B->beta; B->hi = 1.0;
  } else {
    if (ess < B->target) B->hi = B->mid; else B->lo = B->mid;
  }
  B->mid = 0.5 * (B->lo + B->hi);
  if (B->mid == B->lo || B->mid == B->hi) B->done = 2;
}
// CTA reduction of the per-thread triples; the last CTA to finish folds the per-CTA triples in
// block order (deterministic) and advances the bisection state.
__device__ __forceinline__ void nb_fold_update(Tri acc, Bisect* __restrict__ B, Tri* __restrict__ part, int mode, int first) {
  __shared__ Tri sm[NB_THREADS / 32];
  __shared__ bool is_last;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    Tri b;
    b.m = __shfl_down_sync(0xffffffffu, acc.m, o); b.s1 = __shfl_down_sync(0xffffffffu, acc.s1, o); b.s2 = __shfl_down_sync(0xffffffffu, acc.s2, o);
    tri_merge(acc, b);
  }
  if ((threadIdx.x & 31) == 0) sm[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int w = 1; w < NB_THREADS / 32; w++) tri_merge(acc, sm[w]);
    part[blockIdx.x] = acc;
    __threadfence();
    unsigned t = atomicAdd(&B->counter, 1u);
    is_last = (t == gridDim.x - 1);
  }
  __syncthreads();
  if (!is_last) return;
  __threadfence();
  // last block: fold the per-block triples in block order (deterministic), then update the state
  Tri t; t.m = d_neg_inf(); t.s1 = 0.0; t.s2 = 0.0;
  const int nb = gridDim.x;
  const int per = (nb + NB_THREADS - 1) / NB_THREADS;
  for (int i = threadIdx.x * per; i < min(nb, (threadIdx.x + 1) * per); i++) {
    Tri b;  // written by other blocks: read through L2
    b.m = __ldcg(&part[i].m); b.s1 = __ldcg(&part[i].s1); b.s2 = __ldcg(&part[i].s2);
    tri_merge(t, b);
  }
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {  // ordered (left-to-right) tree
    Tri b;
    b.m = __shfl_down_sync(0xffffffffu, t.m, o); b.s1 = __shfl_down_sync(0xffffffffu, t.s1, o); b.s2 = __shfl_down_sync(0xffffffffu, t.s2, o);
    if (((threadIdx.x & 31) & (2 * o - 1)) == 0) tri_merge(t, b);
  }
  __syncthreads();
  if ((threadIdx.x & 31) == 0) sm[threadIdx.x >> 5] = t;
  __syncthreads();
  if (threadIdx.x != 0) return;
  for (int w = 1; w < NB_THREADS / 32; w++) tri_merge(t, sm[w]);
  part[gridDim.x] = t;  // the folded (max, sum e, sum e^2): read by the sharded driver (fc_smc_shard_tri)
  B->counter = 0;
  const double ninf = d_neg_inf();
  double lse1 = (t.m == ninf && t.s1 == 0.0) ? ninf : (t.s1 == 0.0 ? ninf : t.m + log(t.s1));
  double lse2 = (t.m == ninf && t.s2 == 0.0) ? ninf : (t.s2 == 0.0 ? ninf : 2.0 * t.m + log(t.s2));
  B->lse1 = lse1;
  if (mode != 0) return;
  double ess = (!isfinite(lse1) || !isfinite(lse2)) ? B->n : exp(2.0 * lse1 - lse2);
  nb_update(B, ess, first);
}

// mode 0: bisection probe at B->mid (first = 1: the ess_at(1.0) probe); mode 1: lse only at d_fixed
__global__ void __launch_bounds__(NB_THREADS) k_nb_eval(const double* __restrict__ log_w, double lw_const, const double* __restrict__ ll, Bisect* __restrict__ B,
                                                        Tri* __restrict__ part, int64_t P, int mode, int first, double d_fixed) {
  if (mode == 0 && B->done) return;
  const double d = mode == 0 ? B->mid - B->beta : d_fixed;
  Tri acc; acc.m = d_neg_inf(); acc.s1 = 0.0; acc.s2 = 0.0;
  const int64_t nthreads = (int64_t)gridDim.x * NB_THREADS;
  const int64_t P2 = P >> 1;  // double2 units
  for (int64_t v0 = (int64_t)blockIdx.x * NB_THREADS + threadIdx.x; v0 < P2; v0 += nthreads * 4) {
    double lv[8];
#pragma unroll
    for (int k = 0; k < 4; k++) {
      int64_t v = v0 + (int64_t)k * nthreads;
      if (v < P2) {
        double2 a = __ldg((const double2*)ll + v);
        double2 w = log_w ? __ldg((const double2*)log_w + v) : make_double2(lw_const, lw_const);
        lv[2 * k] = w.x + d * a.x; lv[2 * k + 1] = w.y + d * a.y;
      } else lv[2 * k] = lv[2 * k + 1] = d_neg_inf();
    }
    tri_push8(acc, lv);
  }
  if ((P & 1) && blockIdx.x == 0 && threadIdx.x == 0) {
    double lv[8];
#pragma unroll
    for (int k = 0; k < 8; k++) lv[k] = d_neg_inf();
    lv[0] = (log_w ? log_w[P - 1] : lw_const) + d * ll[P - 1];
    tri_push8(acc, lv);
  }
  nb_fold_update(acc, B, part, mode, first);
}

// ---- uniform log_w (the state adaptive_smc is in whenever next_beta runs: initially and after
// every resample): lv_i = c + d ll_i, so max lv = c + d max(ll) and the pass needs no running max.
// k_nb_prepass (once per next_beta call): max(ll) and the "some ll is NaN / +inf" flag. ----
__global__ void __launch_bounds__(NB_THREADS) k_nb_prepass(const double* __restrict__ ll, int64_t P, Bisect* __restrict__ B, double* __restrict__ part) {
  double mx = d_neg_inf();
  int bad = 0;
  for (int64_t i = (int64_t)blockIdx.x * NB_THREADS + threadIdx.x; i < P; i += (int64_t)gridDim.x * NB_THREADS) {
    double v = __ldg(ll + i);
    bad |= (v != v) || v == d_pos_inf();
    mx = fmax(mx, v);
  }
  mx = warp_max(mx);
  bad = __any_sync(0xffffffffu, bad);
  __shared__ double sm[NB_THREADS / 32];
  __shared__ int sbad[NB_THREADS / 32];
  __shared__ bool is_last;
  if ((threadIdx.x & 31) == 0) { sm[threadIdx.x >> 5] = mx; sbad[threadIdx.x >> 5] = bad; }
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int w = 1; w < NB_THREADS / 32; w++) { mx = fmax(mx, sm[w]); bad |= sbad[w]; }
    part[blockIdx.x] = bad ? d_pos_inf() : mx;  // +inf doubles as the bad marker
    __threadfence();
    is_last = atomicAdd(&B->counter, 1u) == gridDim.x - 1;
  }
  __syncthreads();
  if (!is_last || threadIdx.x != 0) return;
  __threadfence();
  double m = d_neg_inf();
  for (int i = 0; i < (int)gridDim.x; i++) m = fmax(m, __ldcg(part + i));
  B->counter = 0;
  B->bad = m == d_pos_inf();
  B->max_ll = m;
}
// exp(x) for x <= 0: fast_exp_tab (fc_fastmath.cuh: 2^k * 2^(j/32) * exp(r), 32-entry table in shared
// memory, degree-6 polynomial, <= 2 ulp) without the upper clamp -- 13 FP64 instructions instead of
// the 19 of the degree-13 series; these sweeps are bound by the FP64 pipe, not by HBM. Below -708
// it returns ~3e-308 instead of 0, negligible against a sum that contains exp(0) = 1 for the max.
__device__ __forceinline__ double exp_nonpos(double x, const double* __restrict__ tab) {
  x = fmax(x, -708.0);
  const double C32 = 46.16624130844683;
  const double L_HI = 0.02166084938653512, L_LO = 5.9631716539705866e-12;
  double t = fma(x, C32, 6755399441055744.0);
  const int ti = __double2loint(t);
  t -= 6755399441055744.0;
  double r = fma(t, -L_HI, x);
  r = fma(t, -L_LO, r);
  double q = 1.388888888888889e-03;
  q = fma(q, r, 8.333333333333333e-03);
  q = fma(q, r, 4.1666666666666664e-02);
  q = fma(q, r, 1.6666666666666666e-01);
  q = fma(q, r, 0.5);
  const double p = fma(r * r, q, r);
  const double T = tab[ti & 31];
  const double v = fma(T, p, T);
  return __hiloint2double(__double2hiint(v) + ((ti >> 5) << 20), __double2loint(v));
}
__global__ void __launch_bounds__(NB_THREADS) k_nb_eval_uni(double lw_const, const double* __restrict__ ll, Bisect* __restrict__ B, Tri* __restrict__ part, int64_t P,
                                                            int mode, int first, double d_fixed) {
  if (mode == 0 && B->done) return;
  __shared__ double s_et[32];
  fast_exp_tab_init(s_et, threadIdx.x);
  __syncthreads();
  if (B->bad) {  // every lse is non-finite: ess_at(.) = n (smc.rs:595-601); the reweight normaliser is NaN
    if (blockIdx.x == 0 && threadIdx.x == 0) {
      B->lse1 = __longlong_as_double(0x7ff8000000000000LL);
      if (mode == 0) nb_update(B, B->n, first);
    }
    return;
  }
  const double d = mode == 0 ? B->mid - B->beta : d_fixed;
  const double m = fma(d, B->max_ll, lw_const);  // the same expression as lv of the max element: x <= 0, == 0 there
  Tri acc; acc.m = m; acc.s1 = 0.0; acc.s2 = 0.0;
  double t1 = 0.0, t2 = 0.0;  // second accumulator pair (shorter dependency chains)
  if (m > d_neg_inf()) {
    const int64_t nthreads = (int64_t)gridDim.x * NB_THREADS;
    const int64_t P2 = P >> 1;
    // software pipeline: the next trip's four 16-byte loads are in flight while this trip's eight exp
    // are evaluated (ncu profiles/ncu_c4_nb_eval_r1_v12.txt: 60 % of the warp samples waited on the
    // loads, DRAM at 33 % -- every warp loaded, then every warp computed)
    double2 a[4], nx[4];
    const int64_t vstart = (int64_t)blockIdx.x * NB_THREADS + threadIdx.x;
#pragma unroll
    for (int k = 0; k < 4; k++) {
      int64_t v = vstart + (int64_t)k * nthreads;
      nx[k] = v < P2 ? __ldg((const double2*)ll + v) : make_double2(d_neg_inf(), d_neg_inf());
    }
    for (int64_t v0 = vstart; v0 < P2; v0 += nthreads * 4) {
#pragma unroll
      for (int k = 0; k < 4; k++) {
        a[k] = nx[k];
        int64_t v = v0 + nthreads * 4 + (int64_t)k * nthreads;
        nx[k] = v < P2 ? __ldg((const double2*)ll + v) : make_double2(d_neg_inf(), d_neg_inf());
      }
#pragma unroll
      for (int k = 0; k < 4; k++) {
        const double e0 = exp_nonpos(fma(d, a[k].x, lw_const) - m, s_et), e1 = exp_nonpos(fma(d, a[k].y, lw_const) - m, s_et);
        acc.s1 += e0; acc.s2 = fma(e0, e0, acc.s2);
        t1 += e1; t2 = fma(e1, e1, t2);
      }
    }
    if ((P & 1) && blockIdx.x == 0 && threadIdx.x == 0) {
      const double e0 = exp_nonpos(fma(d, ll[P - 1], lw_const) - m, s_et);
      acc.s1 += e0; acc.s2 = fma(e0, e0, acc.s2);
    }
    acc.s1 += t1; acc.s2 += t2;
  }
  nb_fold_update(acc, B, part, mode, first);
}

// K candidate increments in one pass over ll (uniform log_w): the sharded driver evaluates three
// bisection levels (7 candidates) per round, so a next_beta search costs ~22 collectives instead of
// ~55. Per-block partial sums [block][2K] (sum e, sum e^2 per candidate, shift m_k = lwc + d_k max ll);
// the host folds the blocks in order.
#define NB_MULTI_MAX 8
struct MultiArgs { double d[NB_MULTI_MAX]; int K; };
__global__ void __launch_bounds__(NB_THREADS) k_nb_multi(double lw_const, const double* __restrict__ ll, const Bisect* __restrict__ B, int64_t P,
                                                         const __grid_constant__ MultiArgs A, double* __restrict__ part) {
  const int K = A.K;
  __shared__ double s_et[32];
  fast_exp_tab_init(s_et, threadIdx.x);
  __syncthreads();
  double m[NB_MULTI_MAX], s1[NB_MULTI_MAX], s2[NB_MULTI_MAX];
#pragma unroll
  for (int k = 0; k < NB_MULTI_MAX; k++) { m[k] = fma(A.d[k], B->max_ll, lw_const); s1[k] = 0.0; s2[k] = 0.0; }
  for (int64_t i = (int64_t)blockIdx.x * NB_THREADS + threadIdx.x; i < P; i += (int64_t)gridDim.x * NB_THREADS) {
    const double a = __ldg(ll + i);
#pragma unroll
    for (int k = 0; k < NB_MULTI_MAX; k++) {
      if (k < K) {
        const double e = m[k] > d_neg_inf() ? exp_nonpos(fma(A.d[k], a, lw_const) - m[k], s_et) : 0.0;
        s1[k] += e;
        s2[k] = fma(e, e, s2[k]);
      }
    }
  }
  __shared__ double sm[NB_THREADS / 32][2 * NB_MULTI_MAX];
#pragma unroll
  for (int k = 0; k < NB_MULTI_MAX; k++) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) { s1[k] += __shfl_down_sync(0xffffffffu, s1[k], o); s2[k] += __shfl_down_sync(0xffffffffu, s2[k], o); }
    if ((threadIdx.x & 31) == 0) { sm[threadIdx.x >> 5][2 * k] = s1[k]; sm[threadIdx.x >> 5][2 * k + 1] = s2[k]; }
  }
  __syncthreads();
  if (threadIdx.x < 2 * NB_MULTI_MAX) {
    double t = 0.0;
    for (int w = 0; w < NB_THREADS / 32; w++) t += sm[w][threadIdx.x];
    part[(int64_t)blockIdx.x * 2 * NB_MULTI_MAX + threadIdx.x] = t;
  }
}

// normalised weights after the reweight: log_w' = (log_w + d ll) - log_norm, w = exp(log_w')
__global__ void k_reweight(const double* __restrict__ log_w, double lw_const, const double* __restrict__ ll, double d, const Bisect* __restrict__ B,
                           double* __restrict__ out_log, double* __restrict__ out_w, int64_t P, double n_unif) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= P) return;
  const double log_norm = B->lse1;
  if (!isfinite(log_norm)) {  // normalize_particles' uniform fallback (smc.rs:719-753); n_unif = population size
    if (out_log) out_log[i] = -log(n_unif);
    out_w[i] = 1.0 / n_unif;
    return;
  }
  double v = ((log_w ? log_w[i] : lw_const) + d * ll[i]) - log_norm;
  if (out_log) out_log[i] = v;
  out_w[i] = exp(v);
}

// gather particle planes by ancestor index (resample_particles, smc.rs:326-349)
template <class T>
__global__ void __launch_bounds__(256) k_gather(const T* __restrict__ src, T* __restrict__ dst, const uint64_t* __restrict__ idx, int64_t rows, int64_t P, int64_t n_out = -1, int64_t dst_stride = 0) {
  int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (n_out >= 0) {  // n_out offspring written plane-major with stride dst_stride (sharded resampling)
    if (j >= n_out) return;
    const int64_t a2 = (int64_t)idx[j];
    for (int64_t r = (int64_t)blockIdx.y; r < rows; r += gridDim.y) dst[r * dst_stride + j] = src[r * P + a2];
    return;
  }
  if (j >= P) return;
  const int64_t a = (int64_t)idx[j];
  // rows r = blockIdx.y*8 + k (+ 8*gridDim.y ...): 8 independent loads in flight per thread
  for (int64_t r0 = (int64_t)blockIdx.y * 8; r0 < rows; r0 += (int64_t)gridDim.y * 8) {
    T v[8];
#pragma unroll
    for (int k = 0; k < 8; k++) if (r0 + k < rows) v[k] = src[(r0 + k) * P + a];
#pragma unroll
    for (int k = 0; k < 8; k++) if (r0 + k < rows) dst[(r0 + k) * P + j] = v[k];
  }
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
  fc::DevBuf<fc::Tri> tri;
  bool log_w_uniform = true;  // log_w == -ln n everywhere (initially and after every resample)
  fc::ResampleWork* rw = nullptr;
  double beta = 0.0, log_evidence = 0.0;
  uint64_t steps = 0, moves = 0;
  bool finished = false;
  double n_global = 0.0;  // population size the uniform weight refers to (0: this session's own n)
  bool shard_prepass_ok = false, shard_ll_bad = false;  // max(ll) / bad flag in `bis` describe the current ll (fc_smc_shard_tri)
};

namespace fc {

// one resident wave: 4 CTAs of 256 threads per SM (56 registers), never more than RDB partials
static int nb_blocks(int64_t n) {
  static int sms = 0;
  if (!sms) { int dev = 0; cudaGetDevice(&dev); cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev); if (sms <= 0) sms = 148; }
  return (int)std::max<int64_t>(1, std::min<int64_t>(std::min<int64_t>(RDB, 4 * sms), cdiv(n, NB_THREADS * 8)));
}

static int32_t next_beta_dev(cudaStream_t st, const double* lw, double lwc, const double* ll, int64_t n, double beta, double target_ess, Bisect* d_bis, Tri* d_tri,
                             uint64_t* launches, double* out) {
  const int nb = nb_blocks(n);
  Bisect hb;
  hb.beta = beta; hb.lo = beta; hb.hi = 1.0; hb.mid = 1.0; hb.target = target_ess; hb.n = (double)n; hb.lse1 = 0.0; hb.done = 0; hb.counter = 0;
  hb.max_ll = 0.0; hb.bad = 0; hb._pad = 0;
  FC_CUDA(cudaMemcpyAsync(d_bis, &hb, sizeof(hb), cudaMemcpyHostToDevice, st));
  if (!lw) { k_nb_prepass<<<nb, NB_THREADS, 0, st>>>(ll, n, d_bis, (double*)d_tri); if (launches) *launches += 1; }
  for (int it = 0; it < 65; it++) {
    if (lw) k_nb_eval<<<nb, NB_THREADS, 0, st>>>(lw, lwc, ll, d_bis, d_tri, n, 0, it == 0 ? 1 : 0, 0.0);
    else k_nb_eval_uni<<<nb, NB_THREADS, 0, st>>>(lwc, ll, d_bis, d_tri, n, 0, it == 0 ? 1 : 0, 0.0);
    if (launches) *launches += 1;
    if (it == 0 || it == 56) {  // early exits: ess_at(1) >= target, or the interval has collapsed
      FC_CUDA(cudaMemcpyAsync(&hb, d_bis, sizeof(hb), cudaMemcpyDeviceToHost, st));
      FC_CUDA(cudaStreamSynchronize(st));
      if (hb.done == 1) { *out = 1.0; return FC_OK; }
      if (hb.done == 2) break;
    }
  }
  FC_KERNEL_CHECK();
  FC_CUDA(cudaMemcpyAsync(&hb, d_bis, sizeof(hb), cudaMemcpyDeviceToHost, st));
  FC_CUDA(cudaStreamSynchronize(st));
  *out = std::fmin(std::fmax(hb.hi, beta + 1e-9), 1.0);  // smc.rs:621
  return FC_OK;
}
static int32_t smc_next_beta(fc_smc_session* S, double target_ess, double* out) {
  return next_beta_dev(S->P->stream, S->log_w_uniform ? nullptr : S->log_w.p, -std::log((double)S->n), S->ll.p, S->n, S->beta, target_ess, S->bis.p, S->tri.p,
                       &S->P->n_kernel_launches, out);
}

// reweight (smc.rs:512-531): log_norm = lse(log_w + d ll); w = exp(log_w + d ll - log_norm).
// log_w itself is only materialised when it will be read again (the final step).
static int32_t smc_reweight(fc_smc_session* S, double d_beta, bool write_log_w, double* log_norm) {
  cudaStream_t st = S->P->stream;
  int64_t n = S->n;
  const double* lw = S->log_w_uniform ? nullptr : S->log_w.p;
  const double lwc = -std::log((double)n);
  FC_CUDA(cudaMemsetAsync(&S->bis.p->counter, 0, sizeof(unsigned), st));
  // (max_ll / bad in the Bisect block are those of this step's next_beta pre-pass: same ll)
  if (lw) k_nb_eval<<<nb_blocks(n), NB_THREADS, 0, st>>>(lw, lwc, S->ll.p, S->bis.p, S->tri.p, n, 1, 0, d_beta);
  else k_nb_eval_uni<<<nb_blocks(n), NB_THREADS, 0, st>>>(lwc, S->ll.p, S->bis.p, S->tri.p, n, 1, 0, d_beta);
  k_reweight<<<(unsigned)cdiv(n, 256), 256, 0, st>>>(lw, lwc, S->ll.p, d_beta, S->bis.p, write_log_w ? S->comb.p : nullptr, S->w.p, n, (double)n);
  FC_KERNEL_CHECK();
  S->P->n_kernel_launches += 2;
  Bisect hb;
  FC_CUDA(cudaMemcpyAsync(&hb, S->bis.p, sizeof(hb), cudaMemcpyDeviceToHost, st));
  FC_CUDA(cudaStreamSynchronize(st));
  *log_norm = hb.lse1;
  if (write_log_w) { std::swap(S->log_w.p, S->comb.p); std::swap(S->log_w.n, S->comb.n); S->log_w_uniform = false; }
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
  double log_norm;
  FC_TRY(smc_reweight(S, d_beta, beta_new >= 1.0, &log_norm));
  S->log_evidence += log_norm;
  S->beta = beta_new;
  if (S->beta < 1.0) {
    // resample (smc.rs:534-541)
    int64_t nu = S->cfg.resampling_method == FC_RESAMPLE_SYSTEMATIC ? 1 : n;
    FC_CUDA(S->us.reserve(nu + 1));
    k_uniform_one<<<(unsigned)cdiv(cdiv(nu, 2), 256), 256, 0, st>>>(S->us.p, S->seed, 0x1000000000000000ULL | S->steps, nu);
    FC_TRY(resample_dev(S->cfg.resampling_method, S->w.p, n, S->us.p, S->idx.p, *S->rw, st, &P->n_kernel_launches));
    dim3 gg(nbk, 1);
    if (P->d_cont) { gg.y = (unsigned)std::min<int64_t>(cdiv(P->d_cont, 8), 64); k_gather<double><<<gg, 256, 0, st>>>(S->q.p, S->q2.p, S->idx.p, P->d_cont, n); std::swap(S->q.p, S->q2.p); std::swap(S->q.n, S->q2.n); }
    // large populations: one CTA row walks all planes (the ancestor index is read once, not once per plane group: 2.89 -> 2.23 ms at 2^24 x 80 planes)
    if (P->d_disc) { gg.y = n >= (1 << 20) ? 1u : (unsigned)std::min<int64_t>(cdiv(P->d_disc, 8), 64); k_gather<disc_t><<<gg, 256, 0, st>>>(S->qd.p, S->qd2.p, S->idx.p, P->d_disc, n); std::swap(S->qd.p, S->qd2.p); std::swap(S->qd.n, S->qd2.n); }
    gg.y = 1;
    k_gather<double><<<gg, 256, 0, st>>>(S->lp3.p, S->lp3b.p, S->idx.p, 3, n);
    std::swap(S->lp3.p, S->lp3b.p); std::swap(S->lp3.n, S->lp3b.n);
    S->log_w_uniform = true;  // weight 1/n, log_weight ln(1/n) (smc.rs:338-345); materialised on demand
    FC_KERNEL_CHECK();
    P->n_kernel_launches += 4;
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
  (void)neg_ln_n;
  return FC_OK;
}

// standalone probe of next_beta on host arrays (parity test against the oracle)
int32_t next_beta_host(const double* h_log_w, const double* h_ll, size_t n, double beta, double target_ess, double* out) {
  DevBuf<double> lw, ll;
  DevBuf<Bisect> bis;
  DevBuf<Tri> tri;
  FC_CUDA(lw.reserve(n)); FC_CUDA(ll.reserve(n)); FC_CUDA(bis.reserve(1)); FC_CUDA(tri.reserve(RDB + 1));
  bool uniform = true;  // the state adaptive_smc calls next_beta in: takes the single-stream path
  for (size_t i = 1; i < n && uniform; i++) uniform = h_log_w[i] == h_log_w[0];
  if (!uniform) FC_CUDA(cudaMemcpy(lw.p, h_log_w, n * 8, cudaMemcpyHostToDevice));
  FC_CUDA(cudaMemcpy(ll.p, h_ll, n * 8, cudaMemcpyHostToDevice));
  return next_beta_dev(nullptr, uniform ? nullptr : lw.p, h_log_w[0], ll.p, (int64_t)n, beta, target_ess, bis.p, tri.p, nullptr, out);
}

// log_w as an array (callers that read it: fc_smc_get, fc_smc_device_state)
static int32_t smc_materialise_log_w(fc_smc_session* S) {
  if (!S->log_w_uniform) return FC_OK;
  k_fill2<<<(unsigned)cdiv(S->n, 256), 256, 0, S->P->stream>>>(S->log_w.p, -std::log(S->n_global > 0.0 ? S->n_global : (double)S->n), S->n);
  FC_KERNEL_CHECK();
  S->P->n_kernel_launches += 1;
  return FC_OK;
}

}  // namespace fc

extern "C" {

int32_t fc_next_beta(const double* h_log_w, const double* h_ll, size_t n, double beta, double target_ess, double* out) {  // smc.rs:588-622
  if (!h_log_w || !h_ll || !out || n == 0) { fc::set_error("NULL / empty argument"); return FC_ERR_INVALID_ARGUMENT; }
  return fc::next_beta_host(h_log_w, h_ll, n, beta, target_ess, out);
}

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
  FC_CUDA(S->idx.reserve(N)); FC_CUDA(S->n_accept.reserve(N)); FC_CUDA(S->bis.reserve(1)); FC_CUDA(S->tri.reserve(RDB + 1));
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
      s->log_w_uniform = false;
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
  FC_TRY(smc_materialise_log_w(s));
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
  if (d_log_w) { FC_TRY(fc::smc_materialise_log_w(s)); *d_log_w = s->log_w.p; }
  if (d_loglik) *d_loglik = s->ll.p;
  return FC_OK;
}

// ---- standalone pieces of the SMC public surface, on the particles resident in a session ----

// smc_prior_particles' weighting (smc.rs:779-787, FG-03): log_weight = log-likelihood of the particle
int32_t fc_smc_prior_weights(fc_smc_session* s) {
  using namespace fc;
  if (!s) { set_error("session is NULL"); return FC_ERR_INVALID_ARGUMENT; }
  FC_CUDA(cudaSetDevice(s->P->device));
  FC_CUDA(cudaMemcpyAsync(s->log_w.p, s->ll.p, s->n * 8, cudaMemcpyDeviceToDevice, s->P->stream));
  s->log_w_uniform = false;
  return FC_OK;
}

// normalize_particles (smc.rs:719-754): weight = exp(log_weight - lse), uniform if every log_weight is
// -inf; log_weight is left as it is. Returns ESS = 1 / sum w^2 (smc.rs:230-233) in *ess (may be NULL).
int32_t fc_smc_normalize(fc_smc_session* s, double* log_norm_out, double* ess) {
  using namespace fc;
  if (!s) { set_error("session is NULL"); return FC_ERR_INVALID_ARGUMENT; }
  fc_program* p = s->P;
  FC_CUDA(cudaSetDevice(p->device));
  cudaStream_t st = p->stream;
  const int64_t n = s->n;
  unsigned nbk = (unsigned)cdiv(n, 256);
  FC_TRY(smc_materialise_log_w(s));
  double log_norm;
  FC_TRY(dev_log_sum_exp(s->log_w.p, n, s->scratch.p, &log_norm, st));
  if (std::isfinite(log_norm)) k_sub_exp<<<nbk, 256, 0, st>>>(s->log_w.p, log_norm, s->comb.p, s->w.p, n);
  else k_fill2<<<nbk, 256, 0, st>>>(s->w.p, 1.0 / (double)n, n);
  FC_KERNEL_CHECK();
  p->n_kernel_launches += 5;
  if (log_norm_out) *log_norm_out = log_norm;
  if (ess) {
    FC_TRY(dev_reduce(2, s->w.p, n, 0.0, s->scratch.p, s->scratch.p + RDB, st));
    double ss;
    FC_CUDA(cudaMemcpyAsync(&ss, s->scratch.p + RDB, 8, cudaMemcpyDeviceToHost, st));
    FC_CUDA(cudaStreamSynchronize(st));
    *ess = 1.0 / ss;
  }
  return FC_OK;
}

// resample_particles (smc.rs:326-349): indices by `method` from the normalised weights, ancestors
// copied, weight = 1/n and log_weight = ln(1/n). h_uniforms (1 for systematic, n otherwise) may be
// NULL: the session's Philox stream is used. h_idx (n entries) may be NULL.
int32_t fc_smc_resample_particles(fc_smc_session* s, int32_t method, const double* h_uniforms, uint64_t* h_idx) {
  using namespace fc;
  if (!s) { set_error("session is NULL"); return FC_ERR_INVALID_ARGUMENT; }
  if (method < FC_RESAMPLE_MULTINOMIAL || method > FC_RESAMPLE_STRATIFIED) { set_error("bad resampling method %d", method); return FC_ERR_INVALID_ARGUMENT; }
  fc_program* P = s->P;
  FC_TRY(fc_smc_normalize(s, nullptr, nullptr));
  cudaStream_t st = P->stream;
  const int64_t n = s->n;
  unsigned nbk = (unsigned)cdiv(n, 256);
  int64_t nu = method == FC_RESAMPLE_SYSTEMATIC ? 1 : n;
  FC_CUDA(s->us.reserve(nu + 1));
  s->steps += 1;
  if (h_uniforms) FC_CUDA(cudaMemcpyAsync(s->us.p, h_uniforms, nu * 8, cudaMemcpyHostToDevice, st));
  else k_uniform_one<<<(unsigned)cdiv(cdiv(nu, 2), 256), 256, 0, st>>>(s->us.p, s->seed, 0x1000000000000000ULL | s->steps, nu);
  FC_TRY(resample_dev(method, s->w.p, n, s->us.p, s->idx.p, *s->rw, st, &P->n_kernel_launches));
  dim3 gg(nbk, 1);
  if (P->d_cont) { gg.y = (unsigned)std::min<int64_t>(cdiv(P->d_cont, 8), 64); k_gather<double><<<gg, 256, 0, st>>>(s->q.p, s->q2.p, s->idx.p, P->d_cont, n); std::swap(s->q.p, s->q2.p); std::swap(s->q.n, s->q2.n); }
  if (P->d_disc) { gg.y = (unsigned)std::min<int64_t>(cdiv(P->d_disc, 8), 64); k_gather<disc_t><<<gg, 256, 0, st>>>(s->qd.p, s->qd2.p, s->idx.p, P->d_disc, n); std::swap(s->qd.p, s->qd2.p); std::swap(s->qd.n, s->qd2.n); }
  gg.y = 1;
  k_gather<double><<<gg, 256, 0, st>>>(s->lp3.p, s->lp3b.p, s->idx.p, 3, n);
  std::swap(s->lp3.p, s->lp3b.p); std::swap(s->lp3.n, s->lp3b.n);
  k_loglik<<<nbk, 256, 0, st>>>(s->lp3.p, s->ll.p, n);
  FC_KERNEL_CHECK();
  P->n_kernel_launches += 5;
  s->log_w_uniform = true;
  s->shard_prepass_ok = false;
  if (h_idx) FC_CUDA(cudaMemcpyAsync(h_idx, s->idx.p, n * 8, cudaMemcpyDeviceToHost, st));
  FC_CUDA(cudaStreamSynchronize(st));
  return FC_OK;
}

// rejuvenate_particles (smc.rs:698-713): `steps` pi_beta-invariant single-site MH moves per particle,
// weights untouched (FG-13). n_accepted (may be NULL) = accepted moves summed over particles.
int32_t fc_smc_rejuvenate_particles(fc_smc_session* s, double beta, uint64_t steps) {
  using namespace fc;
  if (!s) { set_error("session is NULL"); return FC_ERR_INVALID_ARGUMENT; }
  if (!(beta >= 0.0 && beta <= 1.0)) { set_error("beta must be in [0, 1], got %g", beta); return FC_ERR_INVALID_ARGUMENT; }
  fc_program* P = s->P;
  FC_CUDA(cudaSetDevice(P->device));
  for (uint64_t r = 0; r < steps; r++) {
    FC_TRY(s->eng.step(s->q.p, s->qd.p, s->lp3.p, s->total.p, s->n_accept.p, s->moves, beta, true, 0.44, 0.7));
    FC_TRY(s->eng.shared_adapt(0.44, 0.7));
    s->moves += 1;
  }
  k_loglik<<<(unsigned)cdiv(s->n, 256), 256, 0, P->stream>>>(s->lp3.p, s->ll.p, s->n);
  FC_KERNEL_CHECK();
  P->n_kernel_launches += 1;
  s->shard_prepass_ok = false;
  FC_CUDA(cudaStreamSynchronize(P->stream));
  return FC_OK;
}

// ---- one shard of a multi-GPU tempered SMC. The per-rank device steps; the few
// scalars between them travel through torch.distributed / NCCL in fugue_b200/dist.py. ----

// this session holds particles [first, first + n) of a population of n_total: uniform weight 1/n_total
int32_t fc_smc_shard_begin(fc_smc_session* s, uint64_t n_total) {
  using namespace fc;
  if (!s || n_total < (uint64_t)s->n) { set_error("bad shard arguments"); return FC_ERR_INVALID_ARGUMENT; }
  s->n_global = (double)n_total;
  return FC_OK;
}

// (max, sum exp(lv - max), sum exp(2 (lv - max))) of lv_i = log_w_i + d_beta * ll_i over this shard:
// the three numbers ess_at(beta + d_beta) and the reweight log-sum-exp need (smc.rs:589-602,
// 512-518); shards merge them exactly like the per-CTA partials of the single-GPU kernel.
int32_t fc_smc_shard_tri(fc_smc_session* s, double d_beta, double* out3) {
  using namespace fc;
  if (!s || !out3) { set_error("NULL argument"); return FC_ERR_INVALID_ARGUMENT; }
  FC_CUDA(cudaSetDevice(s->P->device));
  cudaStream_t st = s->P->stream;
  const int64_t n = s->n;
  const double lwc = -std::log(s->n_global > 0.0 ? s->n_global : (double)n);
  const double* lw = s->log_w_uniform ? nullptr : s->log_w.p;
  const int nb = nb_blocks(n);
  FC_CUDA(cudaMemsetAsync(&s->bis.p->counter, 0, sizeof(unsigned), st));
  if (!lw) {
    // uniform log_w (the state every next_beta search runs in): single-stream pass, the shift is
    // lwc + d_beta * max(ll) from a pre-pass that is redone only when ll has changed
    if (!s->shard_prepass_ok) {
      Bisect hb;
      std::memset(&hb, 0, sizeof(hb));
      hb.n = s->n_global > 0.0 ? s->n_global : (double)n;
      FC_CUDA(cudaMemcpyAsync(s->bis.p, &hb, sizeof(hb), cudaMemcpyHostToDevice, st));
      k_nb_prepass<<<nb, NB_THREADS, 0, st>>>(s->ll.p, n, s->bis.p, (double*)s->tri.p);
      FC_KERNEL_CHECK();
      FC_CUDA(cudaMemcpyAsync(&hb, s->bis.p, sizeof(hb), cudaMemcpyDeviceToHost, st));
      FC_CUDA(cudaStreamSynchronize(st));
      s->shard_ll_bad = hb.bad != 0;
      s->shard_prepass_ok = true;
      s->P->n_kernel_launches += 1;
    }
    if (s->shard_ll_bad) {  // some ll is NaN / +inf: every log-sum-exp is non-finite (smc.rs:595-601)
      out3[0] = out3[1] = out3[2] = __builtin_nan("");
      return FC_OK;
    }
    k_nb_eval_uni<<<nb, NB_THREADS, 0, st>>>(lwc, s->ll.p, s->bis.p, s->tri.p, n, 1, 0, d_beta);
  } else {
    k_nb_eval<<<nb, NB_THREADS, 0, st>>>(lw, lwc, s->ll.p, s->bis.p, s->tri.p, n, 1, 0, d_beta);
  }
  FC_KERNEL_CHECK();
  s->P->n_kernel_launches += 1;
  Tri t;
  FC_CUDA(cudaMemcpyAsync(&t, s->tri.p + nb, sizeof(Tri), cudaMemcpyDeviceToHost, st));
  FC_CUDA(cudaStreamSynchronize(st));
  out3[0] = t.m; out3[1] = t.s1; out3[2] = t.s2;
  return FC_OK;
}

// the same for up to 8 increments at once (uniform log_w only: the state every next_beta search runs
// in): out[3k..3k+3) = (max, sum e, sum e^2) of candidate k. Returns FC_ERR_UNSUPPORTED_OP if log_w is
// not uniform (the caller falls back to fc_smc_shard_tri).
int32_t fc_smc_shard_tri_multi(fc_smc_session* s, const double* d_betas, int32_t K, double* out) {
  using namespace fc;
  if (!s || !d_betas || !out || K < 1 || K > NB_MULTI_MAX) { set_error("bad arguments"); return FC_ERR_INVALID_ARGUMENT; }
  if (!s->log_w_uniform) { set_error("log weights are not uniform"); return FC_ERR_UNSUPPORTED_OP; }
  FC_CUDA(cudaSetDevice(s->P->device));
  cudaStream_t st = s->P->stream;
  const int64_t n = s->n;
  const double lwc = -std::log(s->n_global > 0.0 ? s->n_global : (double)n);
  const int nb = nb_blocks(n);
  if (!s->shard_prepass_ok) {  // max(ll) and the bad flag: fc_smc_shard_tri's pre-pass
    double tmp[3];
    FC_TRY(fc_smc_shard_tri(s, 0.0, tmp));
  }
  if (s->shard_ll_bad) {
    for (int k = 0; k < 3 * K; k++) out[k] = __builtin_nan("");
    return FC_OK;
  }
  MultiArgs A;
  A.K = K;
  for (int k = 0; k < NB_MULTI_MAX; k++) A.d[k] = k < K ? d_betas[k] : 0.0;
  FC_CUDA(s->scratch.reserve((int64_t)RDB * 2 * NB_MULTI_MAX + 16));
  k_nb_multi<<<nb, NB_THREADS, 0, st>>>(lwc, s->ll.p, s->bis.p, n, A, s->scratch.p);
  FC_KERNEL_CHECK();
  s->P->n_kernel_launches += 1;
  std::vector<double> h((size_t)nb * 2 * NB_MULTI_MAX);
  Bisect hb;
  FC_CUDA(cudaMemcpyAsync(h.data(), s->scratch.p, h.size() * 8, cudaMemcpyDeviceToHost, st));
  FC_CUDA(cudaMemcpyAsync(&hb, s->bis.p, sizeof(hb), cudaMemcpyDeviceToHost, st));
  FC_CUDA(cudaStreamSynchronize(st));
  for (int k = 0; k < K; k++) {
    double a = 0.0, b = 0.0;
    for (int blk = 0; blk < nb; blk++) { a += h[(size_t)blk * 2 * NB_MULTI_MAX + 2 * k]; b += h[(size_t)blk * 2 * NB_MULTI_MAX + 2 * k + 1]; }
    out[3 * k] = std::fma(d_betas[k], hb.max_ll, lwc);
    out[3 * k + 1] = a;
    out[3 * k + 2] = b;
  }
  return FC_OK;
}

// reweight with the GLOBAL normaliser (smc.rs:512-531): log_w = (log_w + d_beta ll) - log_norm,
// w = exp(log_w); advances beta / evidence / step count of the shard.
int32_t fc_smc_shard_reweight(fc_smc_session* s, double beta_new, double log_norm) {
  using namespace fc;
  if (!s) { set_error("session is NULL"); return FC_ERR_INVALID_ARGUMENT; }
  FC_CUDA(cudaSetDevice(s->P->device));
  cudaStream_t st = s->P->stream;
  const int64_t n = s->n;
  const double n_unif = s->n_global > 0.0 ? s->n_global : (double)n;
  const double lwc = -std::log(n_unif);
  const double* lw = s->log_w_uniform ? nullptr : s->log_w.p;
  const double d_beta = beta_new - s->beta;
  FC_CUDA(cudaMemcpyAsync(&s->bis.p->lse1, &log_norm, 8, cudaMemcpyHostToDevice, st));
  FC_CUDA(cudaStreamSynchronize(st));
  k_reweight<<<(unsigned)cdiv(n, 256), 256, 0, st>>>(lw, lwc, s->ll.p, d_beta, s->bis.p, s->comb.p, s->w.p, n, n_unif);
  FC_KERNEL_CHECK();
  s->P->n_kernel_launches += 1;
  std::swap(s->log_w.p, s->comb.p); std::swap(s->log_w.n, s->comb.n);
  s->log_w_uniform = false;
  s->log_evidence += log_norm;
  s->beta = beta_new;
  s->steps += 1;
  if (s->beta >= 1.0) s->finished = true;
  return FC_OK;
}

// exact running sum of this shard's weights from the carry-in of the ranks before it
int32_t fc_smc_shard_scan(fc_smc_session* s, double carry_in, double* end_value) {
  using namespace fc;
  if (!s || !end_value) { set_error("NULL argument"); return FC_ERR_INVALID_ARGUMENT; }
  FC_CUDA(cudaSetDevice(s->P->device));
  return shard_scan(s->w.p, s->n, carry_in, *s->rw, s->P->stream, &s->P->n_kernel_launches, end_value);
}

// the offspring of global output slots [jlo, jhi) (ancestors on this shard, after fc_smc_shard_scan),
// packed plane-major for migration: d_f64 [(d_cont + 3)][cnt] (values, then the three accumulators),
// d_i32 [d_disc][cnt]. With d_f64 == d_i32 == NULL the offspring stay on this rank and are written
// straight to local positions [local_offset, local_offset + cnt) of the new population.
// h_idx (cnt local ancestor indices) may be NULL.
int32_t fc_smc_shard_offspring(fc_smc_session* s, uint64_t n_total, double u_raw, uint64_t jlo, uint64_t jhi, double* d_f64, int32_t* d_i32, uint64_t* h_idx, uint64_t local_offset) {
  using namespace fc;
  if (!s || jhi < jlo) { set_error("bad arguments"); return FC_ERR_INVALID_ARGUMENT; }
  fc_program* P = s->P;
  FC_CUDA(cudaSetDevice(P->device));
  cudaStream_t st = P->stream;
  const int64_t n = s->n, cnt = (int64_t)(jhi - jlo);
  if (cnt == 0) return FC_OK;
  if (cnt > n) FC_CUDA(s->idx.reserve(cnt));
  FC_TRY(shard_ancestors(*s->rw, n, (double)n_total, u_raw, (int64_t)jlo, cnt, s->idx.p, st));
  dim3 gg((unsigned)cdiv(cnt, 256), 1);
  if (!d_f64 && !d_i32) {
    // offspring that stay on this rank: straight into the double buffers at their final position
    // (global slot jlo is local position jlo - first slot of this rank = local_offset)
    const int64_t off = (int64_t)local_offset;
    if (off < 0 || off + cnt > n) { set_error("local offspring out of range"); return FC_ERR_INVALID_ARGUMENT; }
    if (P->d_cont) { gg.y = (unsigned)std::min<int64_t>(P->d_cont, 64); k_gather<double><<<gg, 256, 0, st>>>(s->q.p, s->q2.p + off, s->idx.p, P->d_cont, n, cnt, n); }
    gg.y = 3; k_gather<double><<<gg, 256, 0, st>>>(s->lp3.p, s->lp3b.p + off, s->idx.p, 3, n, cnt, n);
    if (P->d_disc) { gg.y = (unsigned)std::min<int64_t>(P->d_disc, 64); k_gather<disc_t><<<gg, 256, 0, st>>>(s->qd.p, s->qd2.p + off, s->idx.p, P->d_disc, n, cnt, n); }
  } else {
    if (P->d_cont && d_f64) { gg.y = (unsigned)std::min<int64_t>(P->d_cont, 64); k_gather<double><<<gg, 256, 0, st>>>(s->q.p, d_f64, s->idx.p, P->d_cont, n, cnt, cnt); }
    if (d_f64) { gg.y = 3; k_gather<double><<<gg, 256, 0, st>>>(s->lp3.p, d_f64 + P->d_cont * cnt, s->idx.p, 3, n, cnt, cnt); }
    if (P->d_disc && d_i32) { gg.y = (unsigned)std::min<int64_t>(P->d_disc, 64); k_gather<disc_t><<<gg, 256, 0, st>>>(s->qd.p, d_i32, s->idx.p, P->d_disc, n, cnt, cnt); }
  }
  FC_KERNEL_CHECK();
  P->n_kernel_launches += 4;
  if (h_idx) FC_CUDA(cudaMemcpyAsync(h_idx, s->idx.p, cnt * 8, cudaMemcpyDeviceToHost, st));
  FC_CUDA(cudaStreamSynchronize(st));
  return FC_OK;
}

// install `cnt` migrated particles (one source rank's chunk, packed as fc_smc_shard_offspring packs
// them) at local positions [offset, offset + cnt). After the last chunk call fc_smc_shard_resampled.
int32_t fc_smc_shard_install(fc_smc_session* s, uint64_t offset, uint64_t cnt, const double* d_f64, const int32_t* d_i32) {
  using namespace fc;
  if (!s || offset + cnt > (uint64_t)s->n) { set_error("bad arguments"); return FC_ERR_INVALID_ARGUMENT; }
  fc_program* P = s->P;
  FC_CUDA(cudaSetDevice(P->device));
  cudaStream_t st = P->stream;
  const int64_t n = s->n;
  if (cnt == 0) return FC_OK;
  // the new population is assembled in the double buffers and swapped in by fc_smc_shard_resampled
  if (P->d_cont) FC_CUDA(cudaMemcpy2DAsync(s->q2.p + offset, n * 8, d_f64, cnt * 8, cnt * 8, P->d_cont, cudaMemcpyDeviceToDevice, st));
  FC_CUDA(cudaMemcpy2DAsync(s->lp3b.p + offset, n * 8, d_f64 + P->d_cont * cnt, cnt * 8, cnt * 8, 3, cudaMemcpyDeviceToDevice, st));
  if (P->d_disc) FC_CUDA(cudaMemcpy2DAsync(s->qd2.p + offset, n * 4, d_i32, cnt * 4, cnt * 4, P->d_disc, cudaMemcpyDeviceToDevice, st));
  return FC_OK;
}

// this shard's log weights as they are (normalised by the GLOBAL log-sum-exp after a reweight) and
// the weights exp(log_w) the resampler sees (valid after fc_smc_shard_reweight) -> host; either may be NULL
int32_t fc_smc_shard_weights(fc_smc_session* s, double* h_log_w, double* h_w) {
  using namespace fc;
  if (!s) { set_error("session is NULL"); return FC_ERR_INVALID_ARGUMENT; }
  FC_CUDA(cudaSetDevice(s->P->device));
  if (h_log_w) {
    FC_TRY(smc_materialise_log_w(s));
    FC_CUDA(cudaMemcpyAsync(h_log_w, s->log_w.p, s->n * 8, cudaMemcpyDeviceToHost, s->P->stream));
  }
  if (h_w) FC_CUDA(cudaMemcpyAsync(h_w, s->w.p, s->n * 8, cudaMemcpyDeviceToHost, s->P->stream));
  FC_CUDA(cudaStreamSynchronize(s->P->stream));
  return FC_OK;
}

// the migrated population replaces the old one; weights are uniform 1/n_total again (smc.rs:538-541)
int32_t fc_smc_shard_resampled(fc_smc_session* s) {
  using namespace fc;
  if (!s) { set_error("session is NULL"); return FC_ERR_INVALID_ARGUMENT; }
  fc_program* P = s->P;
  FC_CUDA(cudaSetDevice(P->device));
  cudaStream_t st = P->stream;
  std::swap(s->q.p, s->q2.p); std::swap(s->q.n, s->q2.n);
  std::swap(s->qd.p, s->qd2.p); std::swap(s->qd.n, s->qd2.n);
  std::swap(s->lp3.p, s->lp3b.p); std::swap(s->lp3.n, s->lp3b.n);
  k_loglik<<<(unsigned)cdiv(s->n, 256), 256, 0, st>>>(s->lp3.p, s->ll.p, s->n);
  FC_KERNEL_CHECK();
  P->n_kernel_launches += 1;
  s->log_w_uniform = true;
  s->shard_prepass_ok = false;
  FC_CUDA(cudaStreamSynchronize(st));
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
