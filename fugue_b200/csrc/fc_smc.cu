// fc_smc.cu -- adaptive (likelihood-tempered) SMC with the population resident in HBM, on one B200
// or sharded rank-major over the GPUs of one box.
//
// Device restatement of src/inference/smc.rs: adaptive_smc (:455-581), next_beta (:588-622),
// tempered_single_site_mh (:631-688), resample + clone (:326-349, 534-541), normalize_particles
// (:719-753) and smc_prior_particles (:764-790).
//
// Population container. One contiguous RECORD per particle: its f64 sites (sorted address order), the
// three Trace accumulators, the sufficient statistics of a mixture plate (when the program has one), then
// its discrete sites in program order (one BIT per site when every discrete site is two-valued and the
// statistics are carried, else one byte when every support fits, else four), padded to 16 bytes -- a particle
// is cloned as a whole by resampling (smc.rs:326-349) and read as a whole by a rejuvenation move, so both are
// streams of 16-byte vector accesses over consecutive records (C4: 128 B per particle). Only the
// log-likelihood lives in a dense array of its own (the tempering passes stream over it). The cached
// accumulators travel with the particle, so a move scores the proposal only (the reference re-scores the
// current state as well: 2 model runs per move, smc.rs:659-676).
//
// Kernels. For programs made of scalar ops plus two-valued discrete plates and Normal observation
// plates (BASELINE configs[3], examples/mixture_models.rs:77-111) the three per-particle steps are
// ONE kernel each: prior draw + statistics + score (k_smcx_prior), propose + score + accept
// (k_smcx_move; with the statistics the proposal is scored in O(1) from a by-value flattened program),
// ancestor search + state copy (k_smcx_gather). Other programs run the generic scoring engine
// (fc_score.cu) on the same container.
//
// Sharding. One process per GPU; the population buffers and a small mailbox live in one
// cudaIpc-exported slab per rank. Every exchange step of the tempering loop happens ON THE DEVICE
// over peer memory (NVLink): the per-rank sums of a weight pass are stored into every peer's mailbox
// by the last CTA of that pass, which then folds all ranks' sums in rank order and advances the
// replicated bisection state -- the global log-sum-exp is the tail of the weight kernel, not a
// collective call; the exact running-sum carry travels rank to rank through the same mailboxes; and
// the resampling gather writes offspring straight into the destination rank's population buffers.
//
// Documented deviation: the reference threads ONE DiminishingAdaptation through all particles
// sequentially (smc.rs:482,551) -- a cross-particle serial dependency. Here every particle of a
// sweep uses the same per-site scale and the scale is updated once per sweep from the sweep's
// aggregate acceptance rate with gain 2 * sweep^-gamma (per rank when sharded). Any positive scale
// leaves pi_beta invariant, so only efficiency is affected, not the target.
#include <unistd.h>
#include <algorithm>
#include <cmath>
#include <cstdlib>
#include "fc_mh.cuh"
#include "fc_prior.cuh"
#include "fc_scalar.cuh"
#include "fc_fastmath.cuh"

namespace fc {
int32_t dev_reduce(int op, const double* d_x, int64_t n, double m, double* d_scratch, double* d_out, cudaStream_t st);
int32_t dev_log_sum_exp(const double* d_x, int64_t n, double* d_scratch, double* h_out, cudaStream_t st);
struct ResampleWork;
ResampleWork* resample_work_new();
void resample_work_free(ResampleWork*);
int32_t resample_dev(int method, const double* d_w, int64_t n, const double* d_us, uint64_t* d_idx, ResampleWork& W, cudaStream_t st, uint64_t* launches);
int32_t scan_enqueue_head(const double* d_w, int64_t n, double* d_S, void* d_work, cudaStream_t st, uint64_t* launches);
int32_t scan_enqueue_rounds(const double* d_w, int64_t n, double* d_S, void* d_work, int rounds, cudaStream_t st, uint64_t* launches);
int scan_expected_rounds(int64_t n);
int32_t scan_finish(const double* d_w, int64_t n, double* d_S, void* d_work, cudaStream_t st, uint64_t* launches, double* h_end);
size_t scan_work_bytes(int64_t n);

#define RDB 1024  // must match RD_BLOCKS of fc_resample.cu

// =============================================================================================
// peer-memory exchange
// =============================================================================================
#define MB_MAXW 8   // ranks (GPUs of one box)
#define MB_VALS 8   // doubles per rank and exchange
struct MBox {
  unsigned long long flag[2][MB_MAXW];      // [sequence parity][source rank] = sequence number
  double val[2][MB_MAXW][MB_VALS];
  unsigned long long end_flag[2][MB_MAXW];  // exact running-sum end value of rank r for resample #k
  double end_val[2][MB_MAXW];
};
struct Comm {
  MBox* box[MB_MAXW];        // box[r]: rank r's mailbox as mapped into this process (box[rank]: local)
  int rank, world;
  unsigned long long* seq;   // local device counter of completed exchanges (identical on every rank)
  int* err;                  // local: set to 1 when a peer did not answer in time
};
__device__ __forceinline__ unsigned long long ld_acquire_sys(const unsigned long long* p) {
  unsigned long long v;
  asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_release_sys(unsigned long long* p, unsigned long long v) {
  asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ unsigned long long global_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}
#define COMM_TIMEOUT_NS 20000000000ULL
__device__ __forceinline__ bool wait_flag(const unsigned long long* p, unsigned long long want, int* err) {
  const unsigned long long t0 = global_ns();
  while (ld_acquire_sys(p) < want) {
    if (global_ns() - t0 > COMM_TIMEOUT_NS) { *err = 1; return false; }
  }
  return true;
}
// All-gather of nv (<= MB_VALS) doubles per rank, executed by one whole CTA (>= 32 threads):
// s_mine[0..nv) in shared memory -> s_all[r][0..nv) for every rank r. Two mailbox slots alternate by
// sequence parity: a rank cannot finish exchange s+1 before every peer has contributed to s+1, which
// a peer does only after it has read exchange s, so slot s&1 is never overwritten while in use.
__device__ __forceinline__ void comm_allgather(const Comm& cm, const double* s_mine, int nv, double (*s_all)[MB_VALS]) {
  __shared__ unsigned long long s_seq;
  if (threadIdx.x == 0) { s_seq = *cm.seq + 1; *cm.seq = s_seq; }
  __syncthreads();
  const unsigned long long seq = s_seq;
  const int par = (int)(seq & 1), t = threadIdx.x;
  if (t < cm.world) {
    MBox* d = cm.box[t];
    for (int i = 0; i < nv; i++) *(volatile double*)&d->val[par][cm.rank][i] = s_mine[i];
    __threadfence_system();
    st_release_sys(&d->flag[par][cm.rank], seq);
    MBox* me = cm.box[cm.rank];
    wait_flag(&me->flag[par][t], seq, cm.err);
    for (int i = 0; i < nv; i++) s_all[t][i] = *(volatile double*)&me->val[par][t][i];
  }
  __syncthreads();
}
__global__ void k_comm_barrier(const __grid_constant__ Comm cm) {
  __shared__ double s_mine[MB_VALS], s_all[MB_MAXW][MB_VALS];
  if (threadIdx.x == 0) s_mine[0] = 0.0;
  __syncthreads();
  comm_allgather(cm, s_mine, 1, s_all);
}

// =============================================================================================
// small elementwise kernels
// =============================================================================================
__global__ void k_loglik(const double* __restrict__ lp3, double* __restrict__ ll, int64_t P) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < P) ll[i] = lp3[P + i] + lp3[2 * P + i];  // particle_log_likelihood (smc.rs:381-383)
}
__global__ void k_fill2(double* p, double v, int64_t n) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) p[i] = v;
}
// combined = log_w + d_beta * ll   (smc.rs:512-517)
__global__ void k_combine(const double* __restrict__ log_w, double lw_const, const double* __restrict__ ll, double d_beta, double* __restrict__ out, int64_t P) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < P) out[i] = (log_w ? log_w[i] : lw_const) + d_beta * ll[i];
}
// out_log = x - sub, out_exp = exp(x - sub); a non-finite normaliser gives the uniform fallback
// (smc.rs:565-575, 719-753)
__global__ void k_sub_exp(const double* __restrict__ x, const double* __restrict__ d_sub, double n_unif, double* __restrict__ out_log, double* __restrict__ out_exp, int64_t P) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= P) return;
  const double sub = *d_sub;
  if (!isfinite(sub)) {
    if (out_log) out_log[i] = -log(n_unif);
    if (out_exp) out_exp[i] = 1.0 / n_unif;
    return;
  }
  double v = x[i] - sub;
  if (out_log) out_log[i] = v;
  if (out_exp) out_exp[i] = exp(v);
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

// =============================================================================================
// next_beta (smc.rs:588-622) and the reweight log-sum-exp (smc.rs:512-524)
// =============================================================================================
// ess_at(b) = exp(2 lse(lv) - lse(2 lv)), lv_i = log_w_i + (b - beta) ll_i. One launch per evaluation:
// a single streaming pass keeps (sum e, sum e^2) with e = exp(lv - max) -- sum exp(2 lv - 2 max) is
// sum e^2, so one exp per particle serves both log-sum-exps -- the last CTA folds the per-CTA sums in
// block order, exchanges the rank's sums with the peers, folds those in rank order and advances the
// search state, which is therefore replicated bit for bit on every rank. The host only enqueues.
//
// The search. The reference bisects [beta, 1] 64 times, one ess_at per iteration. ess_at is
// non-increasing in b for uniform log_w (d log ESS / d b = 2 (E_b[ll] - E_2b[ll]) <= 0, the tilted mean
// being non-decreasing), so an iteration whose midpoint lies at or beyond a point where ess_at has
// ALREADY been evaluated needs no evaluation: its branch is implied. The kernel therefore first
// localises the crossing with a safeguarded Newton iteration on log ESS in log(b - beta) (the
// derivative sums come from the same pass), certifies a bracket [a, b] a few ulps wide by actual
// evaluations on both sides, and then replays the reference loop -- same midpoints 0.5 * (lo + hi),
// same comparisons -- evaluating ess_at only for midpoints strictly inside (a, b). The returned beta is
// the reference loop's. By default the search ends as soon as Newton has converged (relative step
// <= 1e-9, so the crossing is known to the rounding noise of the sums) and ALL remaining midpoints are
// decided by their side of the crossing: ~7 passes instead of ~57. fc_debug_set("nb_plain", 1) evaluates
// every midpoint as the reference does; ("nb_plain", 2) certifies a bracket by evaluations on both sides
// and evaluates the midpoints strictly inside it (~20 passes).
struct Bisect {
  double beta, target, n;   // n: global population size (the ESS reported when an lse is not finite)
  double lo, hi;            // the reference loop's interval (smc.rs:611-619)
  double a, b;              // certified: ess_at(a) >= target (a == beta: as the reference assumes), ess_at(b) < target
  double mid;               // the point the next launch evaluates
  double lse1;              // lse(lv) of the last evaluation (the reweight normaliser when mid == beta_new)
  double max_ll;            // global max of ll (uniform log_w: max lv = lwc + d max ll)
  double guess;             // first search point (0: none): a multiple of the previous step's increment
  double prev_x, delta;     // search bookkeeping
  int bad;                  // some ll is NaN or +inf => every lse is non-finite => ESS := n (smc.rs:595-601)
  int done;                 // 1: ess_at(1) >= target; 2: the loop has reached its fixed point / 64 iterations
  int phase;                // 0: ess_at(1.0) probe, 1: Newton search, 2: two-sided polish, 3: replay
  int it;                   // iterations of the reference loop performed
  int n_search, plain, evals;
  unsigned counter;
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
// the reference's bisection update for one ESS evaluation at B->mid (smc.rs:605-619)
__device__ __forceinline__ void nb_ref_update(Bisect* B, double ess) {
  if (ess < B->target) { B->hi = B->mid; B->b = fmin(B->b, B->mid); }
  else { B->lo = B->mid; B->a = fmax(B->a, B->mid); }
  B->it += 1;
}
// walk the reference loop as far as the certified bracket decides it; stops at the first midpoint that
// needs an evaluation (left in B->mid) or at the loop's end
__device__ __forceinline__ void nb_replay(Bisect* B) {
  while (B->it < 64) {
    const double mid = 0.5 * (B->lo + B->hi);
    if (mid == B->lo || mid == B->hi) { B->done = 2; return; }  // fixed point of the remaining iterations
    if (mid <= B->a) { B->lo = mid; B->it += 1; continue; }     // ess_at(mid) >= ess_at(a) >= target
    if (mid >= B->b) { B->hi = mid; B->it += 1; continue; }     // ess_at(mid) <= ess_at(b) <  target
    B->mid = mid;
    return;
  }
  B->done = 2;
}
// state machine step after ess_at(B->mid) has been evaluated: s1 = sum e, s2 = sum e^2, t1 = sum ll e,
// t2 = sum ll e^2 over the whole population (shift m), ess as the reference computes it
__device__ void nb_advance(Bisect* B, double ess, double s1, double s2, double t1, double t2) {
  B->evals += 1;
  const double x = B->mid;
  if (B->phase == 0) {  // the ess_at(1.0) probe (smc.rs:605-608)
    if (ess >= B->target) { B->hi = 1.0; B->done = 1; return; }
    B->lo = B->beta; B->hi = 1.0; B->a = B->beta; B->b = 1.0; B->it = 0;
    if (B->plain == 1) { B->phase = 3; nb_replay(B); return; }
    B->phase = 1; B->n_search = 0;
    if (B->guess > B->beta && B->guess < 1.0) { B->prev_x = x; B->mid = B->guess; return; }
  } else if (B->phase == 3) {
    nb_ref_update(B, ess);
    nb_replay(B);
    return;
  } else {  // search phases: certify, then choose the next point
    if (ess < B->target) B->b = fmin(B->b, x); else B->a = fmax(B->a, x);
  }
  if (B->phase == 1) {
    B->n_search += 1;
    const double d = x - B->beta, da = B->a - B->beta, db = B->b - B->beta;
    // Newton on g(d) = ln ESS(d) - ln target in t = ln d: dg/dd = 2 (t1/s1 - t2/s2)
    const double g = log(ess) - log(B->target);
    const double gp = 2.0 * (t1 / s1 - t2 / s2);
    double dn = d * exp(-g / (d * gp));
    if (!(gp < 0.0) || !isfinite(dn) || !(dn > da) || !(dn < db)) dn = da > 0.0 ? sqrt(da * db) : db * 0.0625;
    const double xn = B->beta + dn;
    if (B->plain == 0 && fabs(dn - d) <= 1e-9 * d && xn > B->a && xn < B->b) {
      // Newton has converged quadratically: xn is the crossing to within the rounding noise of the sums
      // themselves (the next correction would be ~1e-18 relative). Every remaining midpoint of the reference
      // loop is decided by which side of xn it lies on: for midpoints farther than the noise that is
      // monotonicity, for nearer ones it is the first-order model of ESS around xn, whose error (second
      // order, ~1e-26 relative) is far below the ~1e-13 evaluation noise that decides them otherwise.
      B->a = xn; B->b = xn; B->phase = 3;
      nb_replay(B);
      return;
    }
    const bool converged = fabs(dn - d) <= 1e-12 * d || B->n_search >= 24 || !(xn > B->a) || !(xn < B->b);
    if (!converged) { B->prev_x = x; B->mid = xn; return; }
    // two-sided polish around the best point: certify the side that is still far away
    B->phase = 2;
    B->delta = ldexp(fmax(x, 1e-300), -44);
  }
  if (B->phase == 2) {
    B->n_search += 1;
    // the evaluated points so far have moved a or b; step from the near side towards the far one
    const bool near_a = (x == B->a);
    const double far_gap = near_a ? (B->b - x) : (x - B->a);
    if (far_gap <= 8.0 * B->delta || B->n_search >= 40) { B->phase = 3; nb_replay(B); return; }
    double xn = near_a ? x + B->delta : x - B->delta;
    B->delta *= 4.0;
    if (!(xn > B->a) || !(xn < B->b)) { B->phase = 3; nb_replay(B); return; }
    B->mid = xn;
    return;
  }
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

// ---- general log_w (standalone next_beta on arbitrary weights; single rank): plain bisection ----
__device__ __forceinline__ void nb_fold_update_general(Tri acc, Bisect* __restrict__ B, Tri* __restrict__ part, int mode) {
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
  B->counter = 0;
  const double ninf = d_neg_inf();
  double lse1 = (t.m == ninf && t.s1 == 0.0) ? ninf : (t.s1 == 0.0 ? ninf : t.m + log(t.s1));
  double lse2 = (t.m == ninf && t.s2 == 0.0) ? ninf : (t.s2 == 0.0 ? ninf : 2.0 * t.m + log(t.s2));
  B->lse1 = lse1;
  if (mode != 0) return;
  double ess = (!isfinite(lse1) || !isfinite(lse2)) ? B->n : exp(2.0 * lse1 - lse2);
  nb_advance(B, ess, 1.0, 1.0, 0.0, 0.0);  // plain = 1: the derivative sums are not used
}
// mode 0: probe at B->mid; mode 1: lse only at d_fixed
__global__ void __launch_bounds__(NB_THREADS) k_nb_eval(const double* __restrict__ log_w, const double* __restrict__ ll, Bisect* __restrict__ B,
                                                        Tri* __restrict__ part, int64_t P, int mode, double d_fixed) {
  if (mode == 0 && B->done) return;
  const double d = mode == 0 ? B->mid - B->beta : d_fixed;
  Tri acc; acc.m = d_neg_inf(); acc.s1 = 0.0; acc.s2 = 0.0;
  const int64_t nthreads = (int64_t)gridDim.x * NB_THREADS;
  for (int64_t i0 = (int64_t)blockIdx.x * NB_THREADS + threadIdx.x; i0 < P; i0 += nthreads * 8) {
    double lv[8];
#pragma unroll
    for (int k = 0; k < 8; k++) {
      int64_t i = i0 + (int64_t)k * nthreads;
      lv[k] = i < P ? __ldg(log_w + i) + d * __ldg(ll + i) : d_neg_inf();
    }
    tri_push8(acc, lv);
  }
  nb_fold_update_general(acc, B, part, mode);
}

// ---- uniform log_w (the state adaptive_smc is in whenever next_beta runs: initially and after every
// resample): lv_i = c + d ll_i, so max lv = c + d max(ll) and the pass needs no running max.
// k_nb_prepass (once per next_beta call): GLOBAL max(ll) and the "some ll is NaN / +inf" flag. ----
__global__ void __launch_bounds__(NB_THREADS) k_nb_prepass(const double* __restrict__ ll, int64_t P, Bisect* __restrict__ B, double* __restrict__ part, const __grid_constant__ Comm cm) {
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
  if (!is_last) return;
  __threadfence();
  __shared__ double s_mine[MB_VALS], s_all[MB_MAXW][MB_VALS];
  if (threadIdx.x == 0) {
    double m = d_neg_inf();
    for (int i = 0; i < (int)gridDim.x; i++) m = fmax(m, __ldcg(part + i));
    s_mine[0] = m;
    B->counter = 0;
  }
  __syncthreads();
  comm_allgather(cm, s_mine, 1, s_all);
  if (threadIdx.x == 0) {
    double m = d_neg_inf();
    for (int r = 0; r < cm.world; r++) m = fmax(m, s_all[r][0]);
    B->bad = m == d_pos_inf();
    B->max_ll = m;
  }
}
// one pass over ll at increment d (mode 0: d = B->mid - B->beta, search state advanced; mode 1: lse only
// at d_fixed). part: per-CTA (s1, s2, t1, t2).
template <bool DERIV>
__device__ __forceinline__ void nb_pass(double lw_const, const double* __restrict__ ll, int64_t P, double d, double m, const double* __restrict__ s_et, double (&out)[4]) {
  double s1 = 0.0, s2 = 0.0, t1 = 0.0, t2 = 0.0, u1 = 0.0, u2 = 0.0, v1 = 0.0, v2 = 0.0;  // two accumulator sets (shorter dependency chains)
  const int64_t nthreads = (int64_t)gridDim.x * NB_THREADS;
  const int64_t P2 = P >> 1;
  // software pipeline: the next trip's four 16-byte loads are in flight while this trip's eight exp are
  // evaluated (ncu profiles/ncu_c4_nb_eval_r1_v12.txt: 60 % of the warp samples waited on the loads,
  // DRAM at 33 % -- every warp loaded, then every warp computed)
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
      s1 += e0; s2 = fma(e0, e0, s2);
      u1 += e1; u2 = fma(e1, e1, u2);
      if (DERIV) {  // ll = -inf contributes e ~ 3e-308: keep its product out of the derivative sums
        const double l0 = a[k].x > -1e300 ? a[k].x * e0 : 0.0, l1 = a[k].y > -1e300 ? a[k].y * e1 : 0.0;
        t1 += l0; t2 = fma(l0, e0, t2);
        v1 += l1; v2 = fma(l1, e1, v2);
      }
    }
  }
  if ((P & 1) && blockIdx.x == 0 && threadIdx.x == 0) {
    const double l = ll[P - 1];
    const double e0 = exp_nonpos(fma(d, l, lw_const) - m, s_et);
    s1 += e0; s2 = fma(e0, e0, s2);
    if (DERIV) { const double l0 = l > -1e300 ? l * e0 : 0.0; t1 += l0; t2 = fma(l0, e0, t2); }
  }
  out[0] = s1 + u1; out[1] = s2 + u2; out[2] = t1 + v1; out[3] = t2 + v2;
}
__global__ void __launch_bounds__(NB_THREADS) k_nb_eval_uni(double lw_const, const double* __restrict__ ll, Bisect* __restrict__ B, double* __restrict__ part, int64_t P,
                                                            int mode, double d_fixed, const __grid_constant__ Comm cm) {
  if (mode == 0 && B->done) return;
  __shared__ double s_et[32];
  fast_exp_tab_init(s_et, threadIdx.x);
  __syncthreads();
  if (B->bad) {  // every lse is non-finite: ess_at(.) = n (smc.rs:595-601); the reweight normaliser is NaN
    if (blockIdx.x == 0 && threadIdx.x == 0) {
      B->lse1 = __longlong_as_double(0x7ff8000000000000LL);
      if (mode == 0) nb_advance(B, B->n, 1.0, 1.0, 0.0, 0.0);  // phase 0: ess_at(1.0) = n >= target -> 1.0
    }
    return;
  }
  const double d = mode == 0 ? B->mid - B->beta : d_fixed;
  const double m = fma(d, B->max_ll, lw_const);  // the same expression as lv of the max element: x <= 0, == 0 there
  const bool deriv = mode == 0 && (B->phase == 1 || (B->phase == 0 && B->plain != 1));
  double acc[4] = {0.0, 0.0, 0.0, 0.0};
  if (m > d_neg_inf()) {
    if (deriv) nb_pass<true>(lw_const, ll, P, d, m, s_et, acc);
    else nb_pass<false>(lw_const, ll, P, d, m, s_et, acc);
  }
  // CTA reduction; the last CTA folds the per-CTA sums in block order (deterministic)
  __shared__ double sm[NB_THREADS / 32][4];
  __shared__ bool is_last;
#pragma unroll
  for (int k = 0; k < 4; k++) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) acc[k] += __shfl_down_sync(0xffffffffu, acc[k], o);
  }
  if ((threadIdx.x & 31) == 0) { sm[threadIdx.x >> 5][0] = acc[0]; sm[threadIdx.x >> 5][1] = acc[1]; sm[threadIdx.x >> 5][2] = acc[2]; sm[threadIdx.x >> 5][3] = acc[3]; }
  __syncthreads();
  if (threadIdx.x < 4) {
    double t = 0.0;
    for (int w = 0; w < NB_THREADS / 32; w++) t += sm[w][threadIdx.x];
    part[(int64_t)blockIdx.x * 4 + threadIdx.x] = t;
  }
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) is_last = atomicAdd(&B->counter, 1u) == gridDim.x - 1;
  __syncthreads();
  if (!is_last) return;
  __threadfence();
  __shared__ double s_mine[MB_VALS], s_all[MB_MAXW][MB_VALS];
  if (threadIdx.x < 4) {
    double t = 0.0;
    for (int i = 0; i < (int)gridDim.x; i++) t += __ldcg(part + (int64_t)i * 4 + threadIdx.x);
    s_mine[threadIdx.x] = t;
  }
  if (threadIdx.x == 0) B->counter = 0;
  __syncthreads();
  comm_allgather(cm, s_mine, 4, s_all);
  if (threadIdx.x != 0) return;
  double s1 = 0.0, s2 = 0.0, t1 = 0.0, t2 = 0.0;
  for (int r = 0; r < cm.world; r++) { s1 += s_all[r][0]; s2 += s_all[r][1]; t1 += s_all[r][2]; t2 += s_all[r][3]; }
  const double ninf = d_neg_inf();
  const double lse1 = (m == ninf || s1 == 0.0) ? ninf : m + log(s1);
  const double lse2 = (m == ninf || s2 == 0.0) ? ninf : 2.0 * m + log(s2);
  B->lse1 = lse1;
  if (mode != 0) return;
  const double ess = (!isfinite(lse1) || !isfinite(lse2)) ? B->n : exp(2.0 * lse1 - lse2);
  nb_advance(B, ess, s1, s2, t1, t2);
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

// ---- global log_sum_exp of an array over all ranks (numerical.rs:15-38): max pass, then sum pass; each
// ends with the mailbox exchange. out[0] = lse, out[1] = global max. ----
__global__ void __launch_bounds__(NB_THREADS) k_glse_max(const double* __restrict__ x, int64_t P, double* __restrict__ part, unsigned* __restrict__ counter, double* __restrict__ out, const __grid_constant__ Comm cm) {
  double mx = d_neg_inf();
  bool nan = false;
  for (int64_t i = (int64_t)blockIdx.x * NB_THREADS + threadIdx.x; i < P; i += (int64_t)gridDim.x * NB_THREADS) { double v = __ldg(x + i); nan |= v != v; mx = fmax(mx, v); }
  mx = warp_max(mx);
  nan = __any_sync(0xffffffffu, nan);
  __shared__ double sm[NB_THREADS / 32];
  __shared__ int snan[NB_THREADS / 32];
  __shared__ bool is_last;
  if ((threadIdx.x & 31) == 0) { sm[threadIdx.x >> 5] = mx; snan[threadIdx.x >> 5] = nan; }
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int w = 1; w < NB_THREADS / 32; w++) { mx = fmax(mx, sm[w]); nan |= snan[w] != 0; }
    part[blockIdx.x] = nan ? d_pos_inf() : mx;  // a NaN element makes the sum pass NaN; flag it through +inf
    __threadfence();
    is_last = atomicAdd(counter, 1u) == gridDim.x - 1;
  }
  __syncthreads();
  if (!is_last) return;
  __threadfence();
  __shared__ double s_mine[MB_VALS], s_all[MB_MAXW][MB_VALS];
  if (threadIdx.x == 0) {
    double m = d_neg_inf();
    for (int i = 0; i < (int)gridDim.x; i++) m = fmax(m, __ldcg(part + i));
    s_mine[0] = m;
    *counter = 0;
  }
  __syncthreads();
  comm_allgather(cm, s_mine, 1, s_all);
  if (threadIdx.x == 0) {
    double m = d_neg_inf();
    for (int r = 0; r < cm.world; r++) m = fmax(m, s_all[r][0]);
    out[1] = m;
  }
}
__global__ void __launch_bounds__(NB_THREADS) k_glse_sum(const double* __restrict__ x, int64_t P, double* __restrict__ part, unsigned* __restrict__ counter, double* __restrict__ out, const __grid_constant__ Comm cm) {
  const double m = out[1];
  double acc = 0.0;
  if (isfinite(m))
    for (int64_t i = (int64_t)blockIdx.x * NB_THREADS + threadIdx.x; i < P; i += (int64_t)gridDim.x * NB_THREADS) acc += exp(__ldg(x + i) - m);
  acc = warp_sum(acc);
  __shared__ double sm[NB_THREADS / 32];
  __shared__ bool is_last;
  if ((threadIdx.x & 31) == 0) sm[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int w = 1; w < NB_THREADS / 32; w++) acc += sm[w];
    part[blockIdx.x] = acc;
    __threadfence();
    is_last = atomicAdd(counter, 1u) == gridDim.x - 1;
  }
  __syncthreads();
  if (!is_last) return;
  __threadfence();
  __shared__ double s_mine[MB_VALS], s_all[MB_MAXW][MB_VALS];
  if (threadIdx.x == 0) {
    double t = 0.0;
    for (int i = 0; i < (int)gridDim.x; i++) t += __ldcg(part + i);
    s_mine[0] = t;
    *counter = 0;
  }
  __syncthreads();
  comm_allgather(cm, s_mine, 1, s_all);
  if (threadIdx.x == 0) {
    double t = 0.0;
    for (int r = 0; r < cm.world; r++) t += s_all[r][0];
    // numerical.rs:15-38: all -inf -> -inf; a non-finite max (NaN element / +inf) is returned as is
    out[0] = m == d_neg_inf() ? m : (!isfinite(m) ? __longlong_as_double(0x7ff8000000000000LL) : (t == 0.0 ? d_neg_inf() : m + log(t)));
  }
}
// sum of squares over all ranks (effective_sample_size, smc.rs:230-233): out[2] = 1 / sum w^2
__global__ void __launch_bounds__(NB_THREADS) k_gsumsq(const double* __restrict__ x, int64_t P, double* __restrict__ part, unsigned* __restrict__ counter, double* __restrict__ out, const __grid_constant__ Comm cm) {
  double acc = 0.0;
  for (int64_t i = (int64_t)blockIdx.x * NB_THREADS + threadIdx.x; i < P; i += (int64_t)gridDim.x * NB_THREADS) { double v = __ldg(x + i); acc += v * v; }
  acc = warp_sum(acc);
  __shared__ double sm[NB_THREADS / 32];
  __shared__ bool is_last;
  if ((threadIdx.x & 31) == 0) sm[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int w = 1; w < NB_THREADS / 32; w++) acc += sm[w];
    part[blockIdx.x] = acc;
    __threadfence();
    is_last = atomicAdd(counter, 1u) == gridDim.x - 1;
  }
  __syncthreads();
  if (!is_last) return;
  __threadfence();
  __shared__ double s_mine[MB_VALS], s_all[MB_MAXW][MB_VALS];
  if (threadIdx.x == 0) {
    double t = 0.0;
    for (int i = 0; i < (int)gridDim.x; i++) t += __ldcg(part + i);
    s_mine[0] = t;
    *counter = 0;
  }
  __syncthreads();
  comm_allgather(cm, s_mine, 1, s_all);
  if (threadIdx.x == 0) {
    double t = 0.0;
    for (int r = 0; r < cm.world; r++) t += s_all[r][0];
    out[2] = 1.0 / t;
  }
}

// =============================================================================================
// the exact running-sum carry from rank to rank (global systematic / stratified resampling)
// =============================================================================================
// rank r starts its exact scan (fc_resample.cu) from the end value of rank r-1 and publishes its own end
// value to every rank; afterwards every rank holds all end values (its slot routing needs them).
__global__ void k_chain_wait(const __grid_constant__ Comm cm, unsigned long long rs, ScanCtl* __restrict__ ctl) {
  double carry = 0.0;
  if (cm.rank > 0) {
    MBox* me = cm.box[cm.rank];
    const int par = (int)(rs & 1);
    wait_flag(&me->end_flag[par][cm.rank - 1], rs, cm.err);
    carry = *(volatile double*)&me->end_val[par][cm.rank - 1];
  }
  ctl->s0 = carry;
}
__global__ void k_chain_publish(const __grid_constant__ Comm cm, unsigned long long rs, const ScanCtl* __restrict__ ctl, int64_t n) {
  const int t = threadIdx.x, par = (int)(rs & 1);
  if (t >= cm.world || ctl->k0 < n) return;  // not complete yet: the host enqueues more rounds, then publishes
  MBox* d = cm.box[t];
  if (ld_acquire_sys(&cm.box[cm.rank]->end_flag[par][cm.rank]) >= rs) return;  // already published
  *(volatile double*)&d->end_val[par][cm.rank] = ctl->s0;
  __threadfence_system();
  st_release_sys(&d->end_flag[par][cm.rank], rs);
}
__global__ void k_chain_collect(const __grid_constant__ Comm cm, unsigned long long rs, double* __restrict__ ends) {
  const int t = threadIdx.x, par = (int)(rs & 1);
  if (t >= cm.world) return;
  MBox* me = cm.box[cm.rank];
  wait_flag(&me->end_flag[par][t], rs, cm.err);
  ends[t] = *(volatile double*)&me->end_val[par][t];
}

// =============================================================================================
// the population container and the fused per-particle kernels
// =============================================================================================
struct Pop {
  uint8_t* rec;   // [n][pitch]: { q[d_cont] | lp3[3] | mixture statistics[5] (when soff) | pad | z[d_disc * zw] | pad }
  double* ll;     // [n] log-likelihood + log-factors of the particle (particle_log_likelihood, smc.rs:381-383)
  int64_t n;
};
struct RecLayout {
  int64_t pitch;    // bytes per record (multiple of 16)
  int64_t zoff;     // byte offset of the discrete sites (multiple of 16)
  int64_t d_cont, d_disc;
  int32_t zw;       // bytes per discrete site: 1 or 4; 0 = ONE BIT per site (every discrete site two-valued)
  int32_t soff;     // byte offset of the mixture plate's sufficient statistics (5 doubles), 0 = none
  __host__ __device__ __forceinline__ double* q(uint8_t* r) const { return (double*)r; }
  __host__ __device__ __forceinline__ double* lp3(uint8_t* r) const { return (double*)r + d_cont; }
  __host__ __device__ __forceinline__ uint8_t* z(uint8_t* r) const { return r + zoff; }
};
struct ZAcc {  // discrete value of a program slot for one particle
  const uint8_t* rec;
  int zw;
  __device__ __forceinline__ long long operator()(int dslot) const {
    if (zw == 0) return (long long)((rec[dslot >> 3] >> (dslot & 7)) & 1u);
    return zw == 1 ? (long long)rec[dslot] : (long long)((const int32_t*)rec)[dslot];
  }
};

// plates the fused evaluator understands (lowered from the program's DOps at session create)
enum { FP_TWOVAL_SAMPLE = 0, FP_NORMAL_OBS = 1, FP_TWOVAL_OBS = 2, FP_MIX = 3 };
enum { FM_NONE = 0, FM_SELECT2 = 1, FM_LATENT_X = 2 };
struct FPlate {
  int32_t kind, n;
  int32_t op, op_b;     // index into all_ops (FP_MIX: op = the discrete sample plate, op_b = the Normal observe)
  int32_t m0, m1, ss;   // Normal observe: kinds of the (up to two) element-dependent mean terms; sigma is base + SELECT2
  int32_t mt0, mt1, st; // their term indices
  int32_t stat, _pad;   // FP_MIX: the record carries this plate's per-component sufficient statistics
};
// scalar ops compiled for the fused evaluator: priors with constant parameters of the three families the
// BASELINE models use become a few FP64 instructions; everything else goes through the interpreter
enum { SO_GENERIC = 0, SO_NORMAL_CC = 1, SO_GAMMA_CC = 2, SO_BETA_CC = 3, SO_FACTOR = 4 };
struct SOp {
  int32_t kind, row;     // row: the site's f64 row (sample ops)
  double c0, c1, c2;     // NORMAL: mu, 1/sigma, -ln sigma - 0.5 ln 2pi; GAMMA: k - 1, rate, k ln rate - lgamma k;
};                       // BETA: a - 1, b - 1, -ln B(a, b); FACTOR: the constant log-weight
struct FusedProg {
  const DOp* ops;           // all ops, program order
  const DOp* scalar_ops;
  const SOp* sops;          // one per scalar op, same order
  const FPlate* plates;
  const int32_t* crow;
  const fc_expr_node* nodes;
  int32_t n_ops, n_scalar, n_plates, _pad;
  RecLayout L;
};

// The commonest fused program -- scalar priors of the compiled families plus ONE mixture plate whose
// statistics travel in the record -- flattened into a by-value kernel argument: the move kernel then reads
// its whole "program" from the constant bank (no descriptor loads, no term-kind dispatch), and its
// arithmetic is the interpreter's, operation for operation (same results bit for bit).
struct CParam {            // link(sum of up to 3 chain-constant terms), as chain_const_param evaluates it
  double c[3];
  int32_t row[3];          // >= 0: c * q[row]; -1: the constant c; -2: unused
  int32_t link, apply;
  double lo, hi;
};
#define CM_MAX_SOP 8
#define CM_HOT_MAX 8      // 16-byte chunks of a record's head staged per particle by the compiled move kernel
#define CM_SROW (2 * CM_HOT_MAX + 1)
struct CompiledMix {
  int32_t on, n_sop;
  SOp sop[CM_MAX_SOP];
  int32_t dist, n;         // the two-valued plate: FC_BERNOULLI or FC_CATEGORICAL [p, 1-p]; plate length
  int32_t m0, ss;          // Normal observe: mean / sigma picked by z
  int32_t row_ma, row_mb, row_sa, row_sb, soff, hot_chunks;  // hot_chunks: ceil((soff + 40) / 16)
  CParam p0, mean, sigma;
};
__device__ __forceinline__ double cparam_eval(const CParam& P, const double* __restrict__ q) {
  double s = 0.0;
#pragma unroll
  for (int t = 0; t < 3; t++) {
    if (P.row[t] == -1) s += P.c[t];
    else if (P.row[t] >= 0) s += P.c[t] * q[P.row[t]];
  }
  if (P.apply) { double dl; s = apply_link(P.link, s, P.lo, P.hi, dl); }
  return s;
}

// In the fused kernels a particle's f64 sites are the first d_cont doubles of its record, so the scoring
// helpers written for the [row][chain] layout are called with (q = record, C = 1, c = 0).
__device__ __forceinline__ double chain_const_param(const DOp& op, int j, const double* __restrict__ q, bool apply) {
  double s = 0.0;
  for (int t = 0; t < 3; t++) {
    const DTerm& tm = op.p[j].t[t];
    if (tm.kind == FC_TERM_CONST) s += tm.c;
    else if (tm.kind == FC_TERM_LATENT) s += tm.c * q[tm.row];
  }
  if (apply && op.p[j].invariant) { double dl; s = apply_link(op.p[j].link, s, op.p[j].lo, op.p[j].hi, dl); }
  return s;
}
struct NormalSetup {  // per-particle constants of a Normal observation plate (same arithmetic as k_plate_vec)
  double base0, va0, vb0, va1, inv_a, inv_b, c_a, c_b;  // c = -ln sigma - 0.5 ln 2 pi of the branch
};
__device__ __forceinline__ NormalSetup normal_setup(const DOp& op, const FPlate& fp, const double* __restrict__ q) {
  NormalSetup S;
  S.base0 = chain_const_param(op, 0, q, true);
  const double base1 = chain_const_param(op, 1, q, true);
  S.va0 = S.vb0 = S.va1 = 0.0;
  if (fp.m0) { const DTerm& T0 = op.p[0].t[fp.mt0]; S.va0 = q[T0.row]; if (fp.m0 == FM_SELECT2) S.vb0 = q[T0.row_b]; }
  if (fp.m1) { const DTerm& T1 = op.p[0].t[fp.mt1]; S.va1 = q[T1.row]; }
  double sa = base1, sb = base1;
  if (fp.ss) { const DTerm& TS = op.p[1].t[fp.st]; sa = base1 + q[TS.row]; sb = base1 + q[TS.row_b]; }
  // an invalid sigma scores -inf through inv = 0, ln = +inf (distribution.rs:189-208 guards)
  const bool oka = sa > 0.0 && d_fin(sa), okb = sb > 0.0 && d_fin(sb);
  S.inv_a = oka ? 1.0 / sa : 0.0; S.c_a = oka ? -log(sa) - 0.5 * FC_LN_2PI : d_neg_inf();
  S.inv_b = okb ? 1.0 / sb : 0.0; S.c_b = okb ? -log(sb) - 0.5 * FC_LN_2PI : d_neg_inf();
  return S;
}
// one Normal observation: -0.5 z^2 - ln sigma - 0.5 ln 2 pi. A NaN (0 * inf from an invalid sigma, or a
// non-finite value / mean) is the reference's -inf; the caller maps a NaN SUM to -inf once, which is the
// same thing because a Normal log-density is never +inf.
__device__ __forceinline__ double normal_elem(const NormalSetup& S, const FPlate& fp, int zm, int zs, double lx0, double lx1, double x) {
  double mean = S.base0;
  if (fp.m0 == FM_SELECT2) mean += zm ? S.vb0 : S.va0;
  if (fp.m0 == FM_LATENT_X) mean += S.va0 * lx0;
  if (fp.m1 == FM_LATENT_X) mean += S.va1 * lx1;
  const bool useb = fp.ss && zs;
  const double z = (x - mean) * (useb ? S.inv_b : S.inv_a);
  return fma(-0.5 * z, z, useb ? S.c_b : S.c_a);
}

// Sufficient statistics of a mixture plate for one particle. The discrete sites never move in a tempered
// rejuvenation move (smc.rs:639-645) and are cloned with the particle by resampling, so the assignment z of
// a particle is FROZEN from the prior draw on, and with it, per component c: how many observations it owns
// (n_c), their mean m_c and M2_c = sum (x_i - m_c)^2. The component's share of the observation plate is then
//   sum_{z_i = c} [ -0.5 ((x_i - mu_c) / sigma_c)^2 - ln sigma_c - 0.5 ln 2 pi ]
//     = -0.5 (M2_c + n_c (m_c - mu_c)^2) / sigma_c^2 + n_c (-ln sigma_c - 0.5 ln 2 pi)
// (centred form: no cancellation), i.e. a move scores its proposal in O(1) instead of one Normal term per
// observation, and never reads the z bytes. st[0] = n_1 (NaN: a site outside {0, 1}, statistics unusable),
// st[1..2] = m_0, M2_0, st[3..4] = m_1, M2_1.
__device__ __noinline__ void mix_stats(const FusedProg& F, const FPlate& fp, const uint8_t* __restrict__ zrec, double* __restrict__ st) {
  const int z0 = (int)F.ops[fp.op].dslot0;
  const double* __restrict__ xs = (const double*)F.ops[fp.op_b].value_col.ptr;
  const ZAcc za{zrec, F.L.zw};
  double s0 = 0.0, s1 = 0.0;
  int n1 = 0;
  bool other = false;
#pragma unroll 1
  for (int i = 0; i < fp.n; i++) {
    const unsigned zi = (unsigned)za(z0 + i);
    const double x = __ldg(xs + i);
    other |= zi > 1u;
    if (zi) { n1++; s1 += x; } else s0 += x;
  }
  const int n0 = fp.n - n1;
  const double m0 = n0 ? s0 / (double)n0 : 0.0, m1 = n1 ? s1 / (double)n1 : 0.0;
  double q0 = 0.0, q1 = 0.0;
#pragma unroll 1
  for (int i = 0; i < fp.n; i++) {
    const unsigned zi = (unsigned)za(z0 + i);
    const double d = __ldg(xs + i) - (zi ? m1 : m0);
    if (zi) q1 = fma(d, d, q1); else q0 = fma(d, d, q0);
  }
  st[0] = other ? __longlong_as_double(0x7ff8000000000000LL) : (double)n1;
  st[1] = m0; st[2] = q0; st[3] = m1; st[4] = q1;
}

// the same from a discrete state held in two registers (one bit per site, at most 128 sites)
__device__ __noinline__ void mix_stats_bits(const FusedProg& F, const FPlate& fp, unsigned long long zr0, unsigned long long zr1, double* __restrict__ st) {
  const int z0 = (int)F.ops[fp.op].dslot0;
  const double* __restrict__ xs = (const double*)F.ops[fp.op_b].value_col.ptr;
  auto zbit = [&](int bit) -> unsigned { return (unsigned)(((bit < 64 ? zr0 >> bit : zr1 >> (bit - 64))) & 1ULL); };
  double s0 = 0.0, s1 = 0.0;
  int n1 = 0;
#pragma unroll 1
  for (int i = 0; i < fp.n; i++) {
    const unsigned zi = zbit(z0 + i);
    const double x = __ldg(xs + i);
    if (zi) { n1++; s1 += x; } else s0 += x;
  }
  const int n0 = fp.n - n1;
  const double m0 = n0 ? s0 / (double)n0 : 0.0, m1 = n1 ? s1 / (double)n1 : 0.0;
  double q0 = 0.0, q1 = 0.0;
#pragma unroll 1
  for (int i = 0; i < fp.n; i++) {
    const unsigned zi = zbit(z0 + i);
    const double d = __ldg(xs + i) - (zi ? m1 : m0);
    if (zi) q1 = fma(d, d, q1); else q0 = fma(d, d, q0);
  }
  st[0] = (double)n1; st[1] = m0; st[2] = q0; st[3] = m1; st[4] = q1;
}

// log_prior / log_likelihood / log_factors of one particle: scalar ops, then every plate, in program
// order; elements of a plate summed in order with one accumulator (the generic kernels' order).
__device__ __forceinline__ void fused_eval(const FusedProg& F, const uint8_t* __restrict__ rec, double (&lp)[3]) {
  const double* __restrict__ q = (const double*)rec;
  const uint8_t* __restrict__ zrec = rec + F.L.zoff;
  double a_prior = 0.0, a_lik = 0.0, a_fac = 0.0;
  ZAcc za{zrec, F.L.zw};
  const double qnan = __longlong_as_double(0x7ff8000000000000LL);
  for (int oi = 0; oi < F.n_scalar; oi++) {
    const SOp so = F.sops[oi];
    if (so.kind == SO_GENERIC) { scalar_ops_eval<false>(F.scalar_ops + oi, 1, F.crow, q, 1, 0, za, nullptr, F.nodes, a_prior, a_lik, a_fac); continue; }
    if (so.kind == SO_FACTOR) { a_fac += so.c0; continue; }
    const double x = q[so.row];
    double v;
    if (so.kind == SO_NORMAL_CC) {        // distribution.rs:189-208
      const double z = (x - so.c0) * so.c1;
      v = fma(-0.5 * z, z, so.c2);
      if (!(v == v)) v = d_neg_inf();     // non-finite x
    } else if (so.kind == SO_GAMMA_CC) {  // distribution.rs:1039-1068; (k - 1) ln x is skipped at k = 1 (+-0)
      if (!(x > 0.0) || !d_fin(x)) v = d_neg_inf();
      else v = (so.c2 + (so.c0 != 0.0 ? so.c0 * log(x) : 0.0)) - so.c1 * x;
    } else {                              // SO_BETA_CC, interior points; the endpoint rules stay in dist_eval
      if (x > 0.0 && x < 1.0) v = ((so.c0 != 0.0 ? so.c0 * log(x) : 0.0) + (so.c1 != 0.0 ? so.c1 * log(1.0 - x) : 0.0)) + so.c2;
      else { const DOp& op = F.scalar_ops[oi]; v = dist_eval<false>(FC_BETA, so.c0 + 1.0, so.c1 + 1.0, 0.0, x, 0, qnan, op.lgc, qnan).lp; }
    }
    a_prior += v;
  }
  lp[0] = 0.0 + a_prior; lp[1] = 0.0 + a_lik; lp[2] = 0.0 + a_fac;
  for (int pi = 0; pi < F.n_plates; pi++) {
    const FPlate fp = F.plates[pi];
    const DOp& op = F.ops[fp.op];
    if (fp.kind == FP_TWOVAL_SAMPLE || fp.kind == FP_TWOVAL_OBS || fp.kind == FP_MIX) {
      const double p0 = chain_const_param(op, 0, q, true);
      const double tv0 = dist_eval<false>(op.dist, p0, 0.0, 0.0, 0.0, 0, qnan).lp;
      const double tv1 = dist_eval<false>(op.dist, p0, 0.0, 0.0, 0.0, 1, qnan).lp;
      const bool bern = op.dist == FC_BERNOULLI;
      const double tv2 = bern ? tv1 : d_neg_inf();  // values >= 2: true for a bool, out of range for [p, 1-p] (distribution.rs:785-791)
      double s = 0.0;
      if (fp.kind == FP_TWOVAL_OBS) {
        for (int i = 0; i < fp.n; i++) {
          const long long xi = dcol_i64(op.value_col, i);
          s += xi == 0 ? tv0 : (xi == 1 ? tv1 : tv2);
        }
        lp[1] += s;
      } else if (fp.kind == FP_TWOVAL_SAMPLE) {
        const int z0 = (int)op.dslot0;
        for (int i = 0; i < fp.n; i++) {
          const long long xi = za(z0 + i);
          s += xi == 0 ? tv0 : (xi == 1 ? tv1 : tv2);
        }
        lp[0] += s;
      } else {
        // FP_MIX: z#i ~ two-valued(p); x#i ~ Normal(mean[z_i], sigma[z_i]) (examples/mixture_models.rs:96-108):
        // one pass over the particle's z bytes serves the discrete prior and the observation plate
        const DOp& ob = F.ops[fp.op_b];
        const NormalSetup S = normal_setup(ob, fp, q);
        if (fp.stat && F.L.soff) {
          const double* __restrict__ stt = (const double*)(rec + F.L.soff);
          const double n1 = stt[0];
          if (n1 == n1) {  // the O(1) path: per-component sufficient statistics (mix_stats)
            const double n0 = (double)fp.n - n1;
            const double ia = S.inv_a, ib = fp.ss ? S.inv_b : S.inv_a, ca = S.c_a, cb = fp.ss ? S.c_b : S.c_a;
            double sl = 0.0;
            if (n0 > 0.0) { const double d = stt[1] - (S.base0 + S.va0); sl += fma(-0.5 * ia * ia, fma(n0 * d, d, stt[2]), n0 * ca); }
            if (n1 > 0.0) { const double d = stt[3] - (S.base0 + S.vb0); sl += fma(-0.5 * ib * ib, fma(n1 * d, d, stt[4]), n1 * cb); }
            lp[0] += (n0 > 0.0 ? n0 * tv0 : 0.0) + (n1 > 0.0 ? n1 * tv1 : 0.0);
            lp[1] += (sl == sl) ? sl : d_neg_inf();
            continue;
          }
        }
        const int z0 = (int)op.dslot0;
        const double* __restrict__ xs = (const double*)ob.value_col.ptr;
        const bool x_f64 = ob.value_col.dtype == FC_COL_F64;
        const double* __restrict__ lxc = fp.m1 == FM_LATENT_X ? (const double*)ob.p[0].t[fp.mt1].col.ptr : nullptr;
        double sl = 0.0;
        const int nfull = (F.L.zw == 1 && (z0 & 15) == 0 && x_f64 && !lxc) ? (fp.n & ~15) : 0;
        int n1 = 0;
        unsigned other = 0;
        if (nfull) {
          // the hot loop of BASELINE configs[3]: 16 sites per 16-byte load. The observation term costs the
          // branch selects and 5 FP64 instructions per element; the discrete prior is a population count
          // (n0 ln p0 + n1 ln p1 instead of 80 sequential adds of the same two values).
          const double ma = S.base0 + S.va0, mb = S.base0 + S.vb0;
          const double ia = S.inv_a, ib = fp.ss ? S.inv_b : S.inv_a, ca = S.c_a, cb = fp.ss ? S.c_b : S.c_a;
          for (int i0 = 0; i0 < nfull; i0 += 16) {
            const uint4 v = *(const uint4*)(zrec + z0 + i0);
            const unsigned wv[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
            for (int k = 0; k < 4; k++) { n1 += __popc(wv[k] & 0x01010101u); other |= wv[k] & 0xfefefefeu; }
#pragma unroll
            for (int b = 0; b < 16; b++) {
              const bool one = ((wv[b >> 2] >> ((b & 3) * 8)) & 0xffu) != 0;
              const double z = (__ldg(xs + i0 + b) - (one ? mb : ma)) * (one ? ib : ia);
              sl += fma(-0.5 * z, z, one ? cb : ca);
            }
          }
          if (other) {  // a byte >= 2: out of range for [p, 1-p] (distribution.rs:785-791), `true` for a bool
            s = 0.0;
            for (int i = 0; i < nfull; i++) { const int zi = (int)za(z0 + i); s += zi == 0 ? tv0 : (zi == 1 ? tv1 : tv2); }
          } else {
            const int n0 = nfull - n1;
            s = (n0 ? (double)n0 * tv0 : 0.0) + (n1 ? (double)n1 * tv1 : 0.0);
          }
        }
        for (int i = nfull; i < fp.n; i++) {
          const int zi = (int)za(z0 + i);
          s += zi == 0 ? tv0 : (zi == 1 ? tv1 : tv2);
          const double x = x_f64 ? __ldg(xs + i) : dcol_f64(ob.value_col, i);
          sl += normal_elem(S, fp, zi != 0, zi != 0, 0.0, lxc ? __ldg(lxc + i) : 0.0, x);
        }
        lp[0] += s;
        lp[1] += (sl == sl) ? sl : d_neg_inf();
      }
    } else {  // FP_NORMAL_OBS
      const NormalSetup S = normal_setup(op, fp, q);
      const int zm0 = fp.m0 == FM_SELECT2 ? op.p[0].t[fp.mt0].dslot : 0;
      const int zs0 = fp.ss ? op.p[1].t[fp.st].dslot : 0;
      const double* __restrict__ lx0c = fp.m0 == FM_LATENT_X ? (const double*)op.p[0].t[fp.mt0].col.ptr : nullptr;
      const double* __restrict__ lx1c = fp.m1 == FM_LATENT_X ? (const double*)op.p[0].t[fp.mt1].col.ptr : nullptr;
      double sl = 0.0;
      for (int i = 0; i < fp.n; i++) {
        const int zm = fp.m0 == FM_SELECT2 ? (za(zm0 + i) != 0) : 0;
        const int zs = fp.ss ? (za(zs0 + i) != 0) : 0;
        sl += normal_elem(S, fp, zm, zs, lx0c ? __ldg(lx0c + i) : 0.0, lx1c ? __ldg(lx1c + i) : 0.0, dcol_f64(op.value_col, i));
      }
      lp[1] += (sl == sl) ? sl : d_neg_inf();
    }
  }
}

// fused_eval for a CompiledMix program (K.on): the same operations in the same order, operands from K
__device__ __forceinline__ void compiled_eval(const CompiledMix& K, const FusedProg& F, const uint8_t* __restrict__ rec, double (&lp)[3]) {
  const double* __restrict__ q = (const double*)rec;
  const double qnan = __longlong_as_double(0x7ff8000000000000LL);
  double a_prior = 0.0, a_fac = 0.0;
#pragma unroll 1
  for (int oi = 0; oi < K.n_sop; oi++) {
    const SOp so = K.sop[oi];
    if (so.kind == SO_FACTOR) { a_fac += so.c0; continue; }
    const double x = q[so.row];
    double v;
    if (so.kind == SO_NORMAL_CC) {
      const double z = (x - so.c0) * so.c1;
      v = fma(-0.5 * z, z, so.c2);
      if (!(v == v)) v = d_neg_inf();
    } else if (so.kind == SO_GAMMA_CC) {
      if (!(x > 0.0) || !d_fin(x)) v = d_neg_inf();
      else v = (so.c2 + (so.c0 != 0.0 ? so.c0 * log(x) : 0.0)) - so.c1 * x;
    } else {
      if (x > 0.0 && x < 1.0) v = ((so.c0 != 0.0 ? so.c0 * log(x) : 0.0) + (so.c1 != 0.0 ? so.c1 * log(1.0 - x) : 0.0)) + so.c2;
      else { const DOp& op = F.scalar_ops[oi]; v = dist_eval<false>(FC_BETA, so.c0 + 1.0, so.c1 + 1.0, 0.0, x, 0, qnan, op.lgc, qnan).lp; }
    }
    a_prior += v;
  }
  lp[0] = 0.0 + a_prior; lp[1] = 0.0 + 0.0; lp[2] = 0.0 + a_fac;
  const double p0 = cparam_eval(K.p0, q);
  double tv0, tv1;
  if (K.dist == FC_BERNOULLI) { tv0 = dist_eval<false>(FC_BERNOULLI, p0, 0.0, 0.0, 0.0, 0, qnan).lp; tv1 = dist_eval<false>(FC_BERNOULLI, p0, 0.0, 0.0, 0.0, 1, qnan).lp; }
  else { tv0 = dist_eval<false>(FC_CATEGORICAL, p0, 0.0, 0.0, 0.0, 0, qnan).lp; tv1 = dist_eval<false>(FC_CATEGORICAL, p0, 0.0, 0.0, 0.0, 1, qnan).lp; }
  // normal_setup
  const double base0 = cparam_eval(K.mean, q), base1 = cparam_eval(K.sigma, q);
  double va0 = 0.0, vb0 = 0.0;
  if (K.m0) { va0 = q[K.row_ma]; vb0 = q[K.row_mb]; }
  double sa = base1, sb = base1;
  if (K.ss) { sa = base1 + q[K.row_sa]; sb = base1 + q[K.row_sb]; }
  const bool oka = sa > 0.0 && d_fin(sa), okb = sb > 0.0 && d_fin(sb);
  const double inv_a = oka ? 1.0 / sa : 0.0, c_a = oka ? -log(sa) - 0.5 * FC_LN_2PI : d_neg_inf();
  const double inv_b = okb ? 1.0 / sb : 0.0, c_b = okb ? -log(sb) - 0.5 * FC_LN_2PI : d_neg_inf();
  const double* __restrict__ stt = (const double*)(rec + K.soff);
  const double n1 = stt[0], n0 = (double)K.n - n1;
  const double ia = inv_a, ib = K.ss ? inv_b : inv_a, ca = c_a, cb = K.ss ? c_b : c_a;
  double sl = 0.0;
  if (n0 > 0.0) { const double d = stt[1] - (base0 + va0); sl += fma(-0.5 * ia * ia, fma(n0 * d, d, stt[2]), n0 * ca); }
  if (n1 > 0.0) { const double d = stt[3] - (base0 + vb0); sl += fma(-0.5 * ib * ib, fma(n1 * d, d, stt[4]), n1 * cb); }
  lp[0] += (n0 > 0.0 ? n0 * tv0 : 0.0) + (n1 > 0.0 ? n1 * tv1 : 0.0);
  lp[1] += (sl == sl) ? sl : d_neg_inf();
}

// ---- prior draw of every sample site (PriorHandler, interpreters.rs:88-104,355-372) in program order; the
// same Philox stream and draw order as the generic k_prior_sample, so both paths give the same population.
// Scored in the same kernel: the record is still in registers / L1. ----
__global__ void __launch_bounds__(128) k_smcx_prior(const __grid_constant__ FusedProg F, Pop pop, uint64_t seed, uint64_t stream0, uint64_t tag) {
  const int64_t c = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= pop.n) return;
  DevRng rng(seed, stream0 + (uint64_t)c, tag);
  uint8_t* rec = pop.rec + c * F.L.pitch;
  double* q = (double*)rec;
  uint8_t* zrec = rec + F.L.zoff;
  // one bit per site: up to 128 discrete sites are assembled in two registers and stored once (read-modify-write
  // on the record's bytes was a chain of dependent global accesses per particle); larger states clear the region
  // and set bits in memory
  const bool zregs = F.L.zw == 0 && F.L.d_disc <= 128;
  unsigned long long zr0 = 0, zr1 = 0;
  if (F.L.zw == 0 && !zregs)
    for (int64_t b = F.L.zoff; b < F.L.pitch; b += 16) *(uint4*)(rec + b) = make_uint4(0u, 0u, 0u, 0u);
  for (int oi = 0; oi < F.n_ops; oi++) {
    const DOp& op = F.ops[oi];
    if (!op.is_sample) continue;
    if (op.kind == FC_OP_SAMPLE) {
      double pv[3] = {0.0, 0.0, 0.0};
      for (int j = 0; j < op.n_params; j++) pv[j] = eval_param_value(op, j, 0, q, nullptr, 0, 1, F.crow, nullptr);  // scalar ops: CONST / LATENT terms only
      double x;
      long long xi;
      op_dist_sample(rng, op, pv[0], pv[1], pv[2], x, xi);
      if (op.cont) q[F.crow[op.cslot0]] = x;
      else if (zregs) { if (xi) { if (op.dslot0 < 64) zr0 |= 1ULL << op.dslot0; else zr1 |= 1ULL << (op.dslot0 - 64); } }
      else if (F.L.zw == 0) { if (xi) zrec[op.dslot0 >> 3] |= (uint8_t)(1u << (op.dslot0 & 7)); }
      else if (F.L.zw == 1) zrec[op.dslot0] = (uint8_t)xi;
      else ((int32_t*)zrec)[op.dslot0] = (int32_t)xi;
    } else {  // two-valued discrete plate with a chain-constant parameter
      const double p0 = chain_const_param(op, 0, q, true);
      const bool bern = op.dist == FC_BERNOULLI;
      const int z0 = (int)op.dslot0, n = (int)op.plate_len;
      // Uniforms come in pairs from one Philox block (DevRng::uniform's own order: a, then the spare b); the
      // loops are kept rolled -- 16 inlined draws per trip thrashed the instruction cache (ncu: 18 of 22
      // stall samples "no instruction", 31 ms) -- and 8 sites are packed per 8-byte store.
      int i = 0;
      if (zregs) {
#pragma unroll 1
        for (; i < n; i++) {
          const bool lt = rng.uniform() < p0;
          const int bit = z0 + i;
          if (bern ? lt : !lt) { if (bit < 64) zr0 |= 1ULL << bit; else zr1 |= 1ULL << (bit - 64); }
        }
      }
      if (F.L.zw == 0 && !zregs) {  // one bit per site: the z region was cleared above; same draws, same order
#pragma unroll 1
        for (; i < n; i += 56) {  // 56 sites per word: shifted by the start bit's offset in its byte it still fits 64 bits
          unsigned long long wd = 0;
          const int m = min(56, n - i);
#pragma unroll 1
          for (int b = 0; b < m; b++) {
            const bool lt = rng.uniform() < p0;
            if (bern ? lt : !lt) wd |= 1ULL << b;
          }
          const int bit0 = z0 + i;
          wd <<= (bit0 & 7);
          uint8_t* dstb = zrec + (bit0 >> 3);
          const int nbytes = ((bit0 & 7) + m + 7) >> 3;
#pragma unroll 1
          for (int k = 0; k < nbytes; k++) dstb[k] |= (uint8_t)(wd >> (8 * k));
        }
        i = n;
      }
      if ((z0 & 7) == 0 && F.L.zw == 1) {
        if (rng.has_spare && n > 0) {  // an odd number of draws so far: realign to whole blocks through the byte path
          const bool lt = rng.uniform() < p0;
          zrec[z0] = (uint8_t)(bern ? (lt ? 1 : 0) : (lt ? 0 : 1));
          i = 1;
          for (; i < n && (i & 7); i++) { const bool l2 = rng.uniform() < p0; zrec[z0 + i] = (uint8_t)(bern ? (l2 ? 1 : 0) : (l2 ? 0 : 1)); }
        }
#pragma unroll 1
        for (; i + 8 <= n && !rng.has_spare; i += 8) {
          unsigned long long wd = 0;
#pragma unroll 1
          for (int b = 0; b < 8; b += 2) {
            double ua, ub;
            rng.ph.uniform2(rng.lo++, rng.hi, ua, ub);
            const unsigned long long za_ = bern ? (ua < p0 ? 1u : 0u) : (ua < p0 ? 0u : 1u), zb_ = bern ? (ub < p0 ? 1u : 0u) : (ub < p0 ? 0u : 1u);  // Bernoulli u < p; Categorical [p, 1-p] CDF search
            wd |= (za_ << (8 * b)) | (zb_ << (8 * b + 8));
          }
          *(unsigned long long*)(zrec + z0 + i) = wd;
        }
      }
#pragma unroll 1
      for (; i < n; i++) {
        const bool lt = rng.uniform() < p0;
        zrec[z0 + i] = (uint8_t)(bern ? (lt ? 1 : 0) : (lt ? 0 : 1));
      }
    }
  }
  if (zregs) {
    *(ulonglong2*)zrec = make_ulonglong2(zr0, zr1);  // the z region is 16-byte aligned and at least 16 bytes long
    for (int64_t b = F.L.zoff + 16; b < F.L.pitch; b += 16) *(uint4*)(rec + b) = make_uint4(0u, 0u, 0u, 0u);
  }
  if (F.L.soff)
    for (int pi = 0; pi < F.n_plates; pi++)
      if (F.plates[pi].kind == FP_MIX && F.plates[pi].stat) {
        if (zregs) mix_stats_bits(F, F.plates[pi], zr0, zr1, (double*)(rec + F.L.soff));
        else mix_stats(F, F.plates[pi], zrec, (double*)(rec + F.L.soff));
      }
  double lp[3];
  fused_eval(F, rec, lp);
  double* l3 = q + F.L.d_cont;
  l3[0] = lp[0]; l3[1] = lp[1]; l3[2] = lp[2];
  pop.ll[c] = lp[1] + lp[2];
}

// ---- one tempered single-site MH move per particle (tempered_single_site_mh, smc.rs:631-688): uniform
// pick among the f64 sites, x' = x + scale * N(0,1), log alpha = d log_prior + beta d (loglik + factors),
// accept rule log alpha >= 0 || u < exp(log alpha). Same random numbers as k_move_propose / k_move_accept. ----
struct MoveX {
  double* scale;          // [d_cont] shared per-site proposal scale
  uint32_t* acc_cnt;      // [d_cont]
  uint32_t* tot_cnt;
  uint64_t seed, chain0, step_tag;
  double beta;
  int32_t adapt;
};
template <bool COMPILED>
__global__ void __launch_bounds__(128, COMPILED ? 8 : 6) k_smcx_move(const __grid_constant__ FusedProg F, const __grid_constant__ CompiledMix K, Pop pop, const __grid_constant__ MoveX M) {
  const int64_t c = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t dc = F.L.d_cont;
  const bool active = c < pop.n && dc > 0;
  int s = -1;
  bool accept = false;
  // COMPILED: the head of the CTA's 128 consecutive records (f64 sites, accumulators, statistics; not z) is
  // staged in shared memory by coalesced 16-byte loads -- one exposed DRAM latency per thread and 4 cache lines
  // per warp-wide load, where per-field reads at a 128-byte stride touched 32 lines each and serialised the
  // misses (ncu: 84 % L1 throughput, 50 % of the stall samples on the long scoreboard). Row stride 17 doubles:
  // a thread's own fields are then conflict-free.
  __shared__ double s_rec[COMPILED ? 128 * CM_SROW : 1];
  if (COMPILED) {
    const int64_t first = (int64_t)blockIdx.x * 128;
    const int live = (int)min((int64_t)128, pop.n - first);
    const int hot = K.hot_chunks, cpp = (int)(F.L.pitch >> 4);
    const uint8_t* __restrict__ base = pop.rec + first * F.L.pitch;
    for (int ch = threadIdx.x; ch < live * hot; ch += 128) {
      const int slot = ch / hot, part = ch - slot * hot;
      const double2 v = *(const double2*)(base + ((int64_t)slot * cpp + part) * 16);
      s_rec[slot * CM_SROW + part * 2] = v.x;
      s_rec[slot * CM_SROW + part * 2 + 1] = v.y;
    }
    __syncthreads();
  }
  if (active) {
    Philox ph(M.seed, M.chain0 + (uint64_t)c);
    double u_site, u_acc, z0, z1;
    ph.uniform2(0, M.step_tag, u_site, u_acc);
    ph.normal2(1, M.step_tag, z0, z1);  // Normal::new(0,1).sample (smc.rs:656)
    int64_t si = (int64_t)(u_site * (double)dc);
    if (si >= dc) si = dc - 1;
    s = (int)si;
    uint8_t* rec = pop.rec + c * F.L.pitch;
    double* q = (double*)rec;
    double* l3 = q + dc;
    if (COMPILED) {
      // the proposal is scored on the staged copy; global memory is written only on acceptance
      double* sq = s_rec + threadIdx.x * CM_SROW;
      const double prop = sq[si] + M.scale[s] * z0;
      sq[si] = prop;
      double lpn[3];
      compiled_eval(K, F, (const uint8_t*)sq, lpn);  // the plate's sites are two-valued draws: the statistics are always usable
      const double c0 = sq[dc], c1 = sq[dc + 1], c2 = sq[dc + 2];
      const double log_alpha = (lpn[0] - c0) + M.beta * ((lpn[1] + lpn[2]) - (c1 + c2));  // smc.rs:679-680
      accept = log_alpha >= 0.0 || u_acc < exp(log_alpha);
      if (accept) {
        q[si] = prop;
        l3[0] = lpn[0]; l3[1] = lpn[1]; l3[2] = lpn[2];
        pop.ll[c] = lpn[1] + lpn[2];
      }
    } else {
      const double cur = q[si];
      q[si] = cur + M.scale[s] * z0;
      double lpn[3];
      fused_eval(F, rec, lpn);
      const double c0 = l3[0], c1 = l3[1], c2 = l3[2];
      const double log_alpha = (lpn[0] - c0) + M.beta * ((lpn[1] + lpn[2]) - (c1 + c2));  // smc.rs:679-680
      accept = log_alpha >= 0.0 || u_acc < exp(log_alpha);
      if (accept) {
        l3[0] = lpn[0]; l3[1] = lpn[1]; l3[2] = lpn[2];
        pop.ll[c] = lpn[1] + lpn[2];
      } else {
        q[si] = cur;
      }
    }
  }
  if (!M.adapt) return;
  // shared per-site counters: aggregate in shared memory, one global atomic per site per CTA
  __shared__ unsigned s_tot[64], s_acc[64];
  if (dc <= 64) {
    for (int i = threadIdx.x; i < 64; i += blockDim.x) { s_tot[i] = 0; s_acc[i] = 0; }
    __syncthreads();
    if (s >= 0) { atomicAdd(&s_tot[s], 1u); if (accept) atomicAdd(&s_acc[s], 1u); }
    __syncthreads();
    for (int i = threadIdx.x; i < (int)dc; i += blockDim.x) {
      if (s_tot[i]) atomicAdd(&M.tot_cnt[i], s_tot[i]);
      if (s_acc[i]) atomicAdd(&M.acc_cnt[i], s_acc[i]);
    }
  } else if (s >= 0) { atomicAdd(&M.tot_cnt[s], 1u); if (accept) atomicAdd(&M.acc_cnt[s], 1u); }
}

// ---- resampling: ancestor search + record copy in one kernel (resample_particles, smc.rs:326-349) ----
// Output slot j (global) takes the particle at which the walk of smc.rs:263-268 / 284-289 stops: the
// first i with S_i >= threshold_j over the exact running sums S of the SOURCE rank's weights (S_0 = its
// carry-in). Ancestors of consecutive slots are non-decreasing, so the source records are read nearly
// in order; the destination (this rank's or a peer's buffer, over NVLink) is written contiguously,
// 16 bytes per thread.
__device__ __forceinline__ int64_t lower_bound_S(const double* __restrict__ S, int64_t n, double t) {  // first i in [0, n] with S[i] >= t, or n + 1
  int64_t lo = 0, hi = n + 1;
  while (lo < hi) {
    int64_t mid = (lo + hi) >> 1;
    if (S[mid] >= t) hi = mid; else lo = mid + 1;
  }
  return lo;
}
__device__ __forceinline__ double resample_threshold(int method, double n_total, double u_raw, uint64_t seed, uint64_t tag, int64_t j) {
  if (method == FC_RESAMPLE_SYSTEMATIC) return __dadd_rn(__ddiv_rn(u_raw, n_total), __ddiv_rn((double)j, n_total));  // u = gen()/n; thr = u + j/n (smc.rs:258,264)
  Philox ph(seed, 0x8000000000000000ULL);
  double a, b;
  ph.uniform2((uint64_t)(j >> 1), tag, a, b);
  return __ddiv_rn(__dadd_rn((double)j, (j & 1) ? b : a), n_total);  // (j + u_j)/n (smc.rs:281-283)
}
struct GatherArgs {
  const double* S;        // source rank's running sums S[0..n_src] (inline walk) or null
  const uint64_t* idx_in; // explicit ancestors (local indices), used when S is null
  const double* us;       // explicit per-slot uniforms of a stratified resample (indexed by global slot), or null
  int64_t n_src;
  double n_total, u_raw;
  int32_t method, wrap0;  // wrap0: this launch is the single u == 0 wrap slot (ancestor = last particle)
  uint64_t seed, tag;
  int64_t j0, cnt;        // global output slots [j0, j0 + cnt)
  int64_t dst_off;        // destination-local index of slot j0
  Pop src, dst;
  int64_t pitch, ll_off;  // record size; byte offset of (log_lik, log_factors) in a record
  uint64_t* idx_out;      // optional: GLOBAL ancestor index of slot j0 + k -> idx_out[idx_off + k]
  int64_t idx_off, first_src;
};
#define GA_THREADS 256
__global__ void __launch_bounds__(GA_THREADS) k_smcx_gather(const __grid_constant__ GatherArgs A) {
  __shared__ int64_t s_anc[GA_THREADS];
  __shared__ int64_t s_range[2];
  const int64_t k = (int64_t)blockIdx.x * GA_THREADS + threadIdx.x;
  const int64_t kbase = (int64_t)blockIdx.x * GA_THREADS;
  int64_t a = 0;
  const bool walk = A.S && !A.wrap0;
  double thr = 0.0;
  if (walk && k < A.cnt) {
    const int64_t j = A.j0 + k;
    thr = (A.method == FC_RESAMPLE_STRATIFIED && A.us) ? __ddiv_rn(__dadd_rn((double)j, A.us[j]), A.n_total)
                                                       : resample_threshold(A.method, A.n_total, A.u_raw, A.seed, A.tag, j);
  }
  if (walk) {
    // thresholds are non-decreasing in the slot index, so every search of this CTA ends inside the range
    // spanned by its first and last slot: two full searches, then 256 short ones that stay in L1 / L2
    const int64_t klast = min(A.cnt, kbase + GA_THREADS) - 1;
    if (k == kbase) s_range[0] = lower_bound_S(A.S, A.n_src, thr);
    if (k == klast) s_range[1] = lower_bound_S(A.S, A.n_src, thr);
    __syncthreads();
  }
  if (k < A.cnt) {
    if (A.S) {
      if (A.wrap0) a = A.n_src - 1;
      else {
        int64_t i;
        if (thr == thr) {
          int64_t lo = s_range[0], hi = min(s_range[1], A.n_src + 1);  // first i in [lo, hi] with S[i] >= thr
          while (lo < hi) {
            const int64_t mid = (lo + hi) >> 1;
            if (A.S[mid] >= thr) hi = mid; else lo = mid + 1;
          }
          i = lo;
        } else i = 0;
        if (i > A.n_src) i = A.n_src;
        a = (i == 0) ? A.n_src - 1 : (i - 1 < A.n_src - 1 ? i - 1 : A.n_src - 1);  // (i - 1).min(n - 1); i == 0 wraps
        if (!(thr == thr)) a = A.n_src - 1;
      }
    } else a = (int64_t)A.idx_in[k];
    if (A.idx_out) A.idx_out[A.idx_off + k] = (uint64_t)(A.first_src + a);
    const double* l = (const double*)(A.src.rec + a * A.pitch + A.ll_off);  // (log_lik, log_factors) of the ancestor
    A.dst.ll[A.dst_off + k] = l[0] + l[1];
  }
  s_anc[threadIdx.x] = a;
  __syncthreads();
  const int cpp = (int)(A.pitch >> 4);
  const int64_t base = (int64_t)blockIdx.x * GA_THREADS;
  const int64_t live = min((int64_t)GA_THREADS, A.cnt - base);
  uint8_t* __restrict__ dst = A.dst.rec + (A.dst_off + base) * A.pitch;
  for (int64_t ch = threadIdx.x; ch < live * cpp; ch += GA_THREADS) {
    const int slot = (int)(ch / cpp), part = (int)(ch - (int64_t)slot * cpp);
    const uint4 v = *(const uint4*)(A.src.rec + s_anc[slot] * A.pitch + (int64_t)part * 16);
    *(uint4*)(dst + ch * 16) = v;
  }
}

// ---- records <-> the generic engine's [row][particle] arrays ----
__global__ void k_rec_unpack(const uint8_t* __restrict__ rec, RecLayout L, const int32_t* __restrict__ drow, int64_t n, double* __restrict__ q, double* __restrict__ lp3, disc_t* __restrict__ qd) {
  const int64_t c = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= n) return;
  const uint8_t* r = rec + c * L.pitch;
  const double* rq = (const double*)r;
  if (q) for (int64_t s = 0; s < L.d_cont; s++) q[s * n + c] = rq[s];
  if (lp3) for (int a = 0; a < 3; a++) lp3[a * n + c] = rq[L.d_cont + a];
  if (qd) {
    const uint8_t* z = r + L.zoff;
    for (int64_t s = 0; s < L.d_disc; s++) qd[(int64_t)drow[s] * n + c] = L.zw == 0 ? (disc_t)((z[s >> 3] >> (s & 7)) & 1u) : (L.zw == 1 ? (disc_t)z[s] : ((const int32_t*)z)[s]);
  }
}
__global__ void k_rec_pack(uint8_t* __restrict__ rec, RecLayout L, const int32_t* __restrict__ drow, int64_t n, const double* __restrict__ q, const double* __restrict__ lp3, const disc_t* __restrict__ qd,
                           double* __restrict__ ll, int* __restrict__ overflow) {
  const int64_t c = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= n) return;
  uint8_t* r = rec + c * L.pitch;
  double* rq = (double*)r;
  if (q) for (int64_t s = 0; s < L.d_cont; s++) rq[s] = q[s * n + c];
  if (lp3) {
    for (int a = 0; a < 3; a++) rq[L.d_cont + a] = lp3[a * n + c];
    ll[c] = lp3[n + c] + lp3[2 * n + c];
  }
  if (qd) {
    uint8_t* z = r + L.zoff;
    for (int64_t s = 0; s < L.d_disc; s++) {
      const disc_t v = qd[(int64_t)drow[s] * n + c];
      if (L.zw == 0) { if (v < 0 || v > 1) *overflow = 1; if (v) z[s >> 3] |= (uint8_t)(1u << (s & 7)); else z[s >> 3] &= (uint8_t)~(1u << (s & 7)); }
      else if (L.zw == 1) { if (v < 0 || v > 255) *overflow = 1; z[s] = (uint8_t)v; }
      else ((int32_t*)z)[s] = v;
    }
  }
}

}  // namespace fc

// =============================================================================================
// session
// =============================================================================================
#define FC_SMC_HANDLE_MAGIC 0x46435350u
struct SmcHandle {  // FC_SMC_HANDLE_BYTES
  cudaIpcMemHandle_t ipc;  // 64 bytes
  uint32_t magic;
  int32_t rank, device, _pad;
  int64_t n_local, first, pid;
  char _fill[128 - 64 - 16 - 24];
};
static_assert(sizeof(SmcHandle) == FC_SMC_HANDLE_BYTES, "handle size");

struct fc_smc_session {
  fc_program* P = nullptr;
  fc_smc_config cfg;
  uint64_t seed = 0;
  int64_t n = 0, n_total = 0, first = 0;  // first: global index of local particle 0
  uint64_t stream0 = 0;                   // random stream of local particle 0
  int rank = 0, world = 1;
  bool connected = false;
  // slab: mailbox + the two record buffers + the log-likelihood array (exported to the peers)
  char* slab = nullptr;
  size_t slab_bytes = 0;
  size_t off_rec[2] = {0, 0}, off_ll = 0;
  int cur = 0;
  fc::RecLayout L;
  char* peer_slab[MB_MAXW] = {nullptr};
  int64_t peer_n[MB_MAXW] = {0}, bounds[MB_MAXW + 1] = {0};
  fc::Comm cm;
  fc::DevBuf<unsigned long long> seq;
  fc::DevBuf<int> err;
  unsigned long long resamples = 0;  // global resamples started (mailbox sequence of the carry chain)
  // local arrays
  fc::DevBuf<double> log_w, comb, w, S, us, part, scal, ends;
  fc::DevBuf<char> work;
  fc::DevBuf<uint64_t> idx;
  fc::DevBuf<fc::Bisect> bis;
  fc::DevBuf<fc::Tri> tri;
  fc::DevBuf<unsigned> counter;
  bool log_w_uniform = true;  // log_w == -ln n_total everywhere (initially and after every resample)
  fc::ResampleWork* rw = nullptr;
  // fused evaluator
  bool fused = false;
  fc::DevBuf<fc::FPlate> fplates;
  fc::DevBuf<fc::SOp> sops;
  fc::FusedProg F;
  fc::CompiledMix K;   // K.on: the move kernel's by-value program (scalar priors + one mixture plate with statistics)
  // generic engine: [row][particle] copies of the records (valid flags say which are current)
  fc::MoveEngine eng;
  fc::DevBuf<fc::disc_t> qd32;
  fc::DevBuf<double> q_soa, lp3_soa, total;
  fc::DevBuf<uint64_t> n_accept;
  bool soa_valid = false;
  double beta = 0.0, log_evidence = 0.0, prev_dbeta = 0.0;
  uint64_t steps = 0, moves = 0, migrated = 0, ess_evals = 0;
  bool finished = false;
  // optional CUDA-event timing of the phases of a tempering step, on the launching stream
  // (fc_debug_set("smc_timing", 1)): next_beta, reweight, exact scan + carry chain, gather + migration, moves
  std::vector<std::pair<int, std::pair<cudaEvent_t, cudaEvent_t>>> phase_events;
  double phase_ms[5] = {0, 0, 0, 0, 0};
  uint64_t phase_launches[5] = {0, 0, 0, 0, 0};
  bool record = false;               // keep the global ancestors of the slots this rank produces (parity tests)
  int64_t last_jlo = 0, last_jhi = 0;
  bool bis_ll_valid = false;         // max(ll) / bad flag in `bis` describe the current ll

  double* ll() const { return (double*)(slab + off_ll); }
  fc::Pop pop(int which) const {
    fc::Pop p;
    p.rec = (uint8_t*)(slab + off_rec[which]); p.ll = ll(); p.n = n;
    return p;
  }
  fc::Pop peer_pop(int r, int which) const {  // rank r's buffers as mapped here (same layout formula, its own n)
    fc::Pop p;
    size_t orec[2], oll;
    layout(peer_n[r], L.pitch, orec, oll);
    char* b = r == rank ? slab : peer_slab[r];
    p.rec = (uint8_t*)(b + orec[which]); p.ll = (double*)(b + oll); p.n = peer_n[r];
    return p;
  }
  static size_t layout(int64_t n, int64_t pitch, size_t (&orec)[2], size_t& oll) {
    auto up = [](size_t x) { return (x + 255) & ~(size_t)255; };
    size_t o = up(sizeof(fc::MBox));
    for (int k = 0; k < 2; k++) { orec[k] = o; o = up(o + (size_t)pitch * n); }
    oll = o; o = up(o + (size_t)n * 8);
    return o;
  }
};

namespace fc {

static int sm_count_cached() {
  static int sms = 0;
  if (!sms) { int dev = 0; cudaGetDevice(&dev); cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev); if (sms <= 0) sms = 148; }
  return sms;
}
// one resident wave: 4 CTAs of 256 threads per SM, never more than RDB partials
static int nb_blocks(int64_t n) {
  return (int)std::max<int64_t>(1, std::min<int64_t>(std::min<int64_t>(RDB, 4 * sm_count_cached()), cdiv(n, NB_THREADS * 8)));
}
// debug switches (fc_debug_set): "nb_plain" = the reference's plain 64-step bisection without the search
// phase; "smc_generic" = run every program on the generic scoring engine instead of the fused kernels
static int g_nb_plain = -1, g_smc_generic = -1, g_smc_timing = 0;
static bool nb_plain() {
  if (g_nb_plain < 0) g_nb_plain = getenv("FC_NB_PLAIN") ? atoi(getenv("FC_NB_PLAIN")) : 0;
  return g_nb_plain == 1;
}
static int nb_mode() { nb_plain(); return g_nb_plain; }
static int g_smc_stats = 1;  // fc_debug_set("smc_stats", 0): score mixture plates term by term (A/B parity)
static bool smc_stats() { return g_smc_stats != 0; }
static int g_smc_compiled = 1;  // fc_debug_set("smc_compiled", 0): the move kernel interprets the op descriptors
static bool smc_compiled() { return g_smc_compiled != 0; }
static int g_smc_zbits = 1;  // fc_debug_set("smc_zbits", 0): keep one byte per two-valued discrete site
static bool smc_zbits() { return g_smc_zbits != 0; }
static bool smc_generic() {
  if (g_smc_generic < 0) g_smc_generic = getenv("FC_SMC_GENERIC") != nullptr;
  return g_smc_generic != 0;
}

struct PhaseTimer {  // brackets a phase with two events when timing is on
  fc_smc_session* S; int phase; cudaEvent_t e0 = nullptr, e1 = nullptr;
  PhaseTimer(fc_smc_session* S_, int phase_) : S(S_), phase(phase_) {
    if (!g_smc_timing) return;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    cudaEventRecord(e0, S->P->stream);
  }
  void stop() {
    if (!e0) return;
    cudaEventRecord(e1, S->P->stream);
    S->phase_events.push_back({phase, {e0, e1}});
    e0 = nullptr;
  }
  ~PhaseTimer() { stop(); }
};
static void phase_collect(fc_smc_session* S) {  // after a stream synchronisation
  for (auto& pe : S->phase_events) {
    float ms = 0.f;
    if (cudaEventElapsedTime(&ms, pe.second.first, pe.second.second) == cudaSuccess) { S->phase_ms[pe.first] += ms; S->phase_launches[pe.first] += 1; }
    cudaEventDestroy(pe.second.first); cudaEventDestroy(pe.second.second);
  }
  S->phase_events.clear();
}

static int32_t comm_check(fc_smc_session* S) {  // after a stream synchronisation
  int e = 0;
  FC_CUDA(cudaMemcpy(&e, S->err.p, sizeof(int), cudaMemcpyDeviceToHost));
  if (e) { set_error("rank %d: a peer did not answer a device-side exchange within %d s", S->rank, (int)(COMM_TIMEOUT_NS / 1000000000ULL)); return FC_ERR_NCCL; }
  return FC_OK;
}

// ---- next_beta on the session's (uniform) weights ----
static void bisect_init(Bisect& hb, double beta, double target, double n_total, double guess) {
  memset(&hb, 0, sizeof(hb));
  hb.beta = beta; hb.lo = beta; hb.hi = 1.0; hb.a = beta; hb.b = 1.0; hb.mid = 1.0; hb.target = target; hb.n = n_total;
  hb.guess = guess; hb.plain = nb_mode();
}
static int32_t next_beta_uniform(cudaStream_t st, const Comm& cm, double lwc, const double* ll, int64_t n, double n_total, double beta, double target_ess, double guess,
                                 Bisect* d_bis, double* d_part, uint64_t* launches, uint64_t* evals, double* out) {
  const int nb = nb_blocks(n);
  Bisect hb;
  bisect_init(hb, beta, target_ess, n_total, guess);
  FC_CUDA(cudaMemcpyAsync(d_bis, &hb, sizeof(hb), cudaMemcpyHostToDevice, st));
  k_nb_prepass<<<nb, NB_THREADS, 0, st>>>(ll, n, d_bis, d_part, cm);
  int launched = 1;
  // a batch of evaluations (launches after the search has finished return at once), one read-back per batch
  int batch = hb.plain == 1 ? 58 : (hb.plain == 2 ? 30 : 12);
  for (int round = 0; round < 8 && !hb.done; round++) {
    for (int it = 0; it < batch; it++) k_nb_eval_uni<<<nb, NB_THREADS, 0, st>>>(lwc, ll, d_bis, d_part, n, 0, 0.0, cm);
    launched += batch;
    FC_CUDA(cudaMemcpyAsync(&hb, d_bis, sizeof(hb), cudaMemcpyDeviceToHost, st));
    FC_CUDA(cudaStreamSynchronize(st));
    batch = 12;
  }
  FC_KERNEL_CHECK();
  if (launches) *launches += launched;
  if (evals) *evals += hb.evals;
  if (hb.done == 1) { *out = 1.0; return FC_OK; }
  if (!hb.done) { set_error("next_beta: the search did not terminate"); return FC_ERR_CUDA; }
  *out = std::fmin(std::fmax(hb.hi, beta + 1e-9), 1.0);  // smc.rs:621
  return FC_OK;
}
// arbitrary log_w, one rank: the reference's plain bisection
static int32_t next_beta_general(cudaStream_t st, const double* lw, const double* ll, int64_t n, double beta, double target_ess, Bisect* d_bis, Tri* d_tri, uint64_t* launches, double* out) {
  const int nb = nb_blocks(n);
  Bisect hb;
  bisect_init(hb, beta, target_ess, (double)n, 0.0);
  hb.plain = 1;
  FC_CUDA(cudaMemcpyAsync(d_bis, &hb, sizeof(hb), cudaMemcpyHostToDevice, st));
  for (int it = 0; it < 66; it++) {
    k_nb_eval<<<nb, NB_THREADS, 0, st>>>(lw, ll, d_bis, d_tri, n, 0, 0.0);
    if (launches) *launches += 1;
    if (it == 0 || it == 56) {
      FC_CUDA(cudaMemcpyAsync(&hb, d_bis, sizeof(hb), cudaMemcpyDeviceToHost, st));
      FC_CUDA(cudaStreamSynchronize(st));
      if (hb.done) break;
    }
  }
  FC_KERNEL_CHECK();
  FC_CUDA(cudaMemcpyAsync(&hb, d_bis, sizeof(hb), cudaMemcpyDeviceToHost, st));
  FC_CUDA(cudaStreamSynchronize(st));
  if (hb.done == 1) { *out = 1.0; return FC_OK; }
  *out = std::fmin(std::fmax(hb.hi, beta + 1e-9), 1.0);
  return FC_OK;
}

// a one-rank communicator for the standalone entry points
struct LocalComm {
  DevBuf<char> box;
  DevBuf<unsigned long long> seq;
  DevBuf<int> err;
  Comm cm;
  int32_t init() {
    FC_CUDA(box.reserve(sizeof(MBox))); FC_CUDA(seq.reserve(1)); FC_CUDA(err.reserve(1));
    FC_CUDA(cudaMemset(box.p, 0, sizeof(MBox))); FC_CUDA(cudaMemset(seq.p, 0, 8)); FC_CUDA(cudaMemset(err.p, 0, 4));
    memset(&cm, 0, sizeof(cm));
    cm.box[0] = (MBox*)box.p; cm.rank = 0; cm.world = 1; cm.seq = seq.p; cm.err = err.p;
    return FC_OK;
  }
};

// standalone probe of next_beta on host arrays (parity test against the oracle)
int32_t next_beta_host(const double* h_log_w, const double* h_ll, size_t n, double beta, double target_ess, double* out) {
  DevBuf<double> lw, ll, part;
  DevBuf<Bisect> bis;
  DevBuf<Tri> tri;
  FC_CUDA(lw.reserve(n)); FC_CUDA(ll.reserve(n)); FC_CUDA(bis.reserve(1)); FC_CUDA(tri.reserve(RDB + 1)); FC_CUDA(part.reserve(4 * RDB + 8));
  bool uniform = true;  // the state adaptive_smc calls next_beta in: takes the single-stream path
  for (size_t i = 1; i < n && uniform; i++) uniform = h_log_w[i] == h_log_w[0];
  if (!uniform) FC_CUDA(cudaMemcpy(lw.p, h_log_w, n * 8, cudaMemcpyHostToDevice));
  FC_CUDA(cudaMemcpy(ll.p, h_ll, n * 8, cudaMemcpyHostToDevice));
  if (!uniform) return next_beta_general(nullptr, lw.p, ll.p, (int64_t)n, beta, target_ess, bis.p, tri.p, nullptr, out);
  LocalComm lc;
  FC_TRY(lc.init());
  return next_beta_uniform(nullptr, lc.cm, h_log_w[0], ll.p, (int64_t)n, (double)n, beta, target_ess, 0.0, bis.p, part.p, nullptr, nullptr, out);
}

static double lwc_of(const fc_smc_session* S) { return -std::log((double)S->n_total); }

// [row][particle] copies of the current records for the generic engine and the copy-out paths
static int32_t ensure_soa(fc_smc_session* S) {
  fc_program* P = S->P;
  if (S->soa_valid) return FC_OK;
  FC_CUDA(S->q_soa.reserve(std::max<int64_t>(1, P->d_cont * S->n)));
  FC_CUDA(S->lp3_soa.reserve(3 * S->n));
  FC_CUDA(S->qd32.reserve(std::max<int64_t>(1, P->d_disc * S->n)));
  k_rec_unpack<<<(unsigned)cdiv(S->n, 256), 256, 0, P->stream>>>(S->pop(S->cur).rec, S->L, P->d_drow.p, S->n, S->q_soa.p, S->lp3_soa.p, P->d_disc ? S->qd32.p : nullptr);
  FC_KERNEL_CHECK();
  P->n_kernel_launches += 1;
  S->soa_valid = true;
  return FC_OK;
}
// write the generic engine's arrays back into the records (q and the accumulators; discrete sites on request)
static int32_t soa_to_records(fc_smc_session* S, bool with_disc) {
  fc_program* P = S->P;
  k_rec_pack<<<(unsigned)cdiv(S->n, 256), 256, 0, P->stream>>>(S->pop(S->cur).rec, S->L, P->d_drow.p, S->n, S->q_soa.p, S->lp3_soa.p, (with_disc && P->d_disc) ? S->qd32.p : nullptr, S->ll(), S->err.p + 1);
  FC_KERNEL_CHECK();
  P->n_kernel_launches += 1;
  return FC_OK;
}

// log_w as an array (callers that read it: fc_smc_get, fc_smc_device_state, fc_smc_weights)
static int32_t smc_materialise_log_w(fc_smc_session* S) {
  if (!S->log_w_uniform) return FC_OK;
  k_fill2<<<(unsigned)cdiv(S->n, 256), 256, 0, S->P->stream>>>(S->log_w.p, lwc_of(S), S->n);
  FC_KERNEL_CHECK();
  S->P->n_kernel_launches += 1;
  return FC_OK;
}

// global log_sum_exp of a local array (+ optional ESS of exp(x - lse)); results in S->scal[0..2] on the device
static int32_t global_lse(fc_smc_session* S, const double* d_x, double* h_lse) {
  cudaStream_t st = S->P->stream;
  const int nb = nb_blocks(S->n);
  k_glse_max<<<nb, NB_THREADS, 0, st>>>(d_x, S->n, S->part.p, S->counter.p, S->scal.p, S->cm);
  k_glse_sum<<<nb, NB_THREADS, 0, st>>>(d_x, S->n, S->part.p, S->counter.p, S->scal.p, S->cm);
  FC_KERNEL_CHECK();
  S->P->n_kernel_launches += 2;
  if (h_lse) {
    FC_CUDA(cudaMemcpyAsync(h_lse, S->scal.p, 8, cudaMemcpyDeviceToHost, st));
    FC_CUDA(cudaStreamSynchronize(st));
    FC_TRY(comm_check(S));
  }
  return FC_OK;
}

// reweight (smc.rs:512-531): log_norm = lse(log_w + d ll) over ALL ranks; w = exp(log_w + d ll - log_norm).
// log_w itself is only materialised when it will be read again (write_log_w).
static int32_t smc_reweight(fc_smc_session* S, double d_beta, bool write_log_w, double* log_norm) {
  cudaStream_t st = S->P->stream;
  const int64_t n = S->n;
  const double lwc = lwc_of(S);
  if (S->log_w_uniform) {
    Bisect hb;
    if (!S->bis_ll_valid) {  // max(ll) / bad flag: those of this step's next_beta pre-pass when it ran on the same ll
      bisect_init(hb, S->beta, 0.0, (double)S->n_total, 0.0);
      FC_CUDA(cudaMemcpyAsync(S->bis.p, &hb, sizeof(hb), cudaMemcpyHostToDevice, st));
      k_nb_prepass<<<nb_blocks(n), NB_THREADS, 0, st>>>(S->ll(), n, S->bis.p, S->part.p, S->cm);
      S->P->n_kernel_launches += 1;
      S->bis_ll_valid = true;
    }
    k_nb_eval_uni<<<nb_blocks(n), NB_THREADS, 0, st>>>(lwc, S->ll(), S->bis.p, S->part.p, n, 1, d_beta, S->cm);
    k_reweight<<<(unsigned)cdiv(n, 256), 256, 0, st>>>(nullptr, lwc, S->ll(), d_beta, S->bis.p, write_log_w ? S->comb.p : nullptr, S->w.p, n, (double)S->n_total);
    FC_KERNEL_CHECK();
    S->P->n_kernel_launches += 2;
    FC_CUDA(cudaMemcpyAsync(&hb, S->bis.p, sizeof(hb), cudaMemcpyDeviceToHost, st));
    FC_CUDA(cudaStreamSynchronize(st));
    FC_TRY(comm_check(S));
    *log_norm = hb.lse1;
  } else {
    // general weights: combined = log_w + d ll, global lse, then normalise
    unsigned nbk = (unsigned)cdiv(n, 256);
    FC_CUDA(S->us.reserve(n + 1));
    k_combine<<<nbk, 256, 0, st>>>(S->log_w.p, 0.0, S->ll(), d_beta, S->us.p, n);
    FC_TRY(global_lse(S, S->us.p, log_norm));
    k_sub_exp<<<nbk, 256, 0, st>>>(S->us.p, S->scal.p, (double)S->n_total, write_log_w ? S->comb.p : nullptr, S->w.p, n);
    FC_KERNEL_CHECK();
    S->P->n_kernel_launches += 2;
  }
  if (write_log_w) { std::swap(S->log_w.p, S->comb.p); std::swap(S->log_w.n, S->comb.n); S->log_w_uniform = false; }
  return FC_OK;
}

// ---- host replica of the device threshold arithmetic (slot routing of a sharded resample) ----
static void host_philox_uniform2(uint64_t seed, uint64_t stream, uint64_t ctr_lo, uint64_t ctr_hi, double& u0, double& u1) {
  uint64_t z = seed + 0x9E3779B97F4A7C15ULL * (stream + 1);
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ULL;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBULL;
  z ^= z >> 31;
  uint32_t a = (uint32_t)z, b = (uint32_t)(z >> 32);
  uint32_t c0 = (uint32_t)ctr_lo, c1 = (uint32_t)(ctr_lo >> 32), c2 = (uint32_t)ctr_hi, c3 = (uint32_t)(ctr_hi >> 32);
  for (int r = 0; r < 10; r++) {
    uint64_t m0 = (uint64_t)0xD2511F53u * c0, m1 = (uint64_t)0xCD9E8D57u * c2;
    uint32_t hi0 = (uint32_t)(m0 >> 32), lo0 = (uint32_t)m0, hi1 = (uint32_t)(m1 >> 32), lo1 = (uint32_t)m1;
    uint32_t n0 = hi1 ^ c1 ^ a, n1 = lo1, n2 = hi0 ^ c3 ^ b, n3 = lo0;
    c0 = n0; c1 = n1; c2 = n2; c3 = n3;
    a += 0x9E3779B9u; b += 0xBB67AE85u;
  }
  uint64_t x = ((uint64_t)c0 << 32) | c1, y = ((uint64_t)c2 << 32) | c3;
  u0 = (double)(x >> 11) * 0x1.0p-53;
  u1 = (double)(y >> 11) * 0x1.0p-53;
}
static double host_threshold(int method, double n_total, double u_raw, uint64_t seed, uint64_t tag, int64_t j) {
  if (method == FC_RESAMPLE_SYSTEMATIC) {
    volatile double a = u_raw / n_total, b = (double)j / n_total;  // two roundings, then the add (no contraction)
    return a + b;
  }
  double u0, u1;
  host_philox_uniform2(seed, 0x8000000000000000ULL, (uint64_t)(j >> 1), tag, u0, u1);
  volatile double s = (double)j + ((j & 1) ? u1 : u0);
  return s / n_total;
}
// #{ j in [0, N) : threshold_j <= x } (thresholds are non-decreasing in j)
static int64_t count_thresholds_le(int method, int64_t N, double u_raw, uint64_t seed, uint64_t tag, double x) {
  int64_t lo = 0, hi = N;
  while (lo < hi) {
    int64_t mid = (lo + hi) / 2;
    if (host_threshold(method, (double)N, u_raw, seed, tag, mid) <= x) lo = mid + 1; else hi = mid;
  }
  return lo;
}

// Global resample of the rank-major population (smc.rs:255-292, 326-349, 534-541). Weights in S->w are
// normalised by the GLOBAL log-sum-exp. h_uniforms: explicit uniforms (1 raw value for systematic; n
// for stratified / multinomial, single rank only) or NULL = the session's Philox stream.
// h_idx (n entries, optional): GLOBAL ancestor index of every local output slot.
static int32_t smc_resample(fc_smc_session* S, int method, const double* h_uniforms, uint64_t* h_idx) {
  fc_program* P = S->P;
  cudaStream_t st = P->stream;
  const int64_t n = S->n, N = S->n_total;
  const int W = S->world;
  if (method < FC_RESAMPLE_MULTINOMIAL || method > FC_RESAMPLE_STRATIFIED) { set_error("bad resampling method %d", method); return FC_ERR_INVALID_ARGUMENT; }
  if (W > 1 && method == FC_RESAMPLE_MULTINOMIAL) { set_error("multinomial resampling is not available on a sharded population (its ancestors are not monotone in the slot index); use Systematic or Stratified"); return FC_ERR_UNSUPPORTED_OP; }
  if (W > 1 && method == FC_RESAMPLE_STRATIFIED && h_uniforms) { set_error("explicit stratified uniforms are not available on a sharded population"); return FC_ERR_UNSUPPORTED_OP; }
  S->resamples += 1;
  const uint64_t tag = 0x1000000000000000ULL | S->resamples;
  const int from = S->cur, to = S->cur ^ 1;
  GatherArgs G;
  memset(&G, 0, sizeof(G));
  G.n_src = n; G.n_total = (double)N; G.method = method; G.seed = S->seed; G.tag = tag; G.src = S->pop(from);
  G.pitch = S->L.pitch; G.ll_off = (P->d_cont + 1) * 8; G.first_src = S->first;
  const bool want_idx = h_idx || S->record;
  double u_raw = 0.0;
  if (method == FC_RESAMPLE_MULTINOMIAL) {
    // single rank: ancestors by search over the exact running sums (fc_resample.cu), then the copy
    FC_CUDA(S->us.reserve(n + 1));
    if (h_uniforms) FC_CUDA(cudaMemcpyAsync(S->us.p, h_uniforms, n * 8, cudaMemcpyHostToDevice, st));
    else k_uniform_one<<<(unsigned)cdiv(cdiv(n, 2), 256), 256, 0, st>>>(S->us.p, S->seed, tag, n);
    FC_TRY(resample_dev(method, S->w.p, n, S->us.p, S->idx.p, *S->rw, st, &P->n_kernel_launches));
    G.S = nullptr; G.idx_in = S->idx.p; G.idx_out = nullptr; G.j0 = 0; G.cnt = n; G.dst_off = 0; G.dst = S->pop(to);
    S->last_jlo = 0; S->last_jhi = n;
    k_smcx_gather<<<(unsigned)cdiv(n, GA_THREADS), GA_THREADS, 0, st>>>(G);
    FC_KERNEL_CHECK();
    P->n_kernel_launches += 2;
    if (h_idx) FC_CUDA(cudaMemcpyAsync(h_idx, S->idx.p, n * 8, cudaMemcpyDeviceToHost, st));
  } else {
    if (method == FC_RESAMPLE_SYSTEMATIC) {
      if (h_uniforms) u_raw = h_uniforms[0];
      else { double u1; host_philox_uniform2(S->seed, 0x8000000000000000ULL, 0, tag, u_raw, u1); }  // element 0 of k_uniform_one's stream: identical on every rank
    } else if (h_uniforms) {
      FC_CUDA(S->us.reserve(n + 1));
      FC_CUDA(cudaMemcpyAsync(S->us.p, h_uniforms, n * 8, cudaMemcpyHostToDevice, st));
      G.us = S->us.p;
    }
    G.u_raw = u_raw;
    // exact running sums of this rank's weights from the carry-in of the ranks before it
    PhaseTimer t_scan(S, 2);
    FC_CUDA(S->work.reserve(scan_work_bytes(n)));
    ScanCtl hs;
    memset(&hs, 0, sizeof(hs));
    hs.w_mean = 1.0 / (double)N;  // the weights are normalised over the whole population
    FC_CUDA(cudaMemcpyAsync(S->work.p, &hs, sizeof(hs), cudaMemcpyHostToDevice, st));
    k_chain_wait<<<1, 1, 0, st>>>(S->cm, S->resamples, (ScanCtl*)S->work.p);
    FC_TRY(scan_enqueue_head(S->w.p, n, S->S.p, S->work.p, st, &P->n_kernel_launches));
    FC_TRY(scan_enqueue_rounds(S->w.p, n, S->S.p, S->work.p, S->rank == 0 ? scan_expected_rounds(n) : 2, st, &P->n_kernel_launches));  // a later rank: one round, two if its range crosses a binade (scan_finish adds more if ever needed)
    if (W > 1) k_chain_publish<<<1, 32, 0, st>>>(S->cm, S->resamples, (const ScanCtl*)S->work.p, n);  // publishes iff the scan is complete
    FC_TRY(scan_finish(S->w.p, n, S->S.p, S->work.p, st, &P->n_kernel_launches, nullptr));
    P->n_kernel_launches += 1;
    double ends[MB_MAXW] = {0.0};
    if (W > 1) {
      k_chain_publish<<<1, 32, 0, st>>>(S->cm, S->resamples, (const ScanCtl*)S->work.p, n);  // no-op if the first one already did
      k_chain_collect<<<1, 32, 0, st>>>(S->cm, S->resamples, S->ends.p);
      FC_KERNEL_CHECK();
      P->n_kernel_launches += 2;
      FC_CUDA(cudaMemcpyAsync(ends, S->ends.p, W * 8, cudaMemcpyDeviceToHost, st));
      FC_CUDA(cudaStreamSynchronize(st));
      FC_TRY(comm_check(S));
    }
    // slots whose ancestor lives here: carry_in < threshold_j <= end (the last rank also takes the slots
    // whose threshold is never reached); the u == 0 wrap slot (threshold_0 == 0 -> i = 0 -> n - 1) belongs to
    // the last rank's last particle
    const double carry_in = S->rank == 0 ? 0.0 : ends[S->rank - 1];
    auto cnt_le = [&](double x) { return count_thresholds_le(method, N, u_raw, S->seed, tag, x); };
    int64_t jlo, jhi;
    if (W == 1) { jlo = 0; jhi = N; }
    else {
      jlo = cnt_le(carry_in);
      jhi = S->rank == W - 1 ? N : cnt_le(ends[S->rank]);
      if (jhi < jlo) jhi = jlo;
    }
    t_scan.stop();
    PhaseTimer t_ga(S, 3);
    G.S = S->S.p;
    S->last_jlo = jlo; S->last_jhi = jhi;
    if (want_idx) FC_CUDA(S->idx.reserve(std::max<int64_t>(n, jhi - jlo)));
    for (int t = 0; t < W; t++) {
      const int64_t a = std::max(jlo, S->bounds[t]), b = std::min(jhi, S->bounds[t + 1]);
      if (b <= a) continue;
      G.j0 = a; G.cnt = b - a; G.dst_off = a - S->bounds[t]; G.dst = S->peer_pop(t, to);
      G.idx_out = want_idx ? S->idx.p : nullptr;
      G.idx_off = a - jlo;
      G.wrap0 = 0;
      k_smcx_gather<<<(unsigned)cdiv(G.cnt, GA_THREADS), GA_THREADS, 0, st>>>(G);
      P->n_kernel_launches += 1;
      if (t != S->rank) S->migrated += (uint64_t)(b - a);
    }
    if (W > 1 && S->rank == W - 1 && jlo > 0 && count_thresholds_le(method, N, u_raw, S->seed, tag, 0.0) > 0) {
      // threshold_0 == 0 exactly: slot 0 wraps to the very last particle of the population
      G.j0 = 0; G.cnt = 1; G.dst_off = 0; G.dst = S->peer_pop(0, to); G.wrap0 = 1; G.idx_out = nullptr;
      k_smcx_gather<<<1, GA_THREADS, 0, st>>>(G);
      P->n_kernel_launches += 1;
    }
    FC_KERNEL_CHECK();
    if (W > 1) {
      // every rank's offspring have landed in every rank's new buffers
      k_comm_barrier<<<1, 32, 0, st>>>(S->cm);
      P->n_kernel_launches += 1;
    }
    if (h_idx && W == 1) FC_CUDA(cudaMemcpyAsync(h_idx, S->idx.p, n * 8, cudaMemcpyDeviceToHost, st));
  }
  S->cur = to;
  S->soa_valid = false;
  S->bis_ll_valid = false;
  S->log_w_uniform = true;  // weight 1/n, log_weight ln(1/n) (smc.rs:338-345); materialised on demand
  return FC_OK;
}

// The sharded resample's device logic on ONE GPU without any waiting kernel: the session's population is cut
// into `vw` virtual shards that are processed one after the other exactly as real ranks would -- exact scan of
// the shard continuing from the previous shard's end value (head skipped, window from the mean-weight hint),
// host slot routing from the (carry, end) pairs, gather of every (source shard, destination shard) block with
// its record / index offsets. Only the mailbox transport is not exercised (ranks that spin on each other must
// not share a GPU). h_idx: the n global ancestors. Systematic (u_raw) or stratified (session stream) thresholds.
static int32_t smc_resample_virtual(fc_smc_session* S, int vw, int method, double u_raw, uint64_t* h_idx) {
  fc_program* P = S->P;
  cudaStream_t st = P->stream;
  const int64_t n = S->n;
  if (S->world != 1 || vw < 1 || vw > MB_MAXW || (int64_t)vw > n) { set_error("virtual shards: need a single-rank session and 1 <= shards <= 8"); return FC_ERR_INVALID_ARGUMENT; }
  if (method != FC_RESAMPLE_SYSTEMATIC && method != FC_RESAMPLE_STRATIFIED) { set_error("virtual shards: systematic or stratified"); return FC_ERR_INVALID_ARGUMENT; }
  S->resamples += 1;
  const uint64_t tag = 0x1000000000000000ULL | S->resamples;
  const int from = S->cur, to = S->cur ^ 1;
  std::vector<int64_t> b(vw + 1);
  for (int r = 0; r <= vw; r++) b[r] = (n / vw) * r + std::min<int64_t>(r, n % vw);
  // every shard's running sums live in their own (n_r + 2)-slice of one buffer
  DevBuf<double> Sall;
  FC_CUDA(Sall.reserve(n + 2 * vw + 2));
  std::vector<double*> Sr(vw);
  std::vector<double> ends(vw);
  double carry = 0.0;
  for (int r = 0; r < vw; r++) {
    const int64_t nr = b[r + 1] - b[r];
    Sr[r] = Sall.p + b[r] + 2 * r;
    FC_CUDA(S->work.reserve(scan_work_bytes(nr)));
    ScanCtl hs;
    memset(&hs, 0, sizeof(hs));
    hs.s0 = carry;
    hs.w_mean = 1.0 / (double)n;
    FC_CUDA(cudaMemcpyAsync(S->work.p, &hs, sizeof(hs), cudaMemcpyHostToDevice, st));
    FC_TRY(scan_enqueue_head(S->w.p + b[r], nr, Sr[r], S->work.p, st, &P->n_kernel_launches));
    FC_TRY(scan_enqueue_rounds(S->w.p + b[r], nr, Sr[r], S->work.p, r == 0 ? scan_expected_rounds(nr) : 3, st, &P->n_kernel_launches));
    FC_TRY(scan_finish(S->w.p + b[r], nr, Sr[r], S->work.p, st, &P->n_kernel_launches, &ends[r]));
    carry = ends[r];
  }
  FC_CUDA(S->idx.reserve(n));
  GatherArgs G;
  memset(&G, 0, sizeof(G));
  G.n_total = (double)n; G.method = method; G.seed = S->seed; G.tag = tag; G.u_raw = u_raw;
  G.pitch = S->L.pitch; G.ll_off = (P->d_cont + 1) * 8; G.idx_out = S->idx.p;
  Pop pf = S->pop(from), pt = S->pop(to);
  for (int r = 0; r < vw; r++) {
    const int64_t nr = b[r + 1] - b[r];
    const double cin = r == 0 ? 0.0 : ends[r - 1];
    int64_t jlo = count_thresholds_le(method, n, u_raw, S->seed, tag, cin);
    int64_t jhi = r == vw - 1 ? n : count_thresholds_le(method, n, u_raw, S->seed, tag, ends[r]);
    if (jhi < jlo) jhi = jlo;
    G.S = Sr[r]; G.n_src = nr; G.first_src = b[r];
    G.src = pf; G.src.rec = pf.rec + b[r] * S->L.pitch; G.src.ll = pf.ll + b[r]; G.src.n = nr;
    for (int t = 0; t < vw; t++) {
      const int64_t a = std::max(jlo, b[t]), e = std::min(jhi, b[t + 1]);
      if (e <= a) continue;
      G.j0 = a; G.cnt = e - a; G.dst_off = a - b[t]; G.idx_off = a; G.wrap0 = 0;
      G.dst = pt; G.dst.rec = pt.rec + b[t] * S->L.pitch; G.dst.ll = pt.ll + b[t]; G.dst.n = b[t + 1] - b[t];
      k_smcx_gather<<<(unsigned)cdiv(G.cnt, GA_THREADS), GA_THREADS, 0, st>>>(G);
      P->n_kernel_launches += 1;
    }
    if (r == vw - 1 && jlo > 0 && vw > 1 && count_thresholds_le(method, n, u_raw, S->seed, tag, 0.0) > 0) {  // the u == 0 wrap slot
      G.j0 = 0; G.cnt = 1; G.dst_off = 0; G.idx_off = 0; G.wrap0 = 1;
      G.dst = pt; G.dst.n = b[1];
      k_smcx_gather<<<1, GA_THREADS, 0, st>>>(G);
      P->n_kernel_launches += 1;
    }
  }
  FC_KERNEL_CHECK();
  if (h_idx) FC_CUDA(cudaMemcpyAsync(h_idx, S->idx.p, n * 8, cudaMemcpyDeviceToHost, st));
  FC_CUDA(cudaStreamSynchronize(st));
  S->cur = to;
  S->soa_valid = false;
  S->bis_ll_valid = false;
  S->log_w_uniform = true;
  return FC_OK;
}

// `steps` pi_beta-invariant single-site moves per particle (smc.rs:544-554, 698-713)
static int32_t smc_rejuvenate(fc_smc_session* S, double beta, uint64_t steps) {
  fc_program* P = S->P;
  cudaStream_t st = P->stream;
  const int64_t n = S->n;
  Pop pop = S->pop(S->cur);
  if (steps) S->bis_ll_valid = false;
  if (S->fused) {
    for (uint64_t r = 0; r < steps; r++) {
      MoveX M;
      M.scale = S->eng.scale.p; M.acc_cnt = S->eng.acc_cnt.p; M.tot_cnt = S->eng.tot_cnt.p;
      M.seed = S->eng.seed; M.chain0 = S->eng.chain0; M.step_tag = S->moves; M.beta = beta; M.adapt = 1;
      if (S->K.on) k_smcx_move<true><<<(unsigned)cdiv(n, 128), 128, 0, st>>>(S->F, S->K, pop, M);
      else k_smcx_move<false><<<(unsigned)cdiv(n, 128), 128, 0, st>>>(S->F, S->K, pop, M);
      FC_KERNEL_CHECK();
      P->n_kernel_launches += 1;
      FC_TRY(S->eng.shared_adapt(0.44, 0.7));
      S->moves += 1;
    }
    if (steps) S->soa_valid = false;
    return FC_OK;
  }
  // generic scoring engine on [row][particle] copies; the records are rewritten afterwards
  if (steps == 0) return FC_OK;
  FC_TRY(ensure_soa(S));
  for (uint64_t r = 0; r < steps; r++) {
    FC_TRY(S->eng.step(S->q_soa.p, S->qd32.p, S->lp3_soa.p, S->total.p, S->n_accept.p, S->moves, beta, true, 0.44, 0.7));
    FC_TRY(S->eng.shared_adapt(0.44, 0.7));
    S->moves += 1;
  }
  return soa_to_records(S, false);
}

// one tempering step (the body of the while loop, smc.rs:502-559)
static int32_t smc_temper_step(fc_smc_session* S) {
  fc_program* P = S->P;
  const double N = (double)S->n_total;
  const double target_ess = std::fmin(std::fmax(S->cfg.ess_threshold * N, 1.0), N);
  S->steps += 1;
  double beta_new;
  PhaseTimer t_nb(S, 0);
  if (S->log_w_uniform) {
    const double guess = S->prev_dbeta > 0.0 ? S->beta + 4.0 * S->prev_dbeta : 0.0;
    FC_TRY(next_beta_uniform(P->stream, S->cm, lwc_of(S), S->ll(), S->n, N, S->beta, target_ess, guess, S->bis.p, S->part.p, &P->n_kernel_launches, &S->ess_evals, &beta_new));
    FC_TRY(comm_check(S));
    S->bis_ll_valid = true;
  } else {
    if (S->world > 1) { set_error("next_beta on non-uniform weights is not available on a sharded population"); return FC_ERR_UNSUPPORTED_OP; }
    FC_TRY(next_beta_general(P->stream, S->log_w.p, S->ll(), S->n, S->beta, target_ess, S->bis.p, S->tri.p, &P->n_kernel_launches, &beta_new));
  }
  t_nb.stop();
  if (S->steps >= 10000) beta_new = 1.0;
  const double d_beta = beta_new - S->beta;
  double log_norm;
  {
    PhaseTimer t_rw(S, 1);
    FC_TRY(smc_reweight(S, d_beta, beta_new >= 1.0, &log_norm));
  }
  S->log_evidence += log_norm;
  S->prev_dbeta = d_beta;
  S->beta = beta_new;
  if (S->beta < 1.0) {
    FC_TRY(smc_resample(S, S->cfg.resampling_method, nullptr, nullptr));  // smc.rs:534-541
    PhaseTimer t_mv(S, 4);
    FC_TRY(smc_rejuvenate(S, S->beta, S->cfg.rejuvenation_steps));       // smc.rs:544-554
  }
  return FC_OK;
}

// which plates can the fused evaluator take? (the shapes of fc_plate_vec.cu, one particle per thread)
static bool fused_lower(fc_program* P, std::vector<FPlate>& out, int& zw) {
  const bool disabled = smc_generic();
  zw = 1;
  for (const DOp& op : P->all_ops)
    if (op.is_sample && !op.cont) {
      const bool small = op.dist == FC_BERNOULLI || (op.dist == FC_CATEGORICAL && (op.n_probs == 0 || op.n_probs <= 256));
      if (!small) zw = 4;
    }
  if (disabled || zw != 1) return false;
  for (size_t oi = 0; oi < P->all_ops.size(); oi++) {
    const DOp& op = P->all_ops[oi];
    const bool plate = op.kind == FC_OP_PLATE_SAMPLE || op.kind == FC_OP_PLATE_OBSERVE;
    if (!plate) continue;
    if (op.plate_len >= (1LL << 30)) return false;
    FPlate fp;
    memset(&fp, 0, sizeof(fp));
    fp.op = (int32_t)oi; fp.op_b = -1; fp.n = (int32_t)op.plate_len;
    const bool twoval = (op.dist == FC_BERNOULLI || (op.dist == FC_CATEGORICAL && op.n_probs == 0)) && op.n_params == 1 && op.p[0].invariant;
    if (twoval) {
      for (int t = 0; t < 3; t++) {
        const int k = op.p[0].t[t].kind;
        if (!(k == FC_TERM_NONE || k == FC_TERM_CONST || k == FC_TERM_LATENT)) return false;
      }
      fp.kind = op.is_sample ? FP_TWOVAL_SAMPLE : FP_TWOVAL_OBS;
      out.push_back(fp);
      continue;
    }
    if (op.dist != FC_NORMAL || op.is_sample || op.n_params != 2) return false;
    int m[2] = {FM_NONE, FM_NONE}, mt[2] = {0, 0}, nm = 0;
    bool ss = false;
    int st = 0;
    for (int j = 0; j < 2; j++)
      for (int t = 0; t < 3; t++) {
        const DTerm& tm = op.p[j].t[t];
        const int k = tm.kind;
        if (k == FC_TERM_NONE || k == FC_TERM_CONST || k == FC_TERM_LATENT) continue;
        if (op.p[j].link != FC_LINK_IDENTITY) return false;
        if (j == 0 && k == FC_TERM_SELECT2) { if (nm == 2) return false; m[nm] = FM_SELECT2; mt[nm] = t; nm++; }
        else if (j == 0 && k == FC_TERM_LATENT_X && tm.col.dtype == FC_COL_F64) { if (nm == 2) return false; m[nm] = FM_LATENT_X; mt[nm] = t; nm++; }
        else if (j == 1 && k == FC_TERM_SELECT2 && !ss) { ss = true; st = t; }
        else return false;
      }
    if (m[1] == FM_SELECT2) {  // canonical order: a SELECT2 mean term first
      if (m[0] == FM_SELECT2) return false;
      std::swap(m[0], m[1]); std::swap(mt[0], mt[1]);
    }
    fp.kind = FP_NORMAL_OBS; fp.m0 = m[0]; fp.m1 = m[1]; fp.ss = ss ? 1 : 0; fp.mt0 = mt[0]; fp.mt1 = mt[1]; fp.st = st;
    // one pass over z when this plate selects on the discrete plate lowered just before it
    if (!out.empty() && out.back().kind == FP_TWOVAL_SAMPLE && fp.m0 == FM_SELECT2 && out.back().n == fp.n) {
      const DOp& zp_op = P->all_ops[out.back().op];
      const bool same_m = op.p[0].t[fp.mt0].dslot == (int32_t)zp_op.dslot0;
      const bool same_s = !ss || op.p[1].t[fp.st].dslot == (int32_t)zp_op.dslot0;
      if (same_m && same_s) {
        FPlate& z = out.back();
        z.kind = FP_MIX; z.op_b = (int32_t)oi; z.m0 = fp.m0; z.m1 = fp.m1; z.ss = fp.ss; z.mt0 = fp.mt0; z.mt1 = fp.mt1; z.st = fp.st;
        // the first such plate whose mean has no other element-dependent term keeps its statistics in the record
        bool have = false;
        for (const FPlate& o : out) have |= o.stat != 0;
        z.stat = (!have && smc_stats() && fp.m1 == FM_NONE && op.value_col.dtype == FC_COL_F64) ? 1 : 0;
        continue;
      }
    }
    out.push_back(fp);
  }
  return true;
}

static int32_t smc_init_population(fc_smc_session* S) {
  fc_program* p = S->P;
  cudaStream_t st = p->stream;
  const int64_t N = S->n;
  Pop pop = S->pop(S->cur);
  const uint64_t tag = 0x7000000000000000ULL;
  if (S->fused) {
    k_smcx_prior<<<(unsigned)cdiv(N, 128), 128, 0, st>>>(S->F, pop, S->seed, S->stream0, tag);
    FC_KERNEL_CHECK();
    p->n_kernel_launches += 1;
    return FC_OK;
  }
  FC_CUDA(S->q_soa.reserve(std::max<int64_t>(1, p->d_cont * N)));
  FC_CUDA(S->lp3_soa.reserve(3 * N));
  FC_CUDA(S->qd32.reserve(std::max<int64_t>(1, p->d_disc * N)));
  FC_TRY(prior_sample_dev(p, S->q_soa.p, S->qd32.p, N, S->seed, S->stream0, tag));
  FC_TRY(score_dev(p, S->q_soa.p, S->qd32.p, N, S->lp3_soa.p, S->total.p, nullptr));
  FC_TRY(soa_to_records(S, true));
  S->soa_valid = true;
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
  for (int r = 0; r < MB_MAXW; r++)
    if (s->peer_slab[r]) cudaIpcCloseMemHandle(s->peer_slab[r]);
  if (s->slab) cudaFree(s->slab);
  if (s->rw) fc::resample_work_free(s->rw);
  delete s;
  return FC_OK;
}

// smc_prior_particles (smc.rs:764-790) + the state adaptive_smc keeps (smc.rs:470-482) for the shard
// [first_particle, first_particle + n) of a population of n_total particles held by `world` ranks.
// stream0: particle j of this session uses the random stream of global particle stream0 + j.
static int32_t smc_create_impl(fc_program* p, const fc_smc_config* cfg, uint64_t seed, size_t n, uint64_t stream0, uint64_t first_particle, uint64_t n_total, int32_t rank, int32_t world, fc_smc_session** out) {
  using namespace fc;
  if (!p || !cfg || !out) { set_error("NULL argument"); return FC_ERR_INVALID_ARGUMENT; }
  *out = nullptr;
  if (n == 0) { set_error("num_particles must be > 0"); return FC_ERR_INVALID_ARGUMENT; }
  if (world < 1 || world > MB_MAXW || rank < 0 || rank >= world || n_total < n || first_particle + n > n_total) { set_error("bad shard arguments (rank %d of %d, %zu of %llu particles from %llu)", rank, world, n, (unsigned long long)n_total, (unsigned long long)first_particle); return FC_ERR_INVALID_ARGUMENT; }
  FC_CUDA(cudaSetDevice(p->device));
  std::unique_ptr<fc_smc_session> S(new fc_smc_session());
  S->P = p; S->cfg = *cfg; S->seed = seed; S->n = (int64_t)n; S->n_total = (int64_t)n_total; S->first = (int64_t)first_particle; S->stream0 = stream0; S->rank = rank; S->world = world;
  const int64_t N = S->n;
  cudaStream_t st = p->stream;
  std::vector<FPlate> fpl;
  int zw = 1;
  S->fused = fused_lower(p, fpl, zw);
  S->L.d_cont = p->d_cont; S->L.d_disc = p->d_disc; S->L.zw = zw; S->L.soff = 0;
  int64_t head = (p->d_cont + 3) * 8;
  if (S->fused)
    for (const FPlate& f : fpl)
      if (f.stat) { S->L.soff = (int32_t)head; head += 5 * 8; }
  // one bit per discrete site when the record carries the mixture statistics (the moves then never read z) and
  // every discrete sample site is two-valued: C4's record shrinks from 192 to 128 bytes
  if (S->fused && S->L.soff && smc_zbits()) {
    bool two = true;
    for (const DOp& op : p->all_ops)
      if (op.is_sample && !op.cont) two &= op.dist == FC_BERNOULLI || (op.dist == FC_CATEGORICAL && op.n_probs == 0);
    if (two) zw = 0;
  }
  S->L.zw = zw;
  S->L.zoff = ((head + 15) / 16) * 16;
  S->L.pitch = S->L.zoff + (((zw == 0 ? (p->d_disc + 7) / 8 : p->d_disc * zw) + 15) / 16) * 16;
  S->slab_bytes = fc_smc_session::layout(N, S->L.pitch, S->off_rec, S->off_ll);
  FC_CUDA(cudaMalloc((void**)&S->slab, S->slab_bytes));
  FC_CUDA(cudaMemsetAsync(S->slab, 0, S->off_rec[0], st));  // the mailbox
  FC_CUDA(S->seq.reserve(1)); FC_CUDA(S->err.reserve(2)); FC_CUDA(S->counter.reserve(1));
  FC_CUDA(cudaMemsetAsync(S->seq.p, 0, 8, st)); FC_CUDA(cudaMemsetAsync(S->err.p, 0, 8, st)); FC_CUDA(cudaMemsetAsync(S->counter.p, 0, 4, st));
  memset(&S->cm, 0, sizeof(S->cm));
  S->cm.rank = rank; S->cm.world = world; S->cm.seq = S->seq.p; S->cm.err = S->err.p;
  S->cm.box[rank] = (MBox*)S->slab;
  S->peer_n[rank] = N;
  if (world == 1) { S->bounds[0] = 0; S->bounds[1] = N; S->connected = true; }
  for (auto* b : {&S->log_w, &S->comb, &S->w, &S->total}) FC_CUDA(b->reserve(N));
  FC_CUDA(S->S.reserve(N + 2));
  FC_CUDA(S->part.reserve(4 * RDB + 8)); FC_CUDA(S->scal.reserve(8)); FC_CUDA(S->ends.reserve(MB_MAXW));
  FC_CUDA(S->idx.reserve(N)); FC_CUDA(S->n_accept.reserve(N)); FC_CUDA(S->bis.reserve(1)); FC_CUDA(S->tri.reserve(RDB + 1));
  FC_CUDA(cudaMemsetAsync(S->n_accept.p, 0, N * 8, st));
  S->rw = resample_work_new();
  FC_TRY(S->eng.init(p, N, true, seed ^ 0x5bd1e995u, stream0));
  if (S->fused) {
    FC_CUDA(S->fplates.reserve(std::max<size_t>(1, fpl.size())));
    if (!fpl.empty()) FC_CUDA(cudaMemcpyAsync(S->fplates.p, fpl.data(), fpl.size() * sizeof(FPlate), cudaMemcpyHostToDevice, st));
    FC_CUDA(cudaStreamSynchronize(st));
    // compile the scalar ops
    std::vector<SOp> so(p->scalar_ops.size());
    for (size_t i = 0; i < so.size(); i++) {
      const DOp& op = p->scalar_ops[i];
      SOp& o = so[i];
      memset(&o, 0, sizeof(o));
      o.kind = SO_GENERIC;
      if (op.kind == FC_OP_FACTOR) { o.kind = SO_FACTOR; o.c0 = op.value; continue; }
      if (op.kind != FC_OP_SAMPLE || !op.cont) continue;
      auto cparam = [&](int j, double& v) -> bool {  // a constant parameter with the identity link
        if (j >= op.n_params || op.p[j].link != FC_LINK_IDENTITY) return false;
        double acc = 0.0;
        bool first = true;
        for (int t = 0; t < 3; t++) {
          if (op.p[j].t[t].kind == FC_TERM_NONE) continue;
          if (op.p[j].t[t].kind != FC_TERM_CONST) return false;
          acc = first ? op.p[j].t[t].c : acc + op.p[j].t[t].c;
          first = false;
        }
        v = acc;
        return !first && std::isfinite(acc);
      };
      double a, b;
      o.row = p->crow[op.cslot0];
      if (op.dist == FC_NORMAL && cparam(0, a) && cparam(1, b) && b > 0.0) { o.kind = SO_NORMAL_CC; o.c0 = a; o.c1 = 1.0 / b; o.c2 = -std::log(b) - 0.5 * FC_LN_2PI; }
      else if (op.dist == FC_GAMMA && cparam(0, a) && cparam(1, b) && a > 0.0 && b > 0.0) { o.kind = SO_GAMMA_CC; o.c0 = a - 1.0; o.c1 = b; o.c2 = a * std::log(b) - std::lgamma(a); }
      else if (op.dist == FC_BETA && cparam(0, a) && cparam(1, b) && a > 0.0 && b > 0.0) { o.kind = SO_BETA_CC; o.c0 = a - 1.0; o.c1 = b - 1.0; o.c2 = -(std::lgamma(a) + std::lgamma(b) - std::lgamma(a + b)); }
    }
    FC_CUDA(S->sops.reserve(std::max<size_t>(1, so.size())));
    if (!so.empty()) FC_CUDA(cudaMemcpy(S->sops.p, so.data(), so.size() * sizeof(SOp), cudaMemcpyHostToDevice));
    FusedProg& F = S->F;
    F.sops = S->sops.p;
    F.ops = p->d_all_ops.p; F.scalar_ops = p->d_scalar_ops.p; F.plates = S->fplates.p; F.crow = p->d_crow.p; F.nodes = p->d_expr_nodes.p;
    F.n_ops = (int32_t)p->all_ops.size(); F.n_scalar = (int32_t)p->scalar_ops.size(); F.n_plates = (int32_t)fpl.size(); F._pad = 0; F.L = S->L;
    // by-value program of the move kernel: every scalar op compiled, exactly one plate, a mixture with statistics
    CompiledMix& K = S->K;
    memset(&K, 0, sizeof(K));
    bool ok = smc_compiled() && S->L.soff && fpl.size() == 1 && fpl[0].kind == FP_MIX && fpl[0].stat && so.size() <= CM_MAX_SOP;
    for (const SOp& o : so) ok = ok && o.kind != SO_GENERIC;
    if (ok) {
      const FPlate& fp = fpl[0];
      const DOp& zp = p->all_ops[fp.op];
      const DOp& ob = p->all_ops[fp.op_b];
      auto cp = [&](const DOp& op, int j, CParam& out) {
        for (int t = 0; t < 3; t++) {
          const DTerm& tm = op.p[j].t[t];
          out.c[t] = tm.c;
          out.row[t] = tm.kind == FC_TERM_CONST ? -1 : (tm.kind == FC_TERM_LATENT ? (int32_t)tm.row : -2);
        }
        out.link = op.p[j].link; out.apply = op.p[j].invariant ? 1 : 0; out.lo = op.p[j].lo; out.hi = op.p[j].hi;
      };
      K.n_sop = (int32_t)so.size();
      for (size_t i = 0; i < so.size(); i++) K.sop[i] = so[i];
      K.dist = zp.dist; K.n = fp.n; K.m0 = fp.m0 == FM_SELECT2 ? 1 : 0; K.ss = fp.ss; K.soff = S->L.soff;
      cp(zp, 0, K.p0); cp(ob, 0, K.mean); cp(ob, 1, K.sigma);
      if (K.m0) { K.row_ma = (int32_t)ob.p[0].t[fp.mt0].row; K.row_mb = (int32_t)ob.p[0].t[fp.mt0].row_b; }
      if (K.ss) { K.row_sa = (int32_t)ob.p[1].t[fp.st].row; K.row_sb = (int32_t)ob.p[1].t[fp.st].row_b; }
      K.hot_chunks = (S->L.soff + 40 + 15) / 16;
      K.on = (fp.m0 != FM_LATENT_X && K.hot_chunks <= CM_HOT_MAX) ? 1 : 0;
    }
  }
  FC_TRY(smc_init_population(S.get()));
  FC_CUDA(cudaStreamSynchronize(st));
  if (!S->fused && p->d_disc) {
    int e = 0;
    FC_CUDA(cudaMemcpy(&e, S->err.p + 1, 4, cudaMemcpyDeviceToHost));
    if (e) { set_error("discrete prior draw outside the byte range of the packed particle record"); return FC_ERR_UNSUPPORTED_OP; }
  }
  *out = S.release();
  return FC_OK;
}
int32_t fc_smc_create_sharded(fc_program* p, const fc_smc_config* cfg, uint64_t seed, size_t n, uint64_t first_particle, uint64_t n_total, int32_t rank, int32_t world, fc_smc_session** out) {
  return smc_create_impl(p, cfg, seed, n, first_particle, first_particle, n_total, rank, world, out);
}
int32_t fc_smc_create(fc_program* p, const fc_smc_config* cfg, uint64_t seed, size_t n, uint64_t first_particle, fc_smc_session** out) {
  return smc_create_impl(p, cfg, seed, n, first_particle, 0, n, 0, 1, out);
}

// the 128-byte descriptor of this rank's slab; gather every rank's (rank order) and pass them to fc_smc_connect
int32_t fc_smc_export(fc_smc_session* s, void* handle) {
  using namespace fc;
  if (!s || !handle) { set_error("NULL argument"); return FC_ERR_INVALID_ARGUMENT; }
  SmcHandle h;
  memset(&h, 0, sizeof(h));
  FC_CUDA(cudaSetDevice(s->P->device));
  FC_CUDA(cudaIpcGetMemHandle(&h.ipc, s->slab));
  h.magic = FC_SMC_HANDLE_MAGIC; h.rank = s->rank; h.device = s->P->device; h.n_local = s->n; h.first = s->first; h.pid = (int64_t)getpid();
  memcpy(handle, &h, sizeof(h));
  return FC_OK;
}
int32_t fc_smc_connect(fc_smc_session* s, const void* handles) {
  using namespace fc;
  if (!s || !handles) { set_error("NULL argument"); return FC_ERR_INVALID_ARGUMENT; }
  FC_CUDA(cudaSetDevice(s->P->device));
  const SmcHandle* H = (const SmcHandle*)handles;
  int64_t acc = 0;
  for (int r = 0; r < s->world; r++) {
    if (H[r].magic != FC_SMC_HANDLE_MAGIC || H[r].rank != r) { set_error("handle %d is not rank %d's export", r, r); return FC_ERR_INVALID_ARGUMENT; }
    if (H[r].first != acc) { set_error("shards are not contiguous in rank order (rank %d starts at %lld, expected %lld)", r, (long long)H[r].first, (long long)acc); return FC_ERR_INVALID_ARGUMENT; }
    s->bounds[r] = acc;
    s->peer_n[r] = H[r].n_local;
    acc += H[r].n_local;
    if (r == s->rank) continue;
    if (H[r].pid == (int64_t)getpid()) { set_error("ranks of a sharded population must be separate processes (cudaIpc)"); return FC_ERR_UNSUPPORTED_OP; }
    void* ptr = nullptr;
    cudaError_t e = cudaIpcOpenMemHandle(&ptr, H[r].ipc, cudaIpcMemLazyEnablePeerAccess);
    if (e != cudaSuccess) { set_error("cudaIpcOpenMemHandle(rank %d) failed: %s", r, cudaGetErrorString(e)); return FC_ERR_NCCL; }
    s->peer_slab[r] = (char*)ptr;
    s->cm.box[r] = (MBox*)ptr;
  }
  s->bounds[s->world] = acc;
  if (acc != s->n_total) { set_error("shards hold %lld particles, population is %lld", (long long)acc, (long long)s->n_total); return FC_ERR_INVALID_ARGUMENT; }
  s->connected = true;
  return FC_OK;
}

static int32_t smc_ready(fc_smc_session* s) {
  if (!s) { fc::set_error("session is NULL"); return FC_ERR_INVALID_ARGUMENT; }
  if (!s->connected) { fc::set_error("sharded session: call fc_smc_connect first"); return FC_ERR_INVALID_ARGUMENT; }
  FC_CUDA(cudaSetDevice(s->P->device));
  return FC_OK;
}

// One step of the tempering ladder (no-op once beta == 1). With rejuvenation_steps == 0 the single
// "step" is the pure importance-sampling reweight (smc.rs:484-493).
int32_t fc_smc_step(fc_smc_session* s, double* beta_out) {
  using namespace fc;
  FC_TRY(smc_ready(s));
  if (!s->finished) {
    if (s->cfg.rejuvenation_steps == 0) {
      cudaStream_t st = s->P->stream;
      unsigned nbk = (unsigned)cdiv(s->n, 256);
      FC_TRY(smc_materialise_log_w(s));
      k_combine<<<nbk, 256, 0, st>>>(s->log_w.p, 0.0, s->ll(), 1.0, s->comb.p, s->n);
      FC_KERNEL_CHECK();
      s->P->n_kernel_launches += 1;
      double le;
      FC_TRY(global_lse(s, s->comb.p, &le));
      s->log_evidence = le;
      s->beta = 1.0;
      std::swap(s->log_w.p, s->comb.p); std::swap(s->log_w.n, s->comb.n);
      s->log_w_uniform = false;
      s->steps = 1;
    } else {
      FC_TRY(smc_temper_step(s));
    }
    if (s->beta >= 1.0) s->finished = true;
    if (!s->phase_events.empty()) { FC_CUDA(cudaStreamSynchronize(s->P->stream)); phase_collect(s); }
  }
  if (beta_out) *beta_out = s->beta;
  return FC_OK;
}

// final normalisation (smc.rs:565-575: log_sum_exp of log_w ALONE, over all ranks) and copy-out of this
// rank's shard. Any pointer may be NULL.
int32_t fc_smc_get(fc_smc_session* s, double* h_values, int64_t* h_dvalues, double* h_weight, double* h_log_weight, double* log_evidence, uint64_t* n_temper_steps) {
  using namespace fc;
  FC_TRY(smc_ready(s));
  fc_program* p = s->P;
  cudaStream_t st = p->stream;
  const int64_t n = s->n;
  unsigned nbk = (unsigned)cdiv(n, 256);
  Pop pop = s->pop(s->cur);
  if (h_weight || h_log_weight) {
    FC_TRY(smc_materialise_log_w(s));
    FC_TRY(global_lse(s, s->log_w.p, nullptr));
    k_sub_exp<<<nbk, 256, 0, st>>>(s->log_w.p, s->scal.p, (double)s->n_total, s->comb.p, s->w.p, n);
    FC_KERNEL_CHECK();
    p->n_kernel_launches += 1;
    if (h_weight) FC_CUDA(cudaMemcpyAsync(h_weight, s->w.p, n * 8, cudaMemcpyDeviceToHost, st));
    if (h_log_weight) FC_CUDA(cudaMemcpyAsync(h_log_weight, s->comb.p, n * 8, cudaMemcpyDeviceToHost, st));
  }
  if ((h_values && p->d_cont) || (h_dvalues && p->d_disc)) FC_TRY(ensure_soa(s));
  if (h_values && p->d_cont) FC_CUDA(cudaMemcpyAsync(h_values, s->q_soa.p, p->d_cont * n * 8, cudaMemcpyDeviceToHost, st));
  if (h_dvalues && p->d_disc) FC_TRY(disc_to_host(p, s->qd32.p, p->d_disc * n, h_dvalues));
  if (log_evidence) *log_evidence = s->log_evidence;
  if (n_temper_steps) *n_temper_steps = s->steps;
  FC_CUDA(cudaStreamSynchronize(st));
  return comm_check(s);
}

int32_t fc_smc_device_state(fc_smc_session* s, double** d_values, int32_t** d_dvalues, double** d_log_w, double** d_loglik) {
  using namespace fc;
  FC_TRY(smc_ready(s));
  // [row][particle] snapshots of the particle records (valid until the next step / resample / move)
  if (d_values || d_dvalues) FC_TRY(ensure_soa(s));
  if (d_values) *d_values = s->q_soa.p;
  if (d_dvalues) *d_dvalues = s->qd32.p;
  if (d_log_w) { FC_TRY(smc_materialise_log_w(s)); *d_log_w = s->log_w.p; }
  if (d_loglik) *d_loglik = s->ll();
  return FC_OK;
}

// ---- standalone pieces of the SMC public surface, on the particles resident in a session ----

// smc_prior_particles' weighting (smc.rs:779-787, FG-03): log_weight = log-likelihood of the particle
int32_t fc_smc_prior_weights(fc_smc_session* s) {
  using namespace fc;
  FC_TRY(smc_ready(s));
  FC_CUDA(cudaMemcpyAsync(s->log_w.p, s->ll(), s->n * 8, cudaMemcpyDeviceToDevice, s->P->stream));
  s->log_w_uniform = false;
  return FC_OK;
}

// normalize_particles (smc.rs:719-754): weight = exp(log_weight - lse) with the lse over ALL ranks,
// uniform if it is not finite; log_weight is left as it is. Returns ESS = 1 / sum w^2 (smc.rs:230-233)
// in *ess (may be NULL).
int32_t fc_smc_normalize(fc_smc_session* s, double* log_norm_out, double* ess) {
  using namespace fc;
  FC_TRY(smc_ready(s));
  fc_program* p = s->P;
  cudaStream_t st = p->stream;
  const int64_t n = s->n;
  unsigned nbk = (unsigned)cdiv(n, 256);
  FC_TRY(smc_materialise_log_w(s));
  FC_TRY(global_lse(s, s->log_w.p, nullptr));
  k_sub_exp<<<nbk, 256, 0, st>>>(s->log_w.p, s->scal.p, (double)s->n_total, nullptr, s->w.p, n);
  if (ess) k_gsumsq<<<nb_blocks(n), NB_THREADS, 0, st>>>(s->w.p, n, s->part.p, s->counter.p, s->scal.p, s->cm);
  FC_KERNEL_CHECK();
  p->n_kernel_launches += ess ? 2 : 1;
  double h[3];
  FC_CUDA(cudaMemcpyAsync(h, s->scal.p, 24, cudaMemcpyDeviceToHost, st));
  FC_CUDA(cudaStreamSynchronize(st));
  FC_TRY(comm_check(s));
  if (log_norm_out) *log_norm_out = h[0];
  if (ess) *ess = h[2];
  return FC_OK;
}

// one reweight of the tempering loop (smc.rs:507-531): log_w += (beta_new - beta) * loglik, normalised by
// the global log-sum-exp, which is added to the evidence estimate; beta := beta_new
int32_t fc_smc_reweight(fc_smc_session* s, double beta_new, double* log_norm_out) {
  using namespace fc;
  FC_TRY(smc_ready(s));
  if (!(beta_new >= s->beta && beta_new <= 1.0)) { set_error("beta_new must be in [beta, 1], got %g (beta = %g)", beta_new, s->beta); return FC_ERR_INVALID_ARGUMENT; }
  double log_norm;
  FC_TRY(smc_reweight(s, beta_new - s->beta, true, &log_norm));
  s->log_evidence += log_norm;
  s->beta = beta_new;
  s->steps += 1;
  if (log_norm_out) *log_norm_out = log_norm;
  return FC_OK;
}

// this rank's log weights as they are and the normalised weights the resampler would see; either may be NULL
int32_t fc_smc_weights(fc_smc_session* s, double* h_log_w, double* h_w) {
  using namespace fc;
  FC_TRY(smc_ready(s));
  cudaStream_t st = s->P->stream;
  FC_TRY(smc_materialise_log_w(s));
  if (h_log_w) FC_CUDA(cudaMemcpyAsync(h_log_w, s->log_w.p, s->n * 8, cudaMemcpyDeviceToHost, st));
  if (h_w) {
    FC_TRY(fc_smc_normalize(s, nullptr, nullptr));
    FC_CUDA(cudaMemcpyAsync(h_w, s->w.p, s->n * 8, cudaMemcpyDeviceToHost, st));
  }
  FC_CUDA(cudaStreamSynchronize(st));
  return FC_OK;
}

// resample_particles (smc.rs:326-349): indices by `method` from the (globally) normalised weights,
// ancestors copied -- across ranks when sharded --, weight = 1/n and log_weight = ln(1/n).
// h_uniforms (1 raw value for systematic, n otherwise) may be NULL: the session's Philox stream is used.
// h_idx (n entries, single-rank sessions only) may be NULL.
int32_t fc_smc_resample_particles(fc_smc_session* s, int32_t method, const double* h_uniforms, uint64_t* h_idx) {
  using namespace fc;
  FC_TRY(smc_ready(s));
  if (h_idx && s->world > 1) { set_error("ancestor indices are reported for single-rank sessions only"); return FC_ERR_UNSUPPORTED_OP; }
  FC_TRY(fc_smc_normalize(s, nullptr, nullptr));
  s->steps += 1;
  FC_TRY(smc_resample(s, method, h_uniforms, h_idx));
  FC_CUDA(cudaStreamSynchronize(s->P->stream));
  return comm_check(s);
}

// parity hook: resample_particles with the population processed as `shards` virtual ranks, one after the other
// (the sharded resample's scan-with-carry, slot routing and offset gather on one GPU; see smc_resample_virtual)
int32_t fc_smc_resample_virtual_shards(fc_smc_session* s, int32_t shards, int32_t method, double u_raw, uint64_t* h_idx) {
  using namespace fc;
  FC_TRY(smc_ready(s));
  FC_TRY(fc_smc_normalize(s, nullptr, nullptr));
  s->steps += 1;
  return smc_resample_virtual(s, shards, method, u_raw, h_idx);
}

// rejuvenate_particles (smc.rs:698-713): `steps` pi_beta-invariant single-site MH moves per particle,
// weights untouched (FG-13).
int32_t fc_smc_rejuvenate_particles(fc_smc_session* s, double beta, uint64_t steps) {
  using namespace fc;
  FC_TRY(smc_ready(s));
  if (!(beta >= 0.0 && beta <= 1.0)) { set_error("beta must be in [0, 1], got %g", beta); return FC_ERR_INVALID_ARGUMENT; }
  FC_TRY(smc_rejuvenate(s, beta, steps));
  FC_CUDA(cudaStreamSynchronize(s->P->stream));
  return FC_OK;
}

// parity hooks: record the GLOBAL ancestor of every output slot this rank produces in a resample, and
// read them back: slots [jlo, jhi) (rank-major concatenation over ranks = the whole ancestor array)
int32_t fc_smc_record_offspring(fc_smc_session* s, int32_t enable) {
  if (!s) { fc::set_error("session is NULL"); return FC_ERR_INVALID_ARGUMENT; }
  s->record = enable != 0;
  return FC_OK;
}
int32_t fc_smc_last_offspring(fc_smc_session* s, uint64_t* jlo, uint64_t* jhi, uint64_t* h_idx) {
  if (!s) { fc::set_error("session is NULL"); return FC_ERR_INVALID_ARGUMENT; }
  if (jlo) *jlo = (uint64_t)s->last_jlo;
  if (jhi) *jhi = (uint64_t)s->last_jhi;
  if (h_idx && s->last_jhi > s->last_jlo) {
    if (!s->record) { fc::set_error("fc_smc_record_offspring was not enabled"); return FC_ERR_INVALID_ARGUMENT; }
    FC_CUDA(cudaSetDevice(s->P->device));
    FC_CUDA(cudaMemcpy(h_idx, s->idx.p, (size_t)(s->last_jhi - s->last_jlo) * 8, cudaMemcpyDeviceToHost));
  }
  return FC_OK;
}

// Host arithmetic of the sharded resample's slot routing (no device work): the output slots [jlo, jhi) whose
// ancestor lives on a rank whose exact running sum goes from carry_in to end -- the walk of smc.rs:263-268
// stops inside that rank iff carry_in < threshold_j <= end; the last rank also takes the slots whose
// threshold is never reached. `tag` selects the stratified uniform stream (ignored for systematic).
int32_t fc_smc_slot_range(int32_t method, uint64_t n_total, double u_raw, uint64_t seed, uint64_t tag, double carry_in, double end, int32_t is_last, uint64_t* jlo, uint64_t* jhi) {
  if (!jlo || !jhi || (method != FC_RESAMPLE_SYSTEMATIC && method != FC_RESAMPLE_STRATIFIED)) { fc::set_error("bad arguments"); return FC_ERR_INVALID_ARGUMENT; }
  const int64_t a = fc::count_thresholds_le(method, (int64_t)n_total, u_raw, seed, tag, carry_in);
  int64_t b = is_last ? (int64_t)n_total : fc::count_thresholds_le(method, (int64_t)n_total, u_raw, seed, tag, end);
  if (b < a) b = a;
  *jlo = (uint64_t)a; *jhi = (uint64_t)b;
  return FC_OK;
}

// debug switches for A/B parity tests: "nb_plain" (1: next_beta evaluates every bisection midpoint, as the
// reference does), "smc_generic" (1: sessions created afterwards use the generic scoring engine)
int32_t fc_debug_set(const char* key, int32_t value) {
  if (key && !strcmp(key, "nb_plain")) { fc::g_nb_plain = (value >= 0 && value <= 2) ? value : 0; return FC_OK; }
  if (key && !strcmp(key, "smc_generic")) { fc::g_smc_generic = value ? 1 : 0; return FC_OK; }
  if (key && !strcmp(key, "smc_timing")) { fc::g_smc_timing = value ? 1 : 0; return FC_OK; }
  if (key && !strcmp(key, "smc_stats")) { fc::g_smc_stats = value ? 1 : 0; return FC_OK; }
  if (key && !strcmp(key, "smc_zbits")) { fc::g_smc_zbits = value ? 1 : 0; return FC_OK; }
  if (key && !strcmp(key, "hmc_hier")) { fc::g_hmc_hier = value ? 1 : 0; return FC_OK; }
  if (key && !strcmp(key, "smc_compiled")) { fc::g_smc_compiled = value ? 1 : 0; return FC_OK; }
  fc::set_error("unknown debug key");
  return FC_ERR_INVALID_ARGUMENT;
}

// CUDA-event time (ms, on the launching stream) spent in the phases of the tempering steps so far, when
// fc_debug_set("smc_timing", 1) was on: next_beta search, reweight, exact scan + carry chain, gather +
// migration (+ barrier), rejuvenation moves; n[k] = how many times phase k ran
int32_t fc_smc_phase_ms(fc_smc_session* s, double* ms5, uint64_t* n5) {
  if (!s) { fc::set_error("session is NULL"); return FC_ERR_INVALID_ARGUMENT; }
  for (int k = 0; k < 5; k++) { if (ms5) ms5[k] = s->phase_ms[k]; if (n5) n5[k] = s->phase_launches[k]; }
  return FC_OK;
}

// counters of a session: particles this rank sent to other ranks, ESS evaluations of all next_beta searches
int32_t fc_smc_counters(const fc_smc_session* s, uint64_t* migrated, uint64_t* ess_evals, int32_t* fused) {
  if (!s) { fc::set_error("session is NULL"); return FC_ERR_INVALID_ARGUMENT; }
  if (migrated) *migrated = s->migrated;
  if (ess_evals) *ess_evals = s->ess_evals;
  if (fused) *fused = s->fused ? 1 : 0;
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
