// fc_resample.cu -- bit-exact systematic / stratified / multinomial resampling on device, plus
// the weight numerics (log_sum_exp, ESS, normalisation).
//
// Reference: src/inference/smc.rs:255-314 (the three *_indices fns), :230-233 (ESS),
// src/core/numerical.rs:15-38,87-90 (log_sum_exp, normalize_log_probs).
//
// All three reference resamplers walk ONE running sum `cum += w[i]` from left to right in f64.
// The index they return for a threshold t is determined by the SEQUENTIALLY ROUNDED prefix sums
//     S_0 = 0,  S_{k+1} = fl(S_k + w_k).
// A parallel scan re-associates the additions and its prefix sums differ from S_k in the last
// bits, which flips indices whenever a threshold falls between the two. This file
// reproduces S_k bit for bit, in parallel:
//   * while the running sum stays inside one binade [2^e, 2^(e+1)) its ulp U is fixed, the sum
//     is a multiple of U, and fl(S + w) = S + U * r(w) where r(w) = round-to-nearest-even of w/U
//     evaluated against the PARITY of S/U (the parity only matters for exact ties). So each
//     weight maps to a pair of integers (increment if S/U even, increment if odd), and pairs
//     compose associatively: the prefix sums inside a binade are an exact integer scan.
//   * the element at which the sum leaves the binade is found by the scan itself (first prefix
//     that reaches the binade's top); that single addition is done in floating point, which is
//     exactly what the sequential loop does, and the scan restarts in the new binade.
// There are at most ~60 binade changes between the first weight and 1.0, each a few launches.
// Thresholds are then resolved by binary search over S (S and the thresholds are monotone).
#include "fc_internal.cuh"

namespace fc {

// ---------------------------------------------------------------------------------------------
// reductions (deterministic: fixed grid, fixed order)
// ---------------------------------------------------------------------------------------------
#define RD_BLOCKS 1024
#define RD_THREADS 256

template <int OP>  // 0: max(x)   1: sum exp(x - m)   2: sum x^2   3: sum x   4: count of invalid weights
__global__ void k_reduce(const double* __restrict__ x, int64_t n, double m, double* __restrict__ part) {
  double acc = OP == 0 ? d_neg_inf() : 0.0;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    double v = x[i];
    if (OP == 0) acc = fmax(acc, v);
    if (OP == 1) acc += exp(v - m);
    if (OP == 2) acc += v * v;
    if (OP == 3) acc += v;
    if (OP == 4) acc += (v >= 0.0 && isfinite(v)) ? 0.0 : 1.0;
  }
  __shared__ double sm[RD_THREADS];
  sm[threadIdx.x] = acc;
  __syncthreads();
  for (int o = RD_THREADS / 2; o > 0; o >>= 1) {
    if (threadIdx.x < o) sm[threadIdx.x] = OP == 0 ? fmax(sm[threadIdx.x], sm[threadIdx.x + o]) : sm[threadIdx.x] + sm[threadIdx.x + o];
    __syncthreads();
  }
  if (threadIdx.x == 0) part[blockIdx.x] = sm[0];
}
template <int OP>
__global__ void k_reduce_final(const double* __restrict__ part, int nb, double* __restrict__ out) {
  __shared__ double sm[RD_THREADS];
  double acc = OP == 0 ? d_neg_inf() : 0.0;
  for (int i = threadIdx.x; i < nb; i += RD_THREADS) acc = OP == 0 ? fmax(acc, part[i]) : acc + part[i];
  sm[threadIdx.x] = acc;
  __syncthreads();
  for (int o = RD_THREADS / 2; o > 0; o >>= 1) {
    if (threadIdx.x < o) sm[threadIdx.x] = OP == 0 ? fmax(sm[threadIdx.x], sm[threadIdx.x + o]) : sm[threadIdx.x] + sm[threadIdx.x + o];
    __syncthreads();
  }
  if (threadIdx.x == 0) *out = sm[0];
}

// d_scratch: RD_BLOCKS + 4 doubles. Result in d_out (device scalar).
int32_t dev_reduce(int op, const double* d_x, int64_t n, double m, double* d_scratch, double* d_out, cudaStream_t st) {
  int nb = (int)std::max<int64_t>(1, std::min<int64_t>(RD_BLOCKS, cdiv(n, RD_THREADS)));
  switch (op) {
    case 0: k_reduce<0><<<nb, RD_THREADS, 0, st>>>(d_x, n, m, d_scratch); k_reduce_final<0><<<1, RD_THREADS, 0, st>>>(d_scratch, nb, d_out); break;
    case 1: k_reduce<1><<<nb, RD_THREADS, 0, st>>>(d_x, n, m, d_scratch); k_reduce_final<1><<<1, RD_THREADS, 0, st>>>(d_scratch, nb, d_out); break;
    case 2: k_reduce<2><<<nb, RD_THREADS, 0, st>>>(d_x, n, m, d_scratch); k_reduce_final<2><<<1, RD_THREADS, 0, st>>>(d_scratch, nb, d_out); break;
    case 3: k_reduce<3><<<nb, RD_THREADS, 0, st>>>(d_x, n, m, d_scratch); k_reduce_final<3><<<1, RD_THREADS, 0, st>>>(d_scratch, nb, d_out); break;
    default: k_reduce<4><<<nb, RD_THREADS, 0, st>>>(d_x, n, m, d_scratch); k_reduce_final<3><<<1, RD_THREADS, 0, st>>>(d_scratch, nb, d_out); break;
  }
  FC_KERNEL_CHECK();
  return FC_OK;
}

// log_sum_exp (numerical.rs:15-38) of a device array; result on the host.
int32_t dev_log_sum_exp(const double* d_x, int64_t n, double* d_scratch, double* h_out, cudaStream_t st) {
  if (n == 0) { *h_out = -INFINITY; return FC_OK; }
  double* d_res = d_scratch + RD_BLOCKS;
  FC_TRY(dev_reduce(0, d_x, n, 0.0, d_scratch, d_res, st));
  double mx;
  FC_CUDA(cudaMemcpyAsync(&mx, d_res, 8, cudaMemcpyDeviceToHost, st));
  FC_CUDA(cudaStreamSynchronize(st));
  if (mx == -INFINITY) { *h_out = -INFINITY; return FC_OK; }
  if (mx != mx || mx == INFINITY) { *h_out = mx; return FC_OK; }
  FC_TRY(dev_reduce(1, d_x, n, mx, d_scratch, d_res, st));
  double s;
  FC_CUDA(cudaMemcpyAsync(&s, d_res, 8, cudaMemcpyDeviceToHost, st));
  FC_CUDA(cudaStreamSynchronize(st));
  *h_out = (s == 0.0) ? -INFINITY : mx + std::log(s);
  return FC_OK;
}

// ---------------------------------------------------------------------------------------------
// exact sequentially-rounded prefix sums
// ---------------------------------------------------------------------------------------------
// increments are counted in ulps of the binade: a single one is < 2^53, the distance to the binade top
// is <= 2^53, so everything >= 2^55 means "has left the binade" and sums saturate there (SAT_INC is even:
// a saturated left operand picks b.e and stays saturated). 56 bits leave room for the look-back tags.
#define SAT_INC (1LL << 55)
struct Inc2 { long long e, o; };

__device__ __forceinline__ long long sat_add(long long a, long long b) { const long long s = a + b; return s < SAT_INC ? s : SAT_INC; }  // a, b in [0, SAT_INC]
__device__ __forceinline__ Inc2 combine(const Inc2& a, const Inc2& b) {
  Inc2 r;
  r.e = sat_add(a.e, (a.e & 1) ? b.o : b.e);
  r.o = sat_add(a.o, (a.o & 1) ? b.e : b.o);
  return r;
}
// The same composition without the clamp, for the in-thread walks over SC_ITEMS items: operands are at most
// SAT_INC, so 1 + SC_ITEMS of them stay below 2^59; the caller clamps the aggregate once (clamp_inc). A value
// above SAT_INC has long passed the binade top (need <= 2^53), and everything after the first crossing is
// discarded, so which of b.e / b.o an over-the-top left operand picks does not matter. (ncu: the clamps'
// 64-bit compare + select pairs were 47 % of the scan's issued instructions.)
__device__ __forceinline__ Inc2 combine_raw(const Inc2& a, const Inc2& b) {
  Inc2 r;
  r.e = a.e + ((a.e & 1) ? b.o : b.e);
  r.o = a.o + ((a.o & 1) ? b.e : b.o);
  return r;
}
__device__ __forceinline__ Inc2 clamp_inc(Inc2 a) { a.e = a.e < SAT_INC ? a.e : SAT_INC; a.o = a.o < SAT_INC ? a.o : SAT_INC; return a; }
// base = bottom of the binade (0 for the subnormal/first-normal region), U = its ulp,
// lim = weights >= lim certainly leave the binade.
// iu1 * iu2 = 1 / U as two powers of two (1 / U itself overflows for the smallest binades): the quotient
// by the ulp is an exact scaling, not a division (two FP64 divisions per weight made the scan FP64-bound)
__device__ __forceinline__ Inc2 make_inc(double w, double base, double U, double lim, double iu1, double iu2) {
  Inc2 r;
  if (!(w < lim)) { r.e = r.o = SAT_INC; return r; }
  double b1 = base + U;
  double te = __dadd_rn(__dadd_rn(base, w), -base);
  double to = __dadd_rn(__dadd_rn(b1, w), -b1);
  r.e = (long long)__dmul_rn(__dmul_rn(te, iu1), iu2);
  r.o = (long long)__dmul_rn(__dmul_rn(to, iu1), iu2);
  return r;
}

#define SC_THREADS 256
#define SC_ITEMS 8
#define SC_TILE (SC_THREADS * SC_ITEMS)

__device__ __forceinline__ Inc2 block_exclusive_scan(Inc2 v, Inc2* total) {
  // exclusive scan of per-thread aggregates across the block under `combine`
  __shared__ Inc2 warp_agg[SC_THREADS / 32];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  Inc2 incl = v;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    Inc2 other;
    other.e = __shfl_up_sync(0xffffffffu, incl.e, o);
    other.o = __shfl_up_sync(0xffffffffu, incl.o, o);
    if (lane >= o) incl = combine(other, incl);
  }
  if (lane == 31) warp_agg[warp] = incl;
  __syncthreads();
  Inc2 excl;
  excl.e = __shfl_up_sync(0xffffffffu, incl.e, 1);
  excl.o = __shfl_up_sync(0xffffffffu, incl.o, 1);
  if (lane == 0) { excl.e = 0; excl.o = 0; }
  // prefix over the warps' aggregates: every warp scans the SC_THREADS / 32 values itself with three shuffle
  // steps (a serial walk per thread was 15 combines = a fifth of the scan's issued instructions)
  constexpr int NW = SC_THREADS / 32;
  Inc2 wv; wv.e = 0; wv.o = 0;
  if (lane < NW) wv = warp_agg[lane];
#pragma unroll
  for (int o = 1; o < NW; o <<= 1) {
    Inc2 other;
    other.e = __shfl_up_sync(0xffffffffu, wv.e, o);
    other.o = __shfl_up_sync(0xffffffffu, wv.o, o);
    if (lane >= o && lane < NW) wv = combine(other, wv);
  }
  Inc2 pre;  // inclusive prefix of warp - 1 (identity for warp 0)
  pre.e = __shfl_sync(0xffffffffu, wv.e, warp > 0 ? warp - 1 : 0);
  pre.o = __shfl_sync(0xffffffffu, wv.o, warp > 0 ? warp - 1 : 0);
  if (warp == 0) { pre.e = 0; pre.o = 0; }
  if (total) {
    total->e = __shfl_sync(0xffffffffu, wv.e, NW - 1);
    total->o = __shfl_sync(0xffffffffu, wv.o, NW - 1);
  }
  __syncthreads();
  return combine(pre, excl);
}

// ---- device-driven windows -------------------------------------------------------------------
// One ROUND handles the window [k0, k0 + len) of the current binade with ONE launch; everything a
// round needs (k0, the running sum s0, the window policy) lives in a control block on the device, so
// the host enqueues a batch of rounds back to back and reads the block once. (Round 1 of this design
// was host-driven: ~27 windows x (4 launches + a blocking copy) = 2.1 ms per resample at n = 2^24.)
//   k_scan_round  one pass over the window (chained scan with look-back): per-element prefix sums -> S,
//                 first crossing; the last CTA does the one floating add that changes binade and sets up
//                 the next window
struct Binade { double base, U, lim, T, iu1, iu2; long long need; int parity; };
__device__ __forceinline__ Binade binade_of(double s0) {
  Binade b;
  const double two_m1021 = ldexp(1.0, -1021);
  int ue;  // U = 2^ue
  if (s0 < two_m1021) { b.base = 0.0; ue = -1074; b.T = two_m1021; }
  else {
    int e;
    frexp(s0, &e);  // s0 = m * 2^e, m in [0.5, 1)
    b.base = ldexp(1.0, e - 1); ue = e - 1 - 52; b.T = ldexp(1.0, e);
  }
  b.U = ldexp(1.0, ue);
  if (ue >= -1000) { b.iu1 = ldexp(1.0, -ue); b.iu2 = 1.0; }
  else { b.iu1 = ldexp(1.0, 600); b.iu2 = ldexp(1.0, -ue - 600); }
  b.lim = (b.base == 0.0) ? b.T : b.base;
  b.need = (long long)__dmul_rn(__dmul_rn(b.T - s0, b.iu1), b.iu2);  // exact: both multiples of U, quotient <= 2^53
  b.parity = (int)(((long long)__dmul_rn(__dmul_rn(s0 - b.base, b.iu1), b.iu2)) & 1LL);
  return b;
}
#define SC_GRID_MAX 1184  // 8 CTAs per SM on 148 SMs

// the first few hundred elements cross a binade every doubling of the count: one thread walks them
// exactly as the reference does (the loads do not depend on the add chain, so this is ~2 us)
// window from the mean weight seen so far: the elements the running sum still needs to leave its binade,
// plus a quarter (a window without a crossing doubles)
__device__ __forceinline__ long long scan_window(double s0, double mean) {
  const Binade B = binade_of(s0);
  double e = mean > 0.0 ? 1.25 * (B.T - s0) / mean : 2048.0;
  if (!(e >= 2048.0)) e = 2048.0;
  if (e > 67108864.0) e = 67108864.0;
  return (long long)e;
}
// ---- one round = the window [k0, k0 + len) of the current binade, as reduce -> scan -> apply ----------------------
// (The first version of this round was ONE kernel, a chained scan with decoupled look-back. ncu launch lists showed it
// at ~0.6 TB/s on multi-million-element windows whatever its instruction count: tiles spin on their predecessors'
// descriptors and take tickets from one atomic. Three plain kernels with no inter-CTA waiting do the same work faster
// and cannot hang: the weights are read twice, the second time mostly from L2.)
//   k_scan_agg    per tile of 2048 weights: the Inc2 aggregate (increment pair for an even / odd start)
//   k_scan_tiles  one CTA: exclusive scan of the tile aggregates under `combine`; the first tile whose inclusive prefix
//                 reaches the binade top (later tiles need no work)
//   k_scan_apply  per tile: prefix sums from the tile's exclusive prefix -> S, first crossing; the last CTA does the
//                 one floating add that changes binade and sets up the next window
__global__ void __launch_bounds__(SC_THREADS, 4) k_scan_agg(const double* __restrict__ w, int64_t n, const ScanCtl* __restrict__ ctl, Inc2* __restrict__ agg) {
  const long long k0 = ctl->k0;
  if (k0 >= n || ctl->serial) return;
  const Binade B = binade_of(ctl->s0);
  const int64_t len = min((long long)(n - k0), (long long)ctl->win);
  const int nt = (int)((len + SC_TILE - 1) / SC_TILE);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  __shared__ Inc2 s_w[SC_THREADS / 32];
  for (int t = blockIdx.x; t < nt; t += gridDim.x) {
    const int t0 = t * SC_TILE + (int)threadIdx.x * SC_ITEMS;
    const int nval = (int)max(0LL, min((long long)SC_ITEMS, (long long)len - t0));
    const double* __restrict__ wp = w + k0 + t0;
    double wv[SC_ITEMS];
#pragma unroll
    for (int i = 0; i < SC_ITEMS; i++) wv[i] = i < nval ? wp[i] : 0.0;
    Inc2 a; a.e = 0; a.o = 0;
#pragma unroll
    for (int i = 0; i < SC_ITEMS; i++) {
      Inc2 it; it.e = 0; it.o = 0;
      if (i < nval) it = make_inc(wv[i], B.base, B.U, B.lim, B.iu1, B.iu2);
      a = combine_raw(a, it);
    }
    a = clamp_inc(a);
    // ordered reduction over the CTA: lane 31 of each warp ends with the warp's aggregate
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      Inc2 other;
      other.e = __shfl_up_sync(0xffffffffu, a.e, o);
      other.o = __shfl_up_sync(0xffffffffu, a.o, o);
      if (lane >= o) a = combine(other, a);
    }
    if (lane == 31) s_w[warp] = a;
    __syncthreads();
    if (threadIdx.x == 0) {
      Inc2 tot = s_w[0];
      for (int q = 1; q < SC_THREADS / 32; q++) tot = combine(tot, s_w[q]);
      agg[t] = tot;
    }
    __syncthreads();
  }
}
#define ST_THREADS 1024
__global__ void __launch_bounds__(ST_THREADS) k_scan_tiles(int64_t n, ScanCtl* __restrict__ ctl, const Inc2* __restrict__ agg, Inc2* __restrict__ excl) {
  const long long k0 = ctl->k0;
  if (k0 >= n || ctl->serial) return;
  const Binade B = binade_of(ctl->s0);
  const int64_t len = min((long long)(n - k0), (long long)ctl->win);
  const int nt = (int)((len + SC_TILE - 1) / SC_TILE);
  const int per = (nt + ST_THREADS - 1) / ST_THREADS;
  const int b0 = threadIdx.x * per, b1 = min(nt, b0 + per);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  __shared__ Inc2 s_w[ST_THREADS / 32];
  __shared__ unsigned s_cross;
  if (threadIdx.x == 0) s_cross = 0xffffffffu;
  Inc2 a; a.e = 0; a.o = 0;
  for (int t = b0; t < b1; t++) a = combine(a, agg[t]);
  Inc2 incl = a;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    Inc2 other;
    other.e = __shfl_up_sync(0xffffffffu, incl.e, o);
    other.o = __shfl_up_sync(0xffffffffu, incl.o, o);
    if (lane >= o) incl = combine(other, incl);
  }
  if (lane == 31) s_w[warp] = incl;
  __syncthreads();
  // prefix over the 32 warp aggregates: every warp scans them itself with five shuffle steps
  Inc2 wv = s_w[lane];
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    Inc2 other;
    other.e = __shfl_up_sync(0xffffffffu, wv.e, o);
    other.o = __shfl_up_sync(0xffffffffu, wv.o, o);
    if (lane >= o) wv = combine(other, wv);
  }
  Inc2 pre;
  pre.e = __shfl_sync(0xffffffffu, wv.e, warp > 0 ? warp - 1 : 0);
  pre.o = __shfl_sync(0xffffffffu, wv.o, warp > 0 ? warp - 1 : 0);
  if (warp == 0) { pre.e = 0; pre.o = 0; }
  Inc2 ex;
  ex.e = __shfl_up_sync(0xffffffffu, incl.e, 1);
  ex.o = __shfl_up_sync(0xffffffffu, incl.o, 1);
  if (lane == 0) { ex.e = 0; ex.o = 0; }
  Inc2 run = combine(pre, ex);  // exclusive prefix of this thread's first tile
  unsigned cross = 0xffffffffu;
  for (int t = b0; t < b1; t++) {
    excl[t] = run;
    run = combine(run, agg[t]);
    if (cross == 0xffffffffu && (B.parity ? run.o : run.e) >= B.need) cross = (unsigned)t;
  }
  if (cross != 0xffffffffu) atomicMin(&s_cross, cross);
  __syncthreads();
  if (threadIdx.x == 0) ctl->cnt_a = s_cross;  // the first tile in which the sum reaches the binade top (0xffffffff: none)
}
__global__ void __launch_bounds__(SC_THREADS, 3) k_scan_apply(const double* __restrict__ w, int64_t n, ScanCtl* __restrict__ ctl, const Inc2* __restrict__ excl, double* __restrict__ S) {
  const long long k0 = ctl->k0;
  if (k0 >= n || ctl->serial) return;
  const double s0 = ctl->s0;
  const Binade B = binade_of(s0);
  const long long win = ctl->win;
  const int64_t len = min((long long)(n - k0), win);
  const int nt_all = (int)((len + SC_TILE - 1) / SC_TILE);
  const unsigned ct = ctl->cnt_a;
  const int nt = ct == 0xffffffffu ? nt_all : min(nt_all, (int)ct + 1);  // tiles after the crossing tile are not needed
  int cross = 0x7fffffff;
  for (int t = blockIdx.x; t < nt; t += gridDim.x) {
    const int t0 = t * SC_TILE + (int)threadIdx.x * SC_ITEMS;
    const int nval = (int)max(0LL, min((long long)SC_ITEMS, (long long)len - t0));
    const double* __restrict__ wp = w + k0 + t0;
    double* __restrict__ Sp = S + k0 + t0 + 1;
    double wv[SC_ITEMS];
#pragma unroll
    for (int i = 0; i < SC_ITEMS; i++) wv[i] = i < nval ? wp[i] : 0.0;
    Inc2 items[SC_ITEMS];
    Inc2 a; a.e = 0; a.o = 0;
#pragma unroll
    for (int i = 0; i < SC_ITEMS; i++) {
      items[i].e = 0; items[i].o = 0;
      if (i < nval) items[i] = make_inc(wv[i], B.base, B.U, B.lim, B.iu1, B.iu2);
      a = combine_raw(a, items[i]);
    }
    a = clamp_inc(a);
    const Inc2 pre = block_exclusive_scan(a, nullptr);
    // the start of every item is known from here on (the window starts at parity B.parity): follow the one actual path
    const Inc2 run = combine(excl[t], pre);
    long long p = B.parity ? run.o : run.e;  // ulps gained since the window start
#pragma unroll
    for (int i = 0; i < SC_ITEMS; i++) {
      p += ((B.parity + (int)p) & 1) ? items[i].o : items[i].e;
      if (i < nval) {
        if (p >= B.need) cross = min(cross, t0 + i);
        else Sp[i] = __dadd_rn(s0, __dmul_rn((double)p, B.U));  // exact: on the U grid below the binade top
      }
    }
  }
  if (cross != 0x7fffffff) atomicMin((unsigned long long*)&ctl->first_cross, (unsigned long long)cross);
  __shared__ bool is_last;
  __syncthreads();
  if (threadIdx.x == 0) { __threadfence(); is_last = atomicAdd(&ctl->cnt_b, 1u) == gridDim.x - 1; }
  __syncthreads();
  if (!is_last || threadIdx.x != 0) return;
  __threadfence();
  // advance: the one addition that changes binade is a plain IEEE add, as the reference does.
  // Window policy: after a crossing the next window is the number of elements the sum still needs to leave
  // the new binade at the mean weight seen so far, plus a quarter; a window without a crossing doubles.
  if (*(volatile int*)&ctl->serial) { ctl->cnt_a = 0; ctl->cnt_b = 0; return; }  // k0 / s0 untouched: the sequential tail finishes from there
  const long long fc = *(volatile long long*)&ctl->first_cross;
  long long nk0; double ns0; bool crossed;
  if (fc < len) {
    const long long k = k0 + fc;
    ns0 = __dadd_rn(__ldcg(&S[k]), w[k]);
    S[k + 1] = ns0;
    nk0 = k + 1;
    crossed = true;
  } else {
    nk0 = k0 + len;
    ns0 = __ldcg(&S[k0 + len]);
    crossed = !(ns0 < B.T);
  }
  long long nw = crossed ? scan_window(ns0, (ns0 - ctl->c0) / (double)nk0) : 2 * win;
  if (nw < 2048) nw = 2048;
  if (nw > (1LL << 26)) nw = 1LL << 26;
  ctl->k0 = nk0; ctl->s0 = ns0; ctl->win = nw;
  ctl->first_cross = 0x7fffffffffffffffLL;
  ctl->cnt_a = 0; ctl->cnt_b = 0;
  ctl->round += 1;
  if (ns0 != ns0 || ns0 == d_pos_inf()) ctl->serial = 1;  // cannot happen with finite weights unless the sum overflows
}
// ---- the front of a scan in ONE single-CTA launch ------------------------------------------------------------
// (a) The first SC_HEAD elements, where the running sum climbs through a binade every doubling of the count: the
//     weights are staged in shared memory by coalesced loads, one thread walks them exactly as the reference does
//     (from shared memory: ~10 cycles per add instead of an L2 round trip), S is written back coalesced.
// (b) Every following window of at most SC_SMALL_MAX elements: the CTA scans its tiles one after the other with a
//     running carry (no descriptors, no look-back), finds the first crossing, does the one floating add that
//     changes binade and sizes the next window -- the same arithmetic as k_scan_round + its last-CTA epilogue.
// ncu on the multi-launch form: a 32 us one-thread head and eight 13 us launches whose work is one tile, per
// resample. A scan that continues a running sum well inside a binade (a later rank of a sharded population)
// skips (a): its first window is sized from the caller's mean-weight hint and is normally too large for (b).
#define SC_HEAD 512
#define SC_SMALL_MAX 16384
__global__ void __launch_bounds__(SC_THREADS) k_scan_front(const double* __restrict__ w, int64_t n, double* __restrict__ S, ScanCtl* __restrict__ ctl) {
  __shared__ double s_w[SC_HEAD];
  __shared__ double s_S[SC_HEAD + 1];
  __shared__ long long sh_k0, sh_win, sh_cross;
  __shared__ double sh_s0, sh_c0;
  __shared__ int sh_serial;
  __shared__ Inc2 s_carry;
  const int tid = threadIdx.x;
  const double c0 = ctl->s0, hint = ctl->w_mean;
  const int64_t m = (c0 > 0.0 && hint > 0.0) ? 0 : (SC_HEAD < n ? SC_HEAD : n);
  for (int64_t i = tid; i < m; i += SC_THREADS) s_w[i] = w[i];
  __syncthreads();
  if (tid == 0) {
    double cum = c0;
    s_S[0] = cum;
    for (int64_t i = 0; i < m; i++) { cum = __dadd_rn(cum, s_w[i]); s_S[i + 1] = cum; }
    sh_k0 = m; sh_s0 = cum; sh_c0 = c0;
    sh_win = scan_window(cum, m > 0 ? (cum - c0) / (double)m : hint);
    sh_serial = (cum != cum || cum == d_pos_inf()) ? 1 : 0;
  }
  __syncthreads();
  for (int64_t i = tid; i <= m; i += SC_THREADS) S[i] = s_S[i];
  for (int it = 0; it < 64; it++) {
    __syncthreads();
    const long long k0 = sh_k0, win = sh_win;
    const double s0 = sh_s0;
    if (k0 >= n || sh_serial || win > SC_SMALL_MAX) break;
    const Binade B = binade_of(s0);
    const int64_t len = min((long long)(n - k0), win);
    const int nt = (int)((len + SC_TILE - 1) / SC_TILE);
    if (tid == 0) { s_carry.e = 0; s_carry.o = 0; sh_cross = 0x7fffffffffffffffLL; }
    __syncthreads();
    for (int t = 0; t < nt; t++) {
      const int t0 = t * SC_TILE + tid * SC_ITEMS;
      const int nval = (int)max(0LL, min((long long)SC_ITEMS, (long long)len - t0));
      const double* __restrict__ wp = w + k0 + t0;
      double* __restrict__ Sp = S + k0 + t0 + 1;
      Inc2 items[SC_ITEMS];
      Inc2 a; a.e = 0; a.o = 0;
#pragma unroll
      for (int i = 0; i < SC_ITEMS; i++) {
        items[i].e = 0; items[i].o = 0;
        if (i < nval) items[i] = make_inc(wp[i], B.base, B.U, B.lim, B.iu1, B.iu2);
        a = combine_raw(a, items[i]);
      }
      a = clamp_inc(a);
      Inc2 total;
      const Inc2 pre = block_exclusive_scan(a, &total);
      const Inc2 run = combine(s_carry, pre);
      long long p = B.parity ? run.o : run.e;
      int cross = 0x7fffffff;
#pragma unroll
      for (int i = 0; i < SC_ITEMS; i++) {
        p += ((B.parity + (int)p) & 1) ? items[i].o : items[i].e;
        if (i < nval) {
          if (p >= B.need) cross = min(cross, t0 + i);
          else Sp[i] = __dadd_rn(s0, __dmul_rn((double)p, B.U));
        }
      }
      if (cross != 0x7fffffff) atomicMin((unsigned long long*)&sh_cross, (unsigned long long)cross);
      __syncthreads();  // everyone has read s_carry
      if (tid == 0) s_carry = combine(s_carry, total);
      __syncthreads();
      if (sh_cross != 0x7fffffffffffffffLL) break;  // uniform: read after the barrier
    }
    __syncthreads();
    if (tid == 0) {  // advance (k_scan_round's epilogue)
      const long long fc = sh_cross;
      long long nk0; double ns0; bool crossed;
      if (fc < len) {
        const long long k = k0 + fc;
        ns0 = __dadd_rn(*(volatile double*)&S[k], w[k]);
        S[k + 1] = ns0;
        nk0 = k + 1;
        crossed = true;
      } else {
        nk0 = k0 + len;
        ns0 = *(volatile double*)&S[k0 + len];
        crossed = !(ns0 < B.T);
      }
      long long nw = crossed ? scan_window(ns0, (ns0 - sh_c0) / (double)nk0) : 2 * win;
      if (nw < 2048) nw = 2048;
      if (nw > (1LL << 26)) nw = 1LL << 26;
      sh_k0 = nk0; sh_s0 = ns0; sh_win = nw;
      if (ns0 != ns0 || ns0 == d_pos_inf()) sh_serial = 1;
    }
  }
  __syncthreads();
  if (tid == 0) {
    ctl->c0 = sh_c0; ctl->k0 = sh_k0; ctl->s0 = sh_s0; ctl->win = sh_win; ctl->serial = sh_serial;
    ctl->first_cross = 0x7fffffffffffffffLL; ctl->cnt_a = 0; ctl->cnt_b = 0; ctl->round = 1;
  }
}
// correctness fallback (weights with negatives / non-finite values): one thread, the reference loop
__global__ void k_scan_serial(const double* __restrict__ w, int64_t from, int64_t n, double* __restrict__ S) {
  double cum = S[from];
  for (int64_t i = from; i < n; i++) { cum = __dadd_rn(cum, w[i]); S[i + 1] = cum; }
}
// finishes a scan whose control block says `serial` (from its k0 on)
__global__ void k_scan_serial_tail(const double* __restrict__ w, int64_t n, double* __restrict__ S, ScanCtl* __restrict__ ctl) {
  if (!ctl->serial) return;
  double cum = ctl->s0;
  for (int64_t i = ctl->k0; i < n; i++) { cum = __dadd_rn(cum, w[i]); S[i + 1] = cum; }
  ctl->k0 = n; ctl->s0 = cum; ctl->serial = 0;
}

size_t scan_work_bytes(int64_t n) { return 64 + 2 * (size_t)(cdiv(std::max<int64_t>(n, 1), SC_TILE) + 8) * sizeof(Inc2); }  // control block, tile aggregates, tile prefixes
static int scan_grid() {
  static int g = 0;
  if (!g) {
    int dev = 0, sms = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    g = std::min(SC_GRID_MAX, 8 * (sms > 0 ? sms : 148));
  }
  return g;
}
// enqueue `rounds` rounds (no host synchronisation); d_work: ScanCtl (64 B) followed by the tile aggregates
int32_t scan_enqueue_rounds(const double* d_w, int64_t n, double* d_S, void* d_work, int rounds, cudaStream_t st, uint64_t* launches) {
  ScanCtl* ctl = (ScanCtl*)d_work;
  const int64_t tiles = cdiv(std::max<int64_t>(n, 1), SC_TILE);
  Inc2* agg = (Inc2*)((char*)d_work + 64);
  Inc2* excl = agg + tiles + 8;
  const int g = (int)std::max<int64_t>(1, std::min<int64_t>(scan_grid(), tiles));
  for (int r = 0; r < rounds; r++) {
    k_scan_agg<<<g, SC_THREADS, 0, st>>>(d_w, n, ctl, agg);
    k_scan_tiles<<<1, ST_THREADS, 0, st>>>(n, ctl, agg, excl);
    k_scan_apply<<<g, SC_THREADS, 0, st>>>(d_w, n, ctl, excl, d_S);
  }
  k_scan_serial_tail<<<1, 1, 0, st>>>(d_w, n, d_S, ctl);
  FC_KERNEL_CHECK();
  if (launches) *launches += 3 * rounds + 1;
  return FC_OK;
}
// start a scan whose carry-in is already in ctl->s0 on the device (written by the host copy below or
// by the sharded driver's carry kernel)
int32_t scan_enqueue_head(const double* d_w, int64_t n, double* d_S, void* d_work, cudaStream_t st, uint64_t* launches) {
  k_scan_front<<<1, SC_THREADS, 0, st>>>(d_w, n, d_S, (ScanCtl*)d_work);
  FC_KERNEL_CHECK();
  if (launches) *launches += 1;
  return FC_OK;
}
// rounds a scan of n elements is expected to need: one per binade the running sum passes after the head
int scan_expected_rounds(int64_t n) {
  int r = 2;  // the front kernel has already passed the binades below SC_SMALL_MAX elements
  for (int64_t k = SC_SMALL_MAX / 2; k < n; k *= 2) r++;
  return r;
}
// wait until the scan has consumed all n elements, enqueueing more rounds if it has not
int32_t scan_finish(const double* d_w, int64_t n, double* d_S, void* d_work, cudaStream_t st, uint64_t* launches, double* h_end) {
  ScanCtl hs;
  for (int guard = 0; guard < 4096; guard++) {
    FC_CUDA(cudaMemcpyAsync(&hs, d_work, sizeof(hs), cudaMemcpyDeviceToHost, st));
    FC_CUDA(cudaStreamSynchronize(st));
    if (hs.k0 >= n) { if (h_end) *h_end = hs.s0; return FC_OK; }
    FC_TRY(scan_enqueue_rounds(d_w, n, d_S, d_work, 8, st, launches));
  }
  set_error("exact_prefix_sums: no progress");
  return FC_ERR_CUDA;
}

// S[0..n] := sequentially rounded prefix sums of w[0..n). d_work: >= 64 B + cdiv(n, SC_TILE) Inc2.
// carry = the running sum the scan starts from (0 for a whole population; the exact end value of
// the previous rank for one shard of a rank-major global population).
// assume_valid: the caller guarantees finite, non-negative weights (skips the check pass).
int32_t exact_prefix_sums(const double* d_w, int64_t n, double* d_S, void* d_work, double* d_scratch, cudaStream_t st, uint64_t* launches, double carry, bool assume_valid = false) {
  if (n == 0) {
    FC_CUDA(cudaMemcpyAsync(d_S, &carry, 8, cudaMemcpyHostToDevice, st));
    FC_CUDA(cudaStreamSynchronize(st));
    return FC_OK;
  }
  if (!assume_valid) {  // validity: all weights finite and >= 0, else the serial fallback
    double bad = 0.0;
    FC_TRY(dev_reduce(4, d_w, n, 0.0, d_scratch, d_scratch + RD_BLOCKS, st));
    FC_CUDA(cudaMemcpyAsync(&bad, d_scratch + RD_BLOCKS, 8, cudaMemcpyDeviceToHost, st));
    FC_CUDA(cudaStreamSynchronize(st));
    if (launches) *launches += 2;
    if (bad != 0.0) {
      FC_CUDA(cudaMemcpyAsync(d_S, &carry, 8, cudaMemcpyHostToDevice, st));
      k_scan_serial<<<1, 1, 0, st>>>(d_w, 0, n, d_S);
      FC_KERNEL_CHECK();
      FC_CUDA(cudaStreamSynchronize(st));
      if (launches) *launches += 1;
      return FC_OK;
    }
  }
  ScanCtl hs;
  memset(&hs, 0, sizeof(hs));
  hs.s0 = carry;
  FC_CUDA(cudaMemcpyAsync(d_work, &hs, sizeof(hs), cudaMemcpyHostToDevice, st));
  FC_TRY(scan_enqueue_head(d_w, n, d_S, d_work, st, launches));
  FC_TRY(scan_enqueue_rounds(d_w, n, d_S, d_work, scan_expected_rounds(n), st, launches));
  return scan_finish(d_w, n, d_S, d_work, st, launches, nullptr);
}

// ---------------------------------------------------------------------------------------------
// threshold search
// ---------------------------------------------------------------------------------------------
// first i in [0, n] with S[i] >= t, or n + 1 if none
__device__ __forceinline__ int64_t lower_bound_S(const double* __restrict__ S, int64_t n, double t) {
  int64_t lo = 0, hi = n + 1;
  while (lo < hi) {
    int64_t mid = (lo + hi) >> 1;
    if (S[mid] >= t) hi = mid; else lo = mid + 1;
  }
  return lo;
}
// systematic (smc.rs:255-272) / stratified (smc.rs:275-292): the shared walk stops at the first i
// with !(cum < thr) i.e. S_i >= thr (or i == n); the index is (i - 1).min(n - 1), and i == 0
// wraps in release builds to n - 1.
__global__ void k_resample_walk(const double* __restrict__ S, int64_t n, int method, const double* __restrict__ us, uint64_t* __restrict__ idx) {
  int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= n) return;
  double nf = (double)n;
  double thr = method == FC_RESAMPLE_SYSTEMATIC ? __dadd_rn(__ddiv_rn(us[0], nf), __ddiv_rn((double)j, nf))  // u = gen()/n; thr = u + j/n
                                                : __ddiv_rn(__dadd_rn((double)j, us[j]), nf);                 // (j + u_j)/n
  int64_t i = lower_bound_S(S, n, thr);
  if (i > n) i = n;
  int64_t k = (i == 0) ? n - 1 : (i - 1 < n - 1 ? i - 1 : n - 1);
  if (!(thr == thr)) k = n - 1;  // NaN threshold: `cum < thr` is false at once => i = 0
  idx[j] = (uint64_t)k;
}
// multinomial (smc.rs:295-314): per draw, first i with u <= cum (cum = S_{i+1}), default n - 1
__global__ void k_resample_multinomial(const double* __restrict__ S, int64_t n, const double* __restrict__ us, uint64_t* __restrict__ idx) {
  int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= n) return;
  double u = us[j];
  int64_t lo = 1, hi = n + 1;  // over S[1..n]
  while (lo < hi) {
    int64_t mid = (lo + hi) >> 1;
    if (S[mid] >= u) hi = mid; else lo = mid + 1;
  }
  idx[j] = (uint64_t)((lo <= n && u == u) ? lo - 1 : n - 1);
}

struct ResampleWork {
  DevBuf<double> S, scratch;
  DevBuf<char> work;
  int32_t reserve(int64_t n) {
    FC_CUDA(S.reserve(n + 2));
    FC_CUDA(scratch.reserve(RD_BLOCKS + 8));
    FC_CUDA(work.reserve(scan_work_bytes(n)));
    return FC_OK;
  }
};

int32_t resample_dev(int method, const double* d_w, int64_t n, const double* d_us, uint64_t* d_idx, ResampleWork& W, cudaStream_t st, uint64_t* launches) {
  if (n == 0) return FC_OK;
  if (method < FC_RESAMPLE_MULTINOMIAL || method > FC_RESAMPLE_STRATIFIED) { set_error("bad resampling method %d", method); return FC_ERR_INVALID_ARGUMENT; }
  FC_TRY(W.reserve(n));
  FC_TRY(exact_prefix_sums(d_w, n, W.S.p, W.work.p, W.scratch.p, st, launches, 0.0));
  unsigned nb = (unsigned)cdiv(n, 256);
  if (method == FC_RESAMPLE_MULTINOMIAL) k_resample_multinomial<<<nb, 256, 0, st>>>(W.S.p, n, d_us, d_idx);
  else k_resample_walk<<<nb, 256, 0, st>>>(W.S.p, n, method, d_us, d_idx);
  FC_KERNEL_CHECK();
  if (launches) *launches += 1;
  return FC_OK;
}

ResampleWork* resample_work_new() { return new ResampleWork(); }
void resample_work_free(ResampleWork* w) { delete w; }

static cudaStream_t g_util_stream = nullptr;
static int32_t util_stream(cudaStream_t* st) {
  if (!g_util_stream) FC_CUDA(cudaStreamCreateWithFlags(&g_util_stream, cudaStreamNonBlocking));
  *st = g_util_stream;
  return FC_OK;
}

__global__ void k_exp_shift(const double* __restrict__ x, double lse, double* __restrict__ out, int64_t n) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) out[i] = exp(x[i] - lse);
}

}  // namespace fc

extern "C" {

// systematic / stratified / multinomial indices with the uniforms passed explicitly
// (1 raw uniform for systematic, n for the others). Host buffers.
int32_t fc_resample(int32_t method, const double* h_w, size_t n, const double* h_uniforms, uint64_t* h_idx) {
  using namespace fc;
  if (n == 0) return FC_OK;
  if (!h_w || !h_uniforms || !h_idx) { set_error("NULL argument"); return FC_ERR_INVALID_ARGUMENT; }
  cudaStream_t st;
  FC_TRY(util_stream(&st));
  size_t nu = method == FC_RESAMPLE_SYSTEMATIC ? 1 : n;
  DevBuf<double> w, us;
  DevBuf<uint64_t> idx;
  FC_CUDA(w.reserve(n)); FC_CUDA(us.reserve(nu)); FC_CUDA(idx.reserve(n));
  FC_CUDA(cudaMemcpyAsync(w.p, h_w, n * 8, cudaMemcpyHostToDevice, st));
  FC_CUDA(cudaMemcpyAsync(us.p, h_uniforms, nu * 8, cudaMemcpyHostToDevice, st));
  ResampleWork W;
  FC_TRY(resample_dev(method, w.p, (int64_t)n, us.p, idx.p, W, st, nullptr));
  FC_CUDA(cudaMemcpyAsync(h_idx, idx.p, n * 8, cudaMemcpyDeviceToHost, st));
  FC_CUDA(cudaStreamSynchronize(st));
  return FC_OK;
}

int32_t fc_resample_dev(int32_t method, const double* d_w, size_t n, const double* d_uniforms, uint64_t* d_idx, void* stream) {
  using namespace fc;
  static thread_local ResampleWork W;  // grows to the largest n seen on this thread
  return resample_dev(method, d_w, (int64_t)n, d_uniforms, d_idx, W, (cudaStream_t)stream, nullptr);
}

// the sequentially rounded prefix sums themselves (S[0..n]), for tests and multi-GPU carries
int32_t fc_exact_prefix_sums(const double* h_w, size_t n, double* h_S) {
  using namespace fc;
  cudaStream_t st;
  FC_TRY(util_stream(&st));
  DevBuf<double> w;
  ResampleWork W;
  FC_CUDA(w.reserve(n));
  FC_TRY(W.reserve((int64_t)n));
  if (n) FC_CUDA(cudaMemcpyAsync(w.p, h_w, n * 8, cudaMemcpyHostToDevice, st));
  FC_TRY(exact_prefix_sums(w.p, (int64_t)n, W.S.p, W.work.p, W.scratch.p, st, nullptr, 0.0));
  FC_CUDA(cudaMemcpyAsync(h_S, W.S.p, (n + 1) * 8, cudaMemcpyDeviceToHost, st));
  FC_CUDA(cudaStreamSynchronize(st));
  return FC_OK;
}

int32_t fc_log_sum_exp(const double* h_x, size_t n, double* out) {  // numerical.rs:15-38
  using namespace fc;
  if (n == 0) { *out = -INFINITY; return FC_OK; }
  cudaStream_t st;
  FC_TRY(util_stream(&st));
  DevBuf<double> x, scratch;
  FC_CUDA(x.reserve(n)); FC_CUDA(scratch.reserve(RD_BLOCKS + 8));
  FC_CUDA(cudaMemcpyAsync(x.p, h_x, n * 8, cudaMemcpyHostToDevice, st));
  return dev_log_sum_exp(x.p, (int64_t)n, scratch.p, out, st);
}

int32_t fc_ess(const double* h_w, size_t n, double* out) {  // smc.rs:230-233
  using namespace fc;
  cudaStream_t st;
  FC_TRY(util_stream(&st));
  DevBuf<double> x, scratch;
  FC_CUDA(x.reserve(n)); FC_CUDA(scratch.reserve(RD_BLOCKS + 8));
  if (n) FC_CUDA(cudaMemcpyAsync(x.p, h_w, n * 8, cudaMemcpyHostToDevice, st));
  FC_TRY(dev_reduce(2, x.p, (int64_t)n, 0.0, scratch.p, scratch.p + RD_BLOCKS, st));
  double s;
  FC_CUDA(cudaMemcpyAsync(&s, scratch.p + RD_BLOCKS, 8, cudaMemcpyDeviceToHost, st));
  FC_CUDA(cudaStreamSynchronize(st));
  *out = 1.0 / s;
  return FC_OK;
}

int32_t fc_normalize_log_probs(const double* h_x, size_t n, double* h_out) {  // numerical.rs:87-90
  using namespace fc;
  if (n == 0) return FC_OK;
  cudaStream_t st;
  FC_TRY(util_stream(&st));
  DevBuf<double> x, scratch;
  FC_CUDA(x.reserve(n)); FC_CUDA(scratch.reserve(RD_BLOCKS + 8));
  FC_CUDA(cudaMemcpyAsync(x.p, h_x, n * 8, cudaMemcpyHostToDevice, st));
  double lse;
  FC_TRY(dev_log_sum_exp(x.p, (int64_t)n, scratch.p, &lse, st));
  k_exp_shift<<<(unsigned)cdiv((int64_t)n, 256), 256, 0, st>>>(x.p, lse, x.p, (int64_t)n);
  FC_KERNEL_CHECK();
  FC_CUDA(cudaMemcpyAsync(h_out, x.p, n * 8, cudaMemcpyDeviceToHost, st));
  FC_CUDA(cudaStreamSynchronize(st));
  return FC_OK;
}
}
