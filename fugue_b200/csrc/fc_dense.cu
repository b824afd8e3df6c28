// fc_dense.cu -- dense GLM plate: value + analytic gradient of
//     sum_i log p(y_i | link(off_c + sum_k X[i][k] * beta[k][c]), other params_c)     for C chains.
//
// Replaces, for a whole block of chains at once, the reference's per-observation walk
// (examples/classification.rs:118-130 builds eta_i left to right, then Bernoulli::log_prob,
// src/core/distribution.rs:598-619) that hmc.rs:304-329 repeats 2*d times per force.
//
// B200 mapping (sm_100a):
//   * one CTA = 128 chains x a contiguous range of observation rows; 8 warps, warp w owns
//     chains [16w, 16w+16) in both contractions, so the residual tile never leaves the warp.
//   * X is streamed once per CTA in tiles of 32 rows by 1-D bulk TMA (cp.async.bulk ->
//     UBLKCP) into a 2-stage shared-memory ring guarded by mbarriers; the tile of 32 rows is
//     contiguous in the row-major matrix, so one bulk copy moves it.
//   * both contractions  eta = X_tile * beta  and  g += X_tile^T * r  run on the FP64 tensor
//     path (mma.sync.m8n8k4.f64 -> DMMA); beta lives in shared memory for the CTA's lifetime,
//     the D x 128 gradient tile lives in registers (52 doubles / thread at D = 100).
//   * the link + log-density + residual is evaluated on the accumulator fragments in between.
//   * per-CTA partial sums go to [split][acc][C]; fc_score.cu's k_finish folds them in fixed
//     order (no atomics => bit-reproducible run to run).
// FP64 has no tcgen05 path; DMMA is the tensor-core route that meets the 1e-10 gate.
#include "fc_dist.cuh"
#include "fc_fastmath.cuh"
#include <cstdlib>
#include "fc_internal.cuh"

namespace fc {

#define DN_BC 128     // chains per CTA
#define DN_TN 32      // observation rows per tile
#define DN_LDB 132    // beta smem row stride (doubles), == 4 mod 16 -> conflict-free B fragments
#define DN_LDW(NT) ((NT) == 2 ? 20 : 12)  // per-warp residual tile row stride (doubles), == 4 mod 8
#define DN_NW(NT) (16 / (NT))             // warps per CTA: each owns NT n-tiles of 8 chains
#define DN_THREADS(NT) (32 * DN_NW(NT))
#define DN_STAGES 2

enum { GLM_GENERIC = 0, GLM_BERN_SIGMOID = 1, GLM_POISSON_ID = 2 };

struct DenseArgs {
  DOp op;
  AccMap acc;
  const double* X;      // [N][K] row-major
  const int32_t* crow;  // program slot -> sorted row
  const double* q;
  double* part;
  int64_t C, N;
  int64_t rows_per_split;  // multiple of DN_TN * sub
  int32_t K, dj, dt;       // dense width, dense param index, dense term index
  int32_t n_acc;           // accumulators per split in `part`
  int32_t n_red;
  int32_t sub;             // 32-row sub-tiles per pipeline stage (1 for wide matrices; > 1 amortises the stage hand-over when K is small)
};

__device__ __forceinline__ void dmma884(double& d0, double& d1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};" : "+d"(d0), "+d"(d1) : "d"(a), "d"(b));
}
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  uint32_t done;
  do {
    asm volatile("{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\nselp.u32 %0, 1, 0, p;\n}\n" : "=r"(done) : "r"(smem_u32(bar)), "r"(parity) : "memory");
  } while (!done);
}
__device__ __forceinline__ void tma_bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}

__device__ __forceinline__ void load_value(const DCol& c, int64_t i, bool cont, double& x, long long& xi) {
  x = 0.0;
  xi = 0;
  switch (c.dtype) {
    case FC_COL_F64: { double v = __ldg((const double*)c.ptr + i); if (cont) x = v; else xi = (long long)v; } break;
    case FC_COL_U8: { long long v = __ldg((const uint8_t*)c.ptr + i); if (cont) x = (double)v; else xi = v; } break;
    case FC_COL_U32: { long long v = __ldg((const uint32_t*)c.ptr + i); if (cont) x = (double)v; else xi = v; } break;
    default: { long long v = __ldg((const long long*)c.ptr + i); if (cont) x = (double)v; else xi = v; } break;
  }
}

__device__ __forceinline__ double link_apply(int link, double s, double lo, double hi, double& dl) {
  switch (link) {
    case FC_LINK_EXP: { double e = exp(s); dl = e; return e; }
    case FC_LINK_SIGMOID_CLAMPED: {
      double p = 1.0 / (1.0 + exp(-s));
      dl = (p > lo && p < hi) ? p * (1.0 - p) : 0.0;
      return p < lo ? lo : (p > hi ? hi : p);
    }
    case FC_LINK_CLAMP: dl = (s > lo && s < hi) ? 1.0 : 0.0; return s < lo ? lo : (s > hi ? hi : s);
    default: dl = 1.0; return s;
  }
}


// generic distribution / link (kept out of line: it is instantiated once, not 16x per tile)
template <bool GRAD>
__device__ __noinline__ void glm_generic_elem(const DOp& op, int dj, double sarg, double p0, double p1, double p2, double l0, double l1, double l2,
                                              double x, long long xi, double lg, double& lp, double& w, double* other) {
  double pv[3] = {p0, p1, p2}, dl[3] = {l0, l1, l2};
  double dlk;
  double val = link_apply(op.p[dj].link, sarg, op.p[dj].lo, op.p[dj].hi, dlk);
  pv[dj] = val;
  dl[dj] = dlk;
  DistOut o = dist_eval<GRAD>(op.dist, pv[0], pv[1], pv[2], x, xi, lg);
  lp = o.lp;
  w = 0.0;
  other[0] = other[1] = other[2] = 0.0;
  if (GRAD) {
    double dpj[3] = {o.dp0, o.dp1, o.dp2};
    for (int pj = 0; pj < 3; pj++) {
      double ww = dpj[pj] * dl[pj];
      if (pj == dj) w = ww;
      else other[pj] = ww;
    }
  }
}

// MT = number of 8-row gradient tiles per warp (K padded to 8*MT); GRAD=false skips the
// second contraction (value only: MH / SMC).
// VAL=false (with GRAD) is the force-only pass of an interior leapfrog step: the log-density of the
// element is not evaluated (no log, no clamp), the lp partial is written as 0 and never read.
template <int MODE, bool GRAD, int MT, int NT, bool VAL = true>
__global__ void __launch_bounds__(DN_THREADS(NT), GRAD ? 1 : 2) k_dense_glm(const __grid_constant__ DenseArgs A) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  constexpr int NW = DN_NW(NT), NTHR = DN_THREADS(NT), LDW = DN_LDW(NT), WC = 8 * NT;  // WC = chains per warp
  const int K = A.K;
  const int KP4 = (K + 3) & ~3;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int gid = lane >> 2, tig = lane & 3;  // mma "groupID" / "threadID_in_group"
  const int64_t C = A.C;
  const int64_t cb0 = (int64_t)blockIdx.x * DN_BC;
  const DOp& op = A.op;

  // ---- shared memory carve-up ----
  int* done = (int*)(smem_raw + 64);                                       // per-stage warp arrival counters
  uint64_t* bars = (uint64_t*)smem_raw;                                   // DN_STAGES mbarriers
  double* s_beta = (double*)(smem_raw + 128);                             // [KP4][DN_LDB]
  double* s_w = s_beta + (size_t)KP4 * DN_LDB;                            // [8 warps][DN_TN][DN_LDW]
  double* s_x = s_w + NW * DN_TN * LDW;                                 // [DN_STAGES][DN_TN*K + pad]
  constexpr int SUB = MT == 1 ? 8 : (MT == 2 ? 4 : (MT == 4 ? 2 : 1));  // == dense_sub(K): MT is chosen from K on the host
  constexpr int ST_ROWS = DN_TN * SUB;                                   // rows per pipeline stage
  const int x_stage_elems = ((ST_ROWS * K + 8 + 15) & ~15);               // + zero pad for k >= K reads

  const int64_t row_begin = (int64_t)blockIdx.y * A.rows_per_split;
  const int64_t row_end = min(A.N, row_begin + A.rows_per_split);
  const int n_tiles = (int)((row_end - row_begin + ST_ROWS - 1) / ST_ROWS);  // stages

  __shared__ double s_exp_tab[32];  // 2^(j/32) for fast_exp_tab (made visible by the __syncthreads below)
  __shared__ double s_log_tab[(MODE != GLM_GENERIC && VAL) ? 364 : 1];  // 1/c, ln c for fast_log_tab
  if (MODE == GLM_BERN_SIGMOID) fast_exp_tab_init(s_exp_tab, tid);
  if (MODE != GLM_GENERIC && VAL) fast_log_tab_init(s_log_tab, tid, NTHR);
  if (tid == 0) {
    for (int s = 0; s < DN_STAGES; s++) { mbar_init(&bars[s], 1); done[s] = 0; }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  // beta tile: s_beta[k][cl] = q[crow[slot+k]][cb0+cl]; zero for padded k / chains
  const int dslot = op.p[A.dj].t[A.dt].slot;
  for (int idx = tid; idx < KP4 * DN_BC; idx += NTHR) {
    int k = idx / DN_BC, cl = idx % DN_BC;
    double v = 0.0;
    if (k < K && cb0 + cl < C) v = A.q[(int64_t)A.crow[dslot + k] * C + cb0 + cl];
    s_beta[k * DN_LDB + cl] = v;
  }
  // zero the pad after each X stage once (bulk copies never touch it)
  for (int s = 0; s < DN_STAGES; s++)
    for (int idx = ST_ROWS * K + tid; idx < x_stage_elems; idx += NTHR) s_x[(size_t)s * x_stage_elems + idx] = 0.0;
  __syncthreads();

  auto issue_tile = [&](int t) {  // called by thread 0 only, full tiles only
    int s = t % DN_STAGES;
    const double* src = A.X + (row_begin + (int64_t)t * ST_ROWS) * K;
    uint32_t bytes = (uint32_t)(ST_ROWS * K * sizeof(double));
    mbar_expect_tx(&bars[s], bytes);
    tma_bulk_g2s(s_x + (size_t)s * x_stage_elems, src, bytes, &bars[s]);
  };
  auto tile_rows = [&](int t) -> int { return (int)min((int64_t)ST_ROWS, row_end - (row_begin + (int64_t)t * ST_ROWS)); };
  // bulk copies need 16-byte sizes; a partial (last) stage is copied cooperatively instead
  auto tile_is_bulk = [&](int t) -> bool { return tile_rows(t) == ST_ROWS && ((ST_ROWS * K) % 2 == 0); };

  if (tid == 0) {
    for (int t = 0; t < DN_STAGES && t < n_tiles; t++)
      if (tile_is_bulk(t)) issue_tile(t);
  }

  // ---- per-thread chain columns: c = cb0 + warp*16 + nt*8 + tig*2 + j ----
  double off[NT][2];                 // CONST/LATENT part of the dense parameter
  double po[3][NT][2], pdl[3][NT][2]; // the other (per-chain) parameters, linked, and dlink
#pragma unroll
  for (int nt = 0; nt < NT; nt++)
#pragma unroll
    for (int j = 0; j < 2; j++) {
      int64_t c = cb0 + warp * WC + nt * 8 + tig * 2 + j;
      bool ok = c < C;
      double s = 0.0;
#pragma unroll
      for (int t = 0; t < 3; t++) {
        const DTerm& tm = op.p[A.dj].t[t];
        if (tm.kind == FC_TERM_CONST) s += tm.c;
        else if (tm.kind == FC_TERM_LATENT && ok) s += tm.c * A.q[(int64_t)tm.row * C + c];
      }
      off[nt][j] = s;
#pragma unroll
      for (int pj = 0; pj < 3; pj++) {
        po[pj][nt][j] = 0.0;
        pdl[pj][nt][j] = 0.0;
        if (MODE == GLM_GENERIC && pj < op.n_params && pj != A.dj) {
          double sp = 0.0;
#pragma unroll
          for (int t = 0; t < 3; t++) {
            const DTerm& tm = op.p[pj].t[t];
            if (tm.kind == FC_TERM_CONST) sp += tm.c;
            else if (tm.kind == FC_TERM_LATENT && ok) sp += tm.c * A.q[(int64_t)tm.row * C + c];
          }
          po[pj][nt][j] = link_apply(op.p[pj].link, sp, op.p[pj].lo, op.p[pj].hi, pdl[pj][nt][j]);
        }
      }
    }

  double g[MT][NT][2];  // gradient tile fragments: k = mt*8+gid, c as above
#pragma unroll
  for (int mt = 0; mt < MT; mt++)
#pragma unroll
    for (int nt = 0; nt < NT; nt++) g[mt][nt][0] = g[mt][nt][1] = 0.0;
  double a_lp[NT][2], a_w[NT][2], a_p[3][NT][2];
#pragma unroll
  for (int nt = 0; nt < NT; nt++)
#pragma unroll
    for (int j = 0; j < 2; j++) {
      a_lp[nt][j] = a_w[nt][j] = 0.0;
      a_p[0][nt][j] = a_p[1][nt][j] = a_p[2][nt][j] = 0.0;
    }

  double* my_w = s_w + warp * DN_TN * LDW;
  const double qnan = __longlong_as_double(0x7ff8000000000000LL);
  const double lnk_lo = op.p[A.dj].lo, lnk_hi = op.p[A.dj].hi;
  const int lnk = op.p[A.dj].link;

  // observed values of one 32-row sub-tile: this thread's 4 rows (mt*8 + gid). They are loaded one
  // sub-tile ahead, so their L1/L2 latency hides behind the previous sub-tile's arithmetic.
  // The value travels as a double (discrete values are exact below 2^53; the generic path converts
  // back), the dtype switch is resolved once per sub-tile, not per row.
#define y_dtype (op.value_col.dtype)
#define y_ptr (op.value_col.ptr)
#define lgam_ptr (op.lgam)
  auto load_y = [&](int64_t rr0, int vrows, double (&yv_)[4], double (&ylg_)[4]) {
    int64_t ri[4];
#pragma unroll
    for (int mt = 0; mt < 4; mt++) { ri[mt] = rr0 + min(mt * 8 + gid, max(vrows - 1, 0)); yv_[mt] = 0.0; ylg_[mt] = qnan; }
    if (vrows <= 0) return;
    if (y_dtype == FC_COL_U32) {
#pragma unroll
      for (int mt = 0; mt < 4; mt++) yv_[mt] = (double)__ldg((const uint32_t*)y_ptr + ri[mt]);
    } else if (y_dtype == FC_COL_U8) {
#pragma unroll
      for (int mt = 0; mt < 4; mt++) yv_[mt] = (double)__ldg((const uint8_t*)y_ptr + ri[mt]);
    } else if (y_dtype == FC_COL_F64) {
#pragma unroll
      for (int mt = 0; mt < 4; mt++) { double v = __ldg((const double*)y_ptr + ri[mt]); yv_[mt] = op.cont ? v : (double)(long long)v; }
    } else {
#pragma unroll
      for (int mt = 0; mt < 4; mt++) yv_[mt] = (double)__ldg((const long long*)y_ptr + ri[mt]);
    }
    if (MODE == GLM_GENERIC && lgam_ptr) {  // the Poisson path subtracts the CTA's lgamma total once (lgam_total)
#pragma unroll
      for (int mt = 0; mt < 4; mt++) ylg_[mt] = __ldg(lgam_ptr + ri[mt]);
    }
  };
#undef y_dtype
#undef y_ptr
#undef lgam_ptr
  // Poisson: sum_i lgamma(y_i + 1) over this CTA's rows does not depend on the chain. It is summed
  // once per CTA (coalesced, fixed tree order) and subtracted from every chain's total in the
  // epilogue, instead of one load + one add per (row, chain) pair in the hot loop.
  double lgam_total = 0.0;
  if (MODE == GLM_POISSON_ID && VAL) {
    __shared__ double s_lg[32];
    double v = 0.0;
    for (int64_t i = row_begin + tid; i < row_end; i += NTHR) v += __ldg(op.lgam + i);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if (lane == 0) s_lg[warp] = v;
    __syncthreads();
    for (int w = 0; w < NW; w++) lgam_total += s_lg[w];
  }
  // K == 8 fast path of contraction 1 (see there): beta fragments for k = 2*tig, 2*tig+1
  const bool k8_path = (MT == 1) && K == 8;
  double b8[NT][2];
#pragma unroll
  for (int nt = 0; nt < NT; nt++) {
    b8[nt][0] = k8_path ? s_beta[(2 * tig) * DN_LDB + warp * WC + nt * 8 + gid] : 0.0;
    b8[nt][1] = k8_path ? s_beta[(2 * tig + 1) * DN_LDB + warp * WC + nt * 8 + gid] : 0.0;
  }
  double nyv[4], nylg[4];
  if (SUB > 1) load_y(row_begin, (int)min((int64_t)DN_TN, row_end - row_begin), nyv, nylg);

  for (int t = 0; t < n_tiles; t++) {
    const int s = t % DN_STAGES;
    double* xs0 = s_x + (size_t)s * x_stage_elems;
    const int64_t st_r0 = row_begin + (int64_t)t * ST_ROWS;
    const int st_valid = tile_rows(t);
    if (SUB == 1) load_y(st_r0, st_valid, nyv, nylg);  // issued before the wait to overlap with it
    if (tile_is_bulk(t)) {
      mbar_wait(&bars[s], (uint32_t)((t / DN_STAGES) & 1));
    } else {
      // cooperative copy + zero fill of the missing rows
      __syncthreads();  // every warp has left stage t-1 (and so t-2, the buffer's previous user)
      const double* src = A.X + st_r0 * K;
      for (int idx = tid; idx < ST_ROWS * K; idx += NTHR) xs0[idx] = (idx < st_valid * K) ? __ldg(src + idx) : 0.0;
      __syncthreads();
    }
   for (int sb = 0; sb < SUB; sb++) {
    const int valid = min(DN_TN, st_valid - sb * DN_TN);
    if (valid <= 0) break;
    double* xs = xs0 + (size_t)sb * DN_TN * K;
    const int64_t r0 = st_r0 + (int64_t)sb * DN_TN;
    double yv[4], ylg[4];  // rows beyond `valid` hold a copy of the last valid row; they are masked below
#pragma unroll
    for (int mt = 0; mt < 4; mt++) { yv[mt] = nyv[mt]; ylg[mt] = nylg[mt]; }
    bool yb[4];  // Bernoulli outcome as a predicate (keeps the per-pair test off the FP64 pipe)
#pragma unroll
    for (int mt = 0; mt < 4; mt++) yb[mt] = yv[mt] != 0.0;
    if (SUB > 1) {
      const int64_t nr0 = r0 + DN_TN;
      load_y(nr0, (int)max((int64_t)0, min((int64_t)DN_TN, row_end - nr0)), nyv, nylg);
    }

    // ---- contraction 1: eta[32][16 per warp] = X_tile[32][K] * beta[K][16] ----
    double eta[4][NT][2];
#pragma unroll
    for (int mt = 0; mt < 4; mt++)
#pragma unroll
      for (int nt = 0; nt < NT; nt++) eta[mt][nt][0] = eta[mt][nt][1] = 0.0;
    if (MT == 1 && k8_path) {
      // K == 8: the contraction index is permuted (step 0 takes k = 2*tig, step 1 takes k = 2*tig+1;
      // a sum over k does not care), so one 16-byte load per row feeds both k-steps. A warp's 32
      // loads of one 8-row group cover 512 contiguous bytes: no bank conflicts (the k4 + tig
      // pattern is 4-way conflicted at a row stride of 8 doubles). beta fragments stay in registers.
#pragma unroll
      for (int mt = 0; mt < 4; mt++) {
        const double2 av = *(const double2*)&xs[(mt * 8 + gid) * 8 + 2 * tig];
#pragma unroll
        for (int nt = 0; nt < NT; nt++) {
          dmma884(eta[mt][nt][0], eta[mt][nt][1], av.x, b8[nt][0]);
          dmma884(eta[mt][nt][0], eta[mt][nt][1], av.y, b8[nt][1]);
        }
      }
    } else
#pragma unroll 5
    for (int k4 = 0; k4 < KP4; k4 += 4) {
      double a[4], b[NT];
#pragma unroll
      for (int mt = 0; mt < 4; mt++) a[mt] = xs[(mt * 8 + gid) * K + k4 + tig];
#pragma unroll
      for (int nt = 0; nt < NT; nt++) b[nt] = s_beta[(k4 + tig) * DN_LDB + warp * WC + nt * 8 + gid];
#pragma unroll
      for (int mt = 0; mt < 4; mt++)
#pragma unroll
        for (int nt = 0; nt < NT; nt++) dmma884(eta[mt][nt][0], eta[mt][nt][1], a[mt], b[nt]);
    }

    // ---- link + log-density + residual on the accumulator fragments ----
#pragma unroll
    for (int mt = 0; mt < 4; mt++) {
      const bool rv = (mt * 8 + gid) < valid;
#pragma unroll
      for (int nt = 0; nt < NT; nt++) {
        double wv[2];
#pragma unroll
        for (int j = 0; j < 2; j++) {
          double sarg = off[nt][j] + eta[mt][nt][j];
          double lp = 0.0, w = 0.0;
          if (MODE == GLM_BERN_SIGMOID) {
            // classification.rs:124-130: p = 1/(1+exp(-eta)); clamp; Bernoulli::log_prob.
            // Same formula, branch-free primitives (host guarantees 1e-300 < lo <= hi < 1).
            double p = fast_rcp(1.0 + fast_exp_tab(-sarg, s_exp_tab));
            bool inside = (p > lnk_lo && p < lnk_hi);
            if (VAL) {
              double pc = fmin(fmax(p, lnk_lo), lnk_hi);
              lp = fast_log_tab(yb[mt] ? pc : 1.0 - pc, s_log_tab);
            }
            if (GRAD) w = inside ? (yb[mt] ? (1.0 - p) : -p) : (p - p);  // clamped: 0; NaN eta: NaN force => divergent (hmc.rs:384-386)
            if (VAL && sarg != sarg) lp = d_neg_inf();  // NaN eta => Bernoulli(NaN) => -inf
          } else if (MODE == GLM_POISSON_ID) {
            // distribution.rs:1237-1258 with lambda = eta. Branch-light: the fast path is evaluated
            // unconditionally on a clamped lambda, the reference's guards are selects, and only an
            // out-of-range lambda takes the (rare) library-function branch.
            const double lam = sarg;
            const double kf = yv[mt];
            // fast-path range as one unsigned compare on the high word: positive, normal, below 512
            // (exponent field in [2, 0x408)). Any in-range cut works: everything else -- including
            // the reference's lambda > 700 && k == 0 special case -- takes the exact branch below.
            const bool ok = ((unsigned)__double2hiint(lam) - 0x00200000u < 0x40600000u) && kf >= 0.0;
            const double lamc = ok ? lam : 1.0;
            if (VAL) lp = kf * fast_log_tab(lamc, s_log_tab) - lam;  // - lgamma(k+1): lgam_total, epilogue
            if (GRAD) w = (!VAL || d_fin(lp)) ? kf * fast_rcp(lamc) - 1.0 : 0.0;  // ok => lp is finite (lgamma column is)
            if (!ok) {
              lp = d_neg_inf(); w = __longlong_as_double(0x7ff8000000000000LL);  // outside the support: no gradient
              if (lam > 0.0 && d_fin(lam) && kf >= 0.0) {  // tiny / huge lambda: exact guards via the library functions
                lp = (lam > 700.0 && kf == 0.0) ? -lam : kf * log(lam) - lam;
                w = (GRAD && d_fin(lp)) ? kf / lam - 1.0 : w;
              }
            }
          } else {
            double op_other[3];
            glm_generic_elem<GRAD>(op, A.dj, sarg, po[0][nt][j], po[1][nt][j], po[2][nt][j], pdl[0][nt][j], pdl[1][nt][j], pdl[2][nt][j], yv[mt], (long long)yv[mt], ylg[mt], lp, w, op_other);
            if (GRAD && rv) {
#pragma unroll
              for (int pj = 0; pj < 3; pj++) a_p[pj][nt][j] += op_other[pj];
            }
          }
          if (rv) { a_lp[nt][j] += lp; a_w[nt][j] += w; }
          else w = 0.0;
          wv[j] = w;
        }
        if (GRAD) *(double2*)&my_w[(mt * 8 + gid) * LDW + nt * 8 + tig * 2] = make_double2(wv[0], wv[1]);
      }
    }

    // ---- contraction 2: g[K][16 per warp] += X_tile^T[K][32] * r[32][16] ----
    if (GRAD) {
      __syncwarp();
#pragma unroll
      for (int i4 = 0; i4 < DN_TN; i4 += 4) {
        double b[NT];
#pragma unroll
        for (int nt = 0; nt < NT; nt++) b[nt] = my_w[(i4 + tig) * LDW + nt * 8 + gid];
        const double* xr = xs + (i4 + tig) * K + gid;
#pragma unroll
        for (int mt = 0; mt < MT; mt++) {
          if (mt * 8 < K) {  // k >= K overshoots by < 8 doubles into the zero pad only
            double a = xr[mt * 8];
#pragma unroll
            for (int nt = 0; nt < NT; nt++) dmma884(g[mt][nt][0], g[mt][nt][1], a, b[nt]);
          }
        }
      }
      __syncwarp();
    }

   }  // sub-tiles of this stage

    // ---- release the stage and refill it ----
    // Warps are NOT barriered against each other: the last warp to leave the stage refills it, so a
    // warp in its link/log-density phase overlaps other warps' DMMA phases (they drift <= 1 tile).
    __syncwarp();
    if (lane == 0) {
      int old = atomicAdd(&done[s], 1);
      if (old == NW - 1) {
        atomicExch(&done[s], 0);
        if (t + DN_STAGES < n_tiles && tile_is_bulk(t + DN_STAGES)) {
          asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
          issue_tile(t + DN_STAGES);
        }
      }
    }
  }

  // ---- write this CTA's partial: lp / reduced scalars (sum over gid via shuffles), then g ----
  const int lp_slot = op.is_sample ? 0 : 1;
  double* part = A.part + (int64_t)blockIdx.y * A.n_acc * C;
#pragma unroll
  for (int nt = 0; nt < NT; nt++)
#pragma unroll
    for (int j = 0; j < 2; j++) {
      double v_lp = a_lp[nt][j], v_w = a_w[nt][j], v_p[3] = {a_p[0][nt][j], a_p[1][nt][j], a_p[2][nt][j]};
#pragma unroll
      for (int o = 4; o < 32; o <<= 1) {
        v_lp += __shfl_xor_sync(0xffffffffu, v_lp, o);
        v_w += __shfl_xor_sync(0xffffffffu, v_w, o);
#pragma unroll
        for (int pj = 0; pj < 3; pj++) v_p[pj] += __shfl_xor_sync(0xffffffffu, v_p[pj], o);
      }
      int64_t c = cb0 + warp * WC + nt * 8 + tig * 2 + j;
      if (gid == 0 && c < C) {
        if (MODE == GLM_POISSON_ID && VAL) v_lp -= lgam_total;  // -inf stays -inf
        for (int a = 0; a < 3; a++) part[(int64_t)a * C + c] = (a == lp_slot) ? v_lp : 0.0;
        if (GRAD) {
          for (int a = 3; a < 3 + A.n_red; a++) {
            double v = 0.0;
#pragma unroll
            for (int pj = 0; pj < 3; pj++)
#pragma unroll
              for (int t = 0; t < 3; t++)
                if (A.acc.a_idx[pj][t] == a) v += ((pj == A.dj) ? v_w : v_p[pj]) * (op.p[pj].t[t].kind == FC_TERM_LATENT ? op.p[pj].t[t].c : 1.0);
            part[(int64_t)a * C + c] = v;
          }
        }
      }
    }
  if (GRAD) {
#pragma unroll
    for (int mt = 0; mt < MT; mt++) {
      int k = mt * 8 + gid;
      if (k < K) {
#pragma unroll
        for (int nt = 0; nt < NT; nt++)
#pragma unroll
          for (int j = 0; j < 2; j++) {
            int64_t c = cb0 + warp * WC + nt * 8 + tig * 2 + j;
            if (c < C) part[(int64_t)(3 + A.n_red + k) * C + c] = g[mt][nt][j];
          }
      }
    }
  }
}

// 32-row sub-tiles per pipeline stage: ~16 KB stages for narrow matrices, 1 from K = 64 up
static int dense_sub(int K) { return K <= 8 ? 8 : (K <= 16 ? 4 : (K <= 32 ? 2 : 1)); }

static size_t dense_smem_bytes(int K, int NT) {
  int KP4 = (K + 3) & ~3;
  size_t x_stage = ((size_t)DN_TN * dense_sub(K) * K + 8 + 15) & ~(size_t)15;
  return 128 + sizeof(double) * ((size_t)KP4 * DN_LDB + DN_NW(NT) * DN_TN * DN_LDW(NT) + DN_STAGES * x_stage);
}

// warps per CTA = 16 / NT: NT = 1 -> 16 warps (8 chains each, <= 128 registers), NT = 2 -> 8 warps
static int g_dense_nt = getenv("FC_DENSE_NT") ? atoi(getenv("FC_DENSE_NT")) : 1;
static int g_dense_sms = 0;
int64_t dense_plate_splits(const fc_program* P, const PlateInfo& pl, int64_t C) {
  if (!g_dense_sms) {
    cudaDeviceGetAttribute(&g_dense_sms, cudaDevAttrMultiProcessorCount, P->device);
    if (g_dense_sms <= 0) g_dense_sms = 148;
  }
  const int K = pl.dop.p[pl.dense_param].t[pl.dense_term].col.width;
  int64_t n_tiles = cdiv(pl.dop.plate_len, (int64_t)DN_TN * dense_sub(K));  // pipeline stages
  int64_t ncb = cdiv(C, DN_BC);
  // grid = ncb x S CTAs, one CTA per SM at a time: make it a whole number of waves when the
  // plate is long enough (148 = 4*37), with at least 8 tiles per CTA.
  auto gcd = [](int64_t a, int64_t b) { while (b) { int64_t t = a % b; a = b; b = t; } return a; };
  int64_t s0 = g_dense_sms / gcd(g_dense_sms, ncb);
  int64_t max_s = std::max<int64_t>(1, n_tiles / 8);
  int64_t S = s0;
  while (ncb * S < 2 * g_dense_sms) S += s0;
  if (S > max_s) S = max_s;
  return std::max<int64_t>(1, S);
}

template <int MODE, bool GRAD, int NT, bool VAL = true>
static int32_t launch_mt(const DenseArgs& A, dim3 grid, size_t smem, cudaStream_t st) {
  int MT = (A.K + 7) / 8;
#define DN_LAUNCH(M)                                                                                          \
  {                                                                                                           \
    FC_CUDA(cudaFuncSetAttribute(k_dense_glm<MODE, GRAD, M, NT, VAL>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)); \
    k_dense_glm<MODE, GRAD, M, NT, VAL><<<grid, DN_THREADS(NT), smem, st>>>(A);                                            \
  }
  // value-only kernels ignore MT except for the sub-tiling it implies (8 / 4 / 2 / 1 sub-tiles per stage)
  if (!GRAD) { if (A.K <= 8) DN_LAUNCH(1) else if (A.K <= 16) DN_LAUNCH(2) else if (A.K <= 32) DN_LAUNCH(4) else DN_LAUNCH(13) }
  else if (MT <= 1) DN_LAUNCH(1)
  else if (MT <= 2) DN_LAUNCH(2)
  else if (MT <= 4) DN_LAUNCH(4)
  else if (MT <= 7) DN_LAUNCH(7)
  else if (MT <= 10) DN_LAUNCH(10)
  else DN_LAUNCH(13)
#undef DN_LAUNCH
  FC_KERNEL_CHECK();
  return FC_OK;
}

int32_t dense_plate_launch(fc_program* P, PlateInfo& pl, const double* d_q, int64_t C, double* d_part, int64_t max_splits, bool grad, int64_t* n_splits, bool need_value) {
  const DOp& op = pl.dop;
  const DTerm& dt = op.p[pl.dense_param].t[pl.dense_term];
  int K = dt.col.width;
  if (K > 104) { set_error("dense GLM plate: matrix width %d > 104 is not supported yet", K); return FC_ERR_UNSUPPORTED_OP; }
  DenseArgs A;
  A.op = op; A.acc = pl.acc; A.X = (const double*)dt.col.ptr; A.crow = P->d_crow.p; A.q = d_q; A.part = d_part;
  A.C = C; A.N = op.plate_len; A.K = K; A.dj = pl.dense_param; A.dt = pl.dense_term;
  A.n_red = pl.acc.n_acc - 3;
  A.n_acc = grad ? 3 + A.n_red + K : 3;
  A.sub = dense_sub(K);
  const int64_t st_rows = (int64_t)DN_TN * A.sub;
  int64_t n_tiles = cdiv(A.N, st_rows);
  int64_t S = std::max<int64_t>(1, std::min(max_splits, n_tiles));
  int64_t tiles_per_split = cdiv(n_tiles, S);
  A.rows_per_split = tiles_per_split * st_rows;
  S = cdiv(A.N, A.rows_per_split);
  *n_splits = S;
  dim3 grid((unsigned)cdiv(C, DN_BC), (unsigned)S);
  const int NT = g_dense_nt;
  size_t smem = dense_smem_bytes(K, NT);
  // single-parameter fast paths (link and distribution fixed at compile time)
  bool only_dense_param = true;
  for (int t = 0; t < 3; t++) {
    int k = op.p[pl.dense_param].t[t].kind;
    (void)k;
  }
  int mode = GLM_GENERIC;
  if (op.dist == FC_BERNOULLI && op.p[0].link == FC_LINK_SIGMOID_CLAMPED && only_dense_param && op.p[0].lo > 1e-300 && op.p[0].hi < 1.0 && op.p[0].lo <= op.p[0].hi) mode = GLM_BERN_SIGMOID;
  if (op.dist == FC_POISSON && op.p[0].link == FC_LINK_IDENTITY && op.lgam) mode = GLM_POISSON_ID;
  int32_t rc;
  cudaEvent_t e0 = nullptr, e1 = nullptr;
  if (P->timing) {
    FC_CUDA(cudaEventCreate(&e0));
    FC_CUDA(cudaEventCreate(&e1));
    FC_CUDA(cudaEventRecord(e0, P->stream));
  }
#define DN_DISPATCH(MODE_)                                                                              \
  rc = NT == 1 ? (grad ? launch_mt<MODE_, true, 1>(A, grid, smem, P->stream) : launch_mt<MODE_, false, 1>(A, grid, smem, P->stream)) \
               : (grad ? launch_mt<MODE_, true, 2>(A, grid, smem, P->stream) : launch_mt<MODE_, false, 2>(A, grid, smem, P->stream));
  if (grad && !need_value && NT == 1 && mode == GLM_BERN_SIGMOID) rc = launch_mt<GLM_BERN_SIGMOID, true, 1, false>(A, grid, smem, P->stream);
  else if (grad && !need_value && NT == 1 && mode == GLM_POISSON_ID) rc = launch_mt<GLM_POISSON_ID, true, 1, false>(A, grid, smem, P->stream);
  else if (mode == GLM_BERN_SIGMOID) { DN_DISPATCH(GLM_BERN_SIGMOID) }
  else if (mode == GLM_POISSON_ID) { DN_DISPATCH(GLM_POISSON_ID) }
  else { DN_DISPATCH(GLM_GENERIC) }
#undef DN_DISPATCH
  if (P->timing) {
    FC_CUDA(cudaEventRecord(e1, P->stream));
    P->timing_events.emplace_back(e0, e1);
  }
  if (rc == FC_OK) P->n_kernel_launches += 1;
  return rc;
}

}  // namespace fc

// accuracy probe for fc_fastmath.cuh (tests/test_gpu_fastmath.py): op 0 exp, 1 log, 2 reciprocal
namespace fc {
__global__ void k_debug_fastmath(int op, const double* __restrict__ x, int64_t n, double* __restrict__ out) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  __shared__ double tab[32];
  __shared__ double ltab[364];
  fast_exp_tab_init(tab, threadIdx.x);
  fast_log_tab_init(ltab, threadIdx.x, blockDim.x);
  __syncthreads();
  if (i >= n) return;
  double v = x[i];
  out[i] = op == 0 ? fast_exp(v) : (op == 1 ? fast_log(v) : (op == 2 ? fast_rcp(v) : (op == 3 ? fast_exp_tab(v, tab) : fast_log_tab(v, ltab))));
}
}  // namespace fc
extern "C" int32_t fc_debug_fastmath(int32_t op, const double* h_x, size_t n, double* h_out) {
  fc::DevBuf<double> x, o;
  FC_CUDA(x.reserve(n)); FC_CUDA(o.reserve(n));
  FC_CUDA(cudaMemcpy(x.p, h_x, n * 8, cudaMemcpyHostToDevice));
  fc::k_debug_fastmath<<<(unsigned)fc::cdiv((int64_t)n, 256), 256>>>(op, x.p, (int64_t)n, o.p);
  FC_KERNEL_CHECK();
  FC_CUDA(cudaMemcpy(h_out, o.p, n * 8, cudaMemcpyDeviceToHost));
  return FC_OK;
}
