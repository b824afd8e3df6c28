// fc_plate.cuh -- the specialised plate kernel (shared by the fc_plate_*.cu instantiation units)
// and the helpers it shares with the generic kernels of fc_score.cu.
#pragma once
#include "fc_dist.cuh"
#include "fc_internal.cuh"

namespace fc {

__device__ __forceinline__ double dcol_f64(const DCol& c, int64_t i) {
  switch (c.dtype) {
    case FC_COL_F64: return __ldg((const double*)c.ptr + i);
    case FC_COL_U8: return (double)__ldg((const uint8_t*)c.ptr + i);
    case FC_COL_U32: return (double)__ldg((const uint32_t*)c.ptr + i);
    default: return (double)__ldg((const long long*)c.ptr + i);
  }
}
__device__ __forceinline__ long long dcol_i64(const DCol& c, int64_t i) {
  switch (c.dtype) {
    case FC_COL_F64: return (long long)__ldg((const double*)c.ptr + i);
    case FC_COL_U8: return __ldg((const uint8_t*)c.ptr + i);
    case FC_COL_U32: return __ldg((const uint32_t*)c.ptr + i);
    default: return __ldg((const long long*)c.ptr + i);
  }
}

// link(s) and d link / d s. SIGMOID_CLAMPED follows examples/classification.rs:124-127
// (p = 1/(1+exp(-s)) then clamp); the derivative is that of the CLAMPED function.
__device__ __forceinline__ double apply_link(int link, double s, double lo, double hi, double& dl) {
  switch (link) {
    case FC_LINK_EXP: { double e = exp(s); dl = e; return e; }
    case FC_LINK_SIGMOID_CLAMPED: {
      double p = 1.0 / (1.0 + exp(-s));
      dl = (p > lo && p < hi) ? p * (1.0 - p) : 0.0;
      return p < lo ? lo : (p > hi ? hi : p);
    }
    case FC_LINK_CLAMP: {
      dl = (s > lo && s < hi) ? 1.0 : 0.0;
      return s < lo ? lo : (s > hi ? hi : s);
    }
    default: dl = 1.0; return s;
  }
}

// dist_eval with the op's lowered extras: the constant normaliser and the K-way Categorical table
// (log probs[x] if in range and > 0, else -inf: distribution.rs:785-791; constant probabilities carry
// no parameter gradient)
template <bool GRAD>
__device__ __forceinline__ DistOut op_dist_eval(const DOp& op, double P0, double P1, double P2, double x, long long xi, double lgam_kp1) {
  if (op.dist == FC_CATEGORICAL && op.n_probs > 0) {
    DistOut g;
    g.dx = g.dp0 = g.dp1 = g.dp2 = 0.0;
    const double pr = (xi >= 0 && xi < (long long)op.n_probs) ? __ldg(op.probs + xi) : 0.0;
    g.lp = pr > 0.0 ? log(pr) : d_neg_inf();
    return g;
  }
  return dist_eval<GRAD>(op.dist, P0, P1, P2, x, xi, lgam_kp1, op.lgc, op.lgc2);
}

#define PL_TX 32
#define PL_TY 8
#define PL_PF 4  // prefetch distance in plate iterations

__device__ __forceinline__ void prefetch_l1(const void* p) { asm volatile("prefetch.global.L1 [%0];" ::"l"(p)); }

struct PlateArgs {
  DOp op;
  AccMap acc;
  const int32_t* crow;
  const int32_t* drow;
  const double* q;
  const disc_t* qd;
  double* grad;
  double* part;           // [n_splits][n_acc][C]
  const int32_t* grp_ptr; // CSR by gather index (or null)
  const int32_t* grp_elems;
  int64_t C;
  int64_t n_units;        // groups if CSR else elements
  int64_t units_per_split;
};

// ---------------------------------------------------------------------------------------
// specialised plate kernel: at most two element-dependent terms whose KINDS are compile-time
// (everything else about the op is hoisted out of the plate loop), Normal on a fast path.
// The generic k_plate above costs ~220 issued instructions per (element, chain); this one ~25.
// Eligibility and dispatch: plate_fast_launch(). Same partial layout and summation structure.
// ---------------------------------------------------------------------------------------
enum { EK_NONE = 0, EK_LATENT_X = 1, EK_GATHER = 2, EK_SELECT2 = 3, EK_PLATE_LATENT = 4 };

struct FastArgs {
  PlateArgs P;
  int32_t ej[2], et[2];  // (param, term) of the two element-dependent terms
};

// raw per-element input of an element-dependent term (the only loads that go to HBM / L1)
template <int KIND>
__device__ __forceinline__ double eterm_load(const PlateArgs& A, const DTerm& tm, int64_t i, int64_t c, int64_t C) {
  if (KIND == EK_LATENT_X) return __ldg((const double*)tm.col.ptr + i);
  if (KIND == EK_SELECT2) return (double)(A.qd[(int64_t)A.drow[tm.dslot + i] * C + c] != 0);
  if (KIND == EK_PLATE_LATENT) return A.q[(int64_t)A.crow[tm.slot + i] * C + c];
  return 0.0;
}
// term value from the raw input; aux = d value / d (reduced latent) or the SELECT2 branch flag
template <int KIND>
__device__ __forceinline__ double eterm_apply(double raw, double va, double vb, double gv, double& aux) {
  if (KIND == EK_LATENT_X) { aux = raw; return va * raw; }
  if (KIND == EK_GATHER) { aux = 1.0; return gv; }
  if (KIND == EK_SELECT2) { aux = raw; return raw != 0.0 ? vb : va; }
  if (KIND == EK_PLATE_LATENT) { aux = 1.0; return raw; }
  aux = 0.0;
  return 0.0;
}

enum { DK_GENERIC = 0, DK_NORMAL = 1, DK_TWOVAL = 2 };
#define PL_UB 8  // elements whose loads are issued together (memory-level parallelism)

// DK: DK_NORMAL = Normal with chain-constant / SELECT2 sigma; DK_TWOVAL = Bernoulli or two-way
// Categorical with a chain-constant parameter (both log-probs evaluated once per chain).
// Value-only launches (MH / SMC: many chains, short plates) run at 4 CTAs per SM so that
// 32 warps x PL_UB independent row loads are in flight per SM.
// TY = plate elements worked on concurrently per CTA (threadIdx.y); TX = 256 / TY chains per CTA.
// TY = 8 (32 chains) suits few chains x long plates; TY = 1 (256 chains, 2 KB contiguous per
// state row and CTA, no shared-memory fold) suits many chains (HMC at >= 1024 chains, MH, SMC).
template <bool GRAD, int DK, int K0, int K1, int TY>
__global__ void __launch_bounds__(256, GRAD ? 2 : 3) k_plate_fast(const __grid_constant__ FastArgs F) {
  constexpr int TX = 256 / TY;
  constexpr bool NORMAL = DK == DK_NORMAL;
  constexpr bool CSR = (K0 == EK_GATHER || K1 == EK_GATHER);
  const PlateArgs& A = F.P;
  const DOp& op = A.op;
  const int64_t C = A.C;
  const int64_t c = (int64_t)blockIdx.x * TX + threadIdx.x;
  const bool active = c < C;
  const int64_t u0 = (int64_t)blockIdx.y * A.units_per_split;
  const int64_t u1 = min(A.n_units, u0 + A.units_per_split);
  const double qnan = __longlong_as_double(0x7ff8000000000000LL);
  const int j0 = F.ej[0], j1 = F.ej[1];
  double lp = 0.0;
  double acc_c[3] = {0.0, 0.0, 0.0};            // sum of w_j: goes to every LATENT term of parameter j
  double acc_e[2] = {0.0, 0.0}, acc_eb[2] = {0.0, 0.0};  // element-term accumulators (b: SELECT2 second branch)
  if (active) {
    // chain-constant part of every parameter
    double base[3] = {0.0, 0.0, 0.0}, dl[3] = {1.0, 1.0, 1.0};
#pragma unroll
    for (int j = 0; j < 3; j++) {
      if (j >= op.n_params) continue;
      double s = 0.0;
#pragma unroll
      for (int t = 0; t < 3; t++) {
        const DTerm& tm = op.p[j].t[t];
        if (tm.kind == FC_TERM_CONST) s += tm.c;
        else if (tm.kind == FC_TERM_LATENT) s += tm.c * A.q[(int64_t)tm.row * C + c];
      }
      base[j] = op.p[j].invariant ? apply_link(op.p[j].link, s, op.p[j].lo, op.p[j].hi, dl[j]) : s;
    }
    const DTerm& T0 = op.p[K0 ? j0 : 0].t[K0 ? F.et[0] : 0];
    const DTerm& T1 = op.p[K1 ? j1 : 0].t[K1 ? F.et[1] : 0];
    double va0 = 0.0, vb0 = 0.0, va1 = 0.0, vb1 = 0.0;
    if (K0 == EK_LATENT_X || K0 == EK_SELECT2) va0 = A.q[(int64_t)T0.row * C + c];
    if (K0 == EK_SELECT2) vb0 = A.q[(int64_t)T0.row_b * C + c];
    if (K1 == EK_LATENT_X || K1 == EK_SELECT2) va1 = A.q[(int64_t)T1.row * C + c];
    if (K1 == EK_SELECT2) vb1 = A.q[(int64_t)T1.row_b * C + c];
    // Normal: sigma handling. Invariant sigma -> constants; SELECT2 sigma -> two constants.
    double n_inv_a = 0.0, n_log_a = 0.0, n_inv_b = 0.0, n_log_b = 0.0;
    bool n_ok_a = false, n_ok_b = false;
    const bool sig_is_e0 = NORMAL && K0 == EK_SELECT2 && j0 == 1, sig_is_e1 = NORMAL && K1 == EK_SELECT2 && j1 == 1;
    if (NORMAL) {
      double sa = base[1], sb = base[1];
      if (sig_is_e0) { sa = base[1] + va0; sb = base[1] + vb0; }
      if (sig_is_e1) { sa = base[1] + va1; sb = base[1] + vb1; }
      n_ok_a = sa > 0.0 && d_fin(sa);
      n_ok_b = sb > 0.0 && d_fin(sb);
      if (n_ok_a) { n_log_a = log(sa); n_inv_a = 1.0 / sa; }
      if (n_ok_b) { n_log_b = log(sb); n_inv_b = 1.0 / sb; }
    }
    double tv_lp0 = 0.0, tv_lp1 = 0.0, tv_d0 = 0.0, tv_d1 = 0.0;
    const bool tv_bern = op.dist == FC_BERNOULLI;
    if (DK == DK_TWOVAL) {
      DistOut o0 = dist_eval<GRAD>(op.dist, base[0], 0.0, 0.0, 0.0, 0, qnan);
      DistOut o1 = dist_eval<GRAD>(op.dist, base[0], 0.0, 0.0, 0.0, 1, qnan);
      tv_lp0 = o0.lp; tv_d0 = o0.dp0; tv_lp1 = o1.lp; tv_d1 = o1.dp0;
    }
    // one element: everything after its loads
    auto body = [&](int64_t i, double raw0, double raw1, double x, long long xi, int64_t xrow, double lg, double gv0, double gv1, double& ga0, double& ga1) {
      double aux0, aux1;
      const double v0 = eterm_apply<K0>(raw0, va0, vb0, gv0, aux0);
      const double v1 = eterm_apply<K1>(raw1, va1, vb1, gv1, aux1);
      double pv0 = base[0], pv1 = base[1], pv2 = base[2];
      if (K0) { pv0 += (j0 == 0) ? v0 : 0.0; pv1 += (j0 == 1) ? v0 : 0.0; pv2 += (j0 == 2) ? v0 : 0.0; }
      if (K1) { pv0 += (j1 == 0) ? v1 : 0.0; pv1 += (j1 == 1) ? v1 : 0.0; pv2 += (j1 == 2) ? v1 : 0.0; }
      double g_lp, g_dx = 0.0, g_dp0 = 0.0, g_dp1 = 0.0, g_dp2 = 0.0;
      if (NORMAL) {
        const bool useb = (sig_is_e0 && aux0 != 0.0) || (sig_is_e1 && aux1 != 0.0);
        const bool ok = useb ? n_ok_b : n_ok_a;
        const double inv = useb ? n_inv_b : n_inv_a, lgs = useb ? n_log_b : n_log_a;
        if (!ok || !d_fin(pv0) || !d_fin(x)) { g_lp = d_neg_inf(); if (GRAD) g_dx = g_dp0 = g_dp1 = qnan; }  // no gradient outside the support: divergent (hmc.rs:384-386)
        else {
          double z = (x - pv0) * inv;
          g_lp = -0.5 * z * z - lgs - 0.5 * FC_LN_2PI;
          if (GRAD) { if (d_fin(g_lp)) { g_dp0 = z * inv; g_dx = -g_dp0; g_dp1 = (z * z - 1.0) * inv; } else g_dx = g_dp0 = g_dp1 = qnan; }
        }
      } else if (DK == DK_TWOVAL) {
        if (xi == 0) { g_lp = tv_lp0; g_dp0 = tv_d0; }
        else if (xi == 1 || tv_bern) { g_lp = tv_lp1; g_dp0 = tv_d1; }
        else g_lp = d_neg_inf();  // two-way Categorical: index out of range (distribution.rs:785-791)
      } else {
        DistOut o = op_dist_eval<GRAD>(op, pv0, pv1, pv2, x, xi, lg);
        g_lp = o.lp; g_dx = o.dx; g_dp0 = o.dp0; g_dp1 = o.dp1; g_dp2 = o.dp2;
      }
      lp += g_lp;
      if (GRAD) {
        if (op.is_sample && op.cont) A.grad[xrow] += g_dx;
        const double w0 = g_dp0 * dl[0], w1 = g_dp1 * dl[1], w2 = g_dp2 * dl[2];
        acc_c[0] += w0; acc_c[1] += w1; acc_c[2] += w2;
        if (K0) {
          const double w = j0 == 0 ? w0 : (j0 == 1 ? w1 : w2);
          if (K0 == EK_LATENT_X) acc_e[0] += w * aux0;
          else if (K0 == EK_SELECT2) { if (aux0 != 0.0) acc_eb[0] += w; else acc_e[0] += w; }
          else if (K0 == EK_GATHER) ga0 += w;
          else if (K0 == EK_PLATE_LATENT && w != 0.0) A.grad[(int64_t)A.crow[T0.slot + i] * C + c] += w;
        }
        if (K1) {
          const double w = j1 == 0 ? w0 : (j1 == 1 ? w1 : w2);
          if (K1 == EK_LATENT_X) acc_e[1] += w * aux1;
          else if (K1 == EK_SELECT2) { if (aux1 != 0.0) acc_eb[1] += w; else acc_e[1] += w; }
          else if (K1 == EK_GATHER) ga1 += w;
          else if (K1 == EK_PLATE_LATENT && w != 0.0) A.grad[(int64_t)A.crow[T1.slot + i] * C + c] += w;
        }
      }
    };
    // the element's own value (site row or data column) and optional lgamma(value + 1)
    auto load_x = [&](int64_t i, double& x, long long& xi, int64_t& xrow, double& lg) {
      x = 0.0; xi = 0; xrow = 0; lg = qnan;
      if (op.is_sample) {
        if (op.cont) { xrow = (int64_t)A.crow[op.cslot0 + i] * C + c; x = A.q[xrow]; }
        else xi = A.qd[(int64_t)A.drow[op.dslot0 + i] * C + c];
      } else {
        if (op.cont) x = dcol_f64(op.value_col, i);
        else xi = dcol_i64(op.value_col, i);
      }
      if (DK == DK_GENERIC && op.lgam) lg = op.lgam[i];
    };
    if (CSR) {
      for (int64_t u = u0 + threadIdx.y; u < u1; u += TY) {
        const int64_t e0 = A.grp_ptr[u], e1 = A.grp_ptr[u + 1];
        double gv0 = 0.0, gv1 = 0.0, ga0 = 0.0, ga1 = 0.0;
        if (K0 == EK_GATHER) gv0 = A.q[(int64_t)A.crow[T0.slot + u] * C + c];
        if (K1 == EK_GATHER) gv1 = A.q[(int64_t)A.crow[T1.slot + u] * C + c];
        for (int64_t e = e0; e < e1; e++) {
          const int64_t i = (int64_t)A.grp_elems[e];
          double x, lg; long long xi; int64_t xrow;
          const double raw0 = eterm_load<K0>(A, T0, i, c, C), raw1 = eterm_load<K1>(A, T1, i, c, C);
          load_x(i, x, xi, xrow, lg);
          body(i, raw0, raw1, x, xi, xrow, lg, gv0, gv1, ga0, ga1);
        }
        if (GRAD) {
          if (K0 == EK_GATHER) A.grad[(int64_t)A.crow[T0.slot + u] * C + c] += ga0;
          if (K1 == EK_GATHER) A.grad[(int64_t)A.crow[T1.slot + u] * C + c] += ga1;
        }
      }
    } else {
      // unit == element. PL_UB elements per trip: all their loads first -- straight-line, no
      // per-element branches (out-of-range slots re-read the last valid element and are masked in
      // the arithmetic), row-table lookups before the state-row loads -- then the arithmetic.
      const bool xs_samp = op.is_sample != 0, xs_cont = op.cont != 0;
      const int32_t* __restrict__ xtab = xs_cont ? A.crow + op.cslot0 : A.drow + op.dslot0;
      const bool same_sel = (K0 == EK_SELECT2 && K1 == EK_SELECT2 && T0.dslot == T1.dslot);
      const int x_dtype = op.value_col.dtype;
      for (int64_t ub = u0 + threadIdx.y; ub < u1; ub += (int64_t)TY * PL_UB) {
        int64_t ic[PL_UB];
        int rx[PL_UB], rr0[PL_UB], rr1[PL_UB];
        double raw0[PL_UB], raw1[PL_UB], x[PL_UB], lg[PL_UB];
        long long xi[PL_UB];
#pragma unroll
        for (int b = 0; b < PL_UB; b++) {
          ic[b] = min(ub + (int64_t)b * TY, u1 - 1);
          rx[b] = rr0[b] = rr1[b] = 0; raw0[b] = raw1[b] = 0.0; x[b] = 0.0; xi[b] = 0; lg[b] = qnan;
        }
        // stage A: row-table lookups (L1-resident)
        if (xs_samp) {
#pragma unroll
          for (int b = 0; b < PL_UB; b++) rx[b] = __ldg(xtab + ic[b]);
        }
        if (K0 == EK_SELECT2 || K0 == EK_PLATE_LATENT) {
          const int32_t* __restrict__ tab = K0 == EK_SELECT2 ? A.drow + T0.dslot : A.crow + T0.slot;
#pragma unroll
          for (int b = 0; b < PL_UB; b++) rr0[b] = __ldg(tab + ic[b]);
        }
        if ((K1 == EK_SELECT2 || K1 == EK_PLATE_LATENT) && !same_sel) {
          const int32_t* __restrict__ tab = K1 == EK_SELECT2 ? A.drow + T1.dslot : A.crow + T1.slot;
#pragma unroll
          for (int b = 0; b < PL_UB; b++) rr1[b] = __ldg(tab + ic[b]);
        }
        // stage B: the state rows / data columns
        if (K0 == EK_SELECT2) {
#pragma unroll
          for (int b = 0; b < PL_UB; b++) raw0[b] = (double)(A.qd[(int64_t)rr0[b] * C + c] != 0);
        } else if (K0 == EK_PLATE_LATENT) {
#pragma unroll
          for (int b = 0; b < PL_UB; b++) raw0[b] = A.q[(int64_t)rr0[b] * C + c];
        } else if (K0 == EK_LATENT_X) {
#pragma unroll
          for (int b = 0; b < PL_UB; b++) raw0[b] = __ldg((const double*)T0.col.ptr + ic[b]);
        }
        if (same_sel) {
#pragma unroll
          for (int b = 0; b < PL_UB; b++) raw1[b] = raw0[b];
        } else if (K1 == EK_SELECT2) {
#pragma unroll
          for (int b = 0; b < PL_UB; b++) raw1[b] = (double)(A.qd[(int64_t)rr1[b] * C + c] != 0);
        } else if (K1 == EK_PLATE_LATENT) {
#pragma unroll
          for (int b = 0; b < PL_UB; b++) raw1[b] = A.q[(int64_t)rr1[b] * C + c];
        } else if (K1 == EK_LATENT_X) {
#pragma unroll
          for (int b = 0; b < PL_UB; b++) raw1[b] = __ldg((const double*)T1.col.ptr + ic[b]);
        }
        if (xs_samp) {
          if (xs_cont) {
#pragma unroll
            for (int b = 0; b < PL_UB; b++) x[b] = A.q[(int64_t)rx[b] * C + c];
          } else {
#pragma unroll
            for (int b = 0; b < PL_UB; b++) xi[b] = A.qd[(int64_t)rx[b] * C + c];
          }
        } else if (xs_cont) {
          if (x_dtype == FC_COL_F64) {
#pragma unroll
            for (int b = 0; b < PL_UB; b++) x[b] = __ldg((const double*)op.value_col.ptr + ic[b]);
          } else {
#pragma unroll
            for (int b = 0; b < PL_UB; b++) x[b] = dcol_f64(op.value_col, ic[b]);
          }
        } else {
#pragma unroll
          for (int b = 0; b < PL_UB; b++) xi[b] = dcol_i64(op.value_col, ic[b]);
        }
        if (DK == DK_GENERIC && op.lgam) {
#pragma unroll
          for (int b = 0; b < PL_UB; b++) lg[b] = __ldg(op.lgam + ic[b]);
        }
        double ga0 = 0.0, ga1 = 0.0;
#pragma unroll
        for (int b = 0; b < PL_UB; b++) {
          const int64_t i = ub + (int64_t)b * TY;
          if (i < u1) body(i, raw0[b], raw1[b], x[b], xi[b], (int64_t)rx[b] * C + c, lg[b], 0.0, 0.0, ga0, ga1);
        }
      }
    }
  }
  __shared__ double sm[TY == 1 ? 1 : TY][TY == 1 ? 1 : TX + 1];
  const int n_acc = GRAD ? A.acc.n_acc : 3;
  const int lp_slot = op.is_sample ? 0 : 1;
  for (int a = 0; a < n_acc; a++) {
    double v = 0.0;
    if (a < 3) v = (a == lp_slot) ? lp : 0.0;
    else {
#pragma unroll
      for (int j = 0; j < 3; j++)
#pragma unroll
        for (int t = 0; t < 3; t++) {
          const bool is0 = K0 && j == j0 && t == F.et[0], is1 = K1 && j == j1 && t == F.et[1];
          if (A.acc.a_idx[j][t] == a) v += is0 ? acc_e[0] : (is1 ? acc_e[1] : acc_c[j] * (op.p[j].t[t].kind == FC_TERM_LATENT ? op.p[j].t[t].c : 1.0));
          if (A.acc.b_idx[j][t] == a) v += is0 ? acc_eb[0] : (is1 ? acc_eb[1] : 0.0);
        }
    }
    if (TY == 1) {
      if (active) A.part[((int64_t)blockIdx.y * n_acc + a) * C + c] = v;
    } else {
      __syncthreads();
      sm[threadIdx.y][threadIdx.x] = v;
      __syncthreads();
      if (threadIdx.y == 0 && active) {
        double s = sm[0][threadIdx.x];
#pragma unroll
        for (int y = 1; y < TY; y++) s += sm[y][threadIdx.x];
        A.part[((int64_t)blockIdx.y * n_acc + a) * C + c] = s;
      }
    }
  }
}

// chain-vectorised value-only kernel (fc_plate_vec.cu): 0 = not eligible, 1 = launched with
// partials [n_splits][3][C], 2 = launched and lp3 already updated (single split)
int plate_vec_launch(const PlateArgs& A, int64_t n_splits, double* d_lp3, cudaStream_t st);

// grouped Normal observation plate, gradient pass, group walk from shared memory (fc_plate_csr.cu)
bool plate_csr_eligible(const DOp& op);
int64_t plate_csr_chain_blocks(int64_t C);
int plate_csr_launch(const PlateArgs& A, int64_t n_splits, bool need_value, cudaStream_t st);

// instantiation units: fc_plate_{v8,v1,g8,g1}.cu (value-only / gradient x TY = 8 / 1)
template <bool GRAD, int TY>
int plate_fast_dispatch(int dk, int k0, int k1, const FastArgs& F, dim3 grid, cudaStream_t st);

#define FC_PLATE_FAST_INSTANTIATE(GRAD_, TY_)                                                                   \
  template <int DK>                                                                                             \
  static int fast_dispatch_dk(int k0, int k1, const FastArgs& F, dim3 grid, cudaStream_t st) {                  \
    dim3 block(256 / TY_, TY_);                                                                                 \
    FAST_CASE_(GRAD_, TY_, EK_NONE, EK_NONE)                                                                    \
    FAST_CASE_(GRAD_, TY_, EK_LATENT_X, EK_NONE)                                                                \
    FAST_CASE_(GRAD_, TY_, EK_GATHER, EK_NONE)                                                                  \
    FAST_CASE_(GRAD_, TY_, EK_SELECT2, EK_NONE)                                                                 \
    FAST_CASE_(GRAD_, TY_, EK_PLATE_LATENT, EK_NONE)                                                            \
    FAST_CASE_(GRAD_, TY_, EK_GATHER, EK_LATENT_X)                                                              \
    FAST_CASE_(GRAD_, TY_, EK_LATENT_X, EK_GATHER)                                                              \
    FAST_CASE_(GRAD_, TY_, EK_SELECT2, EK_SELECT2)                                                              \
    FAST_CASE_(GRAD_, TY_, EK_LATENT_X, EK_LATENT_X)                                                            \
    return 0;                                                                                                   \
  }                                                                                                             \
  template <>                                                                                                   \
  int plate_fast_dispatch<GRAD_, TY_>(int dk, int k0, int k1, const FastArgs& F, dim3 grid, cudaStream_t st) {  \
    if (dk == DK_NORMAL) return fast_dispatch_dk<DK_NORMAL>(k0, k1, F, grid, st);                               \
    if (dk == DK_TWOVAL) return fast_dispatch_dk<DK_TWOVAL>(k0, k1, F, grid, st);                               \
    return fast_dispatch_dk<DK_GENERIC>(k0, k1, F, grid, st);                                                   \
  }
#define FAST_CASE_(G_, T_, A_, B_) \
  if (k0 == A_ && k1 == B_) { k_plate_fast<G_, DK, A_, B_, T_><<<grid, block, 0, st>>>(F); return 1; }

}  // namespace fc
