// fc_score.cu -- batched per-chain log-density and analytic gradient of an op list.
//
// Replaces run(ScoreGivenTrace{base}, model_fn()) (src/runtime/interpreters.rs:138-163,
// src/inference/hmc.rs:264-299) and the finite-difference force grad_log_joint
// (src/inference/hmc.rs:304-329). The three Trace accumulators log_prior / log_likelihood /
// log_factors (src/runtime/trace.rs:168-177) are kept separate and combined as
// (prior + lik) + factors (trace.rs:198-200).
//
// Launch structure per evaluation of C chains (all on the program's stream, no atomics, so
// results are run-to-run reproducible):
//   k_scalar      one thread per chain walks the scalar sample/observe/factor ops in program order
//   k_plate       one launch per plate op: chains on threadIdx.x (coalesced q rows), plate
//                 elements split over threadIdx.y and blockIdx.y; per-chain partial sums of
//                 the log-density and of the reduced gradient rows go to a partial buffer
//   dense kernel  (fc_dense.cu) for plates whose parameter is a dense X*beta contraction
//   k_finish      sums the partials in fixed order into lp3 / grad
#include <cstdlib>
#include "fc_scalar.cuh"

namespace fc {

// ---------------------------------------------------------------------------------------
// scalar ops: one thread per chain, program order
// ---------------------------------------------------------------------------------------
// scalar ops: one thread per chain, program order (interpreter in fc_scalar.cuh)
// ---------------------------------------------------------------------------------------
template <bool GRAD>
__global__ void k_scalar(const DOp* __restrict__ ops, int n_ops, const int32_t* __restrict__ crow, const int32_t* __restrict__ drow,
                         const double* __restrict__ q, const disc_t* __restrict__ qd, int64_t C, double* __restrict__ grad, double* __restrict__ part,
                         const fc_expr_node* __restrict__ nodes) {
  int64_t c = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= C) return;
  double a_prior = 0.0, a_lik = 0.0, a_fac = 0.0;
  auto za = [&](int dslot) -> long long { return qd[(int64_t)drow[dslot] * C + c]; };
  scalar_ops_eval<GRAD>(ops, n_ops, crow, q, C, c, za, grad, nodes, a_prior, a_lik, a_fac);
  part[0 * C + c] = a_prior;
  part[1 * C + c] = a_lik;
  part[2 * C + c] = a_fac;
}

// ---------------------------------------------------------------------------------------
// generic plate op
// ---------------------------------------------------------------------------------------
template <bool GRAD>
__global__ void __launch_bounds__(PL_TX* PL_TY, 2) k_plate(const __grid_constant__ PlateArgs A) {
  const DOp& op = A.op;
  const int64_t C = A.C;
  const int64_t c = (int64_t)blockIdx.x * PL_TX + threadIdx.x;
  const bool active = c < C;
  const int64_t u0 = (int64_t)blockIdx.y * A.units_per_split;
  const int64_t u1 = min(A.n_units, u0 + A.units_per_split);
  const double qnan = __longlong_as_double(0x7ff8000000000000LL);

  double lp = 0.0;
  double acc_a[3][3], acc_b[3][3];
#pragma unroll
  for (int j = 0; j < 3; j++)
#pragma unroll
    for (int t = 0; t < 3; t++) acc_a[j][t] = acc_b[j][t] = 0.0;

  if (active) {
    // ---- everything that is constant along the plate for this chain is loaded / evaluated once:
    //      CONST and LATENT term values, the latent factor of LATENT_X, both SELECT2 branches,
    //      and whole parameters whose terms are all element-invariant ----
    double tva[3][3], tvb[3][3];
#pragma unroll
    for (int j = 0; j < 3; j++)
#pragma unroll
      for (int t = 0; t < 3; t++) {
        tva[j][t] = tvb[j][t] = 0.0;
        if (j >= op.n_params) continue;
        const DTerm& tm = op.p[j].t[t];
        if (tm.kind == FC_TERM_CONST) tva[j][t] = tm.c;
        else if (tm.kind == FC_TERM_LATENT || tm.kind == FC_TERM_LATENT_X || tm.kind == FC_TERM_SELECT2) {
          tva[j][t] = A.q[(int64_t)tm.row * C + c];
          if (tm.kind == FC_TERM_LATENT) tva[j][t] *= tm.c;  // coef * q
          if (tm.kind == FC_TERM_SELECT2) tvb[j][t] = A.q[(int64_t)tm.row_b * C + c];
        }
      }
    double pv_inv[3] = {0.0, 0.0, 0.0}, dl_inv[3] = {0.0, 0.0, 0.0};
#pragma unroll
    for (int j = 0; j < 3; j++) {
      if (j < op.n_params && op.p[j].invariant) {
        const DParam& pr = op.p[j];
        double s = 0.0;
        bool first = true;
#pragma unroll
        for (int t = 0; t < 3; t++) {
          if (pr.t[t].kind == FC_TERM_NONE) continue;
          s = first ? tva[j][t] : s + tva[j][t];
          first = false;
        }
        pv_inv[j] = apply_link(pr.link, s, pr.lo, pr.hi, dl_inv[j]);
      }
    }
    // Normal likelihood/prior: log(sigma) and 1/sigma are memoised on the sigma value (two
    // entries: covers an element-invariant sigma and a SELECT2 sigma that takes two values)
    const bool is_normal = op.dist == FC_NORMAL;
    double mem_sig[2] = {qnan, qnan}, mem_log[2] = {0.0, 0.0}, mem_inv[2] = {0.0, 0.0};
    // two-valued discrete likelihoods with an invariant parameter: both log-probs up front
    const bool two_valued = (op.dist == FC_BERNOULLI || op.dist == FC_CATEGORICAL) && op.p[0].invariant;
    DistOut tv0, tv1;
    if (two_valued) {
      tv0 = dist_eval<GRAD>(op.dist, pv_inv[0], 0.0, 0.0, 0.0, 0, qnan);
      tv1 = dist_eval<GRAD>(op.dist, pv_inv[0], 0.0, 0.0, 0.0, 1, qnan);
    }
    const int32_t* __restrict__ crow = A.crow;
    const int32_t* __restrict__ drow = A.drow;

    for (int64_t u = u0 + threadIdx.y; u < u1; u += PL_TY) {
      int64_t e0 = u, e1 = u + 1;
      if (A.grp_ptr) { e0 = A.grp_ptr[u]; e1 = A.grp_ptr[u + 1]; }
      // The per-unit state rows (own site value, SELECT2 selector, plate/gathered latent) are the
      // only loads that go to HBM; each iteration depends on them, so without help a warp has one
      // 128-byte line in flight. Prefetch the rows PL_PF iterations ahead into L1 (no registers,
      // no dependency) to keep several lines per warp outstanding.
      {
        const int64_t up = u + (int64_t)PL_PF * PL_TY;
        if (up < u1) {
          if (op.is_sample) {
            if (!A.grp_ptr) {
              if (op.cont) prefetch_l1(A.q + (int64_t)crow[op.cslot0 + up] * C + c);
              else prefetch_l1(A.qd + (int64_t)drow[op.dslot0 + up] * C + c);
            }
          }
#pragma unroll
          for (int j = 0; j < 3; j++)
#pragma unroll
            for (int t = 0; t < 3; t++) {
              if (j >= op.n_params) continue;
              const DTerm& tm = op.p[j].t[t];
              if (tm.kind == FC_TERM_GATHER && A.grp_ptr) prefetch_l1(A.q + (int64_t)crow[tm.slot + up] * C + c);
              else if (!A.grp_ptr && tm.kind == FC_TERM_SELECT2) prefetch_l1(A.qd + (int64_t)drow[tm.dslot + up] * C + c);
              else if (!A.grp_ptr && tm.kind == FC_TERM_PLATE_LATENT) prefetch_l1(A.q + (int64_t)crow[tm.slot + up] * C + c);
            }
        }
      }
      double gacc[3][3], gval[3][3];
#pragma unroll
      for (int j = 0; j < 3; j++)
#pragma unroll
        for (int t = 0; t < 3; t++) {
          gacc[j][t] = 0.0;
          gval[j][t] = 0.0;
          // a gathered latent is constant within its CSR group: one load per group
          if (A.grp_ptr && j < op.n_params && op.p[j].t[t].kind == FC_TERM_GATHER) gval[j][t] = A.q[(int64_t)crow[op.p[j].t[t].slot + u] * C + c];
        }
      for (int64_t e = e0; e < e1; e++) {
        const int64_t i = A.grp_elems ? (int64_t)A.grp_elems[e] : e;
        double pv[3] = {pv_inv[0], pv_inv[1], pv_inv[2]}, dl[3] = {dl_inv[0], dl_inv[1], dl_inv[2]};
        double lx[3][3];
        int zsel[3][3];
#pragma unroll
        for (int j = 0; j < 3; j++) {
          if (j >= op.n_params || op.p[j].invariant) continue;
          const DParam& pr = op.p[j];
          double s = 0.0;
          bool first = true;
#pragma unroll
          for (int t = 0; t < 3; t++) {
            const DTerm& tm = pr.t[t];
            if (tm.kind == FC_TERM_NONE) continue;
            double v = tva[j][t];  // CONST / LATENT
            switch (tm.kind) {
              case FC_TERM_LATENT_X: lx[j][t] = dcol_f64(tm.col, i); v = tva[j][t] * lx[j][t]; break;
              case FC_TERM_GATHER: v = A.grp_ptr ? gval[j][t] : A.q[(int64_t)crow[tm.slot + dcol_i64(tm.col, i)] * C + c]; break;
              case FC_TERM_SELECT2: zsel[j][t] = (A.qd[(int64_t)drow[tm.dslot + i] * C + c] == 0) ? 0 : 1; v = zsel[j][t] ? tvb[j][t] : tva[j][t]; break;
              case FC_TERM_PLATE_LATENT: v = A.q[(int64_t)crow[tm.slot + i] * C + c]; break;
              default: break;
            }
            s = first ? v : s + v;
            first = false;
          }
          pv[j] = apply_link(pr.link, s, pr.lo, pr.hi, dl[j]);
        }
        double x = 0.0;
        long long xi = 0;
        int64_t xrow = 0;
        if (op.is_sample) {
          if (op.cont) { xrow = (int64_t)crow[op.cslot0 + i] * C + c; x = A.q[xrow]; }
          else xi = A.qd[(int64_t)drow[op.dslot0 + i] * C + c];
        } else {
          if (op.cont) x = dcol_f64(op.value_col, i);
          else xi = dcol_i64(op.value_col, i);
        }
        DistOut g;
        if (is_normal) {
          const double sig = pv[1];
          int m = (sig == mem_sig[0]) ? 0 : ((sig == mem_sig[1]) ? 1 : -1);
          if (m < 0) {  // miss: evaluate once, replace the older entry
            m = 1;
            mem_sig[1] = mem_sig[0]; mem_log[1] = mem_log[0]; mem_inv[1] = mem_inv[0];
            m = 0;
            mem_sig[0] = sig; mem_log[0] = log(sig); mem_inv[0] = 1.0 / sig;
          }
          g.dx = g.dp0 = g.dp1 = g.dp2 = 0.0;
          if (!(sig > 0.0) || !d_fin(sig) || !d_fin(pv[0]) || !d_fin(x)) { g.lp = d_neg_inf(); if (GRAD) g.dx = g.dp0 = g.dp1 = qnan; }
          else {
            const double inv = m == 0 ? mem_inv[0] : mem_inv[1], lg = m == 0 ? mem_log[0] : mem_log[1];
            double z = (x - pv[0]) * inv;
            g.lp = -0.5 * z * z - lg - 0.5 * FC_LN_2PI;
            if (GRAD) { if (d_fin(g.lp)) { g.dp0 = z * inv; g.dx = -g.dp0; g.dp1 = (z * z - 1.0) * inv; } else g.dx = g.dp0 = g.dp1 = qnan; }
          }
        } else if (two_valued && (xi == 0 || xi == 1)) {
          g = xi ? tv1 : tv0;
        } else {
          g = op_dist_eval<GRAD>(op, pv[0], pv[1], pv[2], x, xi, op.lgam ? op.lgam[i] : qnan);
        }
        lp += g.lp;
        if (GRAD) {
          if (op.is_sample && op.cont) A.grad[xrow] += g.dx;
          const double dpj[3] = {g.dp0, g.dp1, g.dp2};
#pragma unroll
          for (int j = 0; j < 3; j++) {
            if (j >= op.n_params) continue;
            const double w = dpj[j] * dl[j];
            if (w == 0.0) continue;
#pragma unroll
            for (int t = 0; t < 3; t++) {
              const DTerm& tm = op.p[j].t[t];
              switch (tm.kind) {
                case FC_TERM_LATENT: acc_a[j][t] += w * tm.c; break;
                case FC_TERM_LATENT_X: acc_a[j][t] += w * lx[j][t]; break;
                case FC_TERM_SELECT2: if (zsel[j][t]) acc_b[j][t] += w; else acc_a[j][t] += w; break;
                case FC_TERM_GATHER:
                  if (A.grp_ptr) gacc[j][t] += w;
                  break;
                case FC_TERM_PLATE_LATENT: A.grad[(int64_t)crow[tm.slot + i] * C + c] += w; break;
                default: break;
              }
            }
          }
        }
      }
      if (GRAD && A.grp_ptr) {  // this thread is the only writer of group u's gathered rows
#pragma unroll
        for (int j = 0; j < 3; j++)
#pragma unroll
          for (int t = 0; t < 3; t++)
            if (j < op.n_params && op.p[j].t[t].kind == FC_TERM_GATHER) A.grad[(int64_t)crow[op.p[j].t[t].slot + u] * C + c] += gacc[j][t];
      }
    }
  }
  // ---- reduce over threadIdx.y in shared memory (fixed order), write the split's partial ----
  __shared__ double sm[PL_TY][PL_TX + 1];
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
          if (A.acc.a_idx[j][t] == a) v += acc_a[j][t];
          if (A.acc.b_idx[j][t] == a) v += acc_b[j][t];
        }
    }
    __syncthreads();
    sm[threadIdx.y][threadIdx.x] = v;
    __syncthreads();
    if (threadIdx.y == 0 && active) {
      double s = sm[0][threadIdx.x];
#pragma unroll
      for (int y = 1; y < PL_TY; y++) s += sm[y][threadIdx.x];
      A.part[((int64_t)blockIdx.y * n_acc + a) * C + c] = s;
    }
  }
}

// host side: eligibility + dispatch. Returns 1 if launched, 0 if the op needs the generic kernel.
// ty = 8 or 1 (chains per CTA = 256 / ty); grid.x must be cdiv(C, 256 / ty).
static int plate_fast_launch(const PlateArgs& A, bool grad, int ty, dim3 grid, cudaStream_t st) {
  static const bool disabled = getenv("FC_PLATE_GENERIC") != nullptr;
  if (disabled) return 0;
  const DOp& op = A.op;
  FastArgs F;
  F.P = A;
  F.ej[0] = F.ej[1] = F.et[0] = F.et[1] = 0;
  int kinds[2] = {EK_NONE, EK_NONE}, n = 0;
  for (int j = 0; j < op.n_params; j++)
    for (int t = 0; t < 3; t++) {
      int k = op.p[j].t[t].kind, ek = EK_NONE;
      if (k == FC_TERM_LATENT_X) ek = EK_LATENT_X;
      else if (k == FC_TERM_GATHER) ek = EK_GATHER;
      else if (k == FC_TERM_SELECT2) ek = EK_SELECT2;
      else if (k == FC_TERM_PLATE_LATENT) ek = EK_PLATE_LATENT;
      else if (k == FC_TERM_DENSE) return 0;
      if (ek == EK_NONE) continue;
      if (n == 2) return 0;
      if (op.p[j].link != FC_LINK_IDENTITY) return 0;
      if (ek == EK_LATENT_X && op.p[j].t[t].col.dtype != FC_COL_F64) return 0;
      kinds[n] = ek; F.ej[n] = j; F.et[n] = t; n++;
    }
  bool normal = op.dist == FC_NORMAL;
  if (normal) {  // sigma must be chain-constant or a single SELECT2
    int n_sig = 0;
    for (int k = 0; k < n; k++)
      if (F.ej[k] == 1) { n_sig++; if (kinds[k] != EK_SELECT2) normal = false; }
    if (n_sig > 1) normal = false;
  }
  int dk = normal ? DK_NORMAL : DK_GENERIC;
  if (!normal && (op.dist == FC_BERNOULLI || op.dist == FC_CATEGORICAL) && op.p[0].invariant) dk = DK_TWOVAL;
  if (grad) return ty == 1 ? plate_fast_dispatch<true, 1>(dk, kinds[0], kinds[1], F, grid, st) : plate_fast_dispatch<true, 8>(dk, kinds[0], kinds[1], F, grid, st);
  return ty == 1 ? plate_fast_dispatch<false, 1>(dk, kinds[0], kinds[1], F, grid, st) : plate_fast_dispatch<false, 8>(dk, kinds[0], kinds[1], F, grid, st);
}
// can the fast kernel take this op? (same test as plate_fast_launch, without launching)
static bool plate_fast_eligible(const DOp& op) {
  static const bool disabled = getenv("FC_PLATE_GENERIC") != nullptr;
  if (disabled) return false;
  int kinds[2] = {EK_NONE, EK_NONE}, n = 0;
  for (int j = 0; j < op.n_params; j++)
    for (int t = 0; t < 3; t++) {
      int k = op.p[j].t[t].kind, ek = EK_NONE;
      if (k == FC_TERM_LATENT_X) ek = EK_LATENT_X;
      else if (k == FC_TERM_GATHER) ek = EK_GATHER;
      else if (k == FC_TERM_SELECT2) ek = EK_SELECT2;
      else if (k == FC_TERM_PLATE_LATENT) ek = EK_PLATE_LATENT;
      else if (k == FC_TERM_DENSE) return false;
      if (ek == EK_NONE) continue;
      if (n == 2) return false;
      if (op.p[j].link != FC_LINK_IDENTITY) return false;
      if (ek == EK_LATENT_X && op.p[j].t[t].col.dtype != FC_COL_F64) return false;
      kinds[n++] = ek;
    }
  const int k0 = kinds[0], k1 = kinds[1];
  return (k1 == EK_NONE) || (k0 == EK_GATHER && k1 == EK_LATENT_X) || (k0 == EK_LATENT_X && k1 == EK_GATHER) || (k0 == EK_SELECT2 && k1 == EK_SELECT2) ||
         (k0 == EK_LATENT_X && k1 == EK_LATENT_X);
}

// ---------------------------------------------------------------------------------------
// finish: fold one launch's partial segment into lp3 / grad (fixed split order)
// ---------------------------------------------------------------------------------------
struct FinishArgs {
  const double* part;
  int64_t n_splits;
  int32_t n_acc;
  int32_t rows[24 + 128];  // target grad rows of accumulators 3.. (dense plates use up to 128+)
  double* lp3;
  double* grad;
  int64_t C;
};
__global__ void k_finish(const __grid_constant__ FinishArgs A) {
  int64_t c = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  int a = blockIdx.y;
  if (c >= A.C) return;
  double s = 0.0;
  for (int64_t sp = 0; sp < A.n_splits; sp++) s += A.part[(sp * A.n_acc + a) * A.C + c];
  if (a < 3) A.lp3[(int64_t)a * A.C + c] += s;
  else A.grad[(int64_t)A.rows[a - 3] * A.C + c] += s;
}
__global__ void k_total(const double* __restrict__ lp3, double* __restrict__ total, int64_t C) {
  int64_t c = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (c < C) total[c] = (lp3[c] + lp3[C + c]) + lp3[2 * C + c];  // trace.rs:198-200
}

// fast-kernel CTA shape: 256 chains x 1 once there are enough chains to fill whole CTAs
static int g_ty1_min = getenv("FC_PLATE_TY1_MIN") ? atoi(getenv("FC_PLATE_TY1_MIN")) : 16384;
static int plate_ty(const PlateInfo& pl, int64_t C) { return (C >= g_ty1_min && plate_fast_eligible(pl.dop)) ? 1 : PL_TY; }

static bool plate_use_csr(const PlateInfo& pl, bool grad) { return grad && pl.has_gather && plate_csr_eligible(pl.dop); }

static int64_t plate_splits(const PlateInfo& pl, int64_t C, int sm_count, bool grad) {
  const bool csr = plate_use_csr(pl, grad);
  const int ty = csr ? 1 : plate_ty(pl, C);
  int64_t units = pl.has_gather ? pl.n_groups : pl.dop.plate_len;
  int64_t cblocks = csr ? plate_csr_chain_blocks(C) : cdiv(C, 256 / ty);
  int64_t want = cdiv((int64_t)sm_count * 8, cblocks);           // ~8 CTAs per SM
  int64_t max_by_work = std::max<int64_t>(1, units / (ty * 4));  // >= 4 units per thread
  return std::max<int64_t>(1, std::min(want, max_by_work));
}

static int g_sm_count = 0;
static int sm_count() {
  if (!g_sm_count) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&g_sm_count, cudaDevAttrMultiProcessorCount, dev);
    if (g_sm_count <= 0) g_sm_count = 148;
  }
  return g_sm_count;
}

int32_t score_dev(fc_program* P, const double* d_q, const disc_t* d_qd, int64_t C, double* d_lp3, double* d_total, double* d_grad, bool need_value) {
  if (C <= 0) return FC_OK;
  cudaStream_t st = P->stream;
  const bool grad = d_grad != nullptr;
  if (P->d_disc > 0 && !d_qd) { set_error("program has %lld discrete sites but qd is NULL", (long long)P->d_disc); return FC_ERR_TRACE_ADDRESS_NOT_FOUND; }
  // partial buffer size
  int64_t need = 3 * C;
  std::vector<int64_t> splits(P->plates.size());
  for (size_t i = 0; i < P->plates.size(); i++) {
    PlateInfo& pl = *P->plates[i];
    splits[i] = pl.dense ? dense_plate_splits(P, pl, C) : plate_splits(pl, C, sm_count(), grad);
    int64_t n_acc = pl.dense ? (3 + 24 + pl.dop.p[pl.dense_param].t[pl.dense_term].col.width) : pl.acc.n_acc;
    need = std::max(need, splits[i] * n_acc * C);
  }
  FC_CUDA(P->ws.part.reserve(need));
  FC_CUDA(cudaMemsetAsync(d_lp3, 0, 3 * C * sizeof(double), st));
  if (grad && P->d_cont > 0) FC_CUDA(cudaMemsetAsync(d_grad, 0, P->d_cont * C * sizeof(double), st));
  FinishArgs F;
  F.lp3 = d_lp3; F.grad = d_grad; F.C = C; F.part = P->ws.part.p;
  const unsigned fin_bx = (unsigned)cdiv(C, 128);
  if (!P->scalar_ops.empty()) {
    unsigned nb = (unsigned)cdiv(C, 128);
    if (grad) k_scalar<true><<<nb, 128, 0, st>>>(P->d_scalar_ops.p, (int)P->scalar_ops.size(), P->d_crow.p, P->d_drow.p, d_q, d_qd, C, d_grad, P->ws.part.p, P->d_expr_nodes.p);
    else k_scalar<false><<<nb, 128, 0, st>>>(P->d_scalar_ops.p, (int)P->scalar_ops.size(), P->d_crow.p, P->d_drow.p, d_q, d_qd, C, d_grad, P->ws.part.p, P->d_expr_nodes.p);
    FC_KERNEL_CHECK();
    F.n_splits = 1; F.n_acc = 3;
    k_finish<<<dim3(fin_bx, 3), 128, 0, st>>>(F);
    FC_KERNEL_CHECK();
    P->n_kernel_launches += 2;
  }
  for (size_t i = 0; i < P->plates.size(); i++) {
    PlateInfo& pl = *P->plates[i];
    if (pl.dop.plate_len == 0) continue;
    if (pl.dense) {
      int64_t ns = 0;
      FC_TRY(dense_plate_launch(P, pl, d_q, C, P->ws.part.p, splits[i], grad, &ns, need_value));
      const DTerm& dt = pl.dop.p[pl.dense_param].t[pl.dense_term];
      int K = dt.col.width;
      // accumulators: 0..2 lp, then the AccMap's reduced rows, then K dense rows
      int n_red = pl.acc.n_acc - 3;
      F.n_splits = ns;
      F.n_acc = grad ? 3 + n_red + K : 3;
      for (int a = 0; a < n_red; a++) F.rows[a] = pl.acc.rows[a];
      for (int k = 0; k < K; k++) F.rows[n_red + k] = P->crow[dt.slot + k];
      k_finish<<<dim3(fin_bx, F.n_acc), 128, 0, st>>>(F);
      FC_KERNEL_CHECK();
      P->n_kernel_launches += 1;
      continue;
    }
    PlateArgs A;
    A.op = pl.dop; A.acc = pl.acc; A.crow = P->d_crow.p; A.drow = P->d_drow.p;
    A.q = d_q; A.qd = d_qd; A.grad = d_grad; A.part = P->ws.part.p;
    A.grp_ptr = pl.has_gather ? pl.grp_ptr.p : nullptr;
    A.grp_elems = pl.has_gather ? pl.grp_elems.p : nullptr;
    A.C = C;
    A.n_units = pl.has_gather ? pl.n_groups : pl.dop.plate_len;
    A.units_per_split = cdiv(A.n_units, splits[i]);
    int64_t ns = cdiv(A.n_units, A.units_per_split);
    const int ty = plate_ty(pl, C);
    dim3 grid((unsigned)cdiv(C, PL_TX), (unsigned)ns), block(PL_TX, PL_TY), fgrid((unsigned)cdiv(C, 256 / ty), (unsigned)ns);
    int vec = grad ? 0 : plate_vec_launch(A, ns, d_lp3, st);
    if (plate_use_csr(pl, grad)) plate_csr_launch(A, ns, need_value, st);
    else if (!vec && !plate_fast_launch(A, grad, ty, fgrid, st)) {
      if (grad) k_plate<true><<<grid, block, 0, st>>>(A);
      else k_plate<false><<<grid, block, 0, st>>>(A);
    }
    FC_KERNEL_CHECK();
    P->n_kernel_launches += 1;
    if (vec == 2) continue;  // single split: the kernel added straight into lp3
    F.n_splits = ns;
    F.n_acc = grad ? pl.acc.n_acc : 3;
    for (int a = 0; a < pl.acc.n_acc - 3; a++) F.rows[a] = pl.acc.rows[a];
    k_finish<<<dim3(fin_bx, F.n_acc), 128, 0, st>>>(F);
    FC_KERNEL_CHECK();
    P->n_kernel_launches += 1;
  }
  if (d_total) {
    k_total<<<fin_bx, 128, 0, st>>>(d_lp3, d_total, C);
    FC_KERNEL_CHECK();
    P->n_kernel_launches += 1;
  }
  return FC_OK;
}

__global__ void k_i64_to_disc(const long long* __restrict__ in, disc_t* __restrict__ out, int64_t n, int* __restrict__ overflow) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  long long v = in[i];
  if (v > 2147483647LL || v < -2147483648LL) { *overflow = 1; v = v > 0 ? 2147483647LL : -2147483648LL; }
  out[i] = (disc_t)v;
}

// host int64 [d_disc][C] -> device disc_t plane in the workspace
int32_t convert_qd(fc_program* P, const int64_t* h_qd, int64_t C, const disc_t** out) {
  *out = nullptr;
  if (P->d_disc == 0) return FC_OK;
  if (!h_qd) { set_error("program has %lld discrete sites but qd is NULL", (long long)P->d_disc); return FC_ERR_TRACE_ADDRESS_NOT_FOUND; }
  int64_t n = P->d_disc * C;
  FC_CUDA(P->ws.qd64.reserve(n + 1));
  FC_CUDA(P->ws.qd.reserve(n));
  FC_CUDA(cudaMemcpyAsync(P->ws.qd64.p, h_qd, n * 8, cudaMemcpyHostToDevice, P->stream));
  int* flag = (int*)(P->ws.qd64.p + n);
  FC_CUDA(cudaMemsetAsync(flag, 0, 4, P->stream));
  k_i64_to_disc<<<(unsigned)cdiv(n, 256), 256, 0, P->stream>>>((const long long*)P->ws.qd64.p, P->ws.qd.p, n, flag);
  FC_KERNEL_CHECK();
  P->n_kernel_launches += 1;
  int h = 0;
  FC_CUDA(cudaMemcpyAsync(&h, flag, 4, cudaMemcpyDeviceToHost, P->stream));
  FC_CUDA(cudaStreamSynchronize(P->stream));
  if (h) { set_error("discrete site value outside the int32 range of the device planes"); return FC_ERR_UNSUPPORTED_OP; }
  *out = P->ws.qd.p;
  return FC_OK;
}

static int32_t score_host(fc_program* P, const double* h_q, const int64_t* h_qd, size_t C, double* h_lp3[3], double* h_total, double* h_grad) {
  if (!P) { set_error("program is NULL"); return FC_ERR_INVALID_ARGUMENT; }
  if (C == 0) return FC_OK;
  if (P->d_cont > 0 && !h_q) { set_error("q is NULL"); return FC_ERR_INVALID_ARGUMENT; }
  FC_CUDA(cudaSetDevice(P->device));
  int64_t nq = P->d_cont * (int64_t)C;
  FC_CUDA(P->ws.q.reserve(nq));
  FC_CUDA(P->ws.lp3.reserve(3 * C));
  FC_CUDA(P->ws.total.reserve(C));
  if (h_grad) FC_CUDA(P->ws.grad.reserve(nq));
  if (nq) FC_CUDA(cudaMemcpyAsync(P->ws.q.p, h_q, nq * 8, cudaMemcpyHostToDevice, P->stream));
  const disc_t* d_qd = nullptr;
  FC_TRY(convert_qd(P, h_qd, (int64_t)C, &d_qd));
  FC_TRY(score_dev(P, P->ws.q.p, d_qd, (int64_t)C, P->ws.lp3.p, P->ws.total.p, h_grad ? P->ws.grad.p : nullptr));
  for (int a = 0; a < 3; a++)
    if (h_lp3[a]) FC_CUDA(cudaMemcpyAsync(h_lp3[a], P->ws.lp3.p + a * C, C * 8, cudaMemcpyDeviceToHost, P->stream));
  if (h_total) FC_CUDA(cudaMemcpyAsync(h_total, P->ws.total.p, C * 8, cudaMemcpyDeviceToHost, P->stream));
  if (h_grad && nq) FC_CUDA(cudaMemcpyAsync(h_grad, P->ws.grad.p, nq * 8, cudaMemcpyDeviceToHost, P->stream));
  FC_CUDA(cudaStreamSynchronize(P->stream));
  return FC_OK;
}

}  // namespace fc

extern "C" {

int32_t fc_logp(fc_program* p, const double* h_q, const int64_t* h_qd, size_t C, double* h_log_prior, double* h_log_lik, double* h_log_factors) {
  double* lp3[3] = {h_log_prior, h_log_lik, h_log_factors};
  return fc::score_host(p, h_q, h_qd, C, lp3, nullptr, nullptr);
}

int32_t fc_logp_grad(fc_program* p, const double* h_q, const int64_t* h_qd, size_t C, double* h_lp, double* h_grad) {
  double* lp3[3] = {nullptr, nullptr, nullptr};
  if (!h_grad) { fc::set_error("grad is NULL"); return FC_ERR_INVALID_ARGUMENT; }
  return fc::score_host(p, h_q, h_qd, C, lp3, h_lp, h_grad);
}

// device-resident variant: d_qd is an int32 plane [d_disc][C]; d_lp [C]; d_grad [d_cont][C] or NULL.
// Asynchronous on the program's stream (fc_program_stream).
int32_t fc_logp_grad_dev(fc_program* p, const double* d_q, const int32_t* d_qd, size_t C, double* d_lp, double* d_grad) {
  if (!p) { fc::set_error("program is NULL"); return FC_ERR_INVALID_ARGUMENT; }
  FC_CUDA(cudaSetDevice(p->device));
  FC_CUDA(p->ws.lp3.reserve(3 * C));
  return fc::score_dev(p, d_q, d_qd, (int64_t)C, p->ws.lp3.p, d_lp, d_grad);
}
}
