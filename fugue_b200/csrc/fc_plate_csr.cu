// fc_plate_csr.cu -- gradient pass of a grouped Normal observation plate
//     y#i ~ Normal(const + latent + alpha[gid_i] (+ beta * x_i), sigma)        sigma per chain
// (examples/hierarchical_models.rs:7-33, the BASELINE C3 shape) with the group structure walked
// from shared memory.
//
// The generic CSR path of k_plate_fast walks  grp_ptr -> grp_elems -> x[i], y[i]  as a chain of
// dependent, lane-uniform global loads per element (ncu: 65 % of the stall samples are long
// scoreboard, 2.6 TFLOP/s FP64). The element data are chain-independent, so here a CTA of 128 chains
//   1. stages the (x, y) of a chunk of CSR_TILE consecutive CSR elements into shared memory with
//      coalesced loads (each thread loads different elements),
//   2. loads the alpha rows and the gradient rows of CSR_GB groups at once (independent,
//      chain-coalesced loads: memory-level parallelism 2 * CSR_GB per thread),
//   3. walks the elements of those groups from shared memory (broadcast LDS, no global load in
//      the loop), one chain per thread, the group's gradient in a register.
// Control flow is uniform across the CTA (the group / element sequence does not depend on the
// chain), so the chunk reload barrier inside the walk is safe.
// Same partial layout ([split][acc][C]) and same per-element formula as k_plate_fast.
#include "fc_plate.cuh"

namespace fc {

#define CSR_TX 128
#define CSR_TILE 1024
#define CSR_GB 8

struct CsrArgs {
  PlateArgs P;
  int32_t gt;       // term index (in param 0) of the GATHER term
  int32_t xt;       // term index (in param 0) of the LATENT_X term, or -1
};

template <bool HAS_X, bool VAL>
__global__ void __launch_bounds__(CSR_TX, 4) k_plate_csr_normal(const __grid_constant__ CsrArgs F) {
  const PlateArgs& A = F.P;
  const DOp& op = A.op;
  const int64_t C = A.C;
  const int tid = threadIdx.x;
  const int64_t c_raw = (int64_t)blockIdx.x * CSR_TX + tid;
  const bool active = c_raw < C;
  const int64_t c = active ? c_raw : C - 1;  // inactive lanes shadow the last chain (uniform control flow), never store
  const int64_t u0 = (int64_t)blockIdx.y * A.units_per_split;
  const int64_t u1 = min(A.n_units, u0 + A.units_per_split);

  __shared__ double s_x[HAS_X ? CSR_TILE : 1];
  __shared__ double s_y[CSR_TILE];

  // chain constants
  double base0 = 0.0, sig_s = 0.0;
#pragma unroll
  for (int t = 0; t < 3; t++) {
    const DTerm& tm = op.p[0].t[t];
    if (tm.kind == FC_TERM_CONST) base0 += tm.c;
    else if (tm.kind == FC_TERM_LATENT) base0 += tm.c * A.q[(int64_t)tm.row * C + c];
  }
#pragma unroll
  for (int t = 0; t < 3; t++) {
    const DTerm& tm = op.p[1].t[t];
    if (tm.kind == FC_TERM_CONST) sig_s += tm.c;
    else if (tm.kind == FC_TERM_LATENT) sig_s += tm.c * A.q[(int64_t)tm.row * C + c];
  }
  double dl1 = 1.0;
  const double sigma = apply_link(op.p[1].link, sig_s, op.p[1].lo, op.p[1].hi, dl1);
  const bool sig_ok = sigma > 0.0 && d_fin(sigma);
  const double inv = sig_ok ? 1.0 / sigma : 0.0, lgs = sig_ok ? log(sigma) : 0.0;
  const DTerm& TG = op.p[0].t[F.gt];
  const DTerm& TX = op.p[0].t[HAS_X ? F.xt : 0];
  const double va = HAS_X ? A.q[(int64_t)TX.row * C + c] : 0.0;
  const double* __restrict__ xcol = HAS_X ? (const double*)TX.col.ptr : nullptr;
  const int32_t* __restrict__ gtab = A.crow + TG.slot;

  // two interleaved accumulator sets (even / odd element of a group): independent FP64 add chains
  double lp = 0.0, lpb = 0.0, acc_w0 = 0.0, acc_w1 = 0.0, acc_w1b = 0.0, acc_x = 0.0, acc_xb = 0.0;
  const double inv1 = inv * dl1;  // sigma not ok: inv = 0 => every weight below is exactly 0
  const double neg_inf = d_neg_inf();

  // CSR element positions fit in int32 (grp_elems is an int32 table)
  const int e_end = u1 > u0 ? __ldg(A.grp_ptr + u1) : 0;
  int chunk_b = u1 > u0 ? __ldg(A.grp_ptr + u0) : 0, clen = 0;  // empty: first use loads
  auto load_chunk = [&](int b) {
    __syncthreads();  // everyone has finished with the previous chunk
    chunk_b = b;
    clen = min(e_end - b, CSR_TILE);
    for (int k = tid; k < clen; k += CSR_TX) {
      const int64_t i = (int64_t)__ldg(A.grp_elems + b + k);
      if (HAS_X) s_x[k] = __ldg(xcol + i);
      s_y[k] = dcol_f64(op.value_col, i);
    }
    __syncthreads();
  };
  // one element: Normal::log_prob and its parameter derivatives (distribution.rs:138-160), the
  // formula of k_plate_fast. A non-finite mean / observation makes zz non-finite: -inf, no gradient.
  auto elem = [&](int k, double m0, double& lp_, double& ga_, double& w1_, double& ax_) {
    const double y = s_y[k];
    double mean = m0, x = 0.0;
    if (HAS_X) { x = s_x[k]; mean = m0 + va * x; }
    const double z = (y - mean) * inv, zz = z * z;
    const bool fin = d_fin(zz);
    if (VAL) lp_ = fin ? lp_ + (-0.5 * zz - lgs - 0.5 * FC_LN_2PI) : neg_inf;
    const double w0 = fin ? z * inv : 0.0;
    ga_ += w0;
    w1_ += fin ? (zz - 1.0) * inv1 : 0.0;
    if (HAS_X) ax_ += w0 * x;
  };

  for (int64_t ub = u0; ub < u1; ub += CSR_GB) {
    // the batch's group bounds, alpha rows and gradient rows: all loads issued before any use
    int32_t ptr[CSR_GB + 1], row[CSR_GB];
    double gv[CSR_GB], gold[CSR_GB];
#pragma unroll
    for (int b = 0; b <= CSR_GB; b++) ptr[b] = __ldg(A.grp_ptr + min(ub + b, u1));
#pragma unroll
    for (int b = 0; b < CSR_GB; b++) row[b] = __ldg(gtab + min(ub + b, u1 - 1));
#pragma unroll
    for (int b = 0; b < CSR_GB; b++) gv[b] = A.q[(int64_t)row[b] * C + c];
#pragma unroll
    for (int b = 0; b < CSR_GB; b++) gold[b] = A.grad[(int64_t)row[b] * C + c];
#pragma unroll
    for (int b = 0; b < CSR_GB; b++) {
      if (ub + b >= u1) break;
      const double m0 = base0 + gv[b];
      double ga = 0.0, gab = 0.0;
      int k = ptr[b] - chunk_b, kend = ptr[b + 1] - chunk_b;  // positions relative to the staged chunk
      while (k < kend) {                                       // (uniform across the CTA)
        if (k >= clen) { const int nb = chunk_b + k; kend -= k; k = 0; load_chunk(nb); }
        const int kl = min(kend, clen);
        for (; k + 1 < kl; k += 2) {
          elem(k, m0, lp, ga, acc_w1, acc_x);
          elem(k + 1, m0, lpb, gab, acc_w1b, acc_xb);
        }
        if (k < kl) { elem(k, m0, lp, ga, acc_w1, acc_x); k++; }
      }
      ga += gab;
      acc_w0 += ga;
      if (active) A.grad[(int64_t)row[b] * C + c] = gold[b] + ga;
    }
  }
  lp += lpb; acc_w1 += acc_w1b; acc_x += acc_xb;
  if (!sig_ok && u1 > u0 && e_end > __ldg(A.grp_ptr + u0)) lp = neg_inf;  // Normal::new(.., sigma <= 0): -inf

  // partials: same accumulator mapping as k_plate_fast's epilogue
  const int n_acc = A.acc.n_acc;
  const int lp_slot = 1;  // observe -> log_likelihood
  if (active) {
    for (int a = 0; a < n_acc; a++) {
      double v = 0.0;
      if (a < 3) v = (a == lp_slot) ? lp : 0.0;
      else {
#pragma unroll
        for (int j = 0; j < 2; j++)
#pragma unroll
          for (int t = 0; t < 3; t++) {
            if (A.acc.a_idx[j][t] != a) continue;
            if (j == 0 && HAS_X && t == F.xt) v += acc_x;
            else if (j == 0 && t == F.gt) v += 0.0;  // gather rows are written directly
            else v += (j == 0 ? acc_w0 : acc_w1) * (op.p[j].t[t].kind == FC_TERM_LATENT ? op.p[j].t[t].c : 1.0);
          }
      }
      A.part[((int64_t)blockIdx.y * n_acc + a) * C + c_raw] = v;
    }
  }
}

// Normal observation plate from an f64 column, every element-dependent term in the mean (identity
// link): one GATHER and at most one LATENT_X over an f64 column; sigma chain-constant.
bool plate_csr_eligible(const DOp& op) {
  static const bool disabled = getenv("FC_PLATE_NO_CSR") != nullptr;
  if (disabled) return false;
  if (op.dist != FC_NORMAL || op.is_sample || !op.cont || op.n_params != 2) return false;
  if (op.value_col.ptr == nullptr) return false;
  if (op.p[0].link != FC_LINK_IDENTITY || !op.p[1].invariant) return false;
  int ng = 0, nx = 0;
  for (int t = 0; t < 3; t++) {
    int k = op.p[0].t[t].kind;
    if (k == FC_TERM_GATHER) ng++;
    else if (k == FC_TERM_LATENT_X) { nx++; if (op.p[0].t[t].col.dtype != FC_COL_F64) return false; }
    else if (k != FC_TERM_NONE && k != FC_TERM_CONST && k != FC_TERM_LATENT) return false;
  }
  return ng == 1 && nx <= 1;
}

int64_t plate_csr_chain_blocks(int64_t C) { return cdiv(C, (int64_t)CSR_TX); }

int plate_csr_launch(const PlateArgs& A, int64_t n_splits, bool need_value, cudaStream_t st) {
  CsrArgs F;
  F.P = A; F.gt = -1; F.xt = -1;
  for (int t = 0; t < 3; t++) {
    if (A.op.p[0].t[t].kind == FC_TERM_GATHER) F.gt = t;
    if (A.op.p[0].t[t].kind == FC_TERM_LATENT_X) F.xt = t;
  }
  dim3 grid((unsigned)cdiv(A.C, (int64_t)CSR_TX), (unsigned)n_splits);
  if (F.xt >= 0) {
    if (need_value) k_plate_csr_normal<true, true><<<grid, CSR_TX, 0, st>>>(F);
    else k_plate_csr_normal<true, false><<<grid, CSR_TX, 0, st>>>(F);
  } else {
    if (need_value) k_plate_csr_normal<false, true><<<grid, CSR_TX, 0, st>>>(F);
    else k_plate_csr_normal<false, false><<<grid, CSR_TX, 0, st>>>(F);
  }
  return 1;
}

}  // namespace fc
