// fc_hier.cu -- one fused leapfrog step for hierarchical programs with grouped Normal observations
//     alpha#g ~ Normal(loc, scale)                                  g = 0 .. G-1   (loc, scale: chain constants)
//     y#i     ~ Normal(const + latent + alpha[gid_i] (+ beta * x_i), sigma)
// plus any scalar ops (examples/hierarchical_models.rs:7-33, BASELINE configs[2]).
//
// What it replaces. A leapfrog step (hmc.rs:388-401) of the generic path is  memset grad -> k_scalar ->
// k_finish -> k_plate_fast (the alpha plate) -> k_finish -> k_plate_csr_normal -> k_finish -> k_total ->
// k_leap: nine launches that read or write the [G][C] rows of q / grad / p about a dozen times (ncu:
// DRAM 11.6 % busy, 200 launches per transition). The force on alpha_g depends only on alpha_g, its own
// observations and the handful of "global" sites, so here ONE kernel walks the groups of a
// (128 chains x group range) tile, forms the whole force on each alpha row in registers -- prior term +
// observation term, the same two operands in the same order as the generic kernels -- and applies the kick
// and the drift to that row on the spot: q, p, m_inv rows read once, q, p written once. The sums every group
// contributes to the global sites' forces go to per-CTA partials; a one-thread-per-chain tail kernel folds
// them in fixed order, adds the scalar ops (same interpreter as k_scalar) and moves the global rows. Two
// launches per leapfrog step.
//
// The prior term is k_plate_fast's Normal fast path; the observation term is accumulated in residual form (see
// the kernel), so the force agrees with the generic path to rounding, not bit for bit
// (tests/test_gpu_c3.py::test_fused_hierarchical_leapfrog_equals_generic_path).
#include "fc_plate.cuh"
#include "fc_scalar.cuh"

namespace fc {

#define HR_TX 128
#define HR_TILE 1024
#define HR_GB 4
#define HR_NACC 8      // partials per (split, chain): lp_prior, lp_obs, w0, w1, wx, p0, p1, kinetic
#define HR_MAX_GLOB 16

struct HierArgs {
  DOp obs, pri;
  const int32_t* grp_ptr;
  const int32_t* grp_elems;
  const int32_t* crow;
  double* q;              // [d][C] positions, updated in place when drift
  double* p;              // momenta, updated in place
  const double* m_inv;
  const double* eps;      // [C]
  double* g_out;          // force rows (written when store_g)
  int32_t* divergent;
  double* part;           // [n_splits][HR_NACC][C]
  int64_t C, n_groups, units_per_split;
  int32_t gt, xt;         // term indices of GATHER / LATENT_X in the observation mean
  double kick;
  int32_t drift, store_g;
};

template <bool HAS_X, bool VAL>
__global__ void __launch_bounds__(HR_TX, 6) k_hier_step(const __grid_constant__ HierArgs A) {
  const DOp& op = A.obs;
  const DOp& pr = A.pri;
  const int64_t C = A.C;
  const int tid = threadIdx.x;
  const int64_t c_raw = (int64_t)blockIdx.x * HR_TX + tid;
  const bool active = c_raw < C;
  const int64_t c = active ? c_raw : C - 1;  // inactive lanes shadow the last chain (uniform control flow), never store
  const int64_t u0 = (int64_t)blockIdx.y * A.units_per_split;
  const int64_t u1 = min(A.n_groups, u0 + A.units_per_split);

  __shared__ double2 s_xy[HAS_X ? HR_TILE : 1];  // (x, y) of a staged element side by side: one 16-byte load per element
  __shared__ double s_y[HAS_X ? 1 : HR_TILE];

  // ---- chain constants of the observation plate (k_plate_csr_normal's prologue) ----
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
  const DTerm& TG = op.p[0].t[A.gt];
  const DTerm& TX = op.p[0].t[HAS_X ? A.xt : 0];
  const double va = HAS_X ? A.q[(int64_t)TX.row * C + c] : 0.0;
  const double* __restrict__ xcol = HAS_X ? (const double*)TX.col.ptr : nullptr;
  const int32_t* __restrict__ gtab = A.crow + TG.slot;
  const double inv1 = inv * dl1;
  const double neg_inf = d_neg_inf();
  const double qnan = __longlong_as_double(0x7ff8000000000000LL);

  // ---- chain constants of the group-latent plate (k_plate_fast, Normal fast path) ----
  double pb[2] = {0.0, 0.0}, pdl[2] = {1.0, 1.0};
#pragma unroll
  for (int j = 0; j < 2; j++) {
    double s = 0.0;
#pragma unroll
    for (int t = 0; t < 3; t++) {
      const DTerm& tm = pr.p[j].t[t];
      if (tm.kind == FC_TERM_CONST) s += tm.c;
      else if (tm.kind == FC_TERM_LATENT) s += tm.c * A.q[(int64_t)tm.row * C + c];
    }
    pb[j] = apply_link(pr.p[j].link, s, pr.p[j].lo, pr.p[j].hi, pdl[j]);
  }
  const bool p_ok = pb[1] > 0.0 && d_fin(pb[1]);
  double p_log = 0.0, p_inv = 0.0;
  if (p_ok) { p_log = log(pb[1]); p_inv = 1.0 / pb[1]; }
  const bool loc_fin = d_fin(pb[0]);

  const double e = A.eps[c];
  const double ke = A.kick * e;

  // Observation sums in RESIDUAL form: with r_i = y_i - mean_i and one sigma per chain,
  //   sum z_i / sigma = inv^2 sum r_i,   sum (z_i^2 - 1) / sigma = inv (inv^2 sum r_i^2 - n),   sum z_i x_i / sigma = inv^2 sum r_i x_i,
  // so an element costs 5 FP64 instructions (mean, r, and three accumulations) instead of 11, and only sum r_i
  // is per group (ncu on the per-element form: FP64 pipe 39 % busy, 32 issued instructions per element).
  double acc_w0 = 0.0, srr = 0.0, srrb = 0.0, srx = 0.0, srxb = 0.0;
  double lp_pri = 0.0, acc_p0 = 0.0, acc_p1 = 0.0, kin = 0.0;
  const double inv2 = inv * inv;
  bool bad = false;

  const int e_end = u1 > u0 ? __ldg(A.grp_ptr + u1) : 0;
  int chunk_b = u1 > u0 ? __ldg(A.grp_ptr + u0) : 0, clen = 0;
  auto load_chunk = [&](int b) {
    __syncthreads();
    chunk_b = b;
    clen = min(e_end - b, HR_TILE);
    for (int k = tid; k < clen; k += HR_TX) {
      const int64_t i = (int64_t)__ldg(A.grp_elems + b + k);
      if (HAS_X) s_xy[k] = make_double2(__ldg(xcol + i), dcol_f64(op.value_col, i));
      else s_y[k] = dcol_f64(op.value_col, i);
    }
    __syncthreads();
  };
  auto elem = [&](int k, double m0, double& sr_, double& srr_, double& srx_) {
    double r;
    if (HAS_X) { const double2 xy = s_xy[k]; r = xy.y - fma(va, xy.x, m0); srx_ = fma(r, xy.x, srx_); }
    else r = s_y[k] - m0;
    sr_ += r;
    srr_ = fma(r, r, srr_);
  };

  for (int64_t ub = u0; ub < u1; ub += HR_GB) {
    int32_t ptr[HR_GB + 1], row[HR_GB];
    double gv[HR_GB], pv[HR_GB], mv[HR_GB];
#pragma unroll
    for (int b = 0; b <= HR_GB; b++) ptr[b] = __ldg(A.grp_ptr + min(ub + b, u1));
#pragma unroll
    for (int b = 0; b < HR_GB; b++) row[b] = __ldg(gtab + min(ub + b, u1 - 1));
#pragma unroll
    for (int b = 0; b < HR_GB; b++) gv[b] = A.q[(int64_t)row[b] * C + c];
#pragma unroll
    for (int b = 0; b < HR_GB; b++) pv[b] = A.p[(int64_t)row[b] * C + c];
#pragma unroll
    for (int b = 0; b < HR_GB; b++) mv[b] = A.m_inv[(int64_t)row[b] * C + c];
#pragma unroll
    for (int b = 0; b < HR_GB; b++) {
      if (ub + b >= u1) break;
      // observation term of the force on this group's latent
      const double m0 = base0 + gv[b];
      double ga = 0.0, gab = 0.0;
      int k = ptr[b] - chunk_b, kend = ptr[b + 1] - chunk_b;
      while (k < kend) {
        if (k >= clen) { const int nb = chunk_b + k; kend -= k; k = 0; load_chunk(nb); }
        const int kl = min(kend, clen);
        for (; k + 1 < kl; k += 2) {
          elem(k, m0, ga, srr, srx);
          elem(k + 1, m0, gab, srrb, srxb);
        }
        if (k < kl) { elem(k, m0, ga, srr, srx); k++; }
      }
      ga = (ga + gab) * inv2;  // sum of z / sigma over the group's observations (inv = 0 when sigma is invalid: no force)
      acc_w0 += ga;
      // prior term: alpha_g ~ Normal(loc, scale) (distribution.rs:138-160)
      const double x = gv[b];
      double g_lp, g_dx, g_dp0, g_dp1;
      if (!p_ok || !loc_fin || !d_fin(x)) { g_lp = neg_inf; g_dx = g_dp0 = g_dp1 = qnan; }
      else {
        const double z = (x - pb[0]) * p_inv;
        g_lp = -0.5 * z * z - p_log - 0.5 * FC_LN_2PI;
        if (d_fin(g_lp)) { g_dp0 = z * p_inv; g_dx = -g_dp0; g_dp1 = (z * z - 1.0) * p_inv; }
        else g_dx = g_dp0 = g_dp1 = qnan;
      }
      lp_pri += g_lp;
      acc_p0 += g_dp0 * pdl[0];
      acc_p1 += g_dp1 * pdl[1];
      // the force on this row and its leapfrog update (hmc.rs:388-401), as k_leap
      const double force = (0.0 + g_dx) + ga;
      if (!isfinite(force)) bad = true;
      const double pn = pv[b] + ke * force;
      kin += pn * pn * mv[b];
      if (active) {
        const int64_t idx = (int64_t)row[b] * C + c;
        A.p[idx] = pn;
        if (A.drift) A.q[idx] = gv[b] + e * mv[b] * pn;
        if (A.store_g) A.g_out[idx] = force;
      }
    }
  }
  srr += srrb; srx += srxb;
  const double n_el = u1 > u0 ? (double)(e_end - __ldg(A.grp_ptr + u0)) : 0.0;
  const double acc_w1 = inv1 * (inv2 * srr - n_el), acc_x = inv2 * srx;
  double lp = 0.0;
  if (VAL) {
    lp = -0.5 * inv2 * srr - n_el * (lgs + 0.5 * FC_LN_2PI);
    if (!d_fin(srr) || (!sig_ok && n_el > 0.0)) lp = neg_inf;  // a non-finite residual, or Normal::new(.., sigma <= 0): -inf
  }
  if (active) {
    if (bad) atomicOr(&A.divergent[c], 1);
    double* o = A.part + (int64_t)blockIdx.y * HR_NACC * C + c;
    o[0 * C] = lp_pri; o[1 * C] = lp; o[2 * C] = acc_w0; o[3 * C] = acc_w1;
    o[4 * C] = acc_x; o[5 * C] = acc_p0; o[6 * C] = acc_p1; o[7 * C] = kin;
  }
}

// ---- tail: one thread per chain folds the partials (fixed order), adds the scalar ops and moves the global rows ----
struct HierTail {
  DOp obs, pri;
  const DOp* scalar_ops;
  const int32_t* crow;
  const fc_expr_node* nodes;
  double* q;
  double* p;
  const double* m_inv;
  const double* eps;
  double* g_out;
  int32_t* divergent;
  const double* part;
  double* kpart;          // [C]: the whole kinetic sum (k_accept is told one split)
  double* lp3;            // [3][C]
  double* lj;             // [C]
  int64_t C, n_splits;
  int32_t n_scalar, n_glob, gt, xt;
  int32_t glob[HR_MAX_GLOB];  // sorted rows outside the group-latent plate
  double kick;
  int32_t drift, _pad;
};

// CTA = 32 chains x HR_NACC accumulators: thread (cx, a) folds accumulator a of chain cx over the splits (in split
// order: run-to-run reproducible; the loads of 8 splits are issued together), then the first warp finishes its 32
// chains. (One thread per chain doing all of it was 62 us per launch at 1024 chains: 8 CTAs, a serial walk over
// 880 partials per thread -- a quarter of the leapfrog step.)
template <bool VAL>
__global__ void __launch_bounds__(32 * HR_NACC) k_hier_tail(const __grid_constant__ HierTail A) {
  const int64_t C = A.C;
  const int cx = threadIdx.x & 31, ai = threadIdx.x >> 5;
  const int64_t c = (int64_t)blockIdx.x * 32 + cx;
  __shared__ double s_acc[HR_NACC][33];
  {
    double acc = 0.0;
    if (c < C) {
      const double* __restrict__ src = A.part + (int64_t)ai * C + c;
      const int64_t stride = (int64_t)HR_NACC * C;
      int64_t sp = 0;
      for (; sp + 8 <= A.n_splits; sp += 8) {
        double v[8];
#pragma unroll
        for (int u = 0; u < 8; u++) v[u] = src[(sp + u) * stride];
#pragma unroll
        for (int u = 0; u < 8; u++) acc += v[u];
      }
      for (; sp < A.n_splits; sp++) acc += src[sp * stride];
    }
    s_acc[ai][cx] = acc;
  }
  __syncthreads();
  if (ai != 0 || c >= C) return;
  double s[HR_NACC];
#pragma unroll
  for (int a = 0; a < HR_NACC; a++) s[a] = s_acc[a][cx];
  for (int i = 0; i < A.n_glob; i++) A.g_out[(int64_t)A.glob[i] * C + c] = 0.0;
  double a_prior = 0.0, a_lik = 0.0, a_fac = 0.0;
  auto za = [](int) -> long long { return 0; };  // no discrete sites in a fused hierarchical program
  scalar_ops_eval<true>(A.scalar_ops, A.n_scalar, A.crow, A.q, C, c, za, A.g_out, A.nodes, a_prior, a_lik, a_fac);
  // the plates' sums onto the global rows, the group-latent plate first (program order of the generic path)
#pragma unroll
  for (int j = 0; j < 2; j++)
#pragma unroll
    for (int t = 0; t < 3; t++) {
      const DTerm& tm = A.pri.p[j].t[t];
      if (tm.kind == FC_TERM_LATENT) A.g_out[(int64_t)tm.row * C + c] += s[5 + j] * tm.c;
    }
#pragma unroll
  for (int j = 0; j < 2; j++)
#pragma unroll
    for (int t = 0; t < 3; t++) {
      const DTerm& tm = A.obs.p[j].t[t];
      if (tm.kind == FC_TERM_LATENT) A.g_out[(int64_t)tm.row * C + c] += s[2 + j] * tm.c;
      else if (j == 0 && tm.kind == FC_TERM_LATENT_X && t == A.xt) A.g_out[(int64_t)tm.row * C + c] += s[4];
    }
  const double e = A.eps[c];
  double kin = s[7];
  bool bad = false;
  for (int i = 0; i < A.n_glob; i++) {
    const int64_t idx = (int64_t)A.glob[i] * C + c;
    const double gv = A.g_out[idx];
    if (!isfinite(gv)) bad = true;
    const double pv = A.p[idx] + A.kick * e * gv;
    A.p[idx] = pv;
    const double mi = A.m_inv[idx];
    if (A.drift) A.q[idx] += e * mi * pv;
    kin += pv * pv * mi;
  }
  if (bad) atomicOr(&A.divergent[c], 1);
  A.kpart[c] = kin;
  if (VAL) {
    const double l0 = (0.0 + a_prior) + s[0], l1 = (0.0 + a_lik) + s[1], l2 = 0.0 + a_fac;
    A.lp3[0 * C + c] = l0; A.lp3[1 * C + c] = l1; A.lp3[2 * C + c] = l2;
    A.lj[c] = (l0 + l1) + l2;  // Trace::total_log_weight (trace.rs:198-200)
  }
}

// ---- host side ----
int g_hmc_hier = 1;  // fc_debug_set("hmc_hier", 0): sessions created afterwards take the generic leapfrog path
bool hier_detect(fc_program* P, HierPlan* out) {
  static const bool disabled = getenv("FC_NO_HIER_FUSE") != nullptr;
  if (disabled || !g_hmc_hier || P->plates.size() != 2 || P->d_disc != 0) return false;
  int ip = -1, io = -1;
  for (int i = 0; i < 2; i++) {
    const DOp& op = P->plates[i]->dop;
    if (op.kind == FC_OP_PLATE_SAMPLE && op.dist == FC_NORMAL && op.cont && op.n_params == 2 && op.p[0].invariant && op.p[1].invariant) ip = i;
    else if (op.kind == FC_OP_PLATE_OBSERVE && P->plates[i]->has_gather && plate_csr_eligible(op)) io = i;
  }
  if (ip < 0 || io < 0) return false;
  const PlateInfo& pp = *P->plates[ip];
  const PlateInfo& po = *P->plates[io];
  for (int j = 0; j < 2; j++)
    for (int t = 0; t < 3; t++) {
      const int k = pp.dop.p[j].t[t].kind;
      if (!(k == FC_TERM_NONE || k == FC_TERM_CONST || k == FC_TERM_LATENT)) return false;
    }
  int gt = -1, xt = -1;
  for (int t = 0; t < 3; t++) {
    if (po.dop.p[0].t[t].kind == FC_TERM_GATHER) gt = t;
    if (po.dop.p[0].t[t].kind == FC_TERM_LATENT_X) xt = t;
  }
  if (gt < 0 || po.dop.p[0].t[gt].slot != pp.dop.cslot0 || po.n_groups != pp.dop.plate_len || po.n_groups <= 0) return false;
  // the rows outside the group-latent plate
  std::vector<char> in_plate((size_t)P->d_cont, 0);
  for (int64_t g = 0; g < pp.dop.plate_len; g++) in_plate[(size_t)P->crow[pp.dop.cslot0 + g]] = 1;
  std::vector<int32_t> glob;
  for (int64_t r = 0; r < P->d_cont; r++) if (!in_plate[(size_t)r]) glob.push_back((int32_t)r);
  if (glob.size() > HR_MAX_GLOB) return false;
  // a latent parameter of either plate must be a global row (not another group latent)
  auto globals_only = [&](const DOp& op) {
    for (int j = 0; j < 2; j++)
      for (int t = 0; t < 3; t++) {
        const DTerm& tm = op.p[j].t[t];
        if ((tm.kind == FC_TERM_LATENT || tm.kind == FC_TERM_LATENT_X) && in_plate[(size_t)tm.row]) return false;
      }
    return true;
  };
  if (!globals_only(pp.dop) || !globals_only(po.dop)) return false;
  for (const DOp& so : P->scalar_ops)
    if (so.is_sample && so.cont && in_plate[(size_t)P->crow[so.cslot0]]) return false;
  out->pri = ip; out->obs = io; out->gt = gt; out->xt = xt;
  out->n_glob = (int32_t)glob.size();
  for (size_t i = 0; i < glob.size(); i++) out->glob[i] = glob[i];
  return true;
}

// one leapfrog step from the positions in q: force at q, then p += kick * eps * force and (drift) q += eps * M^-1 p,
// in place. store_g: the force goes to g_out; need_value: lp3 / lj of the positions the force was taken at.
int32_t hier_step(fc_program* P, const HierPlan& H, double* q, double* p, const double* m_inv, const double* eps, double* g_out, int32_t* divergent,
                  double* kpart, double* lp3, double* lj, int64_t C, double kick, bool drift, bool store_g, bool need_value) {
  cudaStream_t st = P->stream;
  const PlateInfo& pp = *P->plates[H.pri];
  const PlateInfo& po = *P->plates[H.obs];
  int sms = 148, dev = 0;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  const int64_t cblocks = cdiv(C, (int64_t)HR_TX);
  const int64_t want = cdiv((int64_t)sms * 6, cblocks);
  const int64_t n_splits0 = std::max<int64_t>(1, std::min<int64_t>(want, std::max<int64_t>(1, po.n_groups / 4)));
  const int64_t ups = cdiv(po.n_groups, n_splits0);
  const int64_t n_splits = cdiv(po.n_groups, ups);
  FC_CUDA(P->ws.part.reserve(n_splits * HR_NACC * C));
  HierArgs A;
  A.obs = po.dop; A.pri = pp.dop;
  A.grp_ptr = po.grp_ptr.p; A.grp_elems = po.grp_elems.p; A.crow = P->d_crow.p;
  A.q = q; A.p = p; A.m_inv = m_inv; A.eps = eps; A.g_out = g_out; A.divergent = divergent; A.part = P->ws.part.p;
  A.C = C; A.n_groups = po.n_groups; A.units_per_split = ups; A.gt = H.gt; A.xt = H.xt;
  A.kick = kick; A.drift = drift ? 1 : 0; A.store_g = store_g ? 1 : 0;
  dim3 grid((unsigned)cblocks, (unsigned)n_splits);
  if (H.xt >= 0) {
    if (need_value) k_hier_step<true, true><<<grid, HR_TX, 0, st>>>(A);
    else k_hier_step<true, false><<<grid, HR_TX, 0, st>>>(A);
  } else {
    if (need_value) k_hier_step<false, true><<<grid, HR_TX, 0, st>>>(A);
    else k_hier_step<false, false><<<grid, HR_TX, 0, st>>>(A);
  }
  FC_KERNEL_CHECK();
  HierTail T;
  T.obs = po.dop; T.pri = pp.dop;
  T.scalar_ops = P->d_scalar_ops.p; T.crow = P->d_crow.p; T.nodes = P->d_expr_nodes.p;
  T.q = q; T.p = p; T.m_inv = m_inv; T.eps = eps; T.g_out = g_out; T.divergent = divergent; T.part = P->ws.part.p;
  T.kpart = kpart; T.lp3 = lp3; T.lj = lj; T.C = C; T.n_splits = n_splits;
  T.n_scalar = (int32_t)P->scalar_ops.size(); T.n_glob = H.n_glob; T.gt = H.gt; T.xt = H.xt;
  for (int i = 0; i < HR_MAX_GLOB; i++) T.glob[i] = i < H.n_glob ? H.glob[i] : 0;
  T.kick = kick; T.drift = drift ? 1 : 0; T._pad = 0;
  if (need_value) k_hier_tail<true><<<(unsigned)cdiv(C, 32), 32 * HR_NACC, 0, st>>>(T);
  else k_hier_tail<false><<<(unsigned)cdiv(C, 32), 32 * HR_NACC, 0, st>>>(T);
  FC_KERNEL_CHECK();
  P->n_kernel_launches += 2;
  return FC_OK;
}

}  // namespace fc
