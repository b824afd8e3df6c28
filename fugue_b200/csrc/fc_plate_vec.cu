// fc_plate_vec.cu -- chain-vectorised, value-only plate kernel for the "many chains, short plate"
// regime of SMC rejuvenation and batched MH (millions of particles x a plate of tens of elements,
// BASELINE configs[3]).
//
// Replaces the per-observation walk of ScoreGivenTrace (src/runtime/interpreters.rs:138-163) over
// e.g. examples/mixture_models.rs:96-108 (z#i ~ Categorical([p, 1-p]); x#i ~ Normal(mu[z], sigma[z]))
// for 4 consecutive chains per thread: the per-element cost of the scalar kernel (row-table
// lookup, 64-bit address arithmetic, loop control: ~45-90 issued instructions per element and
// chain, ncu profiles/ncu_c4_plate_r1.txt) is paid once per FOUR chains, and every state-row load
// is a 16-byte vector (int4 of discrete values / two double2 of continuous values), PLV_UB rows
// in flight per thread.
//
// Eligibility (host side, plate_vec_launch): value only; no CSR; C % 4 == 0; distribution kind
// Normal (mean = chain-constant + up to two {SELECT2, LATENT_X} terms, sigma chain-constant or one
// SELECT2) or a two-valued discrete likelihood with a chain-constant parameter. Everything else
// stays on k_plate_fast / k_plate. Same formulas, same summation order along the plate as those
// kernels (elements in order, one accumulator per chain).
#include <algorithm>
#include <cstdlib>
#include "fc_plate.cuh"

namespace fc {

#define PLV_UB 4       // plate elements whose loads are issued together
#define PLV_VEC 4      // chains per thread
#define PLV_THREADS 256

enum { VM_NONE = 0, VM_SELECT2 = 1, VM_LATENT_X = 2 };

struct VecArgs {
  PlateArgs P;
  int32_t mt[2];     // term index (within parameter 0) of the two mean terms
  int32_t st;        // term index (within parameter 1) of the sigma SELECT2
  int32_t direct;    // 1: single split, add straight into lp3 (no partial buffer, no k_finish)
  double* lp3;       // [3][C]
};

__device__ __forceinline__ void ld4(const double* p, double (&v)[4]) {
  double2 a = *(const double2*)p, b = *(const double2*)(p + 2);
  v[0] = a.x; v[1] = a.y; v[2] = b.x; v[3] = b.y;
}

// M0 / M1: kinds of the (up to two) element-dependent terms of the mean (parameter 0);
// SS: sigma (parameter 1) is base + SELECT2. For DK_TWOVAL all three are unused.
template <int DK, int M0, int M1, bool SS>
__global__ void __launch_bounds__(PLV_THREADS, 2) k_plate_vec(const __grid_constant__ VecArgs V) {
  const PlateArgs& A = V.P;
  const DOp& op = A.op;
  const int64_t C = A.C;
  const int64_t c0 = ((int64_t)blockIdx.x * PLV_THREADS + threadIdx.x) * PLV_VEC;
  if (c0 >= C) return;  // C % 4 == 0: a thread's four chains are all valid or all out of range
  const int u0 = (int)((int64_t)blockIdx.y * A.units_per_split);
  const int u1 = (int)min(A.n_units, (int64_t)u0 + A.units_per_split);
  const double qnan = __longlong_as_double(0x7ff8000000000000LL);
  const double* __restrict__ q0 = A.q + c0;
  const disc_t* __restrict__ qd0 = A.qd + c0;
  double lp[4] = {0.0, 0.0, 0.0, 0.0};

  // ---- chain constants ----
  double base0[4], base1[4];
#pragma unroll
  for (int j = 0; j < 2; j++) {
    double s[4] = {0.0, 0.0, 0.0, 0.0};
    if (j < op.n_params) {
#pragma unroll
      for (int t = 0; t < 3; t++) {
        const DTerm& tm = op.p[j].t[t];
        if (tm.kind == FC_TERM_CONST) {
#pragma unroll
          for (int k = 0; k < 4; k++) s[k] += tm.c;
        } else if (tm.kind == FC_TERM_LATENT) {
          double v[4];
          ld4(q0 + (int64_t)tm.row * C, v);
#pragma unroll
          for (int k = 0; k < 4; k++) s[k] += tm.c * v[k];
        }
      }
      if (op.p[j].invariant) {
        double dl;
#pragma unroll
        for (int k = 0; k < 4; k++) s[k] = apply_link(op.p[j].link, s[k], op.p[j].lo, op.p[j].hi, dl);
      }
    }
#pragma unroll
    for (int k = 0; k < 4; k++) { if (j == 0) base0[k] = s[k]; else base1[k] = s[k]; }
  }
  const DTerm& T0 = op.p[0].t[M0 ? V.mt[0] : 0];
  const DTerm& T1 = op.p[0].t[M1 ? V.mt[1] : 0];
  const DTerm& TS = op.p[SS ? 1 : 0].t[SS ? V.st : 0];
  double va0[4], vb0[4], va1[4];
#pragma unroll
  for (int k = 0; k < 4; k++) va0[k] = vb0[k] = va1[k] = 0.0;
  if (M0) ld4(q0 + (int64_t)T0.row * C, va0);
  if (M0 == VM_SELECT2) ld4(q0 + (int64_t)T0.row_b * C, vb0);
  if (M1) ld4(q0 + (int64_t)T1.row * C, va1);
  // Normal: per-chain (1/sigma, log sigma) for both SELECT2 branches; an invalid sigma scores -inf
  // through inv = 0, log = +inf (distribution.rs:189-208 guards)
  double inv_a[4], lg_a[4], inv_b[4], lg_b[4];
  double tv0[4], tv1[4];
  const bool tv_bern = op.dist == FC_BERNOULLI;
  if (DK == DK_NORMAL) {
    double sa[4], sb[4];
#pragma unroll
    for (int k = 0; k < 4; k++) sa[k] = sb[k] = base1[k];
    if (SS) {
      double a[4], b[4];
      ld4(q0 + (int64_t)TS.row * C, a);
      ld4(q0 + (int64_t)TS.row_b * C, b);
#pragma unroll
      for (int k = 0; k < 4; k++) { sa[k] = base1[k] + a[k]; sb[k] = base1[k] + b[k]; }
    }
#pragma unroll
    for (int k = 0; k < 4; k++) {
      const bool oka = sa[k] > 0.0 && d_fin(sa[k]), okb = sb[k] > 0.0 && d_fin(sb[k]);
      inv_a[k] = oka ? 1.0 / sa[k] : 0.0; lg_a[k] = oka ? log(sa[k]) : d_pos_inf();
      inv_b[k] = okb ? 1.0 / sb[k] : 0.0; lg_b[k] = okb ? log(sb[k]) : d_pos_inf();
    }
  } else {
#pragma unroll
    for (int k = 0; k < 4; k++) {
      tv0[k] = dist_eval<false>(op.dist, base0[k], 0.0, 0.0, 0.0, 0, qnan).lp;
      tv1[k] = dist_eval<false>(op.dist, base0[k], 0.0, 0.0, 0.0, 1, qnan).lp;
    }
  }

  // ---- the plate: PLV_UB elements per trip, loads first (row-table lookups, then 16-byte state
  // rows / broadcast data), then the arithmetic; out-of-range slots re-read the last valid
  // element and add 0 ----
  const bool xs_samp = op.is_sample != 0, xs_cont = op.cont != 0;
  const int32_t* __restrict__ xtab = xs_cont ? A.crow + op.cslot0 : A.drow + op.dslot0;
  const int32_t* __restrict__ tab0 = A.drow + T0.dslot;  // used when M0 == VM_SELECT2
  const int32_t* __restrict__ tabs = A.drow + TS.dslot;  // used when SS
  const bool sig_shares_sel = SS && M0 == VM_SELECT2 && T0.dslot == TS.dslot;
  const int x_dtype = op.value_col.dtype;
  for (int ub = u0; ub < u1; ub += PLV_UB) {
    int ic[PLV_UB], rx[PLV_UB], r0[PLV_UB], rs[PLV_UB];
    int4 sel0[PLV_UB], sels[PLV_UB], xdv[PLV_UB];
    double lx0[PLV_UB], lx1[PLV_UB], xs[PLV_UB], xv[PLV_UB][4];
    long long xis[PLV_UB];
#pragma unroll
    for (int b = 0; b < PLV_UB; b++) {
      ic[b] = min(ub + b, u1 - 1);
      rx[b] = r0[b] = rs[b] = 0; lx0[b] = lx1[b] = xs[b] = 0.0; xis[b] = 0;
      sel0[b] = sels[b] = xdv[b] = make_int4(0, 0, 0, 0);
      xv[b][0] = xv[b][1] = xv[b][2] = xv[b][3] = 0.0;
    }
    if (xs_samp) {
#pragma unroll
      for (int b = 0; b < PLV_UB; b++) rx[b] = __ldg(xtab + ic[b]);
    }
    if (M0 == VM_SELECT2) {
#pragma unroll
      for (int b = 0; b < PLV_UB; b++) r0[b] = __ldg(tab0 + ic[b]);
    }
    if (SS && !sig_shares_sel) {
#pragma unroll
      for (int b = 0; b < PLV_UB; b++) rs[b] = __ldg(tabs + ic[b]);
    }
    if (M0 == VM_SELECT2) {
#pragma unroll
      for (int b = 0; b < PLV_UB; b++) sel0[b] = *(const int4*)(qd0 + (int64_t)r0[b] * C);
    } else if (M0 == VM_LATENT_X) {
#pragma unroll
      for (int b = 0; b < PLV_UB; b++) lx0[b] = __ldg((const double*)T0.col.ptr + ic[b]);
    }
    if (M1 == VM_LATENT_X) {
#pragma unroll
      for (int b = 0; b < PLV_UB; b++) lx1[b] = __ldg((const double*)T1.col.ptr + ic[b]);
    }
    if (SS) {
      if (sig_shares_sel) {
#pragma unroll
        for (int b = 0; b < PLV_UB; b++) sels[b] = sel0[b];
      } else {
#pragma unroll
        for (int b = 0; b < PLV_UB; b++) sels[b] = *(const int4*)(qd0 + (int64_t)rs[b] * C);
      }
    }
    if (xs_samp) {
      if (xs_cont) {
#pragma unroll
        for (int b = 0; b < PLV_UB; b++) ld4(q0 + (int64_t)rx[b] * C, xv[b]);
      } else {
#pragma unroll
        for (int b = 0; b < PLV_UB; b++) xdv[b] = *(const int4*)(qd0 + (int64_t)rx[b] * C);
      }
    } else if (xs_cont) {
      if (x_dtype == FC_COL_F64) {
#pragma unroll
        for (int b = 0; b < PLV_UB; b++) xs[b] = __ldg((const double*)op.value_col.ptr + ic[b]);
      } else {
#pragma unroll
        for (int b = 0; b < PLV_UB; b++) xs[b] = dcol_f64(op.value_col, ic[b]);
      }
    } else {
#pragma unroll
      for (int b = 0; b < PLV_UB; b++) xis[b] = dcol_i64(op.value_col, ic[b]);
    }
#pragma unroll
    for (int b = 0; b < PLV_UB; b++) {
      const bool valid = ub + b < u1;
      const int s0v[4] = {sel0[b].x, sel0[b].y, sel0[b].z, sel0[b].w};
      const int ssv[4] = {sels[b].x, sels[b].y, sels[b].z, sels[b].w};
      const int xdk[4] = {xdv[b].x, xdv[b].y, xdv[b].z, xdv[b].w};
#pragma unroll
      for (int k = 0; k < 4; k++) {
        double e;
        if (DK == DK_NORMAL) {
          double mean = base0[k];
          if (M0 == VM_SELECT2) mean += (s0v[k] != 0) ? vb0[k] : va0[k];
          if (M0 == VM_LATENT_X) mean += va0[k] * lx0[b];
          if (M1 == VM_LATENT_X) mean += va1[k] * lx1[b];
          const bool useb = SS && ssv[k] != 0;
          const double inv = useb ? inv_b[k] : inv_a[k], lg = useb ? lg_b[k] : lg_a[k];
          const double x = (xs_samp && xs_cont) ? xv[b][k] : xs[b];
          const double z = (x - mean) * inv;
          e = -0.5 * z * z - lg - 0.5 * FC_LN_2PI;
          if (!(e == e)) e = d_neg_inf();  // non-finite mean / value: the reference's guards return -inf
        } else {
          const long long xi = xs_samp ? (long long)xdk[k] : xis[b];
          e = xi == 0 ? tv0[k] : ((xi == 1 || tv_bern) ? tv1[k] : d_neg_inf());
        }
        lp[k] += valid ? e : 0.0;
      }
    }
  }
  const int lp_slot = op.is_sample ? 0 : 1;
  if (V.direct) {
    double* dst = V.lp3 + (int64_t)lp_slot * C + c0;
    double2 a = *(double2*)dst, b2 = *(double2*)(dst + 2);
    a.x += lp[0]; a.y += lp[1]; b2.x += lp[2]; b2.y += lp[3];
    *(double2*)dst = a; *(double2*)(dst + 2) = b2;
  } else {
#pragma unroll
    for (int a = 0; a < 3; a++) {
      double* dst = A.part + ((int64_t)blockIdx.y * 3 + a) * C + c0;
      const bool on = a == lp_slot;
      *(double2*)dst = make_double2(on ? lp[0] : 0.0, on ? lp[1] : 0.0);
      *(double2*)(dst + 2) = make_double2(on ? lp[2] : 0.0, on ? lp[3] : 0.0);
    }
  }
}

// host side: 0 = not eligible (caller falls back), 1 = launched with partials (caller runs
// k_finish over 3 accumulators), 2 = launched in direct mode (lp3 already updated).
int plate_vec_launch(const PlateArgs& A, int64_t n_splits, double* d_lp3, cudaStream_t st) {
  static const bool disabled = getenv("FC_PLATE_NOVEC") != nullptr;
  if (disabled) return 0;
  const DOp& op = A.op;
  if (A.grp_ptr || (A.C & 3) || A.C < 4096 || op.plate_len >= (1LL << 31)) return 0;
  VecArgs V;
  V.P = A; V.mt[0] = V.mt[1] = 0; V.st = 0; V.lp3 = d_lp3; V.direct = n_splits == 1 ? 1 : 0;
  int m[2] = {VM_NONE, VM_NONE}, nm = 0;
  bool ss = false;
  int dk;
  if (op.dist == FC_NORMAL) {
    dk = DK_NORMAL;
    if (op.n_params != 2) return 0;
    for (int j = 0; j < 2; j++)
      for (int t = 0; t < 3; t++) {
        const DTerm& tm = op.p[j].t[t];
        const int k = tm.kind;
        if (k == FC_TERM_NONE || k == FC_TERM_CONST || k == FC_TERM_LATENT) continue;
        if (op.p[j].link != FC_LINK_IDENTITY) return 0;
        if (j == 0 && k == FC_TERM_SELECT2) { if (nm == 2) return 0; m[nm] = VM_SELECT2; V.mt[nm] = t; nm++; }
        else if (j == 0 && k == FC_TERM_LATENT_X && tm.col.dtype == FC_COL_F64) { if (nm == 2) return 0; m[nm] = VM_LATENT_X; V.mt[nm] = t; nm++; }
        else if (j == 1 && k == FC_TERM_SELECT2 && !ss) { ss = true; V.st = t; }
        else return 0;
      }
    // canonical order: a SELECT2 mean term first
    if (m[1] == VM_SELECT2) {
      if (m[0] == VM_SELECT2) return 0;
      std::swap(m[0], m[1]); std::swap(V.mt[0], V.mt[1]);
    }
  } else if ((op.dist == FC_BERNOULLI || op.dist == FC_CATEGORICAL) && op.p[0].invariant) {
    dk = DK_TWOVAL;
    for (int j = 0; j < op.n_params; j++)
      for (int t = 0; t < 3; t++) {
        const int k = op.p[j].t[t].kind;
        if (!(k == FC_TERM_NONE || k == FC_TERM_CONST || k == FC_TERM_LATENT)) return 0;
      }
  } else return 0;
  dim3 grid((unsigned)cdiv(A.C, (int64_t)PLV_THREADS * PLV_VEC), (unsigned)n_splits);
#define VEC_CASE(DK_, A_, B_, S_) \
  if (dk == DK_ && m[0] == A_ && m[1] == B_ && ss == S_) { k_plate_vec<DK_, A_, B_, S_><<<grid, PLV_THREADS, 0, st>>>(V); return V.direct ? 2 : 1; }
  VEC_CASE(DK_TWOVAL, VM_NONE, VM_NONE, false)
  VEC_CASE(DK_NORMAL, VM_NONE, VM_NONE, false)
  VEC_CASE(DK_NORMAL, VM_NONE, VM_NONE, true)
  VEC_CASE(DK_NORMAL, VM_SELECT2, VM_NONE, false)
  VEC_CASE(DK_NORMAL, VM_SELECT2, VM_NONE, true)
  VEC_CASE(DK_NORMAL, VM_LATENT_X, VM_NONE, false)
  VEC_CASE(DK_NORMAL, VM_LATENT_X, VM_NONE, true)
  VEC_CASE(DK_NORMAL, VM_LATENT_X, VM_LATENT_X, false)
  VEC_CASE(DK_NORMAL, VM_LATENT_X, VM_LATENT_X, true)
  VEC_CASE(DK_NORMAL, VM_SELECT2, VM_LATENT_X, false)
  VEC_CASE(DK_NORMAL, VM_SELECT2, VM_LATENT_X, true)
#undef VEC_CASE
  return 0;
}

}  // namespace fc
