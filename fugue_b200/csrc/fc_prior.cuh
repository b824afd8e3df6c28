// fc_prior.cuh -- device prior draws for the 17 distributions and the per-chain prior pass that
// replaces run(PriorHandler{rng}, model_fn()) (src/runtime/interpreters.rs:88-104,355-372):
// every sample site is drawn from its (state-dependent) prior in PROGRAM order.
// Samplers follow the distribution each `sample` fn of src/core/distribution.rs draws from
// (rand_distr 0.4.3 there; here Philox + textbook transforms: distributional parity only,
// SURVEY H5 / 8c "stream parity unpinned").
#pragma once
#include "fc_plate.cuh"

namespace fc {

struct DevRng {
  Philox ph;
  uint64_t hi, lo;
  double spare;
  bool has_spare;
  __device__ DevRng(uint64_t seed, uint64_t stream, uint64_t tag) : ph(seed, stream), hi(tag), lo(0), spare(0.0), has_spare(false) {}
  __device__ double uniform() {  // [0,1)
    if (has_spare) { has_spare = false; return spare; }
    double a, b;
    ph.uniform2(lo++, hi, a, b);
    spare = b; has_spare = true;
    return a;
  }
  __device__ double uniform_open() { return 1.0 - uniform(); }  // (0,1]
  __device__ double normal() {
    double u0 = uniform_open(), u1 = uniform();
    double s, c;
    sincospi(2.0 * u1, &s, &c);
    (void)s;
    return sqrt(-2.0 * log(u0)) * c;
  }
  __device__ double exponential() { return -log(uniform_open()); }
  __device__ double gamma(double k) {  // Marsaglia & Tsang (2000), unit rate
    if (!(k > 0.0) || !isfinite(k)) return __longlong_as_double(0x7ff8000000000000LL);
    // shape 1 is the exponential: one uniform and one log instead of a rejection loop whose trip count is the
    // maximum over the warp's lanes (Gamma(1, .) and Beta(1, 1) priors: a third of the SMC prior kernel's instructions)
    if (k == 1.0) return exponential();
    double boost = 1.0;
    if (k < 1.0) { boost = pow(uniform_open(), 1.0 / k); k += 1.0; }
    double d = k - 1.0 / 3.0, c = 1.0 / sqrt(9.0 * d);
    for (int it = 0; it < 1000; it++) {
      double x = normal(), v = 1.0 + c * x;
      if (v <= 0.0) continue;
      v = v * v * v;
      double u = uniform_open();
      if (u < 1.0 - 0.0331 * x * x * x * x || log(u) < 0.5 * x * x + d * (1.0 - v + log(v))) return d * v * boost;
    }
    return d * boost;
  }
  __device__ long long poisson(double lam) {
    if (!(lam > 0.0) || !isfinite(lam)) return 0;
    if (lam < 10.0) {  // Knuth multiplication
      double L = exp(-lam), p = 1.0;
      long long k = 0;
      do { k++; p *= uniform_open(); } while (p > L);
      return k - 1;
    }
    // PTRS, Hoermann (1993)
    double slam = sqrt(lam), b = 0.931 + 2.53 * slam, a = -0.059 + 0.02483 * b;
    double inv_alpha = 1.1239 + 1.1328 / (b - 3.4), vr = 0.9277 - 3.6224 / (b - 2.0);
    for (int it = 0; it < 1000; it++) {
      double u = uniform() - 0.5, v = uniform_open();
      double us = 0.5 - fabs(u);
      double kf = floor((2.0 * a / us + b) * u + lam + 0.43);
      if (us >= 0.07 && v <= vr) return (long long)kf;
      if (kf < 0.0 || (us < 0.013 && v > us)) continue;
      if (log(v) + log(inv_alpha) - log(a / (us * us) + b) <= -lam + kf * log(lam) - lgamma(kf + 1.0)) return (long long)kf;
    }
    return (long long)lam;
  }
  __device__ long long binomial(unsigned long long n, double p) {
    if (!(p > 0.0)) return 0;
    if (p >= 1.0) return (long long)n;
    bool flip = p > 0.5;
    double pp = flip ? 1.0 - p : p;
    const double qq = 1.0 - pp, s = pp / qq, u0 = uniform();
    unsigned long long k = 0;
    if ((double)n * pp < 30.0) {
      // inversion by sequential search on the pmf recurrence from 0 (expected O(n p) steps; q^n >= e^-42 here)
      double f = pow(qq, (double)n), u = u0;
      while (u > f && k < n) {
        u -= f;
        k++;
        f *= s * (double)(n - k + 1) / (double)k;
        if (f <= 0.0) break;
      }
    } else {
      // large n p: q^n underflows (n >= 1075 at p = 0.5 made every draw 1), so the search starts at the mode with
      // the pmf evaluated in log space and walks outwards, alternating sides (expected O(sqrt(n p q)) steps)
      const unsigned long long m = (unsigned long long)(((double)n + 1.0) * pp);
      const double fm = exp(lgamma((double)n + 1.0) - lgamma((double)m + 1.0) - lgamma((double)(n - m) + 1.0) + (double)m * log(pp) + (double)(n - m) * log(qq));
      double u = u0 - fm, fu = fm, fd = fm;
      unsigned long long ku = m, kd = m;
      k = m;
      for (int it = 0; u > 0.0 && it < 4000000; it++) {
        bool moved = false;
        if (ku < n) { fu *= s * (double)(n - ku) / (double)(ku + 1); ku++; u -= fu; k = ku; moved = true; if (u <= 0.0) break; }
        if (kd > 0) { fd *= (double)kd / (s * (double)(n - kd + 1)); kd--; u -= fd; k = kd; moved = true; }
        if (!moved) { k = m; break; }
      }
    }
    return (long long)(flip ? n - k : k);
  }
};

// value of parameter j of `op` at plate element i for chain c (value only)
__device__ inline double eval_param_value(const DOp& op, int j, int64_t i, const double* q, const disc_t* qd, int64_t c, int64_t C, const int32_t* crow, const int32_t* drow) {
  const DParam& pr = op.p[j];
  double s = 0.0;
  bool first = true;
  for (int t = 0; t < 3; t++) {
    const DTerm& tm = pr.t[t];
    if (tm.kind == FC_TERM_NONE) continue;
    double v = 0.0;
    switch (tm.kind) {
      case FC_TERM_CONST: v = tm.c; break;
      case FC_TERM_LATENT: v = tm.c * q[(int64_t)tm.row * C + c]; break;
      case FC_TERM_LATENT_X: v = q[(int64_t)tm.row * C + c] * ((const double*)tm.col.ptr)[i]; break;
      case FC_TERM_GATHER: {
        long long gi;
        switch (tm.col.dtype) {
          case FC_COL_U8: gi = ((const uint8_t*)tm.col.ptr)[i]; break;
          case FC_COL_U32: gi = ((const uint32_t*)tm.col.ptr)[i]; break;
          default: gi = ((const long long*)tm.col.ptr)[i]; break;
        }
        v = q[(int64_t)crow[tm.slot + gi] * C + c];
      } break;
      case FC_TERM_SELECT2: v = q[(int64_t)((qd[(int64_t)drow[tm.dslot + i] * C + c] == 0) ? tm.row : tm.row_b) * C + c]; break;
      case FC_TERM_PLATE_LATENT: v = q[(int64_t)crow[tm.slot + i] * C + c]; break;
      default: break;
    }
    s = first ? v : s + v;
    first = false;
  }
  switch (pr.link) {
    case FC_LINK_EXP: return exp(s);
    case FC_LINK_SIGMOID_CLAMPED: { double p = 1.0 / (1.0 + exp(-s)); return p < pr.lo ? pr.lo : (p > pr.hi ? pr.hi : p); }
    case FC_LINK_CLAMP: return s < pr.lo ? pr.lo : (s > pr.hi ? pr.hi : s);
    default: return s;
  }
}

// one draw; continuous result in x, discrete in xi
__device__ inline void dist_sample(DevRng& r, int dist, double P0, double P1, double P2, double& x, long long& xi) {
  x = 0.0;
  xi = 0;
  switch (dist) {
    case FC_NORMAL: x = P0 + P1 * r.normal(); break;
    case FC_UNIFORM: x = P0 + (P1 - P0) * r.uniform(); break;
    case FC_LOGNORMAL: x = exp(P0 + P1 * r.normal()); break;
    case FC_EXPONENTIAL: x = r.exponential() / P0; break;
    case FC_BERNOULLI: xi = r.uniform() < P0 ? 1 : 0; break;
    case FC_CATEGORICAL: xi = r.uniform() < P0 ? 0 : 1; break;  // CDF search over [p, 1-p] (distribution.rs:771-784)
    case FC_BETA: { double a = r.gamma(P0), b = r.gamma(P1); x = a / (a + b); } break;
    case FC_GAMMA: x = r.gamma(P0) / P1; break;
    case FC_BINOMIAL: xi = r.binomial((unsigned long long)P0, P1); break;
    case FC_POISSON: xi = r.poisson(P0); break;
    case FC_STUDENT_T: { double z = r.normal(), ch = 2.0 * r.gamma(0.5 * P0); x = P1 + P2 * z / sqrt(ch / P0); } break;
    case FC_CAUCHY: x = P0 + P1 * tan(3.141592653589793 * (r.uniform() - 0.5)); break;
    case FC_LAPLACE: { double u = r.uniform() - 0.5; x = P0 - P1 * (u < 0.0 ? -1.0 : 1.0) * log(1.0 - 2.0 * fabs(u)); } break;  // inverse CDF (distribution.rs:1524-1534)
    case FC_WEIBULL: x = P1 * pow(r.exponential(), 1.0 / P0); break;
    case FC_CHI_SQUARED: x = 2.0 * r.gamma(0.5 * P0); break;
    case FC_INVERSE_GAMMA: x = P1 / r.gamma(P0); break;  // 1 / Gamma(shape, rate) (distribution.rs:1778-1788)
    case FC_DISCRETE_UNIFORM: {
      long long lo = (long long)P0, hi = (long long)P1;
      unsigned long long span = (unsigned long long)hi - (unsigned long long)lo;
      if (span == 0xffffffffffffffffULL) { uint4 b = r.ph.block(r.lo++, r.hi); xi = (long long)(((uint64_t)b.x << 32) | b.y); }
      else { unsigned long long k = (unsigned long long)(r.uniform() * ((double)span + 1.0)); if (k > span) k = span; xi = lo + (long long)k; }
    } break;
    default: break;
  }
}

// one draw from the op's distribution, including the K-way Categorical with a constant table: u ~ U[0,1),
// first index whose running sum reaches u, capped at K - 1 (distribution.rs:755-784)
__device__ inline void op_dist_sample(DevRng& r, const DOp& op, double P0, double P1, double P2, double& x, long long& xi) {
  if (op.dist == FC_CATEGORICAL && op.n_probs > 0) {
    const double u = r.uniform();
    double acc = 0.0;
    int k = 0;
    for (; k < op.n_probs; k++) { acc += op.probs[k]; if (!(acc < u)) break; }
    x = 0.0;
    xi = k < op.n_probs ? k : op.n_probs - 1;
    return;
  }
  dist_sample(r, op.dist, P0, P1, P2, x, xi);
}

}  // namespace fc
