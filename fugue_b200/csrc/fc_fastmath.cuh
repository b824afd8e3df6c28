// fc_fastmath.cuh -- branch-free FP64 exp / log / reciprocal for the observation-streaming
// kernels. CUDA's exp()/log()/division carry slow-path branches (denormals, specials) that cost
// ~350 issued instructions per (observation, chain) in the logistic kernel; these versions
// assume the argument ranges the kernels guarantee and stay within ~1 ulp:
//   fast_exp : x clamped to [-708, 708]                      (Cody-Waite + degree-13 Taylor)
//   fast_log : x a positive NORMAL double                    (fdlibm e_log.c polynomial)
//   fast_exp_tab : the same contract, 32-entry table in shared memory + degree-6 polynomial
//   fast_rcp : x a positive normal double in [2^-1000, 2^1000] (MUFU.RCP64H seed + 2 Newton steps)
// Accuracy is checked on the GPU against the CUDA library functions (tests/test_gpu_fastmath.py).
#pragma once
#include <cuda_runtime.h>

namespace fc {

__device__ __forceinline__ double fast_exp(double x) {
  x = fmin(fmax(x, -708.0), 708.0);
  const double L2E = 1.4426950408889634;
  const double LN2_HI = 6.93147180369123816490e-01, LN2_LO = 1.90821492927058770002e-10;
  double t = fma(x, L2E, 6755399441055744.0);  // round-to-nearest integer in the low mantissa bits
  int k = __double2loint(t);
  t -= 6755399441055744.0;
  double r = fma(t, -LN2_HI, x);
  r = fma(t, -LN2_LO, r);
  double q = 1.6059043836821613e-10;            // 1/13!
  q = fma(q, r, 2.08767569878681e-09);          // 1/12!
  q = fma(q, r, 2.505210838544172e-08);         // 1/11!
  q = fma(q, r, 2.755731922398589e-07);         // 1/10!
  q = fma(q, r, 2.7557319223985893e-06);        // 1/9!
  q = fma(q, r, 2.48015873015873e-05);          // 1/8!
  q = fma(q, r, 1.984126984126984e-04);         // 1/7!
  q = fma(q, r, 1.388888888888889e-03);         // 1/6!
  q = fma(q, r, 8.333333333333333e-03);         // 1/5!
  q = fma(q, r, 4.1666666666666664e-02);        // 1/4!
  q = fma(q, r, 1.6666666666666666e-01);        // 1/3!
  q = fma(q, r, 0.5);
  double p = fma(r * r, q, r) + 1.0;
  return p * __hiloint2double((k + 1023) << 20, 0);  // * 2^k, k in [-1022, 1022]
}

// Table-driven exp for the hottest loop (the sigmoid of the logistic GLM): x * 32/ln2 = 32k + j + f,
// exp(x) = 2^k * 2^(j/32) * exp(r) with |r| <= ln2/64, so a degree-6 polynomial suffices (r^7/7! <
// 4e-18) and the 2^k scaling is an integer add on the exponent field: 12 FP64 instructions instead
// of 22. `tab` = 32 correctly rounded values of 2^(j/32) in SHARED memory (fast_exp_tab_init).
// x is clamped to [-708, 708], so 2^k * [0.98, 2.03) stays a normal double. Error <= ~1 ulp.
__device__ __constant__ long long c_exp2_32[32] = {
    0x3ff0000000000000LL, 0x3ff059b0d3158574LL, 0x3ff0b5586cf9890fLL, 0x3ff11301d0125b51LL,
    0x3ff172b83c7d517bLL, 0x3ff1d4873168b9aaLL, 0x3ff2387a6e756238LL, 0x3ff29e9df51fdee1LL,
    0x3ff306fe0a31b715LL, 0x3ff371a7373aa9cbLL, 0x3ff3dea64c123422LL, 0x3ff44e086061892dLL,
    0x3ff4bfdad5362a27LL, 0x3ff5342b569d4f82LL, 0x3ff5ab07dd485429LL, 0x3ff6247eb03a5585LL,
    0x3ff6a09e667f3bcdLL, 0x3ff71f75e8ec5f74LL, 0x3ff7a11473eb0187LL, 0x3ff82589994cce13LL,
    0x3ff8ace5422aa0dbLL, 0x3ff93737b0cdc5e5LL, 0x3ff9c49182a3f090LL, 0x3ffa5503b23e255dLL,
    0x3ffae89f995ad3adLL, 0x3ffb7f76f2fb5e47LL, 0x3ffc199bdd85529cLL, 0x3ffcb720dcef9069LL,
    0x3ffd5818dcfba487LL, 0x3ffdfc97337b9b5fLL, 0x3ffea4afa2a490daLL, 0x3fff50765b6e4540LL};
// call with the first 32 threads of the CTA, then __syncthreads()
__device__ __forceinline__ void fast_exp_tab_init(double* tab, int tid) {
  if (tid < 32) tab[tid] = __longlong_as_double(c_exp2_32[tid]);
}
__device__ __forceinline__ double fast_exp_tab(double x, const double* __restrict__ tab) {
  x = fmin(fmax(x, -708.0), 708.0);
  const double C32 = 46.16624130844683;                  // 32 / ln 2
  const double L_HI = 0.02166084938653512, L_LO = 5.9631716539705866e-12;  // ln2/32: fdlibm's ln2HI / ln2LO scaled by 2^-5
  double t = fma(x, C32, 6755399441055744.0);
  const int ti = __double2loint(t);
  t -= 6755399441055744.0;
  double r = fma(t, -L_HI, x);                            // exact: |t| < 2^15, L_HI has 32 significant bits
  r = fma(t, -L_LO, r);
  double q = 1.388888888888889e-03;                       // 1/6!
  q = fma(q, r, 8.333333333333333e-03);                   // 1/5!
  q = fma(q, r, 4.1666666666666664e-02);                  // 1/4!
  q = fma(q, r, 1.6666666666666666e-01);                  // 1/3!
  q = fma(q, r, 0.5);
  const double p = fma(r * r, q, r);                      // exp(r) - 1
  const double T = tab[ti & 31];
  const double v = fma(T, p, T);
  return __hiloint2double(__double2hiint(v) + ((ti >> 5) << 20), __double2loint(v));
}

__device__ __forceinline__ double fast_rcp(double d) {
  double y;
  asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(d));
  double e = fma(-d, y, 1.0);   // seed is good to ~2^-20: two Newton steps reach 2^-80,
  y = fma(y, e, y);             // i.e. the result is limited by the final rounding only
  e = fma(-d, y, 1.0);
  return fma(y, e, y);
}

__device__ __forceinline__ double fast_log(double x) {
  const double LN2_HI = 6.93147180369123816490e-01, LN2_LO = 1.90821492927058770002e-10;
  int hi = __double2hiint(x), lo = __double2loint(x);
  int e = (hi >> 20) - 1023;
  hi = (hi & 0x000fffff) | 0x3ff00000;            // mantissa in [1, 2)
  if (hi > 0x3ff6a09e) { hi -= 0x00100000; e += 1; }  // > sqrt(2): halve
  double m = __hiloint2double(hi, lo);
  double f = m - 1.0;
  double s = f * fast_rcp(2.0 + f);
  double z = s * s;
  double R = 1.479819860511658591e-01;
  R = fma(R, z, 1.531383769920937332e-01);
  R = fma(R, z, 1.818357216161805012e-01);
  R = fma(R, z, 2.222219843214978396e-01);
  R = fma(R, z, 2.857142874366239149e-01);
  R = fma(R, z, 3.999999999940941908e-01);
  R = fma(R, z, 6.666666666666735130e-01);
  R *= z;
  double hfsq = 0.5 * f * f;
  double de = (double)e;
  return fma(de, LN2_HI, f - (hfsq - fma(s, hfsq + R, de * LN2_LO)));
}

}  // namespace fc
