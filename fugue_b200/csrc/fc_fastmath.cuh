// fc_fastmath.cuh -- branch-free FP64 exp / log / reciprocal for the observation-streaming
// kernels. CUDA's exp()/log()/division carry slow-path branches (denormals, specials) that cost
// ~350 issued instructions per (observation, chain) in the logistic kernel; these versions
// assume the argument ranges the kernels guarantee and stay within ~1 ulp:
//   fast_exp : x clamped to [-708, 708]                      (Cody-Waite + degree-13 Taylor)
//   fast_log : x a positive NORMAL double                    (fdlibm e_log.c polynomial)
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
