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

// Table-driven log for the log-bound loops (Poisson / Bernoulli log-densities): x = 2^e * m with
// m in [sqrt(1/2), sqrt 2), c = round(256 m) / 256 (182 centres), d = m - c exactly (|d| <= 2^-9),
// r = d / c via a table of reciprocals, log x = e ln2 + ln c + log1p(r) with a degree-7 series
// (|r| < 2.8e-3: the first dropped term is < 1e-19): ~13 FP64 instructions instead of ~23, no
// reciprocal. ln c and 1/c are correctly rounded; c = 1 is a centre, so x near 1 keeps full RELATIVE
// accuracy. `tab` = 364 doubles in SHARED memory (fast_log_tab_init): [0,182) 1/c, [182,364) ln c.
__device__ __constant__ long long c_log_rc[182] = {
    0x3ff6a13cd1537290LL, 0x3ff6816816816817LL, 0x3ff661ec6a5122f9LL, 0x3ff642c8590b2164LL,
    0x3ff623fa77016240LL, 0x3ff6058160581606LL, 0x3ff5e75bb8d015e7LL, 0x3ff5c9882b931057LL,
    0x3ff5ac056b015ac0LL, 0x3ff58ed2308158edLL, 0x3ff571ed3c506b3aLL, 0x3ff5555555555555LL,
    0x3ff5390948f40febLL, 0x3ff51d07eae2f815LL, 0x3ff5015015015015LL, 0x3ff4e5e0a72f0539LL,
    0x3ff4cab88725af6eLL, 0x3ff4afd6a052bf5bLL, 0x3ff49539e3b2d067LL, 0x3ff47ae147ae147bLL,
    0x3ff460cbc7f5cf9aLL, 0x3ff446f86562d9fbLL, 0x3ff42d6625d51f87LL, 0x3ff4141414141414LL,
    0x3ff3fb013fb013fbLL, 0x3ff3e22cbce4a902LL, 0x3ff3c995a47babe7LL, 0x3ff3b13b13b13b14LL,
    0x3ff3991c2c187f63LL, 0x3ff3813813813814LL, 0x3ff3698df3de0748LL, 0x3ff3521cfb2b78c1LL,
    0x3ff33ae45b57bcb2LL, 0x3ff323e34a2b10bfLL, 0x3ff30d190130d190LL, 0x3ff2f684bda12f68LL,
    0x3ff2e025c04b8097LL, 0x3ff2c9fb4d812ca0LL, 0x3ff2b404ad012b40LL, 0x3ff29e4129e4129eLL,
    0x3ff288b01288b013LL, 0x3ff27350b8812735LL, 0x3ff25e22708092f1LL, 0x3ff2492492492492LL,
    0x3ff23456789abcdfLL, 0x3ff21fb78121fb78LL, 0x3ff20b470c67c0d9LL, 0x3ff1f7047dc11f70LL,
    0x3ff1e2ef3b3fb874LL, 0x3ff1cf06ada2811dLL, 0x3ff1bb4a4046ed29LL, 0x3ff1a7b9611a7b96LL,
    0x3ff19453808ca29cLL, 0x3ff1811811811812LL, 0x3ff16e0689427379LL, 0x3ff15b1e5f75270dLL,
    0x3ff1485f0e0acd3bLL, 0x3ff135c81135c811LL, 0x3ff12358e75d3033LL, 0x3ff1111111111111LL,
    0x3ff0fef010fef011LL, 0x3ff0ecf56be69c90LL, 0x3ff0db20a88f4696LL, 0x3ff0c9714fbcda3bLL,
    0x3ff0b7e6ec259dc8LL, 0x3ff0a6810a6810a7LL, 0x3ff0953f39010954LL, 0x3ff0842108421084LL,
    0x3ff073260a47f7c6LL, 0x3ff0624dd2f1a9fcLL, 0x3ff05197f7d73404LL, 0x3ff0410410410410LL,
    0x3ff03091b51f5e1aLL, 0x3ff0204081020408LL, 0x3ff0101010101010LL, 0x3ff0000000000000LL,
    0x3fefe01fe01fe020LL, 0x3fefc07f01fc07f0LL, 0x3fefa11caa01fa12LL, 0x3fef81f81f81f820LL,
    0x3fef6310aca0dbb5LL, 0x3fef44659e4a4271LL, 0x3fef25f644230ab5LL, 0x3fef07c1f07c1f08LL,
    0x3feee9c7f8458e02LL, 0x3feecc07b301ecc0LL, 0x3feeae807aba01ebLL, 0x3fee9131abf0b767LL,
    0x3fee741aa59750e4LL, 0x3fee573ac901e574LL, 0x3fee3a9179dc1a73LL, 0x3fee1e1e1e1e1e1eLL,
    0x3fee01e01e01e01eLL, 0x3fede5d6e3f8868aLL, 0x3fedca01dca01dcaLL, 0x3fedae6076b981dbLL,
    0x3fed92f2231e7f8aLL, 0x3fed77b654b82c34LL, 0x3fed5cac807572b2LL, 0x3fed41d41d41d41dLL,
    0x3fed272ca3fc5b1aLL, 0x3fed0cb58f6ec074LL, 0x3fecf26e5c44bfc6LL, 0x3fecd85689039b0bLL,
    0x3fecbe6d9601cbe7LL, 0x3feca4b3055ee191LL, 0x3fec8b265afb8a42LL, 0x3fec71c71c71c71cLL,
    0x3fec5894d10d4986LL, 0x3fec3f8f01c3f8f0LL, 0x3fec26b5392ea01cLL, 0x3fec0e070381c0e0LL,
    0x3febf583ee868d8bLL, 0x3febdd2b899406f7LL, 0x3febc4fd65883e7bLL, 0x3febacf914c1bad0LL,
    0x3feb951e2b18ff23LL, 0x3feb7d6c3dda338bLL, 0x3feb65e2e3beee05LL, 0x3feb4e81b4e81b4fLL,
    0x3feb37484ad806ceLL, 0x3feb2036406c80d9LL, 0x3feb094b31d922a4LL, 0x3feaf286bca1af28LL,
    0x3feadbe87f94905eLL, 0x3feac5701ac5701bLL, 0x3feaaf1d2f87ebfdLL, 0x3fea98ef606a63beLL,
    0x3fea82e65130e159LL, 0x3fea6d01a6d01a6dLL, 0x3fea574107688a4aLL, 0x3fea41a41a41a41aLL,
    0x3fea2c2a87c51ca0LL, 0x3fea16d3f97a4b02LL, 0x3fea01a01a01a01aLL, 0x3fe9ec8e951033d9LL,
    0x3fe9d79f176b682dLL, 0x3fe9c2d14ee4a102LL, 0x3fe9ae24ea5510daLL, 0x3fe999999999999aLL,
    0x3fe9852f0d8ec0ffLL, 0x3fe970e4f80cb872LL, 0x3fe95cbb0be377aeLL, 0x3fe948b0fcd6e9e0LL,
    0x3fe934c67f9b2ce6LL, 0x3fe920fb49d0e229LL, 0x3fe90d4f120190d5LL, 0x3fe8f9c18f9c18faLL,
    0x3fe8e6527af1373fLL, 0x3fe8d3018d3018d3LL, 0x3fe8bfce8062ff3aLL, 0x3fe8acb90f6bf3aaLL,
    0x3fe899c0f601899cLL, 0x3fe886e5f0abb04aLL, 0x3fe87427bcc092b9LL, 0x3fe8618618618618LL,
    0x3fe84f00c2780614LL, 0x3fe83c977ab2beddLL, 0x3fe82a4a0182a4a0LL, 0x3fe8181818181818LL,
    0x3fe8060180601806LL, 0x3fe7f405fd017f40LL, 0x3fe7e225515a4f1dLL, 0x3fe7d05f417d05f4LL,
    0x3fe7beb3922e017cLL, 0x3fe7ad2208e0ecc3LL, 0x3fe79baa6bb6398bLL, 0x3fe78a4c8178a4c8LL,
    0x3fe77908119ac60dLL, 0x3fe767dce434a9b1LL, 0x3fe756cac201756dLL, 0x3fe745d1745d1746LL,
    0x3fe734f0c541fe8dLL, 0x3fe724287f46debcLL, 0x3fe713786d9c7c09LL, 0x3fe702e05c0b8170LL,
    0x3fe6f26016f26017LL, 0x3fe6e1f76b4337c7LL, 0x3fe6d1a62681c861LL, 0x3fe6c16c16c16c17LL,
    0x3fe6b1490aa31a3dLL, 0x3fe6a13cd1537290LL};
__device__ __constant__ long long c_log_lc[182] = {
    -0x4029cffcf4c553b7LL, -0x402a2a4220a6a0d0LL, -0x402a8408ac372e05LL, -0x402add51f8c75c28LL,
    -0x402b361f61e8d3c4LL, -0x402b8e723d8e3be5LL, -0x402be64bdc2a1739LL, -0x402c3dad88ccce7cLL,
    -0x402c94988941eee9LL, -0x402ceb0e1e2ca31cLL, -0x402d410f83236cacLL, -0x402d969deecb246eLL,
    -0x402debba92f1472cLL, -0x402e40669ca5946bLL, -0x402e94a33453048dLL, -0x402ee8717dd81b84LL,
    -0x402f3bd2989e9d1dLL, -0x402f8ec79fb2a79eLL, -0x402fe151a9d9396fLL, -0x403066e3934c4c87LL,
    -0x40310a521b23001aLL, -0x4031acf10018efeeLL, -0x40324ec24f2b76c0LL, -0x4032efc80d9aa185LL,
    -0x40339004390ff08fLL, -0x40342f78c7c42753LL, -0x4034ce27a8a431c3LL, -0x40356c12c375261dLL,
    -0x4036093bf8f7699cLL, -0x4036a5a52308fe81LL, -0x4037415014c70174LL, -0x4037dc3e9aae5c3eLL,
    -0x403876727abbb38dLL, -0x40390fed748a9544LL, -0x4039a8b14173ecc6LL, -0x403a40bf94abc24eLL,
    -0x403ad81a1b5e4a73LL, -0x403b6ec27ccc4a9fLL, -0x403c04ba5a66d734LL, -0x403c9a034fea6feaLL,
    -0x403d2e9ef3797ec6LL, -0x403dc28ed5b63dfeLL, -0x403e55d481dc08d6LL, -0x403ee8717dd81b84LL,
    -0x403f7a674a61c5f9LL, -0x4040176ec6242a9aLL, -0x404138c67cf5eee0LL, -0x404258d89c7bb95eLL,
    -0x404377a7fe43b4ddLL, -0x404495377252a4e4LL, -0x4045b189bf4e43c8LL, -0x4046cca1a2a6b677LL,
    -0x4047e681d0bf1c10LL, -0x4048ff2cf5153f1fLL, -0x404a16a5b2686e35LL, -0x404b2ceea2df8154LL,
    -0x404c420a582e119cLL, -0x404d55fb5bb8e85bLL, -0x404e68c42eb9aa99LL, -0x404f7a674a61c5f9LL,
    -0x405115ce3ff94784LL, -0x4053348c32224d34LL, -0x4055510d2f04ef04LL, -0x40576b55eb604cbdLL,
    -0x4059836b0d2b44a8LL, -0x405b99512bd21c16LL, -0x405dad0cd072e7c1LL, -0x405fbea27618bbbcLL,
    -0x40639c2d13eb550eLL, -0x4067b6dad7373541LL, -0x406bcd56da67f33fLL, -0x406fdfa9a76ca7b9LL,
    -0x4077dbb75c775d56LL, -0x407fefea8a77218fLL, -0x408ff7faa6a774cbLL, 0x0000000000000000LL,
    0x3f6ff00aa2b10bc0LL, 0x3f7fe02a6b106789LL, 0x3f87dc475f810a77LL, 0x3f8fc0a8b0fc03e4LL,
    0x3f93cea44346a575LL, 0x3f97b91b07d5b11bLL, 0x3f9b9fc027af9198LL, 0x3f9f829b0e783300LL,
    0x3fa1b0d98923d980LL, 0x3fa39e87b9febd60LL, 0x3fa58a5bafc8e4d5LL, 0x3fa77458f632dcfcLL,
    0x3fa95c830ec8e3ebLL, 0x3fab42dd711971bfLL, 0x3fad276b8adb0b52LL, 0x3faf0a30c01162a6LL,
    0x3fb075983598e471LL, 0x3fb16536eea37ae1LL, 0x3fb253f62f0a1417LL, 0x3fb341d7961bd1d1LL,
    0x3fb42edcbea646f0LL, 0x3fb51b073f06183fLL, 0x3fb60658a93750c4LL, 0x3fb6f0d28ae56b4cLL,
    0x3fb7da766d7b12cdLL, 0x3fb8c345d6319b21LL, 0x3fb9ab42462033adLL, 0x3fba926d3a4ad563LL,
    0x3fbb78c82bb0eda1LL, 0x3fbc5e548f5bc743LL, 0x3fbd4313d66cb35dLL, 0x3fbe27076e2af2e6LL,
    0x3fbf0a30c01162a6LL, 0x3fbfec9131dbeabbLL, 0x3fc0671512ca596eLL, 0x3fc0d77e7cd08e59LL,
    0x3fc14785846742acLL, 0x3fc1b72ad52f67a0LL, 0x3fc2266f190a5acbLL, 0x3fc29552f81ff523LL,
    0x3fc303d718e47fd3LL, 0x3fc371fc201e8f74LL, 0x3fc3dfc2b0ecc62aLL, 0x3fc44d2b6ccb7d1eLL,
    0x3fc4ba36f39a55e5LL, 0x3fc526e5e3a1b438LL, 0x3fc59338d9982086LL, 0x3fc5ff3070a793d4LL,
    0x3fc66acd4272ad51LL, 0x3fc6d60fe719d21dLL, 0x3fc740f8f54037a5LL, 0x3fc7ab890210d909LL,
    0x3fc815c0a14357ebLL, 0x3fc87fa06520c911LL, 0x3fc8e928de886d41LL, 0x3fc9525a9cf456b4LL,
    0x3fc9bb362e7dfb83LL, 0x3fca23bc1fe2b563LL, 0x3fca8becfc882f19LL, 0x3fcaf3c94e80bff3LL,
    0x3fcb5b519e8fb5a4LL, 0x3fcbc286742d8cd6LL, 0x3fcc2968558c18c1LL, 0x3fcc8ff7c79a9a22LL,
    0x3fccf6354e09c5dcLL, 0x3fcd5c216b4fbb91LL, 0x3fcdc1bca0abec7dLL, 0x3fce27076e2af2e6LL,
    0x3fce8c0252aa5a60LL, 0x3fcef0adcbdc5936LL, 0x3fcf550a564b7b37LL, 0x3fcfb9186d5e3e2bLL,
    0x3fd00e6c45ad501dLL, 0x3fd0402594b4d041LL, 0x3fd071b85fcd590dLL, 0x3fd0a324e27390e3LL,
    0x3fd0d46b579ab74bLL, 0x3fd1058bf9ae4ad5LL, 0x3fd136870293a8b0LL, 0x3fd1675cababa60eLL,
    0x3fd1980d2dd4236fLL, 0x3fd1c898c16999fbLL, 0x3fd1f8ff9e48a2f3LL, 0x3fd22941fbcf7966LL,
    0x3fd2596010df763aLL, 0x3fd2895a13de86a3LL, 0x3fd2b9303ab89d25LL, 0x3fd2e8e2bae11d31LL,
    0x3fd31871c9544185LL, 0x3fd347dd9a987d55LL, 0x3fd3772662bfd85bLL, 0x3fd3a64c556945eaLL,
    0x3fd3d54fa5c1f710LL, 0x3fd404308686a7e4LL, 0x3fd432ef2a04e814LL, 0x3fd4618bc21c5ec2LL,
    0x3fd49006804009d1LL, 0x3fd4be5f957778a1LL, 0x3fd4ec973260026aLL, 0x3fd51aad872df82dLL,
    0x3fd548a2c3add263LL, 0x3fd5767717455a6cLL, 0x3fd5a42ab0f4cfe2LL, 0x3fd5d1bdbf5809caLL,
    0x3fd5ff3070a793d4LL, 0x3fd62c82f2b9c795LL};
// call with all threads of the CTA (nthreads >= 1), then __syncthreads()
__device__ __forceinline__ void fast_log_tab_init(double* tab, int tid, int nthreads) {
  for (int i = tid; i < 182; i += nthreads) { tab[i] = __longlong_as_double(c_log_rc[i]); tab[182 + i] = __longlong_as_double(c_log_lc[i]); }
}
// Polynomial / splitting constants of fast_log_tab in the CONSTANT bank: an FP64 literal with a
// non-zero low word costs two IMAD.MOV per use when the register budget (64 in the 2-CTA/SM
// value-only dense kernel) keeps the compiler from hoisting it; a c[bank][offset] operand is free.
__device__ __constant__ double c_logk[8] = {
    1.4285714285714285e-01, -1.6666666666666666e-01, 0.2, 3.3333333333333331e-01,
    6.93147180369123816490e-01, 1.90821492927058770002e-10, 256.0, 6755399441055744.0};
__device__ __forceinline__ double fast_log_tab(double x, const double* __restrict__ tab) {
  int hi = __double2hiint(x), lo = __double2loint(x);
  int e = (hi >> 20) - 1023;
  hi = (hi & 0x000fffff) | 0x3ff00000;            // mantissa in [1, 2)
  if (hi > 0x3ff6a09e) { hi -= 0x00100000; e += 1; }  // > sqrt(2): halve
  const double m = __hiloint2double(hi, lo);
  double t = fma(m, c_logk[6], c_logk[7]);         // round(256 m) in the low mantissa bits
  const int i = __double2loint(t) - 181;
  t -= c_logk[7];
  const double d = fma(t, -0.00390625, m);        // m - c, exact
  const double r = d * tab[i];
  // log1p(r) = r + r^2 (-1/2 + r/3 - r^2/4 + r^3/5 - r^4/6 + r^5/7), |r| <= 2^-8: Estrin form (depth 4
  // instead of Horner's 7 -- the kernels that call this run few warps per scheduler)
  const double r2 = r * r;
  const double qa = fma(c_logk[3], r, -0.5);       // -1/2 + r/3
  const double qb = fma(c_logk[2], r, -0.25);      // -1/4 + r/5
  const double qc = fma(c_logk[0], r, c_logk[1]);  // -1/6 + r/7
  const double q = fma(r2 * r2, qc, fma(r2, qb, qa));
  const double p = fma(r2, q, r);
  const double de = (double)e;
  return fma(de, c_logk[4], tab[182 + i] + fma(de, c_logk[5], p));  // e*ln2_hi + (ln c + (e*ln2_lo + log1p r))
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
