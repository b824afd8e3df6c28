// fc_common.cuh -- shared host/device helpers for libfugue_cuda (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>
#include "../../include/fugue_cuda.h"

namespace fc {

// ---- thread-local error message (fc_last_error) ----
void set_error(const char* fmt, ...);
int cuda_fail(cudaError_t e, const char* file, int line);

#define FC_CUDA(x)                                                       \
  do {                                                                   \
    cudaError_t _e = (x);                                                \
    if (_e != cudaSuccess) return fc::cuda_fail(_e, __FILE__, __LINE__); \
  } while (0)
#define FC_TRY(x)                \
  do {                           \
    int32_t _rc = (x);           \
    if (_rc != FC_OK) return _rc; \
  } while (0)
#define FC_KERNEL_CHECK() FC_CUDA(cudaGetLastError())

inline int64_t cdiv(int64_t a, int64_t b) { return (a + b - 1) / b; }

// ---- RAII device buffer ----
template <class T>
struct DevBuf {
  T* p = nullptr;
  size_t n = 0;
  DevBuf() {}
  DevBuf(const DevBuf&) = delete;
  DevBuf& operator=(const DevBuf&) = delete;
  ~DevBuf() { release(); }
  void release() {
    if (p) cudaFree(p);
    p = nullptr;
    n = 0;
  }
  // grows only; contents are NOT preserved
  cudaError_t reserve(size_t count) {
    if (count <= n) return cudaSuccess;
    release();
    cudaError_t e = cudaMalloc((void**)&p, (count ? count : 1) * sizeof(T));
    if (e == cudaSuccess) n = count;
    return e;
  }
};

#ifdef __CUDACC__
// ---- device numerics ----
__device__ __forceinline__ double d_neg_inf() { return __longlong_as_double(0xfff0000000000000LL); }
__device__ __forceinline__ double d_pos_inf() { return __longlong_as_double(0x7ff0000000000000LL); }
__device__ __forceinline__ bool d_fin(double x) { return isfinite(x); }

__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ double warp_max(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmax(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

// ---- Philox4x32-10 counter RNG (Salmon et al. 2011): streams are keyed by (seed, chain) and
// indexed by (iteration, slot), so draws do not depend on launch geometry. The reference uses
// ChaCha12 + ziggurat (rand 0.8.5 / rand_distr 0.4.3); stream parity is not required, only
// distributional parity. ----
struct Philox {
  uint32_t k0, k1;
  __device__ Philox(uint64_t seed, uint64_t stream) {
    uint64_t z = seed + 0x9E3779B97F4A7C15ULL * (stream + 1);  // splitmix64 key derivation
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ULL;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBULL;
    z ^= z >> 31;
    k0 = (uint32_t)z;
    k1 = (uint32_t)(z >> 32);
  }
  __device__ uint4 block(uint64_t ctr_lo, uint64_t ctr_hi) const {
    uint32_t c0 = (uint32_t)ctr_lo, c1 = (uint32_t)(ctr_lo >> 32), c2 = (uint32_t)ctr_hi, c3 = (uint32_t)(ctr_hi >> 32);
    uint32_t a = k0, b = k1;
#pragma unroll
    for (int r = 0; r < 10; r++) {
      uint32_t hi0 = __umulhi(0xD2511F53u, c0), lo0 = 0xD2511F53u * c0;
      uint32_t hi1 = __umulhi(0xCD9E8D57u, c2), lo1 = 0xCD9E8D57u * c2;
      uint32_t n0 = hi1 ^ c1 ^ a, n1 = lo1, n2 = hi0 ^ c3 ^ b, n3 = lo0;
      c0 = n0; c1 = n1; c2 = n2; c3 = n3;
      a += 0x9E3779B9u; b += 0xBB67AE85u;
    }
    return make_uint4(c0, c1, c2, c3);
  }
  // two uniforms in [0,1) with 53 bits each (rng.gen::<f64>() semantics: multiples of 2^-53)
  __device__ void uniform2(uint64_t ctr_lo, uint64_t ctr_hi, double& u0, double& u1) const {
    uint4 r = block(ctr_lo, ctr_hi);
    uint64_t a = ((uint64_t)r.x << 32) | r.y, b = ((uint64_t)r.z << 32) | r.w;
    u0 = (double)(a >> 11) * 0x1.0p-53;
    u1 = (double)(b >> 11) * 0x1.0p-53;
  }
  // two independent standard normals (Box-Muller, both branches)
  __device__ void normal2(uint64_t ctr_lo, uint64_t ctr_hi, double& z0, double& z1) const {
    double u0, u1;
    uniform2(ctr_lo, ctr_hi, u0, u1);
    u0 = 1.0 - u0;  // (0,1]
    double r = sqrt(-2.0 * log(u0));
    double s, c;
    sincospi(2.0 * u1, &s, &c);
    z0 = r * c;
    z1 = r * s;
  }
};
#endif  // __CUDACC__

}  // namespace fc
