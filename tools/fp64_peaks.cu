// tools/fp64_peaks.cu -- measures the FP64 roofs that MEASURED_PEAKS.json does not hold:
// DFMA issue rate, DMMA (mma.sync f64) rate, and DP exp/log/div throughput on this B200.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o fp64_peaks fp64_peaks.cu
#include <cstdio>
#include <cuda_runtime.h>
#define CK(x) do{cudaError_t e=(x); if(e!=cudaSuccess){printf("CUDA error %s at %d\n",cudaGetErrorString(e),__LINE__); return 1;}}while(0)

__global__ void k_dfma(double* out, int iters) {
  double a[16]; double x = 1.0000001, y = 1e-9 * threadIdx.x;
#pragma unroll
  for (int j = 0; j < 16; j++) a[j] = j + threadIdx.x;
  for (int i = 0; i < iters; i++) {
#pragma unroll
    for (int j = 0; j < 16; j++) a[j] = fma(a[j], x, y);
  }
  double s = 0; for (int j = 0; j < 16; j++) s += a[j];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
__global__ void k_dmma884(double* out, int iters) {
  double c[8][2]; double a = 1.0 + 1e-9 * threadIdx.x, b = 1e-3;
#pragma unroll
  for (int j = 0; j < 8; j++) { c[j][0] = j; c[j][1] = -j; }
  for (int i = 0; i < iters; i++) {
#pragma unroll
    for (int j = 0; j < 8; j++)
      asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};" : "+d"(c[j][0]), "+d"(c[j][1]) : "d"(a), "d"(b));
  }
  double s = 0; for (int j = 0; j < 8; j++) s += c[j][0] + c[j][1];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
__global__ void k_dmma16816(double* out, int iters) {
  double c[4][4]; double a[8], b[4];
#pragma unroll
  for (int j = 0; j < 8; j++) a[j] = 1.0 + 1e-9 * (threadIdx.x + j);
#pragma unroll
  for (int j = 0; j < 4; j++) b[j] = 1e-3 * (j + 1);
#pragma unroll
  for (int j = 0; j < 4; j++) for (int k = 0; k < 4; k++) c[j][k] = j + k;
  for (int i = 0; i < iters; i++) {
#pragma unroll
    for (int j = 0; j < 4; j++)
      asm volatile("mma.sync.aligned.m16n8k16.row.col.f64.f64.f64.f64 {%0,%1,%2,%3}, {%4,%5,%6,%7,%8,%9,%10,%11}, {%12,%13,%14,%15}, {%0,%1,%2,%3};"
                   : "+d"(c[j][0]), "+d"(c[j][1]), "+d"(c[j][2]), "+d"(c[j][3])
                   : "d"(a[0]), "d"(a[1]), "d"(a[2]), "d"(a[3]), "d"(a[4]), "d"(a[5]), "d"(a[6]), "d"(a[7]), "d"(b[0]), "d"(b[1]), "d"(b[2]), "d"(b[3]));
  }
  double s = 0; for (int j = 0; j < 4; j++) for (int k = 0; k < 4; k++) s += c[j][k];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
template <int OP> __global__ void k_fn(double* out, int iters) {
  double v[4];
  for (int j = 0; j < 4; j++) v[j] = 0.5 + 1e-3 * (threadIdx.x + j);
  for (int i = 0; i < iters; i++) {
#pragma unroll
    for (int j = 0; j < 4; j++) {
      if (OP == 0) v[j] = exp(-v[j]) + 0.25;
      if (OP == 1) v[j] = log(v[j] + 1.5) + 0.75;
      if (OP == 2) v[j] = 1.0 / (1.0 + v[j]);
      if (OP == 3) { double p = 1.0 / (1.0 + exp(-v[j])); v[j] = log(p) + 1.2; }
      if (OP == 4) v[j] = lgamma(v[j] + 1.5) + 0.6;
    }
  }
  out[blockIdx.x * blockDim.x + threadIdx.x] = v[0] + v[1] + v[2] + v[3];
}
template <class F> float timeit(F f) {
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  f(); cudaDeviceSynchronize();
  cudaEventRecord(e0); f(); cudaEventRecord(e1); cudaEventSynchronize(e1);
  float ms; cudaEventElapsedTime(&ms, e0, e1); return ms;
}
int main() {
  cudaDeviceProp pr; CK(cudaGetDeviceProperties(&pr, 0));
  int sms = pr.multiProcessorCount; int blocks = sms * 8, thr = 256;
  double* out; CK(cudaMalloc(&out, (size_t)blocks * thr * 8));
  printf("{\"gpu\": \"%s\", \"sms\": %d", pr.name, sms);
  for (int rep = 0; rep < 2; rep++) {
    int it = 20000;
    float ms = timeit([&] { k_dfma<<<blocks, thr>>>(out, it); });
    double tf = 2.0 * 16 * it * (double)blocks * thr / ms / 1e9;
    if (rep) printf(", \"dfma_tflops\": %.2f", tf);
    ms = timeit([&] { k_dmma884<<<blocks, thr>>>(out, it); });
    tf = 2.0 * 8 * 8 * 4 * 8.0 * it * (double)blocks * thr / 32 / ms / 1e9;
    if (rep) printf(", \"dmma_m8n8k4_tflops\": %.2f", tf);
    ms = timeit([&] { k_dmma16816<<<blocks, thr>>>(out, it / 4); });
    tf = 2.0 * 16 * 8 * 16 * 4.0 * (it / 4) * (double)blocks * thr / 32 / ms / 1e9;
    if (rep) printf(", \"dmma_m16n8k16_tflops\": %.2f", tf);
    int it2 = 2000;
    const char* nm[5] = {"exp", "log", "div", "sigmoid_log", "lgamma"};
    for (int op = 0; op < 5; op++) {
      float m2;
      if (op == 0) m2 = timeit([&] { k_fn<0><<<blocks, thr>>>(out, it2); });
      if (op == 1) m2 = timeit([&] { k_fn<1><<<blocks, thr>>>(out, it2); });
      if (op == 2) m2 = timeit([&] { k_fn<2><<<blocks, thr>>>(out, it2); });
      if (op == 3) m2 = timeit([&] { k_fn<3><<<blocks, thr>>>(out, it2); });
      if (op == 4) m2 = timeit([&] { k_fn<4><<<blocks, thr>>>(out, it2); });
      if (rep) printf(", \"%s_gops\": %.1f", nm[op], 4.0 * it2 * (double)blocks * thr / m2 / 1e6);
    }
  }
  printf("}\n");
  return 0;
}
