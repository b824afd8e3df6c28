// p2p_probe.cu -- de-risking probe for the peer-memory exchange used by the sharded SMC:
// two processes (fork before CUDA init), one window of device memory each, exchanged with
// cudaIpcGetMemHandle / cudaIpcOpenMemHandle; (1) latency of a device-side flag exchange
// (P2P store + spin, the all-gather primitive), (2) bandwidth of P2P stores into the peer window.
//   ./p2p_probe same   both processes on device 0 (the one-GPU test box: time-sliced contexts)
//   ./p2p_probe        process r on device r
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <unistd.h>
#include <sys/wait.h>
#include <stdint.h>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { fprintf(stderr, "[%d] %s:%d %s\n", rank, __FILE__, __LINE__, cudaGetErrorString(e)); exit(2); } } while (0)

struct Win { unsigned long long flag[2][8]; double payload[2][8][8]; };

__device__ __forceinline__ unsigned long long ld_acq(const unsigned long long* p) {
  unsigned long long v; asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory"); return v;
}
__device__ __forceinline__ void st_rel(unsigned long long* p, unsigned long long v) {
  asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ unsigned long long gtime() { unsigned long long t; asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t)); return t; }

// one exchange: write my payload + flag into every rank's window, wait for every rank's flag
__global__ void k_exchange(Win* mine, Win* peer, int rank, int world, unsigned long long seq, double val, double* out, int* err) {
  const int s = (int)(seq & 1);
  if (threadIdx.x < world) {
    Win* dst = threadIdx.x == rank ? mine : peer;
    dst->payload[s][rank][0] = val;
    __threadfence_system();
    st_rel(&dst->flag[s][rank], seq);
  }
  __syncthreads();
  if (threadIdx.x < world) {
    unsigned long long t0 = gtime();
    while (ld_acq(&mine->flag[s][threadIdx.x]) < seq) {
      if (gtime() - t0 > 5000000000ULL) { *err = 1; break; }
    }
  }
  __syncthreads();
  if (threadIdx.x == 0) { double a = 0; for (int r = 0; r < world; r++) a += mine->payload[s][r][0]; *out = a; }
}
__global__ void k_fill_peer(double* dst, size_t n, double v) {
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) dst[i] = v;
}

int main(int argc, char** argv) {
  bool same = argc > 1 && !strcmp(argv[1], "same");
  int p01[2], p10[2];
  if (pipe(p01) || pipe(p10)) return 1;
  pid_t pid = fork();
  int rank = pid == 0 ? 1 : 0;
  int rfd = rank == 0 ? p10[0] : p01[0], wfd = rank == 0 ? p01[1] : p10[1];
  int ndev = 0; CK(cudaGetDeviceCount(&ndev));
  int dev = same ? 0 : rank % ndev;
  CK(cudaSetDevice(dev));
  const size_t BIG = 256u << 20;
  char* slab; CK(cudaMalloc(&slab, sizeof(Win) + BIG)); CK(cudaMemset(slab, 0, sizeof(Win) + BIG));
  cudaIpcMemHandle_t h, hp; CK(cudaIpcGetMemHandle(&h, slab));
  if (write(wfd, &h, sizeof(h)) != sizeof(h)) return 1;
  if (read(rfd, &hp, sizeof(hp)) != sizeof(hp)) return 1;
  char* peer; CK(cudaIpcOpenMemHandle((void**)&peer, hp, cudaIpcMemLazyEnablePeerAccess));
  double* out; int* err; CK(cudaMalloc(&out, 8)); CK(cudaMalloc(&err, 4)); CK(cudaMemset(err, 0, 4));
  char c = 1; if (write(wfd, &c, 1) != 1 || read(rfd, &c, 1) != 1) return 1;  // both mapped
  cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
  const int NX = same ? 200 : 2000;
  unsigned long long seq = 1;
  for (int i = 0; i < 10; i++) k_exchange<<<1, 32>>>((Win*)slab, (Win*)peer, rank, 2, seq++, rank + 1.0, out, err);
  CK(cudaDeviceSynchronize());
  CK(cudaEventRecord(e0));
  for (int i = 0; i < NX; i++) k_exchange<<<1, 32>>>((Win*)slab, (Win*)peer, rank, 2, seq++, rank + 1.0, out, err);
  CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1));
  float ms; CK(cudaEventElapsedTime(&ms, e0, e1));
  double hout; int herr; CK(cudaMemcpy(&hout, out, 8, cudaMemcpyDeviceToHost)); CK(cudaMemcpy(&herr, err, 4, cudaMemcpyDeviceToHost));
  printf("[rank %d dev %d %s] exchange: %.2f us each over %d (sum=%g expect 3, timeout_flag=%d)\n", rank, dev, same ? "same-device" : "p2p", ms * 1e3 / NX, NX, hout, herr);
  // bandwidth of P2P stores into the peer's window
  double* pd = (double*)(peer + sizeof(Win));
  k_fill_peer<<<592, 256>>>(pd, BIG / 8, 1.0); CK(cudaDeviceSynchronize());
  CK(cudaEventRecord(e0));
  for (int i = 0; i < 5; i++) k_fill_peer<<<592, 256>>>(pd, BIG / 8, 2.0 + rank);
  CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1));
  CK(cudaEventElapsedTime(&ms, e0, e1));
  printf("[rank %d] peer store bandwidth: %.1f GB/s\n", rank, 5.0 * BIG / (ms * 1e-3) / 1e9);
  if (write(wfd, &c, 1) != 1 || read(rfd, &c, 1) != 1) return 1;  // peer finished writing into my window
  double v; CK(cudaMemcpy(&v, slab + sizeof(Win) + BIG - 8, 8, cudaMemcpyDeviceToHost));
  printf("[rank %d] last double of my window written by peer: %g (expect %g)\n", rank, v, 2.0 + (1 - rank));
  CK(cudaIpcCloseMemHandle(peer));
  if (rank == 0) { int st; waitpid(pid, &st, 0); }
  return 0;
}
