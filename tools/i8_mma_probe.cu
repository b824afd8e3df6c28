// tools/i8_mma_probe.cu -- what does the tcgen05 int8 path sustain on the shape an exact-split ("Ozaki") version of
// the C2 contractions would give it? ("beyond the DMMA roof".) Not part of the library.
//
// One CTA per SM, persistent. A tile = D[128 x 256] (int32, TMEM) = A[128 x 128] . B[256 x 128]^T (int8, shared memory,
// K-major, no swizzle: 8 x 16-byte core matrices), i.e. K = 128 = four tcgen05.mma.kind::i8 of K = 32 -- D = 100 padded
// to 128 is ONE tile deep, so every accumulator is drained after four instructions. Modes:
//   0  MMA only (operands stay in shared memory, the accumulator is never read): the instruction-issue ceiling
//   1  MMA, then all four warps read the accumulator back (tcgen05.ld 32x32b, 128 x 256 int32), serialised
//   2  the same with two accumulators: the MMAs of tile t+1 are in flight while tile t is drained
// The first tile's accumulator is written to global memory and checked against a CPU int32 GEMM.
//   (8 warps drain: a warp reaches only TMEM lanes 32 (w % 4) .. + 31; two warps share a lane quarter, half the columns each)
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -o tools/i8_mma_probe tools/i8_mma_probe.cu
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <cuda_runtime.h>

constexpr int TM = 128, TN = 256, TK = 128;
constexpr uint32_t A_SBO = 128, A_LBO = (TM / 8) * 128;  // bytes: next 8-row group; next 16-byte K chunk
constexpr uint32_t B_SBO = 128, B_LBO = (TN / 8) * 128;

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
// shared-memory matrix descriptor (cute::UMMA::SmemDescriptor): start >> 4 [0,14), LBO >> 4 [16,30), SBO >> 4 [32,46),
// version 1 [46,48), base offset 0, layout type SWIZZLE_NONE [61,64)
__device__ __forceinline__ uint64_t make_desc(uint32_t saddr, uint32_t lbo, uint32_t sbo) {
  return (uint64_t)((saddr >> 4) & 0x3fff) | ((uint64_t)((lbo >> 4) & 0x3fff) << 16) | ((uint64_t)((sbo >> 4) & 0x3fff) << 32) | (1ULL << 46);
}
// instruction descriptor (cute::UMMA::InstrDescriptor), kind::i8: D = S32 (2 at [4,6)), A and B signed 8-bit (1 at [7,10)
// and [10,13)), both K-major, N >> 3 at [17,23), M >> 4 at [24,29)
constexpr uint32_t IDESC = (2u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(TN >> 3) << 17) | ((uint32_t)(TM >> 4) << 24);

__device__ __forceinline__ void mma_i8(uint32_t tmem_d, uint64_t da, uint64_t db, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, {%5, %5, %5, %5}, p;\n\t}\n"
      :: "r"(tmem_d), "l"(da), "l"(db), "r"(IDESC), "r"(accumulate), "r"(0u));
}
__device__ __forceinline__ void mma_commit(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" :: "r"(bar) : "memory");
}
__device__ __forceinline__ bool mbar_wait(uint32_t bar, uint32_t parity) {
  for (unsigned spin = 0; spin < (1u << 22); spin++) {
    uint32_t ok;
    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}\n" : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
    if (ok) return true;
  }
  return false;  // never hang the box: the caller flags the error and carries on
}
__device__ __forceinline__ void tmem_ld64(uint32_t taddr, uint32_t (&v)[64]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x64.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, %17, %18, %19, %20, %21, %22, %23, "
      "%24, %25, %26, %27, %28, %29, %30, %31, %32, %33, %34, %35, %36, %37, %38, %39, %40, %41, %42, %43, %44, %45, %46, %47, %48, %49, %50, %51, %52, %53, "
      "%54, %55, %56, %57, %58, %59, %60, %61, %62, %63}, [%64];\n"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]),
        "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]), "=r"(v[16]), "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]),
        "=r"(v[23]), "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31]), "=r"(v[32]), "=r"(v[33]),
        "=r"(v[34]), "=r"(v[35]), "=r"(v[36]), "=r"(v[37]), "=r"(v[38]), "=r"(v[39]), "=r"(v[40]), "=r"(v[41]), "=r"(v[42]), "=r"(v[43]), "=r"(v[44]),
        "=r"(v[45]), "=r"(v[46]), "=r"(v[47]), "=r"(v[48]), "=r"(v[49]), "=r"(v[50]), "=r"(v[51]), "=r"(v[52]), "=r"(v[53]), "=r"(v[54]), "=r"(v[55]),
        "=r"(v[56]), "=r"(v[57]), "=r"(v[58]), "=r"(v[59]), "=r"(v[60]), "=r"(v[61]), "=r"(v[62]), "=r"(v[63])
      : "r"(taddr));
}

__global__ void __launch_bounds__(256, 1) k_probe(const int8_t* __restrict__ A, const int8_t* __restrict__ B, int32_t* __restrict__ D, int tiles, int mode,
                                                  int* __restrict__ err, unsigned* __restrict__ sink) {
  extern __shared__ __align__(1024) uint8_t smem[];
  uint8_t* sA = smem;                 // 16 KB
  uint8_t* sB = smem + TM * TK;       // 32 KB
  __shared__ __align__(8) unsigned long long bar[2];
  __shared__ uint32_t tmem_base_sh;
  const int tid = threadIdx.x, warp = tid >> 5;
  for (int idx = tid; idx < TM * TK; idx += 256) { const int r = idx / TK, k = idx % TK; sA[(k / 16) * A_LBO + (r / 8) * A_SBO + (r % 8) * 16 + (k % 16)] = (uint8_t)A[idx]; }
  for (int idx = tid; idx < TN * TK; idx += 256) { const int r = idx / TK, k = idx % TK; sB[(k / 16) * B_LBO + (r / 8) * B_SBO + (r % 8) * 16 + (k % 16)] = (uint8_t)B[idx]; }
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" :: "r"(smem_u32(&tmem_base_sh)), "r"(512u) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  if (tid == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" :: "r"(smem_u32(&bar[0])) : "memory");
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" :: "r"(smem_u32(&bar[1])) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // the generic-proxy fills above -> visible to the tensor core's reads
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem = tmem_base_sh;
  const uint32_t a0 = smem_u32(sA), b0 = smem_u32(sB);
  const uint32_t bar_a[2] = {smem_u32(&bar[0]), smem_u32(&bar[1])};
  auto issue = [&](int buf) {  // one tile: four K = 32 instructions into accumulator `buf`, then a commit on its barrier
    for (int k4 = 0; k4 < TK / 32; k4++)
      mma_i8(tmem + buf * TN, make_desc(a0 + k4 * 2 * A_LBO, A_LBO, A_SBO), make_desc(b0 + k4 * 2 * B_LBO, B_LBO, B_SBO), k4 > 0 ? 1u : 0u);
    mma_commit(bar_a[buf]);
  };
  unsigned acc = 0;
  bool bad = false;
  // 8 warps: warp w reads TMEM lanes 32 (w % 4) .. + 31 (a warp reaches only its own lane quarter) and the column
  // half w / 4, as two 64-column loads
  auto drain = [&](int buf, bool store) {
    const int lq = warp & 3, ch = warp >> 2;
    for (int c0 = ch * (TN / 2); c0 < (ch + 1) * (TN / 2); c0 += 64) {
      uint32_t v[64];
      tmem_ld64(tmem + buf * TN + ((uint32_t)(lq * 32) << 16) + c0, v);
      asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
      for (int j = 0; j < 64; j++) { acc += v[j]; if (store) D[(size_t)(lq * 32 + (tid & 31)) * TN + c0 + j] = (int32_t)v[j]; }
    }
  };
  uint32_t ph[2] = {0, 0};
  if (mode == 0) {
    if (tid == 0) for (int t = 0; t < tiles; t++) { for (int k4 = 0; k4 < TK / 32; k4++) mma_i8(tmem, make_desc(a0 + k4 * 2 * A_LBO, A_LBO, A_SBO), make_desc(b0 + k4 * 2 * B_LBO, B_LBO, B_SBO), k4 > 0 ? 1u : 0u); }
    if (tid == 0) mma_commit(bar_a[0]);
    if (!mbar_wait(bar_a[0], 0)) bad = true;
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    drain(0, blockIdx.x == 0);
  } else if (mode == 1) {
    for (int t = 0; t < tiles; t++) {
      if (tid == 0) issue(0);
      if (!mbar_wait(bar_a[0], ph[0])) bad = true;
      ph[0] ^= 1;
      asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
      drain(0, blockIdx.x == 0 && t == 0);
      asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
      __syncthreads();
    }
  } else {
    if (tid == 0) issue(0);
    for (int t = 0; t < tiles; t++) {
      const int buf = t & 1;
      if (tid == 0 && t + 1 < tiles) issue(buf ^ 1);  // the other accumulator was drained (and barrier-synchronised) last iteration
      if (!mbar_wait(bar_a[buf], ph[buf])) bad = true;
      ph[buf] ^= 1;
      asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
      drain(buf, blockIdx.x == 0 && t == 0);
      asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
      __syncthreads();
    }
  }
  if (bad) *err = 1;
  if (acc == 0x12345678u) *sink = acc;  // keep the drained values alive
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" :: "r"(tmem), "r"(512u) : "memory");
}

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("{\"error\": \"%s at line %d\"}\n", cudaGetErrorString(e_), __LINE__); return 1; } } while (0)

int main(int argc, char** argv) {
  const int tiles = argc > 1 ? atoi(argv[1]) : 20000;
  std::vector<int8_t> hA(TM * TK), hB(TN * TK);
  srand(1);
  for (auto& v : hA) v = (int8_t)(rand() % 127 - 63);
  for (auto& v : hB) v = (int8_t)(rand() % 127 - 63);
  std::vector<int32_t> ref(TM * TN);
  for (int m = 0; m < TM; m++)
    for (int n = 0; n < TN; n++) { int32_t s = 0; for (int k = 0; k < TK; k++) s += (int32_t)hA[m * TK + k] * (int32_t)hB[n * TK + k]; ref[m * TN + n] = s; }
  int8_t *dA, *dB; int32_t* dD; int* derr; unsigned* dsink;
  CK(cudaMalloc(&dA, hA.size())); CK(cudaMalloc(&dB, hB.size())); CK(cudaMalloc(&dD, ref.size() * 4)); CK(cudaMalloc(&derr, 4)); CK(cudaMalloc(&dsink, 4));
  CK(cudaMemcpy(dA, hA.data(), hA.size(), cudaMemcpyHostToDevice)); CK(cudaMemcpy(dB, hB.data(), hB.size(), cudaMemcpyHostToDevice));
  int sms = 0; CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0));
  const size_t smem = (size_t)TM * TK + (size_t)TN * TK;
  CK(cudaFuncSetAttribute(k_probe, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
  printf("{\"probe\": \"tcgen05.mma.kind::i8 M=128 N=256 K=128 per tile, one CTA per SM\", \"sms\": %d, \"tiles_per_cta\": %d", sms, tiles);
  const char* names[3] = {"mma_only", "mma_then_drain", "mma_overlapping_drain"};
  for (int mode = 0; mode < 3; mode++) {
    CK(cudaMemset(dD, 0, ref.size() * 4)); CK(cudaMemset(derr, 0, 4));
    k_probe<<<sms, 256, smem>>>(dA, dB, dD, 64, mode, derr, dsink);  // warm-up + the checked tile
    CK(cudaDeviceSynchronize());
    std::vector<int32_t> got(ref.size());
    int herr = 0;
    CK(cudaMemcpy(got.data(), dD, ref.size() * 4, cudaMemcpyDeviceToHost)); CK(cudaMemcpy(&herr, derr, 4, cudaMemcpyDeviceToHost));
    size_t wrong = 0;
    for (size_t i = 0; i < ref.size(); i++) wrong += got[i] != ref[i];
    CK(cudaEventRecord(e0));
    k_probe<<<sms, 256, smem>>>(dA, dB, dD, tiles, mode, derr, dsink);
    CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1));
    float ms = 0; CK(cudaEventElapsedTime(&ms, e0, e1));
    CK(cudaMemcpy(&herr, derr, 4, cudaMemcpyDeviceToHost));
    const double ops = 2.0 * TM * TN * TK * (double)tiles * sms;
    printf(", \"%s\": {\"ms\": %.3f, \"tops\": %.1f, \"mismatches_vs_cpu\": %zu, \"barrier_timeout\": %d}", names[mode], ms, ops / (ms * 1e-3) / 1e12, wrong, herr);
  }
  printf("}\n");
  return 0;
}
