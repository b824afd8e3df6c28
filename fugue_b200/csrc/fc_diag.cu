// fc_diag.cu -- convergence diagnostics on device: split-R-hat, classic R-hat, multi-chain ESS
// (Geyer initial-positive + monotone sequence), pooled mean / std.
//
// Device restatement of src/inference/diagnostics.rs:240-304 (split_f64_chains,
// r_hat_from_f64_chains), :331-391 (pooled mean / std of summarize_f64_parameter) and
// src/inference/mcmc_utils.rs:231-339 (autocovariances, ess_from_chains). Every sum runs in the
// reference's order (draws left to right within a chain, chains in index order), and this file is
// compiled with -fmad=false, so results agree with the CPU restatement to the last bits.
//
// Draw layout: x[(s * d + k) * C + c]  (the fc_hmc_run layout [S][d][C]); one parameter alone is
// the d = 1 case [S][C]. Threads map to chains (coalesced), parameters are processed in batches.
#include "fc_internal.cuh"

namespace fc {

struct Draws {
  const double* x;
  int64_t S, d, C;
  __device__ __forceinline__ double at(int64_t s, int64_t k, int64_t c) const { return x[(s * d + k) * C + c]; }
};

// per (param, chain, segment): mean and sum of squared deviations over draws [s0, s0+len)
// segments: split = 1 -> halves [0,h), [h,2h) with h = S/2 (diagnostics.rs:240-260); split = 0 -> whole chain
__global__ void k_seg_stats(Draws D, int64_t k0, int64_t nk, int split, double* __restrict__ mean, double* __restrict__ ss) {
  int64_t c = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  int64_t k = blockIdx.y;
  if (c >= D.C || k >= nk) return;
  int64_t half = D.S / 2;
  int nseg = (split && half > 0) ? 2 : 1;
  int64_t m = (int64_t)nseg * D.C;
  for (int g = 0; g < nseg; g++) {
    int64_t s0 = (nseg == 2) ? g * half : 0, len = (nseg == 2) ? half : D.S;
    double s = 0.0;
    for (int64_t i = 0; i < len; i++) s += D.at(s0 + i, k0 + k, c);
    double mu = s / (double)len;
    double q = 0.0;
    for (int64_t i = 0; i < len; i++) { double dv = D.at(s0 + i, k0 + k, c) - mu; q += dv * dv; }
    mean[k * m + c * nseg + g] = mu;
    ss[k * m + c * nseg + g] = q;
  }
}
// r_hat_from_f64_chains (diagnostics.rs:262-304), one thread per parameter
__global__ void k_rhat_final(const double* __restrict__ mean, const double* __restrict__ ss, int64_t nk, int64_t m, double n, double* __restrict__ out) {
  int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= nk) return;
  if (m < 2) { out[k] = 1.0; return; }
  if (n == 0.0) { out[k] = __longlong_as_double(0x7ff8000000000000LL); return; }
  double mf = (double)m;
  double overall = 0.0;
  for (int64_t j = 0; j < m; j++) overall += mean[k * m + j];
  overall /= mf;
  double sb = 0.0;
  for (int64_t j = 0; j < m; j++) { double dv = mean[k * m + j] - overall; sb += dv * dv; }
  double b = n / (mf - 1.0) * sb;
  double sw = 0.0;
  for (int64_t j = 0; j < m; j++) sw += ss[k * m + j] / (n - 1.0);
  double w = sw / mf;
  double var_plus = ((n - 1.0) / n) * w + (1.0 / n) * b;
  out[k] = sqrt(var_plus / w);
}

// centred draws + chain means (mcmc_utils.rs:275-283)
__global__ void k_center(Draws D, int64_t k0, int64_t nk, double* __restrict__ cen, double* __restrict__ cmean) {
  int64_t c = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  int64_t k = blockIdx.y;
  if (c >= D.C || k >= nk) return;
  double s = 0.0;
  for (int64_t i = 0; i < D.S; i++) s += D.at(i, k0 + k, c);
  double mu = s / (double)D.S;
  cmean[k * D.C + c] = mu;
  for (int64_t i = 0; i < D.S; i++) cen[(k * D.S + i) * D.C + c] = D.at(i, k0 + k, c) - mu;
}
// autocovariances (mcmc_utils.rs:231-244): biased, sum over i < n - lag, divided by n
#define LAG_CHUNK 16
__global__ void k_acov(const double* __restrict__ cen, int64_t S, int64_t C, int64_t nk, int64_t max_lag, double* __restrict__ acov) {
  int64_t c = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  int64_t k = blockIdx.y;
  int64_t l0 = (int64_t)blockIdx.z * LAG_CHUNK;
  if (c >= C || k >= nk) return;
  const double* x = cen + k * S * C + c;
  for (int64_t lag = l0; lag < l0 + LAG_CHUNK && lag <= max_lag; lag++) {
    double a = 0.0;
    for (int64_t i = 0; i + lag < S; i++) a += x[i * C] * x[(i + lag) * C];
    acov[(k * (max_lag + 1) + lag) * C + c] = a / (double)S;
  }
}
// chain-sum of the lag-t autocovariance, chains in index order (mcmc_utils.rs:307-311)
__global__ void k_acov_chainsum(const double* __restrict__ acov, int64_t C, int64_t nk, int64_t max_lag, double* __restrict__ out) {
  int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  int64_t k = blockIdx.y;
  if (t > max_lag || k >= nk) return;
  const double* a = acov + (k * (max_lag + 1) + t) * C;
  double s = 0.0;
  for (int64_t j = 0; j < C; j++) s += a[j];
  out[k * (max_lag + 1) + t] = s;
}
// the rest of ess_from_chains (mcmc_utils.rs:285-339), one thread per parameter
__global__ void k_ess_final(const double* __restrict__ acov, const double* __restrict__ cmean, const double* __restrict__ lagsum, double* __restrict__ rh_scratch,
                            int64_t S, int64_t C, int64_t nk, int64_t max_lag, double* __restrict__ out) {
  int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= nk) return;
  double nf = (double)S, mf = (double)C;
  double mean_var = 0.0;
  for (int64_t j = 0; j < C; j++) mean_var += acov[(k * (max_lag + 1)) * C + j] * nf / (nf - 1.0);
  mean_var /= mf;
  if (mean_var <= 0.0) { out[k] = (double)(C * S); return; }
  double var_plus = mean_var * (nf - 1.0) / nf;
  if (C > 1) {
    double overall = 0.0;
    for (int64_t j = 0; j < C; j++) overall += cmean[k * C + j];
    overall /= mf;
    double b = 0.0;
    for (int64_t j = 0; j < C; j++) { double dv = cmean[k * C + j] - overall; b += dv * dv; }
    var_plus += b / (mf - 1.0);
  }
  const double* ls = lagsum + k * (max_lag + 1);
  double* rh = rh_scratch + k * (max_lag + 1);
  auto rho = [&](int64_t t) { return 1.0 - (mean_var - ls[t] / mf) / var_plus; };
  for (int64_t t = 0; t <= max_lag; t++) rh[t] = 0.0;
  rh[0] = 1.0;
  if (max_lag >= 1) rh[1] = rho(1);
  int64_t t = 1, max_t = max_lag < 1 ? max_lag : 1;
  while (t + 2 <= max_lag) {  // Geyer initial positive sequence
    double re = rho(t + 1), ro = rho(t + 2);
    if (re + ro < 0.0) break;
    rh[t + 1] = re; rh[t + 2] = ro;
    max_t = t + 2;
    t += 2;
  }
  int64_t kk = 1;
  while (kk + 2 <= max_t) {  // initial monotone sequence
    double prev = rh[kk - 1] + rh[kk], cur = rh[kk + 1] + rh[kk + 2];
    if (cur > prev) { double avg = prev / 2.0; rh[kk + 1] = avg; rh[kk + 2] = avg; }
    kk += 2;
  }
  double sum_rho = 0.0;
  for (int64_t i = 0; i <= max_t; i++) sum_rho += rh[i];
  double tau = fmax(-1.0 + 2.0 * sum_rho, 1.0);
  out[k] = (double)(C * S) / tau;
}

// pooled mean and sample std over all chains (diagnostics.rs:346-351); chain-major order
__global__ void k_chain_sum(Draws D, int64_t k0, int64_t nk, const double* __restrict__ pooled_mean, double* __restrict__ out) {
  int64_t c = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  int64_t k = blockIdx.y;
  if (c >= D.C || k >= nk) return;
  double s = 0.0;
  if (!pooled_mean) for (int64_t i = 0; i < D.S; i++) s += D.at(i, k0 + k, c);
  else { double mu = pooled_mean[k]; for (int64_t i = 0; i < D.S; i++) { double dv = D.at(i, k0 + k, c) - mu; s += dv * dv; } }
  out[k * D.C + c] = s;
}
__global__ void k_pool(const double* __restrict__ part, int64_t C, int64_t nk, double denom, int do_sqrt, double* __restrict__ out) {
  int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= nk) return;
  double s = 0.0;
  for (int64_t j = 0; j < C; j++) s += part[k * C + j];
  s /= denom;
  out[k] = do_sqrt ? sqrt(s) : s;
}

constexpr int64_t kMaxGridY = 65535;  // parameters of one batch ride on gridDim.y
struct DiagWork {
  DevBuf<double> x, a, b, cen, acov, lagsum, rh, out;
};

// what: 1 split-rhat, 2 classic rhat, 4 ess, 8 mean+std. d_draws [S][d][C] on device.
static int32_t diagnostics_dev(const double* d_draws, int64_t S, int64_t d, int64_t C, int what, double* d_mean, double* d_std, double* d_rhat, double* d_ess,
                               DiagWork& W, cudaStream_t st) {
  if (d == 0) return FC_OK;
  Draws D{d_draws, S, d, C};
  const unsigned cbx = (unsigned)cdiv(C, 128);
  // ---- R-hat ----
  if (what & 3) {
    int split = (what & 1) ? 1 : 0;
    int64_t half = S / 2;
    int nseg = (split && half > 0) ? 2 : 1;
    int64_t m = nseg * C;
    double n = (nseg == 2) ? (double)half : (double)S;
    int64_t batch = std::max<int64_t>(1, std::min<int64_t>(kMaxGridY, std::min<int64_t>(d, (int64_t)(1 << 24) / std::max<int64_t>(1, m))));
    FC_CUDA(W.a.reserve(batch * m)); FC_CUDA(W.b.reserve(batch * m));
    for (int64_t k0 = 0; k0 < d; k0 += batch) {
      int64_t nk = std::min(batch, d - k0);
      if (S == 0) { FC_CUDA(cudaMemsetAsync(W.a.p, 0, nk * m * 8, st)); FC_CUDA(cudaMemsetAsync(W.b.p, 0, nk * m * 8, st)); }
      else k_seg_stats<<<dim3(cbx, (unsigned)nk), 128, 0, st>>>(D, k0, nk, split, W.a.p, W.b.p);
      k_rhat_final<<<(unsigned)cdiv(nk, 64), 64, 0, st>>>(W.a.p, W.b.p, nk, m, S == 0 ? 0.0 : n, d_rhat + k0);
      FC_KERNEL_CHECK();
    }
  }
  // ---- pooled mean / std ----
  if (what & 8) {
    int64_t batch = std::max<int64_t>(1, std::min<int64_t>(kMaxGridY, std::min<int64_t>(d, (int64_t)(1 << 24) / std::max<int64_t>(1, C))));
    FC_CUDA(W.a.reserve(batch * C));
    for (int64_t k0 = 0; k0 < d; k0 += batch) {
      int64_t nk = std::min(batch, d - k0);
      k_chain_sum<<<dim3(cbx, (unsigned)nk), 128, 0, st>>>(D, k0, nk, nullptr, W.a.p);
      k_pool<<<(unsigned)cdiv(nk, 64), 64, 0, st>>>(W.a.p, C, nk, (double)(S * C), 0, d_mean + k0);
      k_chain_sum<<<dim3(cbx, (unsigned)nk), 128, 0, st>>>(D, k0, nk, d_mean + k0, W.a.p);
      k_pool<<<(unsigned)cdiv(nk, 64), 64, 0, st>>>(W.a.p, C, nk, (double)(S * C - 1), 1, d_std + k0);
      FC_KERNEL_CHECK();
    }
  }
  // ---- multi-chain ESS ----
  if (what & 4) {
    if (C == 0) { FC_CUDA(cudaMemsetAsync(d_ess, 0, d * 8, st)); return FC_OK; }
    if (S < 4) {  // mcmc_utils.rs:219-222: total draw count
      std::vector<double> v(d, (double)std::max<int64_t>(1, S * C));
      FC_CUDA(cudaMemcpyAsync(d_ess, v.data(), d * 8, cudaMemcpyHostToDevice, st));
      FC_CUDA(cudaStreamSynchronize(st));
      return FC_OK;
    }
    int64_t max_lag = std::min<int64_t>(S - 1, 2048);  // mcmc_utils.rs:266
    int64_t per_param = (max_lag + 1 + S) * C;
    int64_t batch = std::max<int64_t>(1, std::min<int64_t>(kMaxGridY, std::min<int64_t>(d, (int64_t)(1 << 26) / std::max<int64_t>(1, per_param))));
    FC_CUDA(W.cen.reserve(batch * S * C)); FC_CUDA(W.acov.reserve(batch * (max_lag + 1) * C));
    FC_CUDA(W.a.reserve(batch * C)); FC_CUDA(W.lagsum.reserve(batch * (max_lag + 1))); FC_CUDA(W.rh.reserve(batch * (max_lag + 1)));
    for (int64_t k0 = 0; k0 < d; k0 += batch) {
      int64_t nk = std::min(batch, d - k0);
      k_center<<<dim3(cbx, (unsigned)nk), 128, 0, st>>>(D, k0, nk, W.cen.p, W.a.p);
      k_acov<<<dim3(cbx, (unsigned)nk, (unsigned)cdiv(max_lag + 1, LAG_CHUNK)), 128, 0, st>>>(W.cen.p, S, C, nk, max_lag, W.acov.p);
      k_acov_chainsum<<<dim3((unsigned)cdiv(max_lag + 1, 128), (unsigned)nk), 128, 0, st>>>(W.acov.p, C, nk, max_lag, W.lagsum.p);
      k_ess_final<<<(unsigned)cdiv(nk, 32), 32, 0, st>>>(W.acov.p, W.a.p, W.lagsum.p, W.rh.p, S, C, nk, max_lag, d_ess + k0);
      FC_KERNEL_CHECK();
    }
  }
  return FC_OK;
}

static cudaStream_t g_diag_stream = nullptr;

// x_is_dev: the draws already live in HBM (fc_hmc_run_dev) and are read in place
static int32_t diagnostics_any(const double* draws, bool x_is_dev, size_t S, size_t d, size_t C, int what, double* h_mean, double* h_std, double* h_rhat, double* h_ess) {
  if (d == 0) return FC_OK;
  if (!g_diag_stream) FC_CUDA(cudaStreamCreateWithFlags(&g_diag_stream, cudaStreamNonBlocking));
  cudaStream_t st = g_diag_stream;
  DiagWork W;
  size_t n = S * d * C;
  const double* x = draws;
  if (!x_is_dev) {
    FC_CUDA(W.x.reserve(n));
    if (n) FC_CUDA(cudaMemcpyAsync(W.x.p, draws, n * 8, cudaMemcpyHostToDevice, st));
    x = W.x.p;
  }
  FC_CUDA(W.out.reserve(4 * d));
  double* o = W.out.p;
  FC_TRY(diagnostics_dev(x, (int64_t)S, (int64_t)d, (int64_t)C, what, o, o + d, o + 2 * d, o + 3 * d, W, st));
  if (h_mean) FC_CUDA(cudaMemcpyAsync(h_mean, o, d * 8, cudaMemcpyDeviceToHost, st));
  if (h_std) FC_CUDA(cudaMemcpyAsync(h_std, o + d, d * 8, cudaMemcpyDeviceToHost, st));
  if (h_rhat) FC_CUDA(cudaMemcpyAsync(h_rhat, o + 2 * d, d * 8, cudaMemcpyDeviceToHost, st));
  if (h_ess) FC_CUDA(cudaMemcpyAsync(h_ess, o + 3 * d, d * 8, cudaMemcpyDeviceToHost, st));
  FC_CUDA(cudaStreamSynchronize(st));
  return FC_OK;
}
static int32_t diagnostics_host(const double* h_draws, size_t S, size_t d, size_t C, int what, double* h_mean, double* h_std, double* h_rhat, double* h_ess) {
  return diagnostics_any(h_draws, false, S, d, C, what, h_mean, h_std, h_rhat, h_ess);
}

}  // namespace fc

extern "C" {

// r_hat_f64: split-R-hat (diagnostics.rs:230-260). h_draws [S][C].
int32_t fc_split_rhat(const double* h_draws, size_t S, size_t C, double* out) {
  if (C < 1) { *out = 1.0; return FC_OK; }  // < 2 chains -> 1.0 (diagnostics.rs:263-265)
  if (S == 0) { *out = C < 2 ? 1.0 : __builtin_nan(""); return FC_OK; }
  return fc::diagnostics_host(h_draws, S, 1, C, 1, nullptr, nullptr, out, nullptr);
}
int32_t fc_classic_rhat(const double* h_draws, size_t S, size_t C, double* out) {
  if (C < 2) { *out = 1.0; return FC_OK; }
  if (S == 0) { *out = __builtin_nan(""); return FC_OK; }
  return fc::diagnostics_host(h_draws, S, 1, C, 2, nullptr, nullptr, out, nullptr);
}
// effective_sample_size_multichain (mcmc_utils.rs:214-224)
int32_t fc_ess_multichain(const double* h_draws, size_t S, size_t C, double* out) {
  if (C == 0) { *out = 0.0; return FC_OK; }
  return fc::diagnostics_host(h_draws, S, 1, C, 4, nullptr, nullptr, nullptr, out);
}
// all parameters at once: h_draws [S][d][C] (the fc_hmc_run layout); outputs of length d (NULL to skip)
int32_t fc_diagnostics(const double* h_draws, size_t S, size_t d, size_t C, double* h_mean, double* h_std, double* h_rhat, double* h_ess) {
  int what = 0;
  if (h_rhat) what |= 1;
  if (h_ess) what |= 4;
  if (h_mean || h_std) what |= 8;
  if (S == 0 || C == 0) {
    for (size_t k = 0; k < d; k++) {
      if (h_mean) h_mean[k] = __builtin_nan("");
      if (h_std) h_std[k] = __builtin_nan("");
      if (h_rhat) h_rhat[k] = __builtin_nan("");
      if (h_ess) h_ess[k] = 0.0;
    }
    return FC_OK;
  }
  return fc::diagnostics_host(h_draws, S, d, C, what, h_mean, h_std, h_rhat, h_ess);
}
// fc_diagnostics on draws that never left the device (d_draws [S][d][C] in HBM, as fc_hmc_run_dev leaves
// them): the same kernels read them in place; only the 4*d results cross to the host. The caller's
// producer stream must have finished (fc_hmc_run_dev synchronises before it returns).
int32_t fc_diagnostics_dev(const double* d_draws, size_t S, size_t d, size_t C, double* h_mean, double* h_std, double* h_rhat, double* h_ess) {
  int what = 0;
  if (h_rhat) what |= 1;
  if (h_ess) what |= 4;
  if (h_mean || h_std) what |= 8;
  if (S == 0 || C == 0) return fc_diagnostics(nullptr, S, d, C, h_mean, h_std, h_rhat, h_ess);
  if (!d_draws) { fc::set_error("d_draws is NULL"); return FC_ERR_INVALID_ARGUMENT; }
  return fc::diagnostics_any(d_draws, true, S, d, C, what, h_mean, h_std, h_rhat, h_ess);
}
// ---- allocation-free split-R-hat --------------------------------------------------------------------------
// No counterpart in the reference: added for callers that cannot call
// cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it). Same kernels as
// fc_diagnostics_dev's R-hat leg (k_seg_stats -> k_rhat_final); their per-(parameter, chain, half) means and
// sums of squares live in the file-scope array below instead of a DevBuf. Every GPU that loads the library
// holds the array (16 MiB) until exit; every call shares it, so calls that overlap on the GPU race.
// Nothing in it has to start at zero: k_seg_stats overwrites every entry k_rhat_final reads.
static __device__ __align__(256) double g_rhat_noalloc_scratch[2 * FC_RHAT_NOALLOC_MAX_SEGMENTS];
int32_t fc_split_rhat_dev_noalloc(const double* d_draws, size_t S, size_t d, size_t C, double* d_rhat, void* stream) {
  using namespace fc;
  if (d == 0) return FC_OK;
  if (!d_rhat || (!d_draws && S > 0 && C > 0)) { set_error("fc_split_rhat_dev_noalloc: NULL device pointer"); return FC_ERR_INVALID_ARGUMENT; }
  const int nseg = (S / 2 > 0) ? 2 : 1;
  const size_t m = (size_t)nseg * C;
  if (m != 0 && d > (size_t)FC_RHAT_NOALLOC_MAX_SEGMENTS / m) {
    set_error("fc_split_rhat_dev_noalloc: d * segments per parameter = %zu * %zu exceeds the static scratch of %zu segments (FC_RHAT_NOALLOC_MAX_SEGMENTS); use fc_diagnostics_dev",
              d, m, (size_t)FC_RHAT_NOALLOC_MAX_SEGMENTS);
    return FC_ERR_INVALID_ARGUMENT;
  }
  // the array's name is a host-side stand-in here: resolve its device address once per device, with that device current
  static double* scratch_of_device[64] = {nullptr};
  int dev = -1;
  FC_CUDA(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 64) { set_error("fc_split_rhat_dev_noalloc: device ordinal out of range"); return FC_ERR_INVALID_ARGUMENT; }
  if (!scratch_of_device[dev]) {
    double* ptr = nullptr;
    FC_CUDA(cudaGetSymbolAddress((void**)&ptr, g_rhat_noalloc_scratch));
    scratch_of_device[dev] = ptr;
  }
  double* mean = scratch_of_device[dev];
  double* ss = mean + FC_RHAT_NOALLOC_MAX_SEGMENTS;
  cudaStream_t st = (cudaStream_t)stream;
  Draws D{d_draws, (int64_t)S, (int64_t)d, (int64_t)C};
  const double n = S == 0 ? 0.0 : (nseg == 2 ? (double)(S / 2) : (double)S);
  for (size_t k0 = 0; k0 < d; k0 += 32768) {  // gridDim.y limit
    const int64_t nk = (int64_t)std::min<size_t>(32768, d - k0);
    if (S > 0 && C > 0) k_seg_stats<<<dim3((unsigned)cdiv((int64_t)C, 128), (unsigned)nk), 128, 0, st>>>(D, (int64_t)k0, nk, 1, mean + k0 * m, ss + k0 * m);
    k_rhat_final<<<(unsigned)cdiv(nk, 64), 64, 0, st>>>(mean + k0 * m, ss + k0 * m, nk, (int64_t)m, n, d_rhat + k0);
    FC_KERNEL_CHECK();
  }
  return FC_OK;
}

// Per-rank partial statistics for multi-GPU R-hat / ESS: split-half means and sums of
// squares per chain, chain means, and the chain-SUM of autocovariances per lag. A rank holds
// C_local chains; the cross-rank combine needs only all-reduces of O(d * (lags + 3)) doubles
// plus the per-chain means (fugue_b200/dist.py).
//   h_seg_mean, h_seg_ss : [d][nseg*C]  (nseg = 2 if S >= 2 else 1; order c0h0, c0h1, c1h0, ...)
//   h_cmean              : [d][C]
//   h_lagsum             : [d][max_lag+1], max_lag = min(S-1, 2048)   (NULL or S < 4: skipped)
static int32_t diag_partials_any(const double* draws, bool x_is_dev, size_t S, size_t d, size_t C, double* h_seg_mean, double* h_seg_ss, double* h_cmean, double* h_lagsum) {
  using namespace fc;
  if (d == 0 || C == 0 || S == 0) return FC_OK;
  if (!draws) { set_error("draws is NULL"); return FC_ERR_INVALID_ARGUMENT; }
  if (!g_diag_stream) FC_CUDA(cudaStreamCreateWithFlags(&g_diag_stream, cudaStreamNonBlocking));
  cudaStream_t st = g_diag_stream;
  DiagWork W;
  size_t n = S * d * C;
  const double* x = draws;
  if (!x_is_dev) {
    FC_CUDA(W.x.reserve(n));
    FC_CUDA(cudaMemcpyAsync(W.x.p, draws, n * 8, cudaMemcpyHostToDevice, st));
    x = W.x.p;
  }
  Draws D{x, (int64_t)S, (int64_t)d, (int64_t)C};
  const unsigned cbx = (unsigned)cdiv((int64_t)C, 128);
  int nseg = (S / 2 > 0) ? 2 : 1;
  int64_t m = (int64_t)nseg * C;
  int64_t max_lag = std::min<int64_t>((int64_t)S - 1, 2048);
  bool do_acov = h_lagsum && S >= 4;
  int64_t per_param = m * 2 + (do_acov ? ((max_lag + 1 + (int64_t)S) * (int64_t)C) : 0);
  int64_t batch = std::max<int64_t>(1, std::min<int64_t>(kMaxGridY, std::min<int64_t>((int64_t)d, (int64_t)(1 << 26) / std::max<int64_t>(1, per_param))));
  FC_CUDA(W.a.reserve(batch * m)); FC_CUDA(W.b.reserve(batch * m));
  if (do_acov) { FC_CUDA(W.cen.reserve(batch * S * C)); FC_CUDA(W.acov.reserve(batch * (max_lag + 1) * C)); FC_CUDA(W.lagsum.reserve(batch * (max_lag + 1))); FC_CUDA(W.out.reserve(batch * C)); }
  for (int64_t k0 = 0; k0 < (int64_t)d; k0 += batch) {
    int64_t nk = std::min<int64_t>(batch, (int64_t)d - k0);
    k_seg_stats<<<dim3(cbx, (unsigned)nk), 128, 0, st>>>(D, k0, nk, 1, W.a.p, W.b.p);
    FC_KERNEL_CHECK();
    if (h_seg_mean) FC_CUDA(cudaMemcpyAsync(h_seg_mean + k0 * m, W.a.p, nk * m * 8, cudaMemcpyDeviceToHost, st));
    if (h_seg_ss) FC_CUDA(cudaMemcpyAsync(h_seg_ss + k0 * m, W.b.p, nk * m * 8, cudaMemcpyDeviceToHost, st));
    if (do_acov) {
      k_center<<<dim3(cbx, (unsigned)nk), 128, 0, st>>>(D, k0, nk, W.cen.p, W.out.p);
      k_acov<<<dim3(cbx, (unsigned)nk, (unsigned)cdiv(max_lag + 1, LAG_CHUNK)), 128, 0, st>>>(W.cen.p, (int64_t)S, (int64_t)C, nk, max_lag, W.acov.p);
      k_acov_chainsum<<<dim3((unsigned)cdiv(max_lag + 1, 128), (unsigned)nk), 128, 0, st>>>(W.acov.p, (int64_t)C, nk, max_lag, W.lagsum.p);
      FC_KERNEL_CHECK();
      if (h_cmean) FC_CUDA(cudaMemcpyAsync(h_cmean + k0 * C, W.out.p, nk * C * 8, cudaMemcpyDeviceToHost, st));
      FC_CUDA(cudaMemcpyAsync(h_lagsum + k0 * (max_lag + 1), W.lagsum.p, nk * (max_lag + 1) * 8, cudaMemcpyDeviceToHost, st));
    }
    FC_CUDA(cudaStreamSynchronize(st));
  }
  return FC_OK;
}
int32_t fc_diag_partials(const double* h_draws, size_t S, size_t d, size_t C, double* h_seg_mean, double* h_seg_ss, double* h_cmean, double* h_lagsum) {
  return diag_partials_any(h_draws, false, S, d, C, h_seg_mean, h_seg_ss, h_cmean, h_lagsum);
}
// the same on draws resident in HBM (fc_hmc_run_dev / fc_hmc_sample_dev): read in place, only the partials cross to the host
int32_t fc_diag_partials_dev(const double* d_draws, size_t S, size_t d, size_t C, double* h_seg_mean, double* h_seg_ss, double* h_cmean, double* h_lagsum) {
  return diag_partials_any(d_draws, true, S, d, C, h_seg_mean, h_seg_ss, h_cmean, h_lagsum);
}
}
