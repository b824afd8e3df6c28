// fc_prior.cu -- the prior pass (PriorHandler replacement) as a kernel; see fc_prior.cuh.
#include "fc_prior.cuh"

namespace fc {

// All sample ops of the program in program order (scalar and plate), one thread per chain.
// `ops` holds every op of the program lowered to DOp (observes and factors are skipped).
__global__ void k_prior_sample(const DOp* __restrict__ ops, int n_ops, const int32_t* __restrict__ crow, const int32_t* __restrict__ drow,
                               double* __restrict__ q, disc_t* __restrict__ qd, int64_t C, uint64_t seed, uint64_t stream0, uint64_t tag) {
  int64_t c = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= C) return;
  DevRng rng(seed, stream0 + (uint64_t)c, tag);
  for (int oi = 0; oi < n_ops; oi++) {
    const DOp& op = ops[oi];
    if (!op.is_sample) continue;
    int64_t len = op.kind == FC_OP_PLATE_SAMPLE ? op.plate_len : 1;
    for (int64_t i = 0; i < len; i++) {
      double pv[3] = {0.0, 0.0, 0.0};
      for (int j = 0; j < op.n_params; j++) pv[j] = eval_param_value(op, j, i, q, qd, c, C, crow, drow);
      double x;
      long long xi;
      op_dist_sample(rng, op, pv[0], pv[1], pv[2], x, xi);
      if (op.cont) q[(int64_t)crow[op.cslot0 + i] * C + c] = x;
      else qd[(int64_t)drow[op.dslot0 + i] * C + c] = (disc_t)xi;
    }
  }
}


int32_t prior_sample_dev(fc_program* P, double* d_q, disc_t* d_qd, int64_t C, uint64_t seed, uint64_t stream0, uint64_t tag) {
  if (C <= 0 || P->all_ops.empty()) return FC_OK;
  k_prior_sample<<<(unsigned)cdiv(C, 128), 128, 0, P->stream>>>(P->d_all_ops.p, (int)P->all_ops.size(), P->d_crow.p, P->d_drow.p, d_q, d_qd, C, seed, stream0, tag);
  FC_KERNEL_CHECK();
  P->n_kernel_launches += 1;
  return FC_OK;
}

}  // namespace fc
