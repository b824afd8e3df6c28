// fc_mh.cuh -- the single-site move engine shared by fc_mh.cu (adaptive_mcmc_chain) and
// fc_smc.cu (tempered rejuvenation).
#pragma once
#include "fc_internal.cuh"

namespace fc {

struct SiteOverride {  // SiteProposal (mh.rs:145-161) per global site; kind -1 = automatic
  int32_t kind, _pad;
  double lower, upper;
};

struct MoveEngine {
  fc_program* P = nullptr;
  int64_t C = 0, n_sites = 0;
  bool smc_mode = false, has_overrides = false;
  uint64_t seed = 0, chain0 = 0;
  DevBuf<int32_t> target;
  DevBuf<double> old_x, lqdiff, lp3_new, total_new, scale, log_scale, sweeps;
  DevBuf<uint32_t> acc_cnt, tot_cnt;
  DevBuf<int8_t> kind;
  DevBuf<SiteOverride> overrides;
  int32_t init(fc_program* P, int64_t C, bool smc_mode, uint64_t seed, uint64_t chain0);
  int32_t set_overrides(const std::vector<SiteOverride>& ov);
  // one move for every chain: propose -> score (value only) -> accept / restore
  int32_t step(double* q, disc_t* qd, double* lp3_cur, double* total_cur, uint64_t* n_accept, uint64_t step_tag, double beta, bool adapt, double target_rate, double gamma);
  int32_t shared_adapt(double target_rate, double gamma);
};

int32_t disc_to_host(fc_program* P, const disc_t* d_qd, int64_t n, int64_t* h_out);

}  // namespace fc
