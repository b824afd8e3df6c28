// fc_internal.cuh -- program object and the internal (device-pointer) scoring interface
// shared by the sampler translation units of libfugue_cuda.
#pragma once
#include <memory>
#include "fc_common.cuh"

namespace fc {

// Storage type of discrete site values in device memory. The C ABI speaks int64
// (ChoiceValue, src/runtime/trace.rs:32-47); on device discrete planes are int32.
typedef int32_t disc_t;

struct DCol {
  const void* ptr;
  int32_t dtype, width;
  int64_t rows;
};

struct DTerm {
  int32_t kind;
  int32_t row, row_b;  // sorted q rows for LATENT / LATENT_X / SELECT2
  int32_t slot;        // program slot base for DENSE / GATHER / PLATE_LATENT
  int32_t dslot;       // SELECT2: program discrete slot base of the referenced plate
  int32_t _pad;
  double c;
  DCol col;
};

struct DParam {
  DTerm t[3];
  int32_t link;
  int32_t invariant;  // 1 if no term depends on the plate element
  double lo, hi;
};

struct DOp {
  int32_t kind, dist, value_type, n_params;
  int32_t is_sample, cont;
  int64_t plate_len;
  int64_t cslot0, dslot0;  // first program slot of a sample op
  DCol value_col;
  double value;
  const double* lgam;  // optional precomputed lgamma(value+1) per element (count data)
  int32_t expr_first, expr_len;  // FACTOR_EXPR: nodes in the program's expression table
  double lgc, lgc2;  // functions of constant parameters, evaluated once at lowering (NaN: evaluate per call): Normal ln sigma; Gamma lgamma(k), ln rate; Beta ln B(a, b)
  const double* probs;  // K-way Categorical with constant probabilities (distribution.rs:679-713): device table, else null
  int32_t n_probs, _pad2;
  DParam p[3];
};

// one accumulation target of a plate launch
struct AccMap {
  int32_t n_acc;      // 3 + number of reduced gradient rows
  int32_t rows[24];   // target sorted row of accumulator 3+i
  int8_t a_idx[3][3]; // (param, term) -> accumulator index (or -1)
  int8_t b_idx[3][3]; // SELECT2 second branch
};

struct PlateInfo {
  int op_index;
  DOp dop;
  AccMap acc;
  bool dense = false;       // handled by the dense GLM kernel
  int dense_param = -1, dense_term = -1;
  bool has_gather = false;  // iterate CSR groups
  int64_t n_groups = 0;
  DevBuf<int32_t> grp_ptr, grp_elems;
  DevBuf<double> lgam;
};

// one sample site in the global BTreeMap order over ALL addresses (continuous and discrete)
struct SiteInfo {
  int32_t is_cont;  // 1: row indexes q, 0: row indexes qd
  int32_t row;      // sorted row within its plane
  int32_t op;       // index into all_ops
  int32_t elem;     // plate element (0 for scalar ops)
  int32_t value_type;
  int32_t dist;
};

struct Workspace {
  DevBuf<double> part;   // partial accumulators of all launches
  DevBuf<double> lp3;    // [3][C]
  DevBuf<double> q, grad, total;
  DevBuf<disc_t> qd;
  DevBuf<int64_t> qd64;
};

}  // namespace fc

struct fc_program {
  std::vector<fc_op> ops;
  std::vector<std::string> names;
  std::vector<fc::DCol> cols;            // device columns
  std::vector<std::unique_ptr<fc::DevBuf<char>>> col_bufs;
  std::vector<std::vector<char>> col_host;  // host copies (CSR build, validation)
  int64_t d_cont = 0, d_disc = 0;
  std::vector<int64_t> cslot0, dslot0;   // per op
  std::vector<int64_t> perm_c, perm_d;   // sorted index -> program slot
  std::vector<int32_t> crow, drow;       // program slot -> sorted row
  std::vector<std::string> site_c, site_d;  // program-order addresses
  std::vector<std::string> sorted_c;
  std::vector<const char*> sorted_c_ptr;
  std::vector<int32_t> site_dist_c;      // per sorted continuous site: prior dist id
  fc::DevBuf<int32_t> d_crow, d_drow;
  std::vector<fc::SiteInfo> sites_all;    // every sample site, global address order (mh.rs:965)
  fc::DevBuf<fc::SiteInfo> d_sites_all;
  std::vector<std::string> sites_all_names;
  std::vector<fc::DOp> all_ops;           // every op, program order (prior pass)
  fc::DevBuf<fc::DOp> d_all_ops;
  std::vector<fc::DOp> scalar_ops;
  fc::DevBuf<fc::DOp> d_scalar_ops;
  std::vector<fc_expr_node> expr_nodes;      // FACTOR_EXPR table, latent slots already as sorted rows
  fc::DevBuf<fc_expr_node> d_expr_nodes;
  std::vector<std::unique_ptr<fc::PlateInfo>> plates;
  cudaStream_t stream = nullptr;
  int device = 0;
  fc::Workspace ws;
  uint64_t n_kernel_launches = 0;  // counted on the host at every launch site
  // optional CUDA-event timing of the dominant (plate) kernels, on the launching stream
  bool timing = false;
  std::vector<std::pair<cudaEvent_t, cudaEvent_t>> timing_events;
};

namespace fc {
// control block of the device-driven exact prefix scan (fc_resample.cu); lives at the start of the
// scan's work buffer, followed (at +64 bytes) by the per-tile aggregates
struct ScanCtl {
  long long k0;           // elements consumed so far
  double s0;              // S[k0]
  long long first_cross;  // window-relative index of the first element that reaches the binade top
  long long win;          // window policy
  unsigned cnt_a, cnt_b;  // "last CTA" counters
  int serial;             // 1: the running sum became non-finite -> finish with the sequential loop
  int round;              // rounds started (tags the look-back descriptors)
  double c0;              // the carry-in the scan started from (mean weight so far = (s0 - c0) / k0)
  double w_mean;          // caller's hint of the mean weight (0: none): sizes the first window of a scan that skips the head
};
// Scores C chains resident on the device. q [d_cont][C] (sorted rows), qd [d_disc][C].
// d_lp3 [3][C] (prior, lik, factors), d_total [C]; d_grad [d_cont][C] or nullptr (value only).
// need_value=false (only with d_grad): the caller will not read d_lp3 / d_total (interior leapfrog steps)
int32_t score_dev(fc_program* p, const double* d_q, const disc_t* d_qd, int64_t C, double* d_lp3, double* d_total, double* d_grad, bool need_value = true);
// dense GLM plate (fc_dense.cu): writes its partial segment, returns the number of splits used
int32_t dense_plate_launch(fc_program* p, PlateInfo& pl, const double* d_q, int64_t C, double* d_part, int64_t max_splits, bool grad, int64_t* n_splits, bool need_value = true);
int64_t dense_plate_splits(const fc_program* p, const PlateInfo& pl, int64_t C);
// PriorHandler replacement: draws every sample site of chains [0,C) in program order;
// chain c uses Philox stream (seed, stream0 + c), counter tag `tag`.
int32_t prior_sample_dev(fc_program* p, double* d_q, disc_t* d_qd, int64_t C, uint64_t seed, uint64_t stream0, uint64_t tag);
int32_t convert_qd(fc_program* p, const int64_t* h_qd, int64_t C, const disc_t** out);
// fused leapfrog step of hierarchical programs (fc_hier.cu): which plates play which part, the rows outside the
// group-latent plate
struct HierPlan { int pri = -1, obs = -1, gt = -1, xt = -1; int32_t n_glob = 0; int32_t glob[16]; };
extern int g_hmc_hier;
bool hier_detect(fc_program* p, HierPlan* out);
int32_t hier_step(fc_program* p, const HierPlan& H, double* q, double* mom, const double* m_inv, const double* eps, double* g_out, int32_t* divergent,
                  double* kpart, double* lp3, double* lj, int64_t C, double kick, bool drift, bool store_g, bool need_value);
}  // namespace fc
