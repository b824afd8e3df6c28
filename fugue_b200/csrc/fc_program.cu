// fc_program.cu -- the flat op list ("program") a model is lowered to, resident on one B200.
//
// Replaces the reference's per-evaluation re-interpretation of the closure-built Model
// (run(handler, model), src/runtime/handler.rs:124-209): the op list is validated once,
// data columns are uploaded once, and the sorted site order HMC uses
// (BTreeMap<Address,_>, src/inference/hmc.rs:238-248; Ord on the address string,
// src/core/address.rs:150-157) is derived here. Structure errors use the reference's codes
// (src/error.rs:40-58); there is no CPU fallback.
#include <algorithm>
#include <cmath>
#include <cstdarg>
#include <map>
#include "fc_internal.cuh"

namespace fc {

static thread_local char g_err[512] = "";
void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}
int cuda_fail(cudaError_t e, const char* file, int line) {
  set_error("CUDA error '%s' at %s:%d", cudaGetErrorString(e), file, line);
  return FC_ERR_CUDA;
}

// addr!(name) / addr!(name, i): src/core/address.rs:175-223 ('\' and '#' are escaped)
static std::string escape_seg(const std::string& s) {
  std::string o;
  for (char ch : s) {
    if (ch == '\\') o += "\\\\";
    else if (ch == '#') o += "\\#";
    else o += ch;
  }
  return o;
}

static size_t dtype_size(int dt) { return dt == FC_COL_F64 ? 8 : dt == FC_COL_U8 ? 1 : dt == FC_COL_U32 ? 4 : 8; }

// Constructor validation of CONSTANT parameters: the `new` fns of src/core/distribution.rs
// (error codes as listed there; latent-valued parameters are checked on device => -inf).
static int32_t validate_const_params(const fc_op& op) {
  double c[3];
  bool is_c[3];
  for (int j = 0; j < 3; j++) {
    const fc_param& pr = op.param[j];
    int nt = 0;
    bool all_const = true;
    double s = 0.0;
    for (int t = 0; t < 3; t++) {
      if (pr.term[t].kind == FC_TERM_NONE) continue;
      nt++;
      if (pr.term[t].kind != FC_TERM_CONST) all_const = false;
      else s += pr.term[t].c;
    }
    is_c[j] = (j < op.n_params) && nt > 0 && all_const && pr.link == FC_LINK_IDENTITY;
    c[j] = s;
  }
  auto fin = [&](int j) { return !is_c[j] || std::isfinite(c[j]); };
  auto pos = [&](int j) { return !is_c[j] || (c[j] > 0.0 && std::isfinite(c[j])); };
  auto prob = [&](int j) { return !is_c[j] || (std::isfinite(c[j]) && c[j] >= 0.0 && c[j] <= 1.0); };
  switch (op.dist) {
    case FC_NORMAL: case FC_LOGNORMAL: case FC_CAUCHY: case FC_LAPLACE:
      if (!fin(0)) return FC_ERR_INVALID_MEAN;
      if (!pos(1)) return FC_ERR_INVALID_VARIANCE;
      break;
    case FC_UNIFORM:
      if (!fin(0) || !fin(1) || (is_c[0] && is_c[1] && !(c[0] < c[1]))) return FC_ERR_INVALID_RANGE;
      break;
    case FC_EXPONENTIAL: case FC_POISSON:
      if (!pos(0)) return FC_ERR_INVALID_RATE;
      break;
    case FC_BERNOULLI: case FC_CATEGORICAL:
      if (!prob(0)) return FC_ERR_INVALID_PROBABILITY;
      break;
    case FC_BETA:
      if (!pos(0) || !pos(1)) return FC_ERR_INVALID_SHAPE;
      break;
    case FC_GAMMA: case FC_INVERSE_GAMMA:
      if (!pos(0)) return FC_ERR_INVALID_SHAPE;
      if (!pos(1)) return FC_ERR_INVALID_RATE;
      break;
    case FC_BINOMIAL:
      if (!prob(1)) return FC_ERR_INVALID_PROBABILITY;
      break;
    case FC_STUDENT_T:
      if (!pos(0)) return FC_ERR_INVALID_SHAPE;
      if (!fin(1)) return FC_ERR_INVALID_MEAN;
      if (!pos(2)) return FC_ERR_INVALID_VARIANCE;
      break;
    case FC_WEIBULL:
      if (!pos(0)) return FC_ERR_INVALID_SHAPE;
      if (!pos(1)) return FC_ERR_INVALID_VARIANCE;
      break;
    case FC_CHI_SQUARED:
      if (!pos(0)) return FC_ERR_INVALID_SHAPE;
      break;
    case FC_DISCRETE_UNIFORM:
      if (is_c[0] && is_c[1] && !(c[0] <= c[1])) return FC_ERR_INVALID_RANGE;
      break;
    default: break;
  }
  return FC_OK;
}

static bool value_type_matches(int dist, int vt) {
  switch (dist) {
    case FC_BERNOULLI: return vt == FC_T_BOOL;
    case FC_CATEGORICAL: return vt == FC_T_USIZE;
    case FC_BINOMIAL: case FC_POISSON: return vt == FC_T_U64;
    case FC_DISCRETE_UNIFORM: return vt == FC_T_I64;
    default: return vt == FC_T_F64;
  }
}

__global__ void k_lgam_col(DCol col, int64_t n, double* out) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  double v;
  switch (col.dtype) {
    case FC_COL_F64: v = ((const double*)col.ptr)[i]; break;
    case FC_COL_U8: v = (double)((const uint8_t*)col.ptr)[i]; break;
    case FC_COL_U32: v = (double)((const uint32_t*)col.ptr)[i]; break;
    default: v = (double)((const int64_t*)col.ptr)[i]; break;
  }
  out[i] = lgamma(v + 1.0);
}

static int64_t host_col_i64(const fc_column& c, int64_t i) {
  switch (c.dtype) {
    case FC_COL_F64: return (int64_t)((const double*)c.data)[i];
    case FC_COL_U8: return ((const uint8_t*)c.data)[i];
    case FC_COL_U32: return ((const uint32_t*)c.data)[i];
    default: return ((const int64_t*)c.data)[i];
  }
}

static int32_t build(fc_program* P, const fc_op* ops, size_t n_ops, const fc_column* cols, size_t n_cols, const fc_expr_node* nodes, size_t n_nodes) {
  if (n_ops > 0 && !ops) { set_error("ops is NULL"); return FC_ERR_INVALID_ARGUMENT; }
  P->ops.assign(ops, ops + n_ops);
  P->names.resize(n_ops);
  for (size_t i = 0; i < n_ops; i++) {
    P->names[i] = ops[i].name ? ops[i].name : "";
    P->ops[i].name = nullptr;
  }
  P->cslot0.assign(n_ops, -1);
  P->dslot0.assign(n_ops, -1);
  // ---- validate ops, assign program slots ----
  for (size_t i = 0; i < n_ops; i++) {
    const fc_op& op = P->ops[i];
    if (op.kind < FC_OP_SAMPLE || op.kind > FC_OP_FACTOR_EXPR) { set_error("op %zu: bad kind %d", i, op.kind); return FC_ERR_INVALID_ARGUMENT; }
    if (op.kind == FC_OP_FACTOR) continue;
    if (op.kind == FC_OP_FACTOR_EXPR) {
      // nodes in SSA order; a factor may only read latents sampled EARLIER in program order
      if (op.expr_len < 1 || op.expr_len > FC_EXPR_MAX_NODES || op.expr_first < 0 || (size_t)op.expr_first + (size_t)op.expr_len > n_nodes || !nodes) {
        set_error("op %zu: factor expression [%d, +%d) outside the node table (%zu nodes, at most %d per expression)", i, op.expr_first, op.expr_len, n_nodes, FC_EXPR_MAX_NODES);
        return FC_ERR_INVALID_ARGUMENT;
      }
      for (int k = 0; k < op.expr_len; k++) {
        const fc_expr_node& nd = nodes[op.expr_first + k];
        if (nd.op < FC_X_CONST || nd.op > FC_X_MAX) { set_error("op %zu: bad expression opcode %d", i, nd.op); return FC_ERR_INVALID_ARGUMENT; }
        if (nd.op == FC_X_LATENT && (nd.slot < 0 || nd.slot >= P->d_cont)) { set_error("op %zu: factor reads continuous slot %d but only %lld are defined", i, nd.slot, (long long)P->d_cont); return FC_ERR_TRACE_ADDRESS_NOT_FOUND; }
        bool binary = nd.op >= FC_X_ADD && nd.op <= FC_X_DIV || nd.op >= FC_X_POW;
        if (nd.op >= FC_X_ADD && (nd.a < 0 || nd.a >= k || (binary && (nd.b < 0 || nd.b >= k)))) { set_error("op %zu: expression node %d refers to a later node", i, k); return FC_ERR_INVALID_ARGUMENT; }
      }
      continue;
    }
    if (op.dist < 0 || op.dist >= FC_N_DISTS) { set_error("op %zu: bad distribution id %d", i, op.dist); return FC_ERR_INVALID_ARGUMENT; }
    static const int np[FC_N_DISTS] = {2, 2, 2, 1, 1, 1, 2, 2, 2, 1, 3, 2, 2, 2, 1, 2, 2};
    const bool kway = op.dist == FC_CATEGORICAL && op.n_params == 0;
    if (kway) {  // Categorical::new(probs) (distribution.rs:679-713)
      if (op.probs_col < 0 || (size_t)op.probs_col >= n_cols || cols[op.probs_col].dtype != FC_COL_F64 || cols[op.probs_col].width != 1) { set_error("op %zu: K-way Categorical needs an f64 vector column of probabilities", i); return FC_ERR_INVALID_ARGUMENT; }
      const fc_column& pc = cols[op.probs_col];
      if (pc.rows <= 0) { set_error("op %zu ('%s'): Probability vector cannot be empty", i, P->names[i].c_str()); return FC_ERR_INVALID_PROBABILITY; }
      double sum = 0.0;
      for (int64_t k = 0; k < pc.rows; k++) sum += ((const double*)pc.data)[k];
      if (std::fabs(sum - 1.0) > 1e-6) { set_error("op %zu ('%s'): Probabilities must sum to 1.0 (sum %.6f)", i, P->names[i].c_str(), sum); return FC_ERR_INVALID_PROBABILITY; }
      for (int64_t k = 0; k < pc.rows; k++) { double v = ((const double*)pc.data)[k]; if (!std::isfinite(v) || v < 0.0) { set_error("op %zu ('%s'): All probabilities must be non-negative and finite", i, P->names[i].c_str()); return FC_ERR_INVALID_PROBABILITY; } }
    }
    if (!kway && op.n_params != np[op.dist]) { set_error("op %zu: distribution %d takes %d parameters, got %d", i, op.dist, np[op.dist], op.n_params); return FC_ERR_INVALID_ARGUMENT; }
    if (!value_type_matches(op.dist, op.value_type)) { set_error("op %zu ('%s'): value type %d does not match distribution %d", i, P->names[i].c_str(), op.value_type, op.dist); return FC_ERR_TYPE_MISMATCH; }
    bool plate = op.kind == FC_OP_PLATE_SAMPLE || op.kind == FC_OP_PLATE_OBSERVE;
    if (plate && op.plate_len < 0) { set_error("op %zu: negative plate length", i); return FC_ERR_INVALID_ARGUMENT; }
    if (op.kind == FC_OP_PLATE_OBSERVE) {
      if (op.value_col < 0 || (size_t)op.value_col >= n_cols) { set_error("op %zu: value column %d out of range", i, op.value_col); return FC_ERR_INVALID_ARGUMENT; }
      if (cols[op.value_col].rows < op.plate_len || cols[op.value_col].width != 1) { set_error("op %zu: value column has %lld rows (width %d), plate needs %lld", i, (long long)cols[op.value_col].rows, cols[op.value_col].width, (long long)op.plate_len); return FC_ERR_INVALID_ARGUMENT; }
    }
    for (int j = 0; j < op.n_params; j++) {
      int nterms = 0;
      for (int t = 0; t < 3; t++) {
        const fc_term& tm = op.param[j].term[t];
        if (tm.kind == FC_TERM_NONE) continue;
        nterms++;
        if (tm.kind < FC_TERM_NONE || tm.kind > FC_TERM_SCALED) { set_error("op %zu: bad term kind %d", i, tm.kind); return FC_ERR_INVALID_ARGUMENT; }
        bool elem_term = tm.kind == FC_TERM_LATENT_X || tm.kind == FC_TERM_DENSE || tm.kind == FC_TERM_GATHER || tm.kind == FC_TERM_SELECT2 || tm.kind == FC_TERM_PLATE_LATENT;
        if (elem_term && !plate) { set_error("op %zu ('%s'): per-element term %d outside a plate", i, P->names[i].c_str(), tm.kind); return FC_ERR_UNSUPPORTED_OP; }
        int64_t span = 1;
        if (tm.kind == FC_TERM_LATENT_X || tm.kind == FC_TERM_DENSE || tm.kind == FC_TERM_GATHER) {
          if (tm.col < 0 || (size_t)tm.col >= n_cols) { set_error("op %zu: term column %d out of range", i, tm.col); return FC_ERR_INVALID_ARGUMENT; }
          const fc_column& c = cols[tm.col];
          if (c.rows < op.plate_len) { set_error("op %zu: term column has %lld rows, plate needs %lld", i, (long long)c.rows, (long long)op.plate_len); return FC_ERR_INVALID_ARGUMENT; }
          if (tm.kind == FC_TERM_DENSE) {
            if (c.dtype != FC_COL_F64) { set_error("op %zu: dense term needs an f64 matrix column", i); return FC_ERR_UNSUPPORTED_OP; }
            span = c.width;
          } else if (c.width != 1) { set_error("op %zu: term column must be a vector", i); return FC_ERR_INVALID_ARGUMENT; }
          if (tm.kind == FC_TERM_LATENT_X && c.dtype != FC_COL_F64) { set_error("op %zu: LATENT_X needs an f64 column", i); return FC_ERR_UNSUPPORTED_OP; }
          if (tm.kind == FC_TERM_GATHER) {
            if (c.dtype == FC_COL_F64) { set_error("op %zu: gather index column must be integer", i); return FC_ERR_INVALID_ARGUMENT; }
            int64_t mx = -1;
            for (int64_t e = 0; e < op.plate_len; e++) {
              int64_t g = host_col_i64(c, e);
              if (g < 0) { set_error("op %zu: negative gather index", i); return FC_ERR_INVALID_ARGUMENT; }
              mx = std::max(mx, g);
            }
            span = mx + 1;
          }
        }
        if (tm.kind == FC_TERM_PLATE_LATENT) span = op.plate_len;
        if (tm.kind == FC_TERM_SELECT2) {
          if (tm.col < 0 || (size_t)tm.col >= i || P->ops[tm.col].kind != FC_OP_PLATE_SAMPLE || P->ops[tm.col].value_type == FC_T_F64 || P->ops[tm.col].plate_len < op.plate_len) {
            set_error("op %zu: SELECT2 must reference an earlier discrete plate of at least the same length", i);
            return FC_ERR_INVALID_ARGUMENT;
          }
          if (tm.slot_b < 0 || tm.slot_b >= P->d_cont) { set_error("op %zu: SELECT2 slot_b %d not defined yet", i, tm.slot_b); return FC_ERR_TRACE_ADDRESS_NOT_FOUND; }
        }
        if (tm.kind != FC_TERM_CONST) {
          // a parameter may only read latents sampled EARLIER in program order
          if (tm.slot < 0 || tm.slot + span > P->d_cont) { set_error("op %zu ('%s'): term reads continuous slots [%d,%lld) but only %lld are defined", i, P->names[i].c_str(), tm.slot, (long long)(tm.slot + span), (long long)P->d_cont); return FC_ERR_TRACE_ADDRESS_NOT_FOUND; }
        }
      }
      if (nterms == 0) { set_error("op %zu: parameter %d has no terms", i, j); return FC_ERR_INVALID_ARGUMENT; }
      int lk = op.param[j].link;
      if (lk < FC_LINK_IDENTITY || lk > FC_LINK_CLAMP) { set_error("op %zu: bad link %d", i, lk); return FC_ERR_INVALID_ARGUMENT; }
    }
    FC_TRY([&]() -> int32_t {
      int32_t rc = validate_const_params(op);
      if (rc != FC_OK) set_error("op %zu ('%s'): invalid constant parameter for distribution %d", i, P->names[i].c_str(), op.dist);
      return rc;
    }());
    if (op.kind == FC_OP_SAMPLE || op.kind == FC_OP_PLATE_SAMPLE) {
      int64_t len = op.kind == FC_OP_SAMPLE ? 1 : op.plate_len;
      bool cont = op.value_type == FC_T_F64;
      (cont ? P->cslot0 : P->dslot0)[i] = cont ? P->d_cont : P->d_disc;
      std::string base = escape_seg(P->names[i]);
      for (int64_t e = 0; e < len; e++) (cont ? P->site_c : P->site_d).push_back(op.kind == FC_OP_PLATE_SAMPLE ? base + "#" + std::to_string(e) : base);
      (cont ? P->d_cont : P->d_disc) += len;
    }
  }
  // ---- sorted (BTreeMap) order + duplicate detection over ALL sample addresses ----
  auto mk = [](const std::vector<std::string>& s, std::vector<int64_t>& perm, std::vector<int32_t>& row) {
    perm.resize(s.size());
    for (size_t i = 0; i < s.size(); i++) perm[i] = (int64_t)i;
    std::sort(perm.begin(), perm.end(), [&](int64_t a, int64_t b) { return s[a] < s[b]; });
    row.resize(s.size());
    for (size_t i = 0; i < s.size(); i++) row[perm[i]] = (int32_t)i;
  };
  mk(P->site_c, P->perm_c, P->crow);
  mk(P->site_d, P->perm_d, P->drow);
  {
    std::vector<const std::string*> all;
    for (auto& s : P->site_c) all.push_back(&s);
    for (auto& s : P->site_d) all.push_back(&s);
    std::sort(all.begin(), all.end(), [](const std::string* a, const std::string* b) { return *a < *b; });
    for (size_t i = 1; i < all.size(); i++)
      if (*all[i] == *all[i - 1]) { set_error("address '%s' is sampled twice (AddressConflict)", all[i]->c_str()); return FC_ERR_ADDRESS_CONFLICT; }
  }
  P->sorted_c.resize(P->d_cont);
  P->sorted_c_ptr.resize(P->d_cont);
  P->site_dist_c.assign(P->d_cont, 0);
  for (int64_t s = 0; s < P->d_cont; s++) P->sorted_c[s] = P->site_c[P->perm_c[s]];
  for (int64_t s = 0; s < P->d_cont; s++) P->sorted_c_ptr[s] = P->sorted_c[s].c_str();
  for (size_t i = 0; i < n_ops; i++)
    if (P->cslot0[i] >= 0) {
      int64_t len = P->ops[i].kind == FC_OP_SAMPLE ? 1 : P->ops[i].plate_len;
      for (int64_t e = 0; e < len; e++) P->site_dist_c[P->crow[P->cslot0[i] + e]] = P->ops[i].dist;
    }
  // ---- device state ----
  FC_CUDA(cudaGetDevice(&P->device));
  FC_CUDA(cudaStreamCreateWithFlags(&P->stream, cudaStreamNonBlocking));
  P->cols.resize(n_cols);
  P->col_bufs.resize(n_cols);
  for (size_t i = 0; i < n_cols; i++) {
    const fc_column& c = cols[i];
    if (c.dtype < FC_COL_F64 || c.dtype > FC_COL_I64 || c.width < 1 || c.rows < 0 || (c.rows > 0 && !c.data)) { set_error("column %zu: bad descriptor", i); return FC_ERR_INVALID_ARGUMENT; }
    size_t bytes = (size_t)c.rows * c.width * dtype_size(c.dtype);
    P->col_bufs[i].reset(new DevBuf<char>());
    FC_CUDA(P->col_bufs[i]->reserve(bytes + 16));
    if (bytes) FC_CUDA(cudaMemcpyAsync(P->col_bufs[i]->p, c.data, bytes, cudaMemcpyHostToDevice, P->stream));
    P->cols[i] = DCol{P->col_bufs[i]->p, c.dtype, c.width, c.rows};
  }
  FC_CUDA(P->d_crow.reserve(P->crow.size()));
  FC_CUDA(P->d_drow.reserve(P->drow.size()));
  if (!P->crow.empty()) FC_CUDA(cudaMemcpyAsync(P->d_crow.p, P->crow.data(), P->crow.size() * 4, cudaMemcpyHostToDevice, P->stream));
  if (!P->drow.empty()) FC_CUDA(cudaMemcpyAsync(P->d_drow.p, P->drow.data(), P->drow.size() * 4, cudaMemcpyHostToDevice, P->stream));
  // ---- lower each op to its device descriptor ----
  for (size_t i = 0; i < n_ops; i++) {
    const fc_op& op = P->ops[i];
    DOp d;
    memset(&d, 0, sizeof(d));
    d.kind = op.kind; d.dist = op.dist; d.value_type = op.value_type; d.n_params = (op.kind == FC_OP_FACTOR || op.kind == FC_OP_FACTOR_EXPR) ? 0 : op.n_params;
    d.expr_first = op.expr_first; d.expr_len = op.kind == FC_OP_FACTOR_EXPR ? op.expr_len : 0;
    d.is_sample = (op.kind == FC_OP_SAMPLE || op.kind == FC_OP_PLATE_SAMPLE);
    d.cont = op.value_type == FC_T_F64;
    d.plate_len = op.plate_len; d.cslot0 = P->cslot0[i]; d.dslot0 = P->dslot0[i];
    d.value = op.value;
    if (op.kind == FC_OP_PLATE_OBSERVE) d.value_col = P->cols[op.value_col];
    bool plate = op.kind == FC_OP_PLATE_SAMPLE || op.kind == FC_OP_PLATE_OBSERVE;
    std::unique_ptr<PlateInfo> pl;
    if (plate) { pl.reset(new PlateInfo()); pl->op_index = (int)i; }
    AccMap am;
    memset(&am, 0, sizeof(am));
    memset(am.a_idx, -1, sizeof(am.a_idx));
    memset(am.b_idx, -1, sizeof(am.b_idx));
    am.n_acc = 3;
    std::map<int32_t, int> row_acc;
    auto acc_of = [&](int32_t row) -> int {
      auto it = row_acc.find(row);
      if (it != row_acc.end()) return it->second;
      int a = am.n_acc++;
      am.rows[a - 3] = row;
      row_acc[row] = a;
      return a;
    };
    int gather_col = -1;
    int n_dense = 0;
    for (int j = 0; j < d.n_params; j++) {
      DParam& dp = d.p[j];
      dp.link = op.param[j].link; dp.lo = op.param[j].lo; dp.hi = op.param[j].hi;
      dp.invariant = 1;
      for (int t = 0; t < 3; t++) {
        const fc_term& tm = op.param[j].term[t];
        DTerm& dt = dp.t[t];
        dt.kind = tm.kind; dt.c = tm.c; dt.slot = tm.slot;
        switch (tm.kind) {
          // on the device a LATENT term is coef * q[row]: coef = 1 for LATENT, c for SCALED
          case FC_TERM_LATENT: dt.c = 1.0; dt.row = P->crow[tm.slot]; if (plate) am.a_idx[j][t] = (int8_t)acc_of(dt.row); break;
          case FC_TERM_SCALED: dt.kind = FC_TERM_LATENT; dt.c = tm.c; dt.row = P->crow[tm.slot]; if (plate) am.a_idx[j][t] = (int8_t)acc_of(dt.row); break;
          case FC_TERM_LATENT_X: dt.row = P->crow[tm.slot]; dt.col = P->cols[tm.col]; dp.invariant = 0; am.a_idx[j][t] = (int8_t)acc_of(dt.row); break;
          case FC_TERM_SELECT2:
            dt.row = P->crow[tm.slot]; dt.row_b = P->crow[tm.slot_b]; dt.dslot = (int32_t)P->dslot0[tm.col]; dp.invariant = 0;
            am.a_idx[j][t] = (int8_t)acc_of(dt.row); am.b_idx[j][t] = (int8_t)acc_of(dt.row_b);
            break;
          case FC_TERM_DENSE: dt.col = P->cols[tm.col]; dp.invariant = 0; n_dense++; pl->dense = true; pl->dense_param = j; pl->dense_term = t; break;
          case FC_TERM_GATHER:
            dt.col = P->cols[tm.col]; dp.invariant = 0;
            if (gather_col >= 0 && gather_col != tm.col) { set_error("op %zu: all GATHER terms of one plate must share one index column", i); return FC_ERR_UNSUPPORTED_OP; }
            gather_col = tm.col;
            break;
          case FC_TERM_PLATE_LATENT: dp.invariant = 0; break;
          default: break;
        }
        if (am.n_acc > 3 + 24) { set_error("op %zu: too many reduced gradient targets", i); return FC_ERR_UNSUPPORTED_OP; }
      }
    }
    if (op.kind == FC_OP_PLATE_OBSERVE && (op.dist == FC_POISSON || op.dist == FC_BINOMIAL) && op.plate_len > 0) {
      // precomputed lgamma(k+1) for count-valued observations (constant per model)
      FC_CUDA(pl->lgam.reserve(op.plate_len));
      k_lgam_col<<<(unsigned)cdiv(op.plate_len, 256), 256, 0, P->stream>>>(d.value_col, op.plate_len, pl->lgam.p);
      FC_KERNEL_CHECK();
      d.lgam = pl->lgam.p;
    }
    // normalisers of constant parameters (lgamma is ~200 FP64 instructions per call on the device;
    // C4 scores Beta(1,1) and two Gamma(1,1) priors for 2^24 particles per temper step)
    if (op.dist == FC_CATEGORICAL && op.n_params == 0 && op.kind != FC_OP_FACTOR && op.kind != FC_OP_FACTOR_EXPR) { d.probs = (const double*)P->cols[op.probs_col].ptr; d.n_probs = (int32_t)P->cols[op.probs_col].rows; }
    d.lgc = std::nan("");
    d.lgc2 = std::nan("");
    {
      auto const_param = [&](int j, double& v) -> bool {
        if (j >= d.n_params || d.p[j].link != FC_LINK_IDENTITY) return false;
        double s = 0.0;
        bool first = true;
        for (int t = 0; t < 3; t++) {
          if (d.p[j].t[t].kind == FC_TERM_NONE) continue;
          if (d.p[j].t[t].kind != FC_TERM_CONST) return false;
          s = first ? d.p[j].t[t].c : s + d.p[j].t[t].c;  // the kernels' own term fold
          first = false;
        }
        v = s;
        return !first && std::isfinite(s) && s > 0.0;
      };
      double a, b;
      if (d.dist == FC_GAMMA && const_param(0, a)) d.lgc = std::lgamma(a);
      if (d.dist == FC_GAMMA && const_param(1, b)) d.lgc2 = std::log(b);
      if (d.dist == FC_NORMAL && const_param(1, b)) d.lgc = std::log(b);
      if (d.dist == FC_BETA && const_param(0, a) && const_param(1, b)) d.lgc = std::lgamma(a) + std::lgamma(b) - std::lgamma(a + b);
    }
    P->all_ops.push_back(d);
    if (!plate) { P->scalar_ops.push_back(d); continue; }
    if (n_dense > 1) { set_error("op %zu: at most one DENSE term per plate", i); return FC_ERR_UNSUPPORTED_OP; }
    if (pl->dense) {
      if (op.kind != FC_OP_PLATE_OBSERVE) { set_error("op %zu: DENSE terms are supported in plate observes only", i); return FC_ERR_UNSUPPORTED_OP; }
      for (int j = 0; j < d.n_params; j++)
        for (int t = 0; t < 3; t++) {
          int k = d.p[j].t[t].kind;
          if (k != FC_TERM_NONE && k != FC_TERM_CONST && k != FC_TERM_LATENT && k != FC_TERM_DENSE) { set_error("op %zu: a dense GLM plate may combine DENSE only with CONST/LATENT terms", i); return FC_ERR_UNSUPPORTED_OP; }
        }
    }
    if (gather_col >= 0) {  // CSR of elements by gather index (stable => element order kept)
      const fc_column& c = cols[gather_col];
      int64_t G = 0;
      std::vector<int64_t> gi(op.plate_len);
      for (int64_t e = 0; e < op.plate_len; e++) { gi[e] = host_col_i64(c, e); G = std::max(G, gi[e] + 1); }
      std::vector<int32_t> ptr(G + 1, 0), elems(op.plate_len);
      for (int64_t e = 0; e < op.plate_len; e++) ptr[gi[e] + 1]++;
      for (int64_t g = 0; g < G; g++) ptr[g + 1] += ptr[g];
      std::vector<int32_t> fill(ptr.begin(), ptr.end() - 1);
      for (int64_t e = 0; e < op.plate_len; e++) elems[fill[gi[e]]++] = (int32_t)e;
      pl->has_gather = true; pl->n_groups = G;
      FC_CUDA(pl->grp_ptr.reserve(G + 1));
      FC_CUDA(pl->grp_elems.reserve(op.plate_len));
      FC_CUDA(cudaMemcpyAsync(pl->grp_ptr.p, ptr.data(), (G + 1) * 4, cudaMemcpyHostToDevice, P->stream));
      if (op.plate_len) FC_CUDA(cudaMemcpyAsync(pl->grp_elems.p, elems.data(), op.plate_len * 4, cudaMemcpyHostToDevice, P->stream));
      FC_CUDA(cudaStreamSynchronize(P->stream));  // ptr/elems are locals
    }
    pl->dop = d;
    pl->acc = am;
    P->plates.push_back(std::move(pl));
  }
  {  // global site table: every sample address in BTreeMap order (adaptive_mcmc_chain's `sites`, mh.rs:965)
    std::vector<std::pair<std::string, SiteInfo>> all;
    for (size_t i = 0; i < n_ops; i++) {
      const fc_op& op = P->ops[i];
      if (op.kind != FC_OP_SAMPLE && op.kind != FC_OP_PLATE_SAMPLE) continue;
      bool cont = op.value_type == FC_T_F64;
      int64_t len = op.kind == FC_OP_SAMPLE ? 1 : op.plate_len;
      for (int64_t e = 0; e < len; e++) {
        int64_t slot = (cont ? P->cslot0[i] : P->dslot0[i]) + e;
        SiteInfo si{cont ? 1 : 0, cont ? P->crow[slot] : P->drow[slot], (int32_t)i, (int32_t)e, op.value_type, op.dist};
        all.emplace_back(cont ? P->site_c[slot] : P->site_d[slot], si);
      }
    }
    std::sort(all.begin(), all.end(), [](const std::pair<std::string, SiteInfo>& a, const std::pair<std::string, SiteInfo>& b) { return a.first < b.first; });
    for (auto& a : all) { P->sites_all_names.push_back(a.first); P->sites_all.push_back(a.second); }
    FC_CUDA(P->d_sites_all.reserve(P->sites_all.size()));
    if (!P->sites_all.empty()) FC_CUDA(cudaMemcpyAsync(P->d_sites_all.p, P->sites_all.data(), P->sites_all.size() * sizeof(SiteInfo), cudaMemcpyHostToDevice, P->stream));
  }
  FC_CUDA(P->d_all_ops.reserve(P->all_ops.size()));
  if (!P->all_ops.empty()) FC_CUDA(cudaMemcpyAsync(P->d_all_ops.p, P->all_ops.data(), P->all_ops.size() * sizeof(DOp), cudaMemcpyHostToDevice, P->stream));
  // expression table of FACTOR_EXPR ops: latent slots resolved to sorted state rows
  P->expr_nodes.assign(nodes, nodes + (nodes ? n_nodes : 0));
  for (auto& nd : P->expr_nodes)
    if (nd.op == FC_X_LATENT) nd.slot = (nd.slot >= 0 && nd.slot < P->d_cont) ? P->crow[nd.slot] : 0;
  FC_CUDA(P->d_expr_nodes.reserve(std::max<size_t>(1, P->expr_nodes.size())));
  if (!P->expr_nodes.empty()) FC_CUDA(cudaMemcpyAsync(P->d_expr_nodes.p, P->expr_nodes.data(), P->expr_nodes.size() * sizeof(fc_expr_node), cudaMemcpyHostToDevice, P->stream));
  FC_CUDA(P->d_scalar_ops.reserve(P->scalar_ops.size()));
  if (!P->scalar_ops.empty()) FC_CUDA(cudaMemcpyAsync(P->d_scalar_ops.p, P->scalar_ops.data(), P->scalar_ops.size() * sizeof(DOp), cudaMemcpyHostToDevice, P->stream));
  FC_CUDA(cudaStreamSynchronize(P->stream));
  return FC_OK;
}

}  // namespace fc

extern "C" {

const char* fc_last_error(void) { return fc::g_err; }

int32_t fc_device_count(int32_t* out) {
  int n = 0;
  FC_CUDA(cudaGetDeviceCount(&n));
  *out = n;
  return FC_OK;
}
int32_t fc_set_device(int32_t device) {
  FC_CUDA(cudaSetDevice(device));
  return FC_OK;
}

int32_t fc_program_create(const fc_op* ops, size_t n_ops, const fc_column* cols, size_t n_cols, fc_program** out) {
  return fc_program_create_ex(ops, n_ops, cols, n_cols, nullptr, 0, out);
}

int32_t fc_program_create_ex(const fc_op* ops, size_t n_ops, const fc_column* cols, size_t n_cols, const fc_expr_node* nodes, size_t n_nodes, fc_program** out) {
  if (!out) { fc::set_error("out is NULL"); return FC_ERR_INVALID_ARGUMENT; }
  *out = nullptr;
  fc_program* P = new fc_program();
  int32_t rc = fc::build(P, ops, n_ops, cols, n_cols, nodes, n_nodes);
  if (rc != FC_OK) { fc_program_destroy(P); return rc; }
  *out = P;
  return FC_OK;
}

int32_t fc_program_destroy(fc_program* p) {
  if (!p) return FC_OK;
  if (p->stream) { cudaStreamSynchronize(p->stream); cudaStreamDestroy(p->stream); }
  for (auto& e : p->timing_events) { cudaEventDestroy(e.first); cudaEventDestroy(e.second); }
  delete p;
  return FC_OK;
}

// A second trace of the same model must produce the same address sequence, kinds, value types
// and distributions (score_given_trace_strict, src/runtime/interpreters.rs:271-283,667-679).
int32_t fc_program_check_structure(const fc_program* p, const fc_op* ops, size_t n_ops) {
  if (n_ops != p->ops.size()) { fc::set_error("model structure changed: %zu ops traced, program has %zu", n_ops, p->ops.size()); return FC_ERR_UNEXPECTED_MODEL_STRUCTURE; }
  for (size_t i = 0; i < n_ops; i++) {
    const fc_op& a = p->ops[i];
    const fc_op& b = ops[i];
    std::string nb = b.name ? b.name : "";
    if (a.kind != b.kind || a.dist != b.dist || a.value_type != b.value_type || a.plate_len != b.plate_len || nb != p->names[i] ||
        (a.kind == FC_OP_FACTOR_EXPR && a.expr_len != b.expr_len)) {
      fc::set_error("model structure changed at op %zu: '%s' (kind %d, dist %d, len %lld) vs traced '%s' (kind %d, dist %d, len %lld)", i, p->names[i].c_str(), a.kind, a.dist, (long long)a.plate_len, nb.c_str(), b.kind, b.dist, (long long)b.plate_len);
      return FC_ERR_UNEXPECTED_MODEL_STRUCTURE;
    }
  }
  return FC_OK;
}

int32_t fc_program_dims(const fc_program* p, int64_t* d_cont, int64_t* d_disc) {
  if (d_cont) *d_cont = p->d_cont;
  if (d_disc) *d_disc = p->d_disc;
  return FC_OK;
}

int32_t fc_program_sites(const fc_program* p, const char* const** names, const int64_t** perm, int64_t* n) {
  if (names) *names = p->sorted_c_ptr.data();
  if (perm) *perm = p->perm_c.data();
  if (n) *n = p->d_cont;
  return FC_OK;
}

int32_t fc_program_stream(const fc_program* p, void** stream) {
  *stream = (void*)p->stream;
  return FC_OK;
}
int32_t fc_program_device(const fc_program* p, int32_t* device) {
  *device = (int32_t)p->device;
  return FC_OK;
}

// CUDA-event timing of the dominant plate kernel (the dense GLM kernel), recorded on the
// launching stream. enable(1) clears previous records.
int32_t fc_program_kernel_timing(fc_program* p, int32_t enable) {
  for (auto& e : p->timing_events) { cudaEventDestroy(e.first); cudaEventDestroy(e.second); }
  p->timing_events.clear();
  p->timing = enable != 0;
  return FC_OK;
}
int32_t fc_program_kernel_time(fc_program* p, double* total_ms, uint64_t* launches) {
  FC_CUDA(cudaStreamSynchronize(p->stream));
  double t = 0.0;
  for (auto& e : p->timing_events) {
    float ms = 0.f;
    FC_CUDA(cudaEventElapsedTime(&ms, e.first, e.second));
    t += ms;
  }
  if (total_ms) *total_ms = t;
  if (launches) *launches = p->timing_events.size();
  return FC_OK;
}

int32_t fc_program_launch_count(const fc_program* p, uint64_t* n) {
  *n = p->n_kernel_launches;
  return FC_OK;
}
}
