/* fugue_cuda.h -- C ABI of the B200-native inference backend for fugue.
 *
 * This is what a Rust `fugue-cuda` crate binds with `extern "C"`.
 * Every entry point names the reference interface it replaces (paths relative to
 * /root/reference). Plain C: pointers + sizes, int32 status codes, no exceptions.
 *
 * Layout conventions
 *   q  : continuous site values, [d_cont][C] doubles, chain-minor (q[s*C + c]); site order is
 *        the SORTED address order (BTreeMap<Address,_> iteration, src/inference/hmc.rs:238-248,
 *        src/core/address.rs:150-157) as returned by fc_program_sites().
 *   qd : discrete site values, [d_disc][C] int64, same conventions.
 *   Pointers named h_* are HOST memory, d_* are DEVICE memory. Functions without a suffix
 *   take host buffers and copy (the reference-facing path); *_dev variants take device pointers.
 */
#ifndef FUGUE_CUDA_H
#define FUGUE_CUDA_H
#include <stddef.h>
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

/* ---- status codes: numerically equal to ErrorCode (src/error.rs:40-58) ---- */
enum {
  FC_OK = 0,
  FC_ERR_INVALID_MEAN = 100, FC_ERR_INVALID_VARIANCE = 101, FC_ERR_INVALID_PROBABILITY = 102,
  FC_ERR_INVALID_RANGE = 103, FC_ERR_INVALID_SHAPE = 104, FC_ERR_INVALID_RATE = 105,
  FC_ERR_INVALID_COUNT = 106,
  FC_ERR_ADDRESS_CONFLICT = 301,            /* src/runtime/interpreters.rs:23-45 */
  FC_ERR_UNEXPECTED_MODEL_STRUCTURE = 302,  /* src/runtime/interpreters.rs:667-679 */
  FC_ERR_TRACE_ADDRESS_NOT_FOUND = 500,
  FC_ERR_TYPE_MISMATCH = 600,
  FC_ERR_UNSUPPORTED_OP = 1000, FC_ERR_CUDA = 1001, FC_ERR_NCCL = 1002, FC_ERR_INVALID_ARGUMENT = 1003
};

/* ---- distribution ids: order of src/core/distribution.rs:125-1936 ---- */
enum {
  FC_NORMAL = 0, FC_UNIFORM = 1, FC_LOGNORMAL = 2, FC_EXPONENTIAL = 3, FC_BERNOULLI = 4,
  FC_CATEGORICAL = 5, FC_BETA = 6, FC_GAMMA = 7, FC_BINOMIAL = 8, FC_POISSON = 9,
  FC_STUDENT_T = 10, FC_CAUCHY = 11, FC_LAPLACE = 12, FC_WEIBULL = 13, FC_CHI_SQUARED = 14,
  FC_INVERSE_GAMMA = 15, FC_DISCRETE_UNIFORM = 16, FC_N_DISTS = 17
};

/* ---- op kinds (Model variants: src/core/model.rs:20-131; plate!: src/macros/mod.rs:82-86) ---- */
enum { FC_OP_SAMPLE = 0, FC_OP_OBSERVE = 1, FC_OP_FACTOR = 2, FC_OP_PLATE_SAMPLE = 3, FC_OP_PLATE_OBSERVE = 4,
       FC_OP_FACTOR_EXPR = 5 /* factor(f(latents)): log-weight computed from the latents (Model::Factor
                                fed by a bind closure, src/core/model.rs:97-104); see fc_expr_node */ };

/* ---- value types (ChoiceValue: src/runtime/trace.rs:32-47) ---- */
enum { FC_T_F64 = 0, FC_T_I64 = 1, FC_T_U64 = 2, FC_T_USIZE = 3, FC_T_BOOL = 4 };

/* ---- parameter expressions: value = link(term0 + term1 + term2), evaluated left to right ---- */
enum {
  FC_TERM_NONE = 0,
  FC_TERM_CONST = 1,      /* c                                                            */
  FC_TERM_LATENT = 2,     /* q[slot]                                                      */
  FC_TERM_LATENT_X = 3,   /* q[slot] * col[i]               (col: f64 vector column)      */
  FC_TERM_DENSE = 4,      /* sum_k q[slot+k]*X[i][k], k=0..K-1 left-to-right from 0.0     */
                          /*   (col: f64 row-major matrix column, K = its width)          */
  FC_TERM_GATHER = 5,     /* q[slot + idx[i]]               (col: index column)           */
  FC_TERM_SELECT2 = 6,    /* z==0 ? q[slot] : q[slot_b]     z = plate-discrete site i of  */
                          /*   PLATE_SAMPLE op `col` (an op index, not a column)          */
  FC_TERM_PLATE_LATENT = 7,/* q[slot + i]  (element i of a continuous plate)              */
  FC_TERM_SCALED = 8      /* c * q[slot]   (`let mu = 2.0 * p - 1.0`, dsl.rs:12-13)        */
};
enum { FC_LINK_IDENTITY = 0, FC_LINK_EXP = 1, FC_LINK_SIGMOID_CLAMPED = 2, FC_LINK_CLAMP = 3 };

typedef struct fc_term {
  int32_t kind;
  int32_t slot;    /* PROGRAM-order continuous slot index (see fc_program_create) */
  int32_t slot_b;
  int32_t col;
  double  c;
} fc_term;

typedef struct fc_param {
  fc_term term[3];
  int32_t link;
  int32_t _pad;
  double  lo, hi;  /* SIGMOID_CLAMPED / CLAMP bounds (examples/classification.rs:127,
                      examples/mixture_models.rs:93) */
} fc_param;

typedef struct fc_op {
  int32_t kind;        /* FC_OP_*                                                     */
  int32_t dist;        /* FC_NORMAL.. ; ignored for FACTOR                            */
  int32_t value_type;  /* FC_T_*                                                      */
  int32_t n_params;    /* 1..3. Categorical: 1 = the two-way form [p, 1-p] (p may depend on
                          latents); 0 = K-way with CONSTANT probabilities taken from the
                          f64 vector column `probs_col` (Categorical::new(vec), validated as
                          src/core/distribution.rs:679-713)                            */
  const char* name;    /* address base; scalar ops use addr!(name), plates addr!(name,i)
                          => "name#i" (src/core/address.rs:208-223)                   */
  int64_t plate_len;   /* PLATE_*: number of elements; else 0                         */
  int32_t value_col;   /* OBSERVE/PLATE_OBSERVE: data column holding the value(s)     */
  int32_t probs_col;   /* K-way Categorical (n_params == 0): column of the K probabilities */
  double  value;       /* OBSERVE (scalar): observed value; FACTOR: log-weight;
                          FACTOR_EXPR: 1.0 = a NaN result counts as -inf (the playground's
                          policy, crates/fugue-wasm/src/dsl.rs:922-926), 0.0 = added as is */
  fc_param param[3];
  int32_t expr_first;  /* FACTOR_EXPR: first node / number of nodes in the program's        */
  int32_t expr_len;    /*   expression table (fc_program_create_ex); the last node is the value */
} fc_op;

/* ---- expression nodes of FACTOR_EXPR ops, in evaluation (SSA) order: node k may only refer to
 *      nodes a, b < k of the same expression (indices relative to expr_first). Arithmetic and
 *      functions are those of the playground's evaluator (crates/fugue-wasm/src/dsl.rs:755-786):
 *      f64 + - * /, exp ln sqrt abs floor sin cos tanh, powf, f64::min / f64::max. ---- */
enum {
  FC_X_CONST = 0,   /* c                                   */
  FC_X_LATENT = 1,  /* q[slot] (program-order continuous slot) */
  FC_X_ADD = 2, FC_X_SUB = 3, FC_X_MUL = 4, FC_X_DIV = 5, FC_X_NEG = 6,
  FC_X_EXP = 7, FC_X_LN = 8, FC_X_SQRT = 9, FC_X_ABS = 10, FC_X_FLOOR = 11, FC_X_SIN = 12,
  FC_X_COS = 13, FC_X_TANH = 14, FC_X_POW = 15, FC_X_MIN = 16, FC_X_MAX = 17
};
#define FC_EXPR_MAX_NODES 64
typedef struct fc_expr_node {
  int32_t op;      /* FC_X_*                        */
  int32_t a, b;    /* operand nodes (unary: a only) */
  int32_t slot;    /* FC_X_LATENT                   */
  double  c;       /* FC_X_CONST                    */
} fc_expr_node;

/* ---- data columns ---- */
enum { FC_COL_F64 = 0, FC_COL_U8 = 1, FC_COL_U32 = 2, FC_COL_I64 = 3 };
typedef struct fc_column {
  int32_t dtype;      /* FC_COL_*                               */
  int32_t width;      /* 1 = vector; K = row-major [rows][K]    */
  int64_t rows;
  const void* data;   /* host pointer; copied at create         */
} fc_column;

typedef struct fc_program fc_program;

/* ---- library ---- */
const char* fc_last_error(void);
int32_t fc_device_count(int32_t* out);
int32_t fc_set_device(int32_t device);

/* ---- program (replaces: run(handler, model_fn()) lowering; src/runtime/handler.rs:124-209) ----
 * Ops are in PROGRAM order. Continuous program slots are numbered in program order, a
 * PLATE_SAMPLE of f64 occupying plate_len consecutive slots. Duplicate addresses ->
 * FC_ERR_ADDRESS_CONFLICT. Columns are copied to the current device. */
int32_t fc_program_create(const fc_op* ops, size_t n_ops, const fc_column* cols, size_t n_cols,
                          fc_program** out);
/* The same with the expression table FACTOR_EXPR ops index into (nodes may be NULL if n_nodes == 0). */
int32_t fc_program_create_ex(const fc_op* ops, size_t n_ops, const fc_column* cols, size_t n_cols,
                             const fc_expr_node* nodes, size_t n_nodes, fc_program** out);
int32_t fc_program_destroy(fc_program* p);
/* Checks a second trace of the same model against the program (address sequence, kinds, types);
 * mismatch -> FC_ERR_UNEXPECTED_MODEL_STRUCTURE (src/runtime/interpreters.rs:271-283). */
int32_t fc_program_check_structure(const fc_program* p, const fc_op* ops, size_t n_ops);
int32_t fc_program_dims(const fc_program* p, int64_t* d_cont, int64_t* d_disc);
/* Sorted site names (BTreeMap order) of the continuous sites; pointers owned by the program.
 * perm[i] = program-order slot of sorted site i. */
int32_t fc_program_sites(const fc_program* p, const char* const** names, const int64_t** perm,
                         int64_t* n);
/* cudaStream_t the program launches on (for CUDA-event timing by the caller). */
int32_t fc_program_stream(const fc_program* p, void** stream);
/* CUDA device ordinal the program's columns and stream live on (where device-resident buffers handed
 * to the *_dev entry points must be allocated). */
int32_t fc_program_device(const fc_program* p, int32_t* device);
/* Number of kernels this library has launched on behalf of the program (bench.py gpu_launches). */
int32_t fc_program_launch_count(const fc_program* p, uint64_t* n);
/* CUDA-event timing (on the launching stream) of the dominant observation-streaming kernel:
 * enable, run, then read the summed duration and the number of timed launches. */
int32_t fc_program_kernel_timing(fc_program* p, int32_t enable);
int32_t fc_program_kernel_time(fc_program* p, double* total_ms, uint64_t* launches);

/* ---- scoring: replaces run(ScoreGivenTrace{base}, model) (src/runtime/interpreters.rs:138-163,
 * src/inference/hmc.rs:264-299). Three accumulators as in Trace (src/runtime/trace.rs:168-177). */
int32_t fc_logp(fc_program* p, const double* h_q, const int64_t* h_qd, size_t C,
                double* h_log_prior, double* h_log_lik, double* h_log_factors);
/* value + analytic gradient (replaces grad_log_joint, src/inference/hmc.rs:304-329);
 * h_grad is [d_cont][C]. */
int32_t fc_logp_grad(fc_program* p, const double* h_q, const int64_t* h_qd, size_t C,
                     double* h_lp, double* h_grad);
/* Device-resident variant: d_qd is the int32 plane [d_disc][C] discrete values live in on the
 * device; d_grad may be NULL (value only). Asynchronous on fc_program_stream(). */
int32_t fc_logp_grad_dev(fc_program* p, const double* d_q, const int32_t* d_qd, size_t C,
                         double* d_lp, double* d_grad);

/* ---- HMC (src/inference/hmc.rs) ---- */
typedef struct fc_hmc_config {   /* HMCConfig, hmc.rs:105-135 */
  uint64_t n_leapfrog;           /* 16    */
  double   target_accept;        /* 0.8   */
  double   init_step_size;       /* NaN = None */
  double   finite_diff_eps;      /* 1e-5 (used only by grad_mode = 1) */
  int32_t  adapt_mass;           /* false */
  int32_t  grad_mode;            /* 0 = analytic (device), 1 = central finite differences */
} fc_hmc_config;
typedef struct fc_hmc_stats {
  uint64_t n_accept, n_divergent;
  double   step_size;            /* current (warmup) or frozen step size */
} fc_hmc_stats;
typedef struct fc_hmc_session fc_hmc_session;

void    fc_hmc_config_default(fc_hmc_config* cfg);
/* HmcSession::new (hmc.rs:667): prior init (or h_q0 / h_qd0 if non-NULL), find_reasonable_epsilon
 * unless cfg->init_step_size is set. C independent chains, device resident. */
int32_t fc_hmc_create(fc_program* p, const fc_hmc_config* cfg, uint64_t seed, size_t C,
                      size_t n_warmup, const double* h_q0, const int64_t* h_qd0,
                      fc_hmc_session** out);
/* Same, for one shard of a multi-GPU job: local chain c uses the random stream of global chain
 * first_chain + c, so the union over ranks equals the single-GPU run. */
int32_t fc_hmc_create_sharded(fc_program* p, const fc_hmc_config* cfg, uint64_t seed, size_t C,
                              size_t n_warmup, const double* h_q0, const int64_t* h_qd0,
                              uint64_t first_chain, fc_hmc_session** out);
int32_t fc_hmc_destroy(fc_hmc_session* s);
/* HmcSession::step / advance (hmc.rs:819-919): n transitions for all C chains. Asynchronous on
 * the program's stream; fc_hmc_sync() or any getter waits. */
int32_t fc_hmc_step(fc_hmc_session* s, size_t n);
int32_t fc_hmc_sync(fc_hmc_session* s);
/* HmcSession::step_recorded (hmc.rs:811-817): one transition, recording LeapfrogPoint{q, h}
 * (hmc.rs:331-343) at the start and after every leapfrog step: h_q [L+1][d][C], h_h [L+1][C]
 * (H = -log pi(q) + 1/2 p^T M^-1 p; NaN once a chain has left the support). Synchronous. */
int32_t fc_hmc_step_recorded(fc_hmc_session* s, double* h_q, double* h_h, uint64_t* n_points);
/* HmcSession::position / step_size (hmc.rs:752-777): copy state to host. Any pointer may be NULL. */
int32_t fc_hmc_get(fc_hmc_session* s, double* h_q, double* h_lp, fc_hmc_stats* h_stats);
/* HmcSession::trace accumulators of the current state (trace.rs:168-177). */
int32_t fc_hmc_trace(fc_hmc_session* s, double* h_log_prior, double* h_log_lik, double* h_log_factors);
/* device pointers of the resident state: q [d][C], log-joint [C], accumulators [3][C] */
int32_t fc_hmc_device_state(fc_hmc_session* s, double** d_q, double** d_lj, double** d_lp3);
int32_t fc_hmc_set_position(fc_hmc_session* s, const double* h_q);
int32_t fc_hmc_set_step_size(fc_hmc_session* s, double eps);      /* hmc.rs:738-744 */
int32_t fc_hmc_set_n_leapfrog(fc_hmc_session* s, uint64_t l);     /* hmc.rs:748-750 */
int32_t fc_hmc_grad_evals(const fc_hmc_session* s, uint64_t* per_chain);
int32_t fc_hmc_iterations(const fc_hmc_session* s, uint64_t* iters, int32_t* warming); /* hmc.rs:780-787 */
/* HmcStepInfo of the last transition (hmc.rs:585-600): accept_prob, accepted, divergent per chain */
int32_t fc_hmc_last_accept_prob(fc_hmc_session* s, double* h_ap, int32_t* h_accepted, int32_t* h_divergent);
/* Parity hooks: supply the StandardNormal draws z [d][C] and the accept uniform u [C] of the next
 * transition / step-size search (hmc.rs:436-441,461,489-495). */
int32_t fc_hmc_set_explicit_randomness(fc_hmc_session* s, const double* h_z, const double* h_u);
int32_t fc_hmc_find_reasonable_epsilon(fc_hmc_session* s, double* h_eps);  /* hmc.rs:479-535 */
/* hmc_chain (hmc.rs:566-583): warmup then n_samples draws; h_draws [n_samples][d][C], h_lp
 * [n_samples][C] (either may be NULL). */
int32_t fc_hmc_run(fc_program* p, const fc_hmc_config* cfg, uint64_t seed, size_t C,
                   size_t n_warmup, size_t n_samples, double* h_draws, double* h_lp,
                   fc_hmc_stats* h_stats);
/* The same run with the draws left in HBM (1024 chains x 500 draws x 10 004
 * parameters = 41 GB never crosses PCIe): d_draws [n_samples][d][C] and d_lp [n_samples][C] are
 * DEVICE pointers (either may be NULL). fc_hmc_sample_dev is the session-level form (n further
 * transitions, every state kept). Returns after the stream has drained. */
int32_t fc_hmc_run_dev(fc_program* p, const fc_hmc_config* cfg, uint64_t seed, size_t C,
                       size_t n_warmup, size_t n_samples, double* d_draws, double* d_lp,
                       fc_hmc_stats* h_stats);
int32_t fc_hmc_sample_dev(fc_hmc_session* s, size_t n, double* d_draws, double* d_lp);

/* ---- single-site MH (src/inference/mh.rs:698-744, 921-1014) ---- */
typedef struct fc_mh_config {
  double target_accept;   /* 0.44 (mh.rs:946) */
  double gamma;           /* 0.7               */
} fc_mh_config;
typedef struct fc_mh_stats { uint64_t n_accept; uint64_t n_steps; } fc_mh_stats;
/* SiteProposal override for an f64 site (mh.rs:145-161), matched by address string */
enum { FC_PROPOSAL_GAUSSIAN = 0, FC_PROPOSAL_LOGSPACE = 1, FC_PROPOSAL_REFLECT = 2, FC_PROPOSAL_PRIOR_RESAMPLE = 3 };
typedef struct fc_site_proposal {
  const char* address;
  int32_t kind;           /* FC_PROPOSAL_* */
  int32_t _pad;
  double  lower, upper;   /* REFLECT bounds */
} fc_site_proposal;
typedef struct fc_mh_session fc_mh_session;
void    fc_mh_config_default(fc_mh_config* cfg);
/* adaptive_mcmc_chain_with_overrides set-up (mh.rs:938-965): prior init, one scoring pass.
 * Chain c uses the random stream of global chain first_chain + c. */
int32_t fc_mh_create(fc_program* p, const fc_mh_config* cfg, uint64_t seed, size_t C, size_t n_warmup,
                     const fc_site_proposal* overrides, size_t n_overrides, uint64_t first_chain,
                     fc_mh_session** out);
int32_t fc_mh_destroy(fc_mh_session* s);
/* n single_site_mh_step()s for every chain (mh.rs:698-744); the first n_warmup steps adapt. */
int32_t fc_mh_step(fc_mh_session* s, size_t n);
/* h_q [d_cont][C], h_qd [d_disc][C], h_lp [C], h_scales [n_sites][C] (DiminishingAdaptation
 * scales, global address order); any pointer may be NULL. */
int32_t fc_mh_get(fc_mh_session* s, double* h_q, int64_t* h_qd, double* h_lp, fc_mh_stats* h_stats,
                  double* h_scales);
int32_t fc_mh_n_sites(const fc_mh_session* s, int64_t* n);
/* adaptive_mcmc_chain (mh.rs:921-1014). Every thin-th sampling step is kept:
 * h_draws [n_kept][d_cont][C], h_ddraws [n_kept][d_disc][C], h_lp [n_kept][C]. */
int32_t fc_mh_run(fc_program* p, const fc_mh_config* cfg, uint64_t seed, size_t C,
                  size_t n_warmup, size_t n_samples, size_t thin, double* h_draws, int64_t* h_ddraws,
                  double* h_lp, fc_mh_stats* h_stats);

/* ---- SMC (src/inference/smc.rs) ---- */
enum { FC_RESAMPLE_MULTINOMIAL = 0, FC_RESAMPLE_SYSTEMATIC = 1, FC_RESAMPLE_STRATIFIED = 2 };
typedef struct fc_smc_config {  /* SMCConfig, smc.rs:172-189 */
  int32_t  resampling_method;   /* Systematic */
  int32_t  _pad;
  double   ess_threshold;       /* 0.5 */
  uint64_t rejuvenation_steps;  /* 0   */
} fc_smc_config;
typedef struct fc_smc_session fc_smc_session;
void    fc_smc_config_default(fc_smc_config* cfg);
/* adaptive_smc (smc.rs:455-581). h_values [d_cont][P], h_dvalues [d_disc][P]; outputs may be NULL.
 * n_temper_steps returns the number of tempering steps taken. */
int32_t fc_smc_run(fc_program* p, const fc_smc_config* cfg, uint64_t seed, size_t P,
                   double* h_values, int64_t* h_dvalues, double* h_weight, double* h_log_weight,
                   double* log_evidence, uint64_t* n_temper_steps);
/* The same, one tempering step at a time, particles resident on the device:
 * create = smc_prior_particles (smc.rs:764-790); step = one pass of the while loop (smc.rs:499-560:
 * next_beta, reweight, evidence, resample, rejuvenate) and returns the new beta (1.0 = finished);
 * get = final normalisation (smc.rs:565-575) + copy-out. Particle j of this call uses the random
 * stream of global particle first_particle + j. */
int32_t fc_smc_create(fc_program* p, const fc_smc_config* cfg, uint64_t seed, size_t P,
                      uint64_t first_particle, fc_smc_session** out);
int32_t fc_smc_step(fc_smc_session* s, double* beta_out);
int32_t fc_smc_get(fc_smc_session* s, double* h_values, int64_t* h_dvalues, double* h_weight,
                   double* h_log_weight, double* log_evidence, uint64_t* n_temper_steps);
int32_t fc_smc_device_state(fc_smc_session* s, double** d_values, int32_t** d_dvalues,
                            double** d_log_w, double** d_loglik);
int32_t fc_smc_destroy(fc_smc_session* s);
/* Standalone pieces of the SMC surface on a session's resident particles:
 *   fc_smc_prior_weights       smc_prior_particles' weighting, log_weight = log-likelihood (smc.rs:779-787)
 *   fc_smc_normalize           normalize_particles (smc.rs:719-754) + effective_sample_size (smc.rs:230-233)
 *   fc_smc_resample_particles  resample_particles (smc.rs:326-349); h_uniforms NULL = session stream
 *   fc_smc_rejuvenate_particles rejuvenate_particles (smc.rs:698-713), weights unchanged */
int32_t fc_smc_prior_weights(fc_smc_session* s);
int32_t fc_smc_normalize(fc_smc_session* s, double* log_norm, double* ess);
int32_t fc_smc_resample_particles(fc_smc_session* s, int32_t method, const double* h_uniforms,
                                  uint64_t* h_idx);
int32_t fc_smc_rejuvenate_particles(fc_smc_session* s, double beta, uint64_t steps);
/* one reweight of the tempering loop (smc.rs:507-531): log_w += (beta_new - beta) * loglik, normalised by
 * the log-sum-exp over the WHOLE population (all ranks), which is added to the evidence estimate and
 * returned in *log_norm (may be NULL); beta := beta_new */
int32_t fc_smc_reweight(fc_smc_session* s, double beta_new, double* log_norm);
/* this rank's log weights as they are, and the globally normalised weights exp(log_w - lse) the resampler
 * sees; either pointer may be NULL */
int32_t fc_smc_weights(fc_smc_session* s, double* h_log_w, double* h_w);
/* particles this rank sent to other ranks in global resamples; ESS evaluations spent by all next_beta
 * searches; whether the program runs on the fused per-particle kernels. Any pointer may be NULL. */
int32_t fc_smc_counters(const fc_smc_session* s, uint64_t* migrated, uint64_t* ess_evals, int32_t* fused);

/* CUDA-event time (ms, on the launching stream) per phase of the tempering steps so far, recorded while
 * fc_debug_set("smc_timing", 1) is on: [0] next_beta search, [1] reweight, [2] exact scan + carry chain,
 * [3] gather + migration (+ barrier), [4] rejuvenation moves; n5[k] = how many times phase k ran */
int32_t fc_smc_phase_ms(fc_smc_session* s, double* ms5, uint64_t* n5);

/* ---- sharded SMC: the population of n_total particles is split rank-major over `world`
 * processes, one per GPU of a box. The exchange steps of the tempering loop -- the global weight
 * log-sum-exp / ESS of every next_beta probe and of the reweight (smc.rs:512-524, 588-602), the exact
 * rank-to-rank carry of the sequentially rounded running sum and the particle migration of a global
 * systematic / stratified resample (smc.rs:255-292, 534-541) -- run ON THE DEVICE over peer memory:
 * every rank exports one cudaIpc slab (mailbox + population buffers), the last CTA of a weight pass
 * stores the rank's sums into every peer's mailbox and folds all ranks' sums in rank order, and the
 * resampling gather writes offspring straight into the destination rank's buffers over NVLink.
 *   create_sharded  rank `rank` holds particles [first_particle, first_particle + n) of n_total
 *   export          this rank's 128-byte slab descriptor; all-gather them in rank order (any host channel)
 *   connect         map the peers' slabs; then fc_smc_step / fc_smc_get / the standalone pieces above work
 *                   on the global population (every rank must make the same sequence of calls)
 * Multinomial resampling is not available sharded (FC_ERR_UNSUPPORTED_OP): its ancestors are not monotone
 * in the slot index. A peer that does not answer within 20 s gives FC_ERR_NCCL. */
#define FC_SMC_HANDLE_BYTES 128
int32_t fc_smc_create_sharded(fc_program* p, const fc_smc_config* cfg, uint64_t seed, size_t n, uint64_t first_particle,
                              uint64_t n_total, int32_t rank, int32_t world, fc_smc_session** out);
int32_t fc_smc_export(fc_smc_session* s, void* handle);
int32_t fc_smc_connect(fc_smc_session* s, const void* handles);
/* host arithmetic of the slot routing (no device work): the output slots [jlo, jhi) whose ancestor lives on
 * a rank whose exact running sum runs from carry_in to end (smc.rs:263-268: the walk stops there iff
 * carry_in < u/N + j/N <= end; the last rank also takes the slots whose threshold is never reached) */
int32_t fc_smc_slot_range(int32_t method, uint64_t n_total, double u_raw, uint64_t seed, uint64_t tag, double carry_in,
                          double end, int32_t is_last, uint64_t* jlo, uint64_t* jhi);
/* parity hooks: record the GLOBAL ancestor index of every output slot this rank produces in a resample
 * (the slots whose ancestor it holds), and read the last resample's back: slots [jlo, jhi), h_idx of
 * jhi - jlo entries (may be NULL). Concatenated over ranks in rank order they are the whole index array
 * the reference's systematic_indices / stratified_indices return. */
int32_t fc_smc_record_offspring(fc_smc_session* s, int32_t enable);
/* resample_particles with a single-rank session's population processed as `shards` (1..8) VIRTUAL ranks one after
 * the other: the sharded resample's exact scan continuing from the previous shard's end value, its host slot
 * routing and its offset gathers, on one GPU and without any kernel that waits on another. h_idx: the n
 * ancestors (may be NULL). method: systematic (u_raw) or stratified (the session's uniform stream). */
int32_t fc_smc_resample_virtual_shards(fc_smc_session* s, int32_t shards, int32_t method, double u_raw, uint64_t* h_idx);
int32_t fc_smc_last_offspring(fc_smc_session* s, uint64_t* jlo, uint64_t* jhi, uint64_t* h_idx);
/* next_beta (smc.rs:588-622) on host arrays: the largest beta' in (beta, 1] whose incremental
 * weights log_w + (beta' - beta) * ll keep ESS >= target_ess (64-step bisection; 1.0 if a full
 * jump already does). Parity probe for the tempering ladder. */
int32_t fc_next_beta(const double* h_log_w, const double* h_ll, size_t n, double beta,
                     double target_ess, double* out);
/* systematic/stratified/multinomial_indices (smc.rs:255-314) with the uniforms passed explicitly:
 * 1 uniform for systematic (the raw rng.gen::<f64>(), before /n), n for the others.
 * Indices are bit-exact w.r.t. the reference's sequential running sum. */
int32_t fc_resample(int32_t method, const double* h_w, size_t n, const double* h_uniforms,
                    uint64_t* h_idx);
int32_t fc_resample_dev(int32_t method, const double* d_w, size_t n, const double* d_uniforms,
                        uint64_t* d_idx, void* stream);
/* S[0..n]: the sequentially rounded running sums cum_weight takes in smc.rs:262-268 */
int32_t fc_exact_prefix_sums(const double* h_w, size_t n, double* h_S);
int32_t fc_log_sum_exp(const double* h_x, size_t n, double* out);        /* numerical.rs:15-38 */
int32_t fc_ess(const double* h_w, size_t n, double* out);                /* smc.rs:230-233     */
int32_t fc_normalize_log_probs(const double* h_x, size_t n, double* h_out); /* numerical.rs:87-90 */

/* ---- diagnostics (src/inference/diagnostics.rs:218-304, src/inference/mcmc_utils.rs:195-339).
 * h_draws is [S][C] for one parameter (draw-major, chain-minor). ---- */
int32_t fc_split_rhat(const double* h_draws, size_t S, size_t C, double* out);
int32_t fc_classic_rhat(const double* h_draws, size_t S, size_t C, double* out);
int32_t fc_ess_multichain(const double* h_draws, size_t S, size_t C, double* out);
/* all parameters at once: h_draws [S][d][C] (the fc_hmc_run layout); out arrays of length d. */
int32_t fc_diagnostics(const double* h_draws, size_t S, size_t d, size_t C, double* h_mean,
                       double* h_std, double* h_rhat, double* h_ess);
/* ... on draws resident in HBM (device pointer, same layout); diagnostics.rs:240-304 and
 * mcmc_utils.rs:231-339 run in place, 4*d doubles come back. */
int32_t fc_diagnostics_dev(const double* d_draws, size_t S, size_t d, size_t C, double* h_mean,
                           double* h_std, double* h_rhat, double* h_ess);
/* Allocation-free split-R-hat of draws resident in HBM. NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA graph
 * capture, tight serving loops, harnesses that forbid it). The same kernels as fc_diagnostics_dev's
 * R-hat leg, with their scratch (per parameter x chain x half: mean and sum of squares) in a
 * file-scope static __device__ array of 2 * FC_RHAT_NOALLOC_MAX_SEGMENTS doubles (16 MiB, held by
 * every GPU that loads the library until exit). d_draws [S][d][C] and d_rhat [d] are DEVICE pointers
 * on the current device; the work is enqueued on `stream` (a cudaStream_t; NULL = the default
 * stream) and the call returns without synchronising. It calls no cudaMalloc / cudaFree or
 * allocating library routine. d * 2 * C > FC_RHAT_NOALLOC_MAX_SEGMENTS -> FC_ERR_INVALID_ARGUMENT
 * (use fc_diagnostics_dev). Every call shares the one array: calls that overlap on the GPU (two
 * streams, two threads) race; calls on one stream are safe back to back. The first call on each
 * device resolves the array's address (cudaGetSymbolAddress): make it before starting a capture. */
#define FC_RHAT_NOALLOC_MAX_SEGMENTS (1u << 20)
int32_t fc_split_rhat_dev_noalloc(const double* d_draws, size_t S, size_t d, size_t C,
                                  double* d_rhat, void* stream);

/* Per-rank partial statistics for multi-GPU R-hat / ESS (chains sharded over ranks): split-half
 * means / sums of squares per chain [d][2C], chain means [d][C], and chain-summed autocovariances
 * per lag [d][min(S-1,2048)+1]. Combined across ranks by all-reduce (fugue_b200/dist.py). */
int32_t fc_diag_partials(const double* h_draws, size_t S, size_t d, size_t C, double* h_seg_mean,
                         double* h_seg_ss, double* h_cmean, double* h_lagsum);
/* ... on draws resident in HBM (device pointer, the fc_hmc_run_dev layout): read in place. */
int32_t fc_diag_partials_dev(const double* d_draws, size_t S, size_t d, size_t C, double* h_seg_mean,
                             double* h_seg_ss, double* h_cmean, double* h_lagsum);

/* switches for A/B parity tests: "nb_plain" = 1: next_beta evaluates every midpoint of the reference's
 * 64-step bisection (no search phase); "smc_generic" = 1: SMC sessions created afterwards run on the
 * generic scoring engine instead of the fused per-particle kernels; "smc_timing" = 1: per-phase
 * CUDA-event timing of the tempering steps (fc_smc_phase_ms); "smc_stats" = 0: sessions created afterwards
 * score mixture plates term by term instead of from the per-particle sufficient statistics; "smc_zbits" = 0:
 * keep one byte (not one bit) per two-valued discrete site in the particle record; "smc_compiled" = 0: the move
 * kernel interprets the op descriptors instead of running the by-value flattened program; "hmc_hier" = 0: HMC
 * sessions created afterwards run hierarchical programs through the generic leapfrog kernels (fc_hier.cu off) */
int32_t fc_debug_set(const char* key, int32_t value);
/* accuracy probe of the kernels' branch-free FP64 primitives: op 0 exp, 1 log, 2 reciprocal */
int32_t fc_debug_fastmath(int32_t op, const double* h_x, size_t n, double* h_out);

#ifdef __cplusplus
}
#endif
#endif
