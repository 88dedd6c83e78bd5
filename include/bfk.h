/* bfk.h — C ABI of libbfk (B200 / sm_100a kernels for the BayesFlow generative-network hot path).
 *
 * The reference (bayesflow v2.0.11) has NO FFI: its extension points are Python objects.  The entry points below are
 * what a `keras.Layer` shim on the torch backend binds (ctypes stub in INTEGRATION.md) to replace, on CUDA tensors:
 *
 *   bfk_velocity        <- FlowMatching.velocity / DiffusionModel.velocity(stochastic_solver=False)
 *                          bayesflow/networks/inference/flow_matching/flow_matching.py:220-236
 *                          bayesflow/networks/inference/diffusion/diffusion_model.py:501-559
 *                          (= TimeMLP.call, bayesflow/networks/subnets/mlp/time_mlp.py:125-147, + output_projector)
 *   bfk_sample_fixed    <- integrate_fixed / integrate_scheduled driven by FlowMatching._inverse/_forward and
 *                          DiffusionModel._inverse: bayesflow/utils/integrate.py:217-288,
 *                          flow_matching.py:254-337, diffusion_model.py:782-869
 *   bfk_adaptive_*      <- integrate_adaptive, bayesflow/utils/integrate.py:291-406
 *   bfk_cond_prologue   <- the conditions half of TimeMLP.call (time_mlp.py:133-139), hoisted per dataset
 *   bfk_eval_table      <- FourierEmbedding.call + FiLM.call (fourier_embedding.py:57-76,
 *                          featurewise_linear_modulation.py:25-33) and the noise-schedule scalars
 *                          (diffusion_model.py:351-405, 466-489, 539-552, 711-717), per evaluation time
 *   bfk_pack            <- weight import: Keras variable order / (in,out) kernels (time_mlp.py:149-183)
 *
 * Conventions: every pointer is a DEVICE pointer unless named host_*; tensors are dense row-major fp32; `stream`
 * is a cudaStream_t passed as void*.  Return value: 0 ok, <0 argument error, >0 cudaError_t.  Nothing throws.
 * Re-entrant; no global state besides the last-error string (thread-local).
 */
#ifndef BFK_H_
#define BFK_H_
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define BFK_MAX_BLOCKS 16
#define BFK_MAX_D 8      /* fused on-chip sampler: state dim <= 8 */
#define BFK_MAX_STAGES 7 /* RK stages incl. the FSAL evaluation */

enum { BFK_ACT_MISH = 0, BFK_ACT_SILU = 1, BFK_ACT_RELU = 2, BFK_ACT_GELU = 3, BFK_ACT_TANH = 4 };
enum { BFK_MERGE_NONE = 0, BFK_MERGE_CONCAT = 1, BFK_MERGE_ADD = 2 };
enum { BFK_NORM_NONE = 0, BFK_NORM_LAYER = 1, BFK_NORM_RMS = 2 };
enum { BFK_PREC_FP16 = 0, BFK_PREC_BF16 = 1 };
enum { BFK_METHOD_EULER = 0, BFK_METHOD_RK45 = 1, BFK_METHOD_TSIT5 = 2 };
/* what the network output means (DiffusionModel prediction_type, diffusion_model.py:351-405); FM = velocity of x */
enum { BFK_MODEL_FM = 0, BFK_MODEL_DM = 1 };
enum { BFK_PRED_VELOCITY = 0, BFK_PRED_NOISE = 1, BFK_PRED_F = 2, BFK_PRED_X = 3, BFK_PRED_SCORE = 4 };
enum { BFK_SCHED_EDM = 0, BFK_SCHED_COSINE = 1 };

typedef struct bfk_net_config {
  int32_t D;              /* state / parameter dim */
  int32_t C;              /* condition dim, 0 = unconditional */
  int32_t W;              /* hidden width (all blocks equal) */
  int32_t L;              /* number of ConditionalDenseBlocks */
  int32_t T;              /* time embedding size: time_embedding_dim+1 (33), or 1 for identity */
  int32_t activation;     /* BFK_ACT_* */
  int32_t merge;          /* BFK_MERGE_* */
  int32_t norm;           /* BFK_NORM_* */
  int32_t residual;       /* 0/1 */
  int32_t film_use_gamma; /* 0/1 */
  int32_t precision;      /* BFK_PREC_*: tensor-core operand type, fp32 accumulate */
  float fourier_scale;
  /* Unequal hidden widths (time_mlp.py:100-120 with widths like (64, 96, 32)): W is then the common PADDED width (a
   * multiple of 64, every weight zero-padded to it) and block_width[l] the true output width of block l, over which its
   * normalisation runs; all zero = every block has width W.  Blocks whose input and output widths differ carry a
   * residual projector (bfk_weights.proj_k / proj_b).  Served by the layer-wise training / density kernels only. */
  int32_t block_width[BFK_MAX_BLOCKS];
} bfk_net_config;

/* fp32 device pointers in the reference's variable order; Dense kernels are (in,out) row-major */
typedef struct bfk_weights {
  const float* freq;                 /* (T-1)/2 */
  const float* x_k; const float* x_b; /* (D,W), (W) */
  const float* c_k; const float* c_b; /* (C,W), (W)   or NULL */
  const float* m_k; const float* m_b; /* (2W|W,W),(W) or NULL */
  const float* dense_k[BFK_MAX_BLOCKS]; const float* dense_b[BFK_MAX_BLOCKS];   /* (W,W),(W) */
  const float* film_g_k[BFK_MAX_BLOCKS]; const float* film_g_b[BFK_MAX_BLOCKS]; /* (T,W),(W) or NULL */
  const float* film_b_k[BFK_MAX_BLOCKS]; const float* film_b_b[BFK_MAX_BLOCKS]; /* (T,W),(W) */
  const float* norm_g[BFK_MAX_BLOCKS]; const float* norm_b[BFK_MAX_BLOCKS];     /* (W),(W) or NULL */
  const float* out_k; const float* out_b; /* (W,D),(D) */
  /* residual projector of block l, skip = x Wp + bp (conditional_dense_block.py:125-127, 152-156), or NULL: identity skip */
  const float* proj_k[BFK_MAX_BLOCKS]; const float* proj_b[BFK_MAX_BLOCKS]; /* (W,W),(W) */
} bfk_weights;

typedef struct bfk_schedule {
  int32_t model;       /* BFK_MODEL_* */
  int32_t prediction;  /* BFK_PRED_* (DM only) */
  int32_t schedule;    /* BFK_SCHED_* (DM only) */
  int32_t exploding;   /* variance_type == "exploding" */
  double sigma_data, sigma_min, sigma_max, rho; /* EDM */
  double log_snr_min, log_snr_max, shift, t_min, t_max; /* cosine (and EDM bounds) */
} bfk_schedule;

int bfk_version(void);
const char* bfk_last_error_string(void);

/* ---- weights -------------------------------------------------------------------------------------------- */
size_t bfk_packed_bytes(const bfk_net_config* cfg);
int bfk_pack(const bfk_net_config* cfg, const bfk_weights* w, void* packed, void* stream);

/* ---- per-call prologues --------------------------------------------------------------------------------- */
/* conditions (G,C) -> pre_add (G,W) [added before the first activation] and post_add (G,W) [added to the merge GEMM] */
int bfk_cond_prologue(const bfk_net_config* cfg, const void* packed, const float* conditions, int64_t G,
                      float* pre_add, float* post_add, void* stream);
size_t bfk_table_stride(const bfk_net_config* cfg); /* floats per evaluation */
/* times (n) -> table (n, stride): FiLM shifts (and scales) of every block + velocity coefficients */
int bfk_eval_table(const bfk_net_config* cfg, const void* packed, const bfk_schedule* sched, const float* times,
                   int64_t n, float* table, void* stream);

/* ---- the hot path ------------------------------------------------------------------------------------------ */
/* One velocity evaluation at a single shared time (table holds 1 evaluation). */
int bfk_velocity(const bfk_net_config* cfg, const void* packed, const float* table, const float* pre_add,
                 const float* post_add, int64_t rows_per_group, const float* target_mask, const float* xz, float* out,
                 int64_t B, void* stream);
/* Fixed / scheduled stepping.  table holds n_steps*stages evaluations in execution order, step_h the n_steps step
 * sizes.  x_in may equal x_out.  Rows keep their whole trajectory on chip. */
int bfk_sample_fixed(const bfk_net_config* cfg, const void* packed, const float* table, const float* step_h,
                     int32_t method, int64_t n_steps, const float* pre_add, const float* post_add,
                     int64_t rows_per_group, const float* target_mask, const float* x_in, float* x_out, int64_t B,
                     void* stream);

/* ---- adaptive stepping (integrate_adaptive, bayesflow/utils/integrate.py:291-406) ---------------------------------
 * The controller state lives on the DEVICE as BFK_CTRL_FLOATS floats (indices below) so that the trial-step loop needs
 * no host round trip: per trial the host enqueues  bfk_adaptive_plan -> bfk_eval_table(times = ctrl+BFK_CTRL_TIMES,
 * n = 6) -> bfk_adaptive_trial  and polls ctrl[BFK_CTRL_DONE] every few trials; once done, the three calls are no-ops.
 * ybuf / kbuf hold two (B,D) slots each: slot ctrl[SEL] is the current state / FSAL slope k1, the other slot receives
 * the candidate.  acc = ctrl+BFK_CTRL_ACC holds {max row-L2 error, max|state|, any-NaN flag, any-non-NaN flag} of the
 * last trial (non-negative floats, so a cross-rank MAX all-reduce of these 4 floats between trial and plan reproduces
 * the reference's GLOBAL step control on sharded rows). */
enum {
  BFK_CTRL_TIME = 0, BFK_CTRL_STEP_SIZE = 1, BFK_CTRL_H = 2, BFK_CTRL_STEP = 3, BFK_CTRL_REJECTED = 4,
  BFK_CTRL_START = 5, BFK_CTRL_STOP = 6, BFK_CTRL_ATOL = 7, BFK_CTRL_RTOL = 8, BFK_CTRL_MIN_STEPS = 9,
  BFK_CTRL_MAX_STEPS = 10, BFK_CTRL_SEL = 11, BFK_CTRL_DONE = 12, BFK_CTRL_HAVE_TRIAL = 13, BFK_CTRL_ALL_NAN = 14,
  BFK_CTRL_TRIALS = 15, BFK_CTRL_MIN_STEP_SIZE = 16, BFK_CTRL_FINAL = 17, BFK_CTRL_SKIP = 18,
  BFK_CTRL_ACC = 20,   /* 4 floats */
  BFK_CTRL_TIMES = 24, /* 8 floats: evaluation times of the planned trial (stages 2..6 and t+h) */
  BFK_CTRL_FLOATS = 32
};
/* write the initial controller state (start/stop/min_steps/max_steps/atol/rtol; step_size = (stop-start)/min_steps) */
int bfk_adaptive_init(float* ctrl, double start_time, double stop_time, int64_t min_steps, int64_t max_steps,
                      double atol, double rtol, void* stream);
/* digest the last trial (accept / reject, new step size: integrate.py:353-378), test the loop condition (:325-335),
 * plan the next trial (:339-342) and write its evaluation times; method selects the stage abscissae */
int bfk_adaptive_plan(float* ctrl, int32_t method, void* stream);
/* after the loop: the final step that lands exactly on stop_time (integrate.py:393-403); then bfk_eval_table +
 * bfk_adaptive_trial once more and bfk_adaptive_commit */
int bfk_adaptive_plan_final(float* ctrl, int32_t method, void* stream);
int bfk_adaptive_commit(float* ctrl, void* stream);
/* one trial step over all rows (k2..k6, candidate state, FSAL k7, error norm): rk45_step / tsit5_step with
 * use_adaptive_step_size=True, integrate.py:57-120, 123-214.  table holds the 6 evaluations planned in ctrl. */
int bfk_adaptive_trial(const bfk_net_config* cfg, const void* packed, const float* table, float* ctrl, int32_t method,
                       const float* pre_add, const float* post_add, int64_t rows_per_group, const float* target_mask,
                       float* ybuf, float* kbuf, int64_t B, void* stream);

/* ---- layer-wise tensor-core GEMM (training forward/backward, widths > 256) ------------------------------------------
 * Operands are fp16 (precision 0) or bf16 (1) in the packed K-major image P[r/RT][k/8][r%RT][k%8] (row tile RT = 128
 * for A, = NT for B; NT in {64,128,256} and N % NT == 0; K % 32 == 0; rows zero-padded to the tile).  These replace
 * the `keras.layers.Dense` matmuls of TimeMLP / ConditionalDenseBlock (time_mlp.py:131-139,
 * conditional_dense_block.py:117) and their autograd transposes (torch_approximator.py:88-99). */
size_t bfk_packed_operand_bytes(int64_t rows, int64_t k, int32_t row_tile);
/* src (R, ld) fp32 row-major, contraction along columns -> packed (rows padded to row_tile, k padded to k_pad) */
int bfk_pack_kmajor(const float* src, int64_t ld, int64_t R, int32_t k_valid, int32_t k_pad, int32_t row_tile,
                    int32_t precision, void* dst, void* stream);
/* src (Kc, ld) fp32 row-major, contraction along ROWS; operand rows are the F columns of src -> packed */
int bfk_pack_kmajor_t(const float* src, int64_t ld, int64_t k_valid, int64_t k_pad, int32_t F, int32_t row_tile,
                      int32_t precision, void* dst, void* stream);
/* C (M, ldc) fp32 = alpha * A * B (+ bias) ; flags: 1 = atomic accumulate (use with k_splits > 1), 2 = add to existing */
int bfk_gemm(const void* A_packed, const void* B_packed, float* C, int64_t ldc, int64_t M, int32_t N, int32_t K,
             int32_t NT, int32_t k_splits, int32_t flags, const float* bias, float alpha, int32_t precision,
             void* stream);

/* ---- training step (FlowMatching.compute_metrics flow_matching.py:132-178, DiffusionModel.compute_metrics
 * diffusion_model.py:142-232; gradients = what torch autograd returns in torch_approximator.py:88-99) ----------------
 * The network sees x (B,D), the per-row subnet time t (B) and conditions (B,C).  The loss head is
 *     loss = loss_scale * sum_b w_b * mean_i L(a_b * pred_bi - target_bi),  L = square (0), abs (1) or the
 *     pseudo-Huber sqrt(r^2 + huber_c^2) - huber_c (2; ConsistencyModel.compute_metrics, consistency_model.py:405-416)
 * which covers FlowMatching (a = 1, target = x1 - x0) and every DiffusionModel prediction/loss type (a_b, target from
 * the noise schedule).  Gradients are ADDED to the tensors in `grads` (same struct as the weights; zero them first),
 * already multiplied by grad_scale (= 1/(B_global*D) for the reference's mean).  grads == NULL: forward (+ loss) only.
 * pred_out (B,D) optional.  dropout: rate of the blocks' Dropout layers (conditional_dense_block.py:120-121), mask
 * derived from seed.  Operands are fp16 with fp32 accumulation. */
typedef struct bfk_train_args {
  const float* x; const float* t; const float* cond; const float* target;
  const float* pred_scale; /* (B) or NULL */
  const float* row_weight; /* (B) or NULL */
  int64_t B;
  int32_t loss;            /* 0 mse, 1 mae, 2 pseudo-Huber */
  int32_t dw_swap;         /* debug: swap LBO/SBO of the MN-major descriptors; must be 0 */
  float dropout;
  float loss_scale, grad_scale;
  uint64_t seed;
  float huber_c;           /* loss == 2 only */
  const uint32_t* seed_dev; /* NULL, or 2 device words (lo, hi) that override `seed`: lets a captured CUDA graph of the step
                               be replayed with a new dropout mask */
  int64_t row_offset;      /* global index of row 0 (data-parallel shard): the dropout mask is a function of the global row */
  const float* elem_mask;  /* (B,D) of 0 / 1 or NULL: loss of mask*target against mask*prediction (drop_target_prob,
                              flow_matching.py:170-175, diffusion_model.py:177-216) */
} bfk_train_args;
size_t bfk_train_workspace_bytes(const bfk_net_config* cfg, int64_t B, int32_t backward);
int bfk_train_step(const bfk_net_config* cfg, const bfk_weights* weights, const bfk_weights* grads,
                   const bfk_train_args* args, float* loss, float* pred_out, void* workspace, size_t workspace_bytes,
                   void* stream);
/* The Dropout keep scales bfk_train_step applies in block `layer` (0-based) for global rows row_offset .. row_offset+B-1:
 * out (B,W) of 0 or 1/(1-rate).  Keras draws its mask from the backend RNG (conditional_dense_block.py:120-121, not
 * reproducible across backends); here it is a counter-based function of (seed, layer, global row, column) so that forward,
 * backward and every data-parallel shard agree.  Exposed for tests and for users who want the mask of a step. */
int bfk_train_dropout_mask(uint64_t seed, const uint32_t* seed_dev, int32_t layer, int64_t row_offset, int64_t B, int32_t W,
                           float rate, float* out, void* stream);
/* Vector-Jacobian product of the velocity network w.r.t. its input, for the density path
 * (FlowMatching._velocity_trace, flow_matching.py:238-252 -> utils/jacobian/jacobian_trace.py:10-55 -> utils/jacobian/vjp.py:44-68;
 * the reference runs D torch.autograd.grad passes per evaluation).  Consumes the activations the preceding
 * bfk_train_step(cfg, weights, NULL, args, NULL, pred, workspace, ...) left in `workspace` (sized with backward = 1).
 * Cotangent: dout (B,D), or the one-hot e_col when dout == NULL.  dx (B,D) is written when non-NULL; trace (B) gets
 * dx[:,col] ADDED when non-NULL.  Touches no parameter gradient. */
int bfk_train_vjp(const bfk_net_config* cfg, const bfk_weights* weights, const bfk_train_args* args, const float* dout,
                  int32_t col, float* dx, float* trace, void* workspace, size_t workspace_bytes, void* stream);
/* DiffusionModel probability-flow velocity as v = coefZ*z + coefP*pred(z, net_time) at time t, in double
 * (diffusion_model.py:351-405, 466-489, 539-552, 711-717; schedules/*.py) - the host-side twin of the table builder. */
int bfk_schedule_coefs(const bfk_schedule* sched, double t, double* coefZ, double* coefP, double* net_time);
/* Reverse-SDE terms at time t (DiffusionModel.velocity with stochastic_solver=True, diffusion_model.py:539-552, and
 * diffusion_term, :561-595): beta = g(t)^2 and fz with f = fz*z; SDE drift = 2*v_ode - fz*z. */
int bfk_schedule_sde_terms(const bfk_schedule* sched, double t, double* fz, double* beta);
/* alpha_t, sigma_t at time t (NoiseSchedule.get_alpha_sigma of get_log_snr(t, training=False):
 * networks/inference/diffusion/schedules/noise_schedule.py, edm_noise_schedule.py:61-75, cosine_noise_schedule.py:60-63). */
int bfk_schedule_alpha_sigma(const bfk_schedule* sched, double t, double* alpha, double* sigma);
/* One annealed-Langevin update on (B, D) rows (utils/integrate.py:924-961 `_apply_corrector`, :1167-1251
 * `integrate_langevin`): score = 2*(fz*x - v_ode)/beta from the probability-flow velocity the network kernels return;
 * out = x + e*score + sqrt(2e)*z with e = 2*a*(b*||z||/max(||score||,1e-8))^2 per row (per_row != 0: a = alpha_t,
 * b = step_size_factor) or e = a (per_row == 0: a = effective_factor*sigma_t^2).  out must not alias the inputs. */
int bfk_sde_langevin_update(const float* x, const float* v_ode, const float* z, int64_t B, int32_t D, double fz, double beta,
                            double a, double b, int32_t per_row, float* out, void* stream);
/* Runge-Kutta linear combination for host-driven integration (utils/integrate.py:47-54 add_scaled, :31-44, :98-105):
 * out[i] = y[i] + h * sum_j coef[j] * k[j][i], j < n (n <= 8), accumulated left to right in fp32 like the reference. */
int bfk_rk_combine(const float* y, const float* const* k, const float* coef, int32_t n, float h, int64_t numel, float* out,
                   void* stream);
/* x = t*x1 + (1-t)*x0, target = x1 - x0 (flow_matching.py:159-168) */
int bfk_fm_interpolate(const float* x0, const float* x1, const float* t, int64_t B, int32_t D, float* x, float* target,
                       void* stream);
/* keras.optimizers.Adam / AdamW with per-variable clipnorm (basic_workflow.py:514-526) over one flat parameter buffer;
 * seg_offsets (n_seg+1, device int64) delimit the variables, seg_sqnorm (n_seg) is scratch.  lr_t = lr*sqrt(1-b2^t)/(1-b1^t). */
int bfk_adam_step(float* params, const float* grads, float* m, float* v, const int64_t* seg_offsets, int32_t n_seg,
                  float* seg_sqnorm, float clipnorm, float lr_t, float beta1, float beta2, float eps, float wd_lr,
                  void* stream);
/* the same step with the step-dependent scalars read from HBM (hyper_dev[0] = lr_t, hyper_dev[1] = weight_decay * lr),
 * so that a captured CUDA graph of the whole training step can be replayed while the learning-rate schedule advances */
int bfk_adam_step_dev(float* params, const float* grads, float* m, float* v, const int64_t* seg_offsets, int32_t n_seg,
                      float* seg_sqnorm, float clipnorm, const float* hyper_dev, float beta1, float beta2, float eps,
                      void* stream);
/* FlowMatching.compute_metrics' sampling step (flow_matching.py:143-163) in one launch: x0 ~ N(0,1) (the bfk_normal_fill
 * stream of `seed`), u ~ U(0,1) per GLOBAL row (row_offset + r), t = u^(1+alpha), x = t*x1 + (1-t)*x0, target = x1 - x0.
 * seed_dev: NULL or 2 device words overriding seed (graph replay).  x0_out optional. */
int bfk_fm_train_batch(const float* x1, int64_t B, int32_t D, int64_t row_offset, uint64_t seed, const uint32_t* seed_dev,
                       float alpha, float* x0_out, float* t_out, float* x, float* target, void* stream);

/* ---- layer-wise sampler: networks the fused on-chip kernel cannot hold (W > 256 or D > 8; BASELINE configs[3]) --------
 * Same contract as bfk_sample_fixed (integrate.py:217-288 driven by flow_matching.py:254-337 / diffusion_model.py:782-869)
 * but taking the fp32 weights directly: one tcgen05 GEMM per Dense layer, rows processed in chunks of `chunk_rows` whose
 * activations stay L2-resident; each chunk runs its whole trajectory.  times: n_steps*stages evaluation times (device,
 * execution order), step_h: n_steps (device).  conditions (G,C): row r uses dataset r / rows_per_group (the Sampler's
 * repeat, samplers.py:51-63, without materialising it).  velocity_only != 0: one evaluation at times[0], x_out = velocity.
 * W % 64 == 0, W <= 1024, D <= 64, fp16 operands. */
size_t bfk_layerwise_workspace_bytes(const bfk_net_config* cfg, int64_t chunk_rows, int64_t n_evals, int64_t G);
int bfk_sample_layerwise(const bfk_net_config* cfg, const bfk_weights* weights, const bfk_schedule* sched,
                         const float* times, const float* step_h, int32_t method, int64_t n_steps,
                         const float* conditions, int64_t G, int64_t rows_per_group, const float* target_mask,
                         const float* x_in, float* x_out, int64_t B, int32_t velocity_only, int64_t chunk_rows,
                         void* workspace, size_t workspace_bytes, void* stream);

/* ---- introspection ------------------------------------------------------------------------------------------- */
int bfk_num_sms(void);
int bfk_fused_supported(const bfk_net_config* cfg); /* 1 if the on-chip sampler handles this config */

/* ---- the approximator edge of the sampling path ---------------------------------------------
 * bfk_normal_fill replaces DiagonalNormal.sample (bayesflow/distributions/diagonal_normal.py:100-105: mean + std *
 * keras.random.normal) for the z of InferenceNetwork.sample (inference_network.py:109-125): out (rows, D) written in
 * HBM, no host copy.  Philox4x32-10 keyed by `seed`, counter = global element index / 4, Box-Muller; element (r, d) of
 * the stream is a function of (seed, (row_offset + r) * D + d) only, so a rank that passes its row_offset draws exactly
 * the rows a single-GPU run would.  mean / stdv are (D) or NULL (0 / 1).  raw_words (rows*D uint32) or NULL: the raw
 * generator output, for known-answer tests.
 * bfk_standardize replaces Standardize.call with transformation_type="location_scale"
 * (bayesflow/networks/helpers/standardization/standardize.py:50-133): forward (x - mean) / std, inverse x * std + mean,
 * std = m2 > 0 ? sqrt(m2 / count) : 1 (:27-37); x, out are (rows, F), may alias.
 * bfk_batch_moments + bfk_moments_merge replace Standardize._update_moments (:135-170): stats (2F+1 floats: mean[F],
 * m2[F], count) are the batch moments of x (rows, F), computed in ONE read of x; bfk_moments_merge folds n_parts such
 * records (n_parts, 2F+1) into the running (mean, m2, count) with the reference's pairwise update, in index order — one
 * record for a single process, world_size records after an all-gather for data-parallel training.
 * workspace: bfk_moments_workspace_floats(rows, F) floats. */
int bfk_normal_fill(float* out, int64_t rows, int32_t D, int64_t row_offset, const float* mean, const float* stdv,
                    uint64_t seed, uint32_t* raw_words, void* stream);
int bfk_standardize(const float* x, float* out, int64_t rows, int32_t F, const float* mean, const float* m2,
                    const float* count, int32_t forward, void* stream);
/* out[i] = a*x[i] + b*y[i] + c*z[i] (y, z may be NULL; out may alias any input): the state updates of the
 * ConsistencyModel multistep sampler (consistency_model.py:300-306: x_n = x + sqrt(t_n^2 - eps^2) * noise) and of its
 * consistency function (:336-342: c_skip(t) * x + c_out(t) * f). */
/* conditions (B, C) with each dataset's row repeated (samplers.py:51-63): out2[0] = rows per dataset (index of the first
 * row that differs from row 0; B if all rows are equal), out2[1] = 0 iff every row equals the first row of its group and
 * out2[0] divides B.  Two streaming launches; replaces a sort-based unique pass over the materialised conditions. */
int bfk_detect_repeats(const float* conditions, int64_t B, int32_t C, int64_t* out2, void* stream);
int bfk_lincomb3(float* out, float a, const float* x, float b, const float* y, float c, const float* z, int64_t numel,
                 void* stream);
/* dst (rows_pad, cols_pad) = src (rows, cols) zero-padded; and the inverse copy of the valid region.  Parameter staging for
 * networks with unequal hidden widths (bfk_net_config.block_width): the kernels see zero-padded weights of one common width. */
int bfk_pad2d(const float* src, int32_t rows, int32_t cols, float* dst, int32_t rows_pad, int32_t cols_pad, void* stream);
int bfk_unpad2d(const float* src, int32_t cols_pad, float* dst, int32_t rows, int32_t cols, void* stream);
size_t bfk_moments_workspace_floats(int64_t rows, int32_t F);
int bfk_batch_moments(const float* x, int64_t rows, int32_t F, float* workspace, float* stats, void* stream);
/* Allocation-free variant of bfk_batch_moments.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that
 * forbid it).  Same inputs, outputs and results as bfk_batch_moments, but the per-block partial moments live in a
 * file-scope `static __device__` array of bfk_static_scratch_floats() = 2^19 floats (2 MiB, held by every GPU that loads
 * the library until the process exits) instead of a caller workspace.  Calls no cudaMalloc, cudaFree or allocating
 * library routine.  Returns -69 with a message (bfk_last_error_string) when bfk_moments_workspace_floats(rows, F) exceeds
 * the array: on a 148-SM B200 any number of rows fits for F <= 147 (1184 blocks x 3F floats), and wider F fits while
 * the launch uses at most 2^19 / (3 F) blocks; beyond that use bfk_batch_moments.  All calls share the array:
 * calls that overlap on the GPU (different streams without ordering) race; back-to-back calls on one stream are safe,
 * and nothing in the array needs clearing between calls (every slot that is read is overwritten first). */
int bfk_batch_moments_static(const float* x, int64_t rows, int32_t F, float* stats, void* stream);
size_t bfk_static_scratch_floats(void);
int bfk_moments_merge(float* mean, float* m2, float* count, const float* parts, int32_t n_parts, int32_t F,
                      void* stream);

/* ---- mini-batch optimal-transport pairing for FlowMatching training ----------------------------
 * Replace bayesflow/utils/optimal_transport: log_sinkhorn.py:14-172 (log-stabilised Sinkhorn-Knopp + categorical
 * assignment, the FlowMatching default: networks/defaults.py:54-63, flow_matching.py:146-157), ot_utils.py:6-34 (costs)
 * and :95-164 (conditional weight search).  All matrices are dense row-major fp32 in device memory.
 *   bfk_ot_cost      M (n,m) = max(|x1_i|^2 + |x2_j|^2 - 2 x1_i.x2_j, 0); cosine != 0: 1 - clip(cos(x1_i, x2_j), -1, 1)
 *   bfk_ot_ratio     *count = #{(i,j): w * C_ij <= M_ii - M_ij} over (n,n)  (get_ratio numerator, ot_utils.py:117-121)
 *   bfk_ot_scale     M <- -(M + w * C) / regularization in place (C may be NULL)
 *   bfk_ot_init      ctrl (8 floats: err, steps, done, atol, max_steps, ...), u = 0 (n), v = 0 (m)
 *   bfk_ot_log_sinkhorn_step  n_steps bodies of the while loop (log_sinkhorn.py:131-137); log marginals log_a_vec (n) /
 *                    log_b_vec (m), or NULL for the uniform -log n / -log m (partial OT passes the augmented ones,
 *                    ot_utils.py:37-92, with S the (n0+1, m0+1) augmented matrix); a no-op once
 *                    ctrl[2] (done) is set, so the host may enqueue several steps between polls of ctrl
 *   bfk_ot_sample    assignments[i] ~ Categorical(softmax_j(S_ij + u_i + v_j)) by Gumbel-max over the Philox4x32-10
 *                    stream keyed by seed (word i*m + j); S has leading dimension ld >= m (partial OT samples the
 *                    leading (n0, m0) block of the augmented matrix); int64 indices into x2. */
int bfk_ot_cost(const float* x1, const float* x2, int64_t n, int64_t m, int32_t D, int32_t cosine, float* M, void* stream);
int bfk_ot_ratio(const float* M, const float* Cm, int64_t n, float w, unsigned long long* count, void* stream);
int bfk_ot_scale(float* M, const float* Cm, float w, float regularization, int64_t numel, void* stream);
size_t bfk_ot_work_floats(int64_t n, int64_t m);
int bfk_ot_init(float* ctrl, float* u, float* v, int64_t n, int64_t m, float atol, int64_t max_steps, void* stream);
int bfk_ot_log_sinkhorn_step(const float* S, int64_t n, int64_t m, const float* log_a_vec, const float* log_b_vec,
                             float* u, float* v, float* ctrl, float* work, int32_t n_steps, void* stream);
/* Plain Sinkhorn-Knopp (utils/optimal_transport/sinkhorn.py:65-186) on the same sweeps: after bfk_ot_init,
 * bfk_ot_plain_mode(S, numel, ctrl) records max(S) and switches bfk_ot_log_sinkhorn_step to the multiplicative algorithm's
 * arithmetic (scalings kept as logs, the reference's `tiny` = 1e-12 terms and its log-relative-change error);
 * bfk_ot_sample_plan draws assignments ~ Categorical(log(plan + 1e-10)) (sinkhorn.py:60-62). */
int bfk_ot_plain_mode(const float* S, int64_t numel, float* ctrl, void* stream);
int bfk_ot_sample_plan(const float* S, int64_t ld, const float* u, const float* v, const float* ctrl, int64_t n, int64_t m,
                       uint64_t seed, int64_t* assignments, void* stream);
int bfk_ot_sample(const float* S, int64_t ld, const float* v, int64_t n, int64_t m, uint64_t seed, int64_t* assignments,
                  void* stream);

#ifdef __cplusplus
}
#endif
#endif /* BFK_H_ */
