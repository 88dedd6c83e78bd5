// libbfk: C-ABI entry points (include/bfk.h) over the sm_100a kernels.  No torch types, no exceptions.
#include <cuda_runtime.h>
#include <stdio.h>
#include <string.h>
#include <stdlib.h>

#include "bfk_sampler.cuh"
#include "bfk_sampler_fold.cuh"
#include "bfk_adaptive.cuh"
#include "bfk_gemm.cuh"

using namespace bfk;

static thread_local char g_err[512] = "";
static int fail(int code, const char* msg) {
  snprintf(g_err, sizeof(g_err), "%s", msg);
  return code;
}
static int cuda_fail(cudaError_t e, const char* where) {
  snprintf(g_err, sizeof(g_err), "%s: %s", where, cudaGetErrorString(e));
  return (int)e;
}
#define CK(call)                                              \
  do {                                                        \
    cudaError_t _e = (call);                                  \
    if (_e != cudaSuccess) return cuda_fail(_e, #call);       \
  } while (0)

extern "C" int bfk_version(void) { return 100; }
extern "C" const char* bfk_last_error_string(void) { return g_err; }

extern "C" int bfk_num_sms(void) {
  int dev = 0, n = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return -1;
  if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess) return -1;
  return n;
}

static int check_cfg(const bfk_net_config* c) {
  if (!c) return fail(-1, "null config");
  if (c->D < 1 || c->W < 32 || c->L < 1 || c->L > BFK_MAX_BLOCKS) return fail(-2, "bad D/W/L");
  if (c->T != 1 && (c->T < 3 || c->T > 65 || (c->T % 2) == 0)) return fail(-3, "time embedding size must be 1 or odd in [3,65]");
  if (c->C < 0) return fail(-4, "bad C");
  if (c->C > 0 && c->merge != BFK_MERGE_CONCAT && c->merge != BFK_MERGE_ADD) return fail(-5, "bad merge mode");
  return 0;
}

typedef void (*sampler_fn)(const SamplerArgs);
static sampler_fn pick_kernel(const bfk_net_config* c, int* threads);

// Which fused kernel serves the reference's default block shape: the chase schedule (bfk_sampler_chase.cuh, default) or
// the two-tile ping-pong kernel it replaced (BFK_SAMPLER=fast, kept for A/B timing).  Read once per process: the packed
// blob and the per-evaluation table are laid out for the kernel that will read them.
// The reference's default block shape (LayerNorm + residual + FiLM shift) runs on the folded single-pass kernel
// (bfk_sampler_fold.cuh); bfk_pack and bfk_eval_table lay the weights and the per-evaluation table out for it.
static bool use_fold(const bfk_net_config* c) { return is_fast(*c) != 0; }

extern "C" int bfk_fused_supported(const bfk_net_config* c) {
  if (check_cfg(c)) return 0;
  if (c->D > BFK_MAX_D) return 0;
  if (c->W > 256 || (c->W % 32) != 0) return 0;
  for (int l = 0; l < c->L; ++l)
    if (c->block_width[l] != 0) return 0;  // unequal hidden widths: the layer-wise training / density kernels only
  if (c->precision != BFK_PREC_FP16 && c->precision != BFK_PREC_BF16) return 0;
  SmemPlan p = use_fold(c) ? make_fold_plan(*c) : make_smem_plan(*c);
  if (p.n_stage < 2) return 0;
  int threads = 0;
  return pick_kernel(c, &threads) != nullptr;  // configurations without an instantiation take the layer-wise path
}

extern "C" size_t bfk_packed_bytes(const bfk_net_config* c) {
  if (check_cfg(c)) return 0;
  return make_layout(*c).total;
}
extern "C" size_t bfk_table_stride(const bfk_net_config* c) {
  if (check_cfg(c)) return 0;
  return (size_t)table_stride(*c);
}

static inline unsigned cdiv(size_t a, size_t b) { return (unsigned)((a + b - 1) / b); }

// cudaFuncSetAttribute is per device: cache "already set" per device id, not per process
static inline bool* attr_flag(bool* flags) {
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) dev = 0;
  return flags + dev;
}


extern "C" int bfk_pack(const bfk_net_config* c, const bfk_weights* w, void* packed, void* stream_) {
  if (int r = check_cfg(c)) return r;
  if (!w || !packed) return fail(-10, "null pointer");
  if ((c->W % 8) != 0) return fail(-11, "W must be a multiple of 8");
  cudaStream_t st = (cudaStream_t)stream_;
  const PackedLayout L = make_layout(*c);
  unsigned char* P = (unsigned char*)packed;
  const int W = c->W, bf = c->precision == BFK_PREC_BF16;
  CK(cudaMemsetAsync(P, 0, L.total, st));
  // GEMM operands
  int g = 0;
  if (L.has_merge) {
    if (!w->m_k || !w->m_b || !w->c_k || !w->c_b) return fail(-12, "conditional net needs c_proj and merge_proj");
    // x half / whole (add); the folded layout carries pre-activations times log2(e) and has XK extra (zero) K rows
    if (use_fold(c))
      pack_gemm_kernel<<<cdiv((size_t)W / 8 * W, 256), 256, 0, st>>>(w->m_k, W, W, bf, (uint16_t*)(P + L.gemm_x), 1.4426950408889634f);
    else
      pack_gemm_kernel<<<cdiv((size_t)W / 8 * W, 256), 256, 0, st>>>(w->m_k, W, W, bf, (uint16_t*)(P + L.gemm));
    ++g;
  }
  const bool chase = use_fold(c);
  float* cv = (float*)(P + L.cvec);
  for (int l = 0; l < c->L; ++l, ++g) {
    if (!w->dense_k[l] || !w->dense_b[l] || !w->film_b_k[l] || !w->film_b_b[l]) return fail(-13, "missing block weights");
    if (chase) {
      // LayerNorm of block l-1 folded into the Dense kernel of block l (see fold_pack_kernel)
      if (!w->norm_g[l] || !w->norm_b[l]) return fail(-15, "LayerNorm weights missing");
      fold_pack_kernel<<<cdiv(W, 128), 128, 0, st>>>(w->dense_k[l], w->dense_b[l], l ? w->norm_g[l - 1] : nullptr,
                                                    l ? w->norm_b[l - 1] : nullptr, W, W, bf,
                                                    (uint16_t*)(P + L.gemm_x + (size_t)g * (W + XK) * W * 2),
                                                    cv + L.cv_blk + (size_t)l * W);
      copy_f32_kernel<<<cdiv(W, 256), 256, 0, st>>>(w->norm_b[l], (float*)(P + L.ln_beta) + (size_t)l * W, W);
    } else {
      pack_gemm_kernel<<<cdiv((size_t)W / 8 * W, 256), 256, 0, st>>>(w->dense_k[l], W, W, bf,
                                                                    (uint16_t*)(P + L.gemm + (size_t)g * W * W * 2));
    }
  }
  // constant vectors (SMEM image)
  pad2d_kernel<<<cdiv((size_t)L.DP * W, 256), 256, 0, st>>>(w->x_k, c->D, W, cv + L.cv_wx, L.DP, W);
  copy_f32_kernel<<<cdiv(W, 256), 256, 0, st>>>(w->x_b, cv + L.cv_bx, W);
  if (chase) scale_f32_kernel<<<cdiv((size_t)(L.DP + 1) * W, 256), 256, 0, st>>>(cv + L.cv_wx, 1.4426950408889634f, (size_t)(L.DP + 1) * W);
  for (int l = 0; l < c->L && !chase; ++l) {
    copy_f32_kernel<<<cdiv(W, 256), 256, 0, st>>>(w->dense_b[l], cv + L.cv_blk + (3 * l + 0) * W, W);
    if (c->norm != BFK_NORM_NONE && w->norm_g[l])
      copy_f32_kernel<<<cdiv(W, 256), 256, 0, st>>>(w->norm_g[l], cv + L.cv_blk + (3 * l + 1) * W, W);
    else
      fill_f32_kernel<<<cdiv(W, 256), 256, 0, st>>>(cv + L.cv_blk + (3 * l + 1) * W, 1.f, W);
    copy_f32_kernel<<<cdiv(W, 256), 256, 0, st>>>(c->norm == BFK_NORM_LAYER ? w->norm_b[l] : nullptr,
                                                  cv + L.cv_blk + (3 * l + 2) * W, W);
  }
  if (is_fast(*c)) {
    fold_proj_kernel<<<1, 32, 0, st>>>(w->out_k, w->out_b, w->norm_g[c->L - 1], w->norm_b[c->L - 1], W, c->D, L.DP,
                                      cv + L.cv_wo, cv + L.cv_bo);
  } else {
    pad2d_kernel<<<cdiv((size_t)W * L.DP, 256), 256, 0, st>>>(w->out_k, W, c->D, cv + L.cv_wo, W, L.DP);
    pad2d_kernel<<<1, 32, 0, st>>>(w->out_b, 1, c->D, cv + L.cv_bo, 1, L.DP);
  }
  // prologue parameters
  if (L.Th > 0) copy_f32_kernel<<<1, 64, 0, st>>>(w->freq, (float*)(P + L.freq), L.Th);
  for (int l = 0; l < c->L; ++l) {
    const size_t n = (size_t)c->T * W;
    copy_f32_kernel<<<cdiv(n, 256), 256, 0, st>>>(w->film_b_k[l], (float*)(P + L.film_b_k) + l * n, n);
    copy_f32_kernel<<<cdiv(W, 256), 256, 0, st>>>(w->film_b_b[l], (float*)(P + L.film_b_b) + (size_t)l * W, W);
    if (c->film_use_gamma) {
      if (!w->film_g_k[l] || !w->film_g_b[l]) return fail(-14, "film_use_gamma set but to_gamma weights missing");
      copy_f32_kernel<<<cdiv(n, 256), 256, 0, st>>>(w->film_g_k[l], (float*)(P + L.film_g_k) + l * n, n);
      copy_f32_kernel<<<cdiv(W, 256), 256, 0, st>>>(w->film_g_b[l], (float*)(P + L.film_g_b) + (size_t)l * W, W);
    }
  }
  if (L.has_merge) {
    const size_t n = (size_t)c->C * W;
    copy_f32_kernel<<<cdiv(n, 256), 256, 0, st>>>(w->c_k, (float*)(P + L.c_k), n);
    copy_f32_kernel<<<cdiv(W, 256), 256, 0, st>>>(w->c_b, (float*)(P + L.c_b), W);
    if (c->merge == BFK_MERGE_CONCAT)
      copy_f32_kernel<<<cdiv((size_t)W * W, 256), 256, 0, st>>>(w->m_k + (size_t)W * W, (float*)(P + L.m_c), (size_t)W * W);
    copy_f32_kernel<<<cdiv(W, 256), 256, 0, st>>>(w->m_b, (float*)(P + L.m_b), W);
  }
  CK(cudaGetLastError());
  return 0;
}

extern "C" int bfk_cond_prologue(const bfk_net_config* c, const void* packed, const float* cond, int64_t G,
                                 float* pre_add, float* post_add, void* stream_) {
  if (int r = check_cfg(c)) return r;
  if (c->C <= 0) return fail(-20, "network has no conditions");
  if (!packed || !cond || !post_add || G <= 0) return fail(-21, "bad arguments");
  if (c->merge == BFK_MERGE_ADD && !pre_add) return fail(-22, "merge='add' needs pre_add");
  const PackedLayout L = make_layout(*c);
  const unsigned grid = (unsigned)(G < 148 * 16 ? G : 148 * 16);
  const size_t sh = (size_t)(c->C + c->W) * 4;
  cond_prologue_kernel<<<grid, 256, sh, (cudaStream_t)stream_>>>(*c, (const unsigned char*)packed, L, cond, G, pre_add, post_add,
                                                                use_fold(c) ? 1.4426950408889634f : 1.f);
  CK(cudaGetLastError());
  return 0;
}

extern "C" int bfk_eval_table(const bfk_net_config* c, const void* packed, const bfk_schedule* sched, const float* times,
                              int64_t n, float* table, void* stream_) {
  if (int r = check_cfg(c)) return r;
  if (!packed || !sched || !times || !table || n <= 0) return fail(-30, "bad arguments");
  const PackedLayout L = make_layout(*c);
  const unsigned grid = (unsigned)(n < 148 * 8 ? n : 148 * 8);
  eval_table_kernel<<<grid, 256, 0, (cudaStream_t)stream_>>>(*c, (const unsigned char*)packed, L, *sched, times, n,
                                                           table_stride(*c), table, use_fold(c) ? 1 : 0);
  CK(cudaGetLastError());
  return 0;
}

// ---------------------------------------------------------------------------------------------------- tableaux
// integrate.py:74-96 (Dormand-Prince) and :139-191 (Tsit5); coefficients are cast to fp32 exactly where torch casts
// the python doubles (scalar * fp32 tensor).
static void fill_tableau(int method, SamplerArgs& a) {
  memset(a.rk_a, 0, sizeof(a.rk_a));
  memset(a.rk_b, 0, sizeof(a.rk_b));
  memset(a.rk_e, 0, sizeof(a.rk_e));
  a.s0 = 0; a.err_kind = 0;
  if (method == BFK_METHOD_EULER) {
    a.S = 1;
    a.rk_b[0] = 1.f;
    return;
  }
  a.S = 6;
  if (method == BFK_METHOD_RK45) {
    const double A[6][5] = {{0},
                            {1.0 / 5},
                            {3.0 / 40, 9.0 / 40},
                            {44.0 / 45, -56.0 / 15, 32.0 / 9},
                            {19372.0 / 6561, -25360.0 / 2187, 64448.0 / 6561, -212.0 / 729},
                            {9017.0 / 3168, -355.0 / 33, 46732.0 / 5247, 49.0 / 176, -5103.0 / 18656}};
    const double Bc[6] = {35.0 / 384, 0.0, 500.0 / 1113, 125.0 / 192, -2187.0 / 6784, 11.0 / 84};
    const double B4[7] = {5179.0 / 57600, 0.0, 7571.0 / 16695, 393.0 / 640, -92097.0 / 339200, 187.0 / 2100, 1.0 / 40};
    for (int i = 0; i < 6; ++i) {
      for (int j = 0; j < 5; ++j) a.rk_a[i][j] = (float)A[i][j];
      a.rk_b[i] = (float)Bc[i];
    }
    for (int i = 0; i < 7; ++i) a.rk_e[i] = (float)B4[i];
    a.err_kind = 0;
  } else {
    const double A[6][5] = {{0},
                            {0.161},
                            {-0.0084806554923570, 0.3354806554923570},
                            {2.897153057105494, -6.359448489975075, 4.362295432869581},
                            {5.325864828439257, -11.74888356406283, 7.495539342889836, -0.09249506636175525},
                            {5.86145544294270, -12.92096931784711, 8.159367898576159, -0.07158497328140100, -0.02826905039406838}};
    const double Bc[6] = {0.09646076681806523, 0.01, 0.4798896504144996, 1.379008574103742, -3.290069515436081, 2.324710524099774};
    const double E[7] = {-0.00178001105222577714, -0.0008164344596567469, 0.007880878010261995, -0.1447110071732629,
                         0.5823571654525552, -0.45808210592918697, 0.015151515151515152};
    for (int i = 0; i < 6; ++i) {
      for (int j = 0; j < 5; ++j) a.rk_a[i][j] = (float)A[i][j];
      a.rk_b[i] = (float)Bc[i];
    }
    for (int i = 0; i < 7; ++i) a.rk_e[i] = (float)E[i];
    a.err_kind = 1;
  }
}

static StageC stage_c(int method) {
  StageC sc{};
  const double rk[6] = {1.0 / 5, 3.0 / 10, 4.0 / 5, 8.0 / 9, 1.0, 1.0};
  const double ts[6] = {0.161, 0.327, 0.9, 0.9800255409045097, 1.0, 1.0};
  for (int i = 0; i < 6; ++i) sc.c[i] = (float)(method == BFK_METHOD_RK45 ? rk[i] : ts[i]);
  return sc;
}

static sampler_fn pick_kernel(const bfk_net_config* c, int* threads) {
  const int dp = dp_of(c->D);
  const bool bf = c->precision == BFK_PREC_BF16;
  if (use_fold(c)) {
    *threads = F_NUM_THREADS;
#define PICKD(ACT_, DP_, BF_) \
  if (c->activation == ACT_ && dp == DP_ && bf == BF_) return sampler_fold_kernel<ACT_, DP_, BF_>;
    PICKD(BFK_ACT_MISH, 2, false)
#ifndef BFK_DEV_BUILD
    PICKD(BFK_ACT_MISH, 4, false) PICKD(BFK_ACT_MISH, 8, false)
    PICKD(BFK_ACT_MISH, 2, true) PICKD(BFK_ACT_MISH, 4, true) PICKD(BFK_ACT_MISH, 8, true)
    PICKD(BFK_ACT_SILU, 2, false) PICKD(BFK_ACT_SILU, 4, false) PICKD(BFK_ACT_SILU, 8, false)
    PICKD(BFK_ACT_RELU, 2, false) PICKD(BFK_ACT_RELU, 4, false) PICKD(BFK_ACT_RELU, 8, false)
    PICKD(BFK_ACT_GELU, 2, false) PICKD(BFK_ACT_GELU, 4, false) PICKD(BFK_ACT_GELU, 8, false)
    PICKD(BFK_ACT_TANH, 2, false) PICKD(BFK_ACT_TANH, 4, false) PICKD(BFK_ACT_TANH, 8, false)
#endif
#undef PICKD
    return nullptr;
  }
  // non-default options (norm None/rms, no residual, FiLM gamma): generic 10-warp kernel
  *threads = NUM_THREADS;
#define PICK(ACT_, DP_, BF_) \
  if (c->activation == ACT_ && dp == DP_ && bf == BF_) return sampler_kernel<ACT_, DP_, BF_>;
  PICK(BFK_ACT_MISH, 2, false)
#ifndef BFK_DEV_BUILD
  PICK(BFK_ACT_MISH, 4, false) PICK(BFK_ACT_MISH, 8, false)
  PICK(BFK_ACT_SILU, 2, false) PICK(BFK_ACT_SILU, 4, false) PICK(BFK_ACT_SILU, 8, false)
  PICK(BFK_ACT_RELU, 2, false) PICK(BFK_ACT_RELU, 4, false) PICK(BFK_ACT_RELU, 8, false)
  PICK(BFK_ACT_GELU, 2, false) PICK(BFK_ACT_GELU, 4, false) PICK(BFK_ACT_GELU, 8, false)
  PICK(BFK_ACT_TANH, 2, false) PICK(BFK_ACT_TANH, 4, false) PICK(BFK_ACT_TANH, 8, false)
#endif
#undef PICK
  return nullptr;
}

static int launch_sampler(const bfk_net_config* c, const void* packed, const float* table, const float* step_h, int method,
                          int64_t n_steps, const float* pre_add, const float* post_add, int64_t rpg, const float* tmask,
                          const float* x_in, float* x_out, int64_t B, int mode, cudaStream_t st, float* ctrl = nullptr,
                          float* ybuf = nullptr, float* kbuf = nullptr) {
  if (int r = check_cfg(c)) return r;
  if (!bfk_fused_supported(c)) return fail(-40, "config not supported by the fused on-chip sampler (need D<=8, W<=256, W%32==0)");
  if (mode == 2) {
    if (!packed || !table || !ctrl || !ybuf || !kbuf || B <= 0) return fail(-41, "bad arguments");
    if (method != BFK_METHOD_RK45 && method != BFK_METHOD_TSIT5)
      return fail(-46, "Adaptive step sizing is not supported for the 'euler' method.");
  } else if (!packed || !table || !step_h || !x_in || !x_out || B <= 0 || n_steps <= 0) {
    return fail(-41, "bad arguments");
  }
  if (method < BFK_METHOD_EULER || method > BFK_METHOD_TSIT5) return fail(-42, "unknown method");
  const PackedLayout L = make_layout(*c);
  if (L.has_merge && !post_add) return fail(-43, "conditional net needs post_add from bfk_cond_prologue");
  if (L.has_merge && c->merge == BFK_MERGE_ADD && !pre_add) return fail(-44, "merge='add' needs pre_add");
  if (rpg <= 0) rpg = 1;
  int threads = 0;
  sampler_fn fn = pick_kernel(c, &threads);
  if (!fn) return fail(-45, "no kernel instantiated for this activation / D / precision");

  SamplerArgs a{};
  a.wg = (const unsigned char*)packed + L.gemm;
  a.cvec = (const float*)((const unsigned char*)packed + L.cvec);
  a.table = table; a.step_h = step_h;
  a.pre_add = (L.has_merge && c->merge == BFK_MERGE_ADD) ? pre_add : nullptr;
  a.post_add = L.has_merge ? post_add : nullptr;
  a.x_in = x_in; a.x_out = x_out; a.tmask = tmask;
  a.B = B; a.rows_per_group = rpg; a.n_steps = n_steps;
  a.D = c->D; a.W = c->W; a.L = c->L; a.NG = L.NG; a.has_merge = L.has_merge;
  a.residual = c->residual; a.norm = c->norm; a.film_gamma = c->film_use_gamma;
  a.mode = mode;
  a.tstride = table_stride(*c); a.cvec_floats = L.cvec_floats;
  a.cv_wx = L.cv_wx; a.cv_bx = L.cv_bx; a.cv_blk = L.cv_blk; a.cv_wo = L.cv_wo; a.cv_bo = L.cv_bo;
  a.idesc = umma_idesc_f16(TILE_M, c->W, c->precision == BFK_PREC_BF16 ? 1 : 0);
  a.plan = use_fold(c) ? make_fold_plan(*c) : make_smem_plan(*c);
  if (use_fold(c)) a.wg = (const unsigned char*)packed + L.gemm_x;
  a.err_flag = nullptr;
  { const char* e = getenv("BFK_DEBUG"); a.dbg = e ? atoi(e) : 0; }
  fill_tableau(method, a);
  if (mode == 1) { a.S = 1; }
  if (mode == 2) { a.S = 7; a.s0 = 1; a.ctrl = ctrl; a.ybuf = ybuf; a.kbuf = kbuf; }

  int dev = 0, sms = 0;
  CK(cudaGetDevice(&dev));
  CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  CK(cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)a.plan.total));
  const long long rows_per_cta = TILES_PER_CTA * TILE_M;
  const long long n_pairs = (B + rows_per_cta - 1) / rows_per_cta;
  const unsigned grid = (unsigned)(n_pairs < sms ? n_pairs : sms);
  fn<<<grid, threads, a.plan.total, st>>>(a);
  CK(cudaGetLastError());
  return 0;
}

extern "C" int bfk_velocity(const bfk_net_config* c, const void* packed, const float* table, const float* pre_add,
                            const float* post_add, int64_t rpg, const float* tmask, const float* xz, float* out, int64_t B,
                            void* stream_) {
  static const float one = 1.f;
  (void)one;
  // step_h is read once per step; any valid device float will do for the single evaluation: reuse the table
  return launch_sampler(c, packed, table, table, BFK_METHOD_EULER, 1, pre_add, post_add, rpg, tmask, xz, out, B, 1,
                        (cudaStream_t)stream_);
}

extern "C" int bfk_sample_fixed(const bfk_net_config* c, const void* packed, const float* table, const float* step_h,
                                int32_t method, int64_t n_steps, const float* pre_add, const float* post_add, int64_t rpg,
                                const float* tmask, const float* x_in, float* x_out, int64_t B, void* stream_) {
  return launch_sampler(c, packed, table, step_h, method, n_steps, pre_add, post_add, rpg, tmask, x_in, x_out, B, 0,
                        (cudaStream_t)stream_);
}

#ifdef BFK_PROFILE
// timing builds only (-DBFK_PROFILE): per-CTA, per-warp cycle counters of the last chase-kernel launch -> host
extern "C" int bfk_debug_read_prof(long long* host_out) {
  CK(cudaDeviceSynchronize());
  CK(cudaMemcpyFromSymbol(host_out, g_bfk_prof, sizeof(long long) * 148 * 18 * 8));
  return 0;
}
#endif

// ---------------------------------------------------------------------------------------------------- adaptive
extern "C" int bfk_adaptive_init(float* ctrl, double start_time, double stop_time, int64_t min_steps, int64_t max_steps,
                                 double atol, double rtol, void* stream_) {
  if (!ctrl) return fail(-50, "null controller");
  if (max_steps <= min_steps) return fail(-51, "Maximum number of steps must be greater than minimum number of steps.");
  // integrate.py:316-321: python doubles cast to fp32 where they meet a tensor
  const float step_size = (float)((stop_time - start_time) / (double)min_steps);
  adaptive_init_kernel<<<1, 32, 0, (cudaStream_t)stream_>>>(ctrl, (float)start_time, (float)stop_time, (float)min_steps,
                                                           (float)max_steps, (float)atol, (float)rtol, step_size);
  CK(cudaGetLastError());
  return 0;
}
static int check_adaptive_method(int method) {
  if (method == BFK_METHOD_EULER) return fail(-46, "Adaptive step sizing is not supported for the 'euler' method.");
  if (method != BFK_METHOD_RK45 && method != BFK_METHOD_TSIT5) return fail(-42, "unknown method");
  return 0;
}
extern "C" int bfk_adaptive_plan(float* ctrl, int32_t method, void* stream_) {
  if (!ctrl) return fail(-50, "null controller");
  if (int r = check_adaptive_method(method)) return r;
  adaptive_plan_kernel<<<1, 32, 0, (cudaStream_t)stream_>>>(ctrl, stage_c(method));
  CK(cudaGetLastError());
  return 0;
}
extern "C" int bfk_adaptive_plan_final(float* ctrl, int32_t method, void* stream_) {
  if (!ctrl) return fail(-50, "null controller");
  if (int r = check_adaptive_method(method)) return r;
  adaptive_plan_final_kernel<<<1, 32, 0, (cudaStream_t)stream_>>>(ctrl, stage_c(method));
  CK(cudaGetLastError());
  return 0;
}
extern "C" int bfk_adaptive_commit(float* ctrl, void* stream_) {
  if (!ctrl) return fail(-50, "null controller");
  adaptive_commit_kernel<<<1, 32, 0, (cudaStream_t)stream_>>>(ctrl);
  CK(cudaGetLastError());
  return 0;
}
extern "C" int bfk_adaptive_trial(const bfk_net_config* c, const void* packed, const float* table, float* ctrl,
                                  int32_t method, const float* pre_add, const float* post_add, int64_t rpg,
                                  const float* tmask, float* ybuf, float* kbuf, int64_t B, void* stream_) {
  return launch_sampler(c, packed, table, nullptr, method, 1, pre_add, post_add, rpg, tmask, nullptr, nullptr, B, 2,
                        (cudaStream_t)stream_, ctrl, ybuf, kbuf);
}

// ---------------------------------------------------------------------------------------------------- layer-wise GEMM
static inline int64_t round_up64(int64_t x, int64_t a) { return (x + a - 1) / a * a; }

extern "C" size_t bfk_packed_operand_bytes(int64_t rows, int64_t k, int32_t row_tile) {
  if (rows <= 0 || k <= 0 || row_tile <= 0) return 0;
  return (size_t)round_up64(rows, row_tile) * (size_t)round_up64(k, 32) * 2;
}

extern "C" int bfk_pack_kmajor(const float* src, int64_t ld, int64_t R, int32_t k_valid, int32_t k_pad, int32_t row_tile,
                               int32_t precision, void* dst, void* stream_) {
  if (!src || !dst || R <= 0 || k_valid <= 0 || k_pad < k_valid || (k_pad % 32) || row_tile <= 0 || (row_tile % 8))
    return fail(-60, "bfk_pack_kmajor: bad arguments");
  const int64_t R_pad = round_up64(R, row_tile);
  const int64_t n = R_pad * (k_pad / 8);
  pack_kmajor_kernel<<<(unsigned)((n + 255) / 256), 256, 0, (cudaStream_t)stream_>>>(src, ld, R, R_pad, k_valid, k_pad, row_tile,
                                                                                   precision == BFK_PREC_BF16, (uint16_t*)dst);
  CK(cudaGetLastError());
  return 0;
}

extern "C" int bfk_pack_kmajor_t(const float* src, int64_t ld, int64_t k_valid, int64_t k_pad, int32_t F, int32_t row_tile,
                                 int32_t precision, void* dst, void* stream_) {
  if (!src || !dst || F <= 0 || k_valid <= 0 || k_pad < k_valid || (k_pad % 32) || row_tile <= 0 || (row_tile % 8))
    return fail(-61, "bfk_pack_kmajor_t: bad arguments");
  const int F_pad = (int)round_up64(F, row_tile);
  const int64_t n = (int64_t)F_pad * (k_pad / 8);
  pack_kmajor_t_kernel<<<(unsigned)((n + 255) / 256), 256, 0, (cudaStream_t)stream_>>>(src, ld, k_valid, k_pad, F, F_pad, row_tile,
                                                                                     precision == BFK_PREC_BF16, (uint16_t*)dst);
  CK(cudaGetLastError());
  return 0;
}

extern "C" int bfk_gemm(const void* A, const void* B, float* Cm, int64_t ldc, int64_t M, int32_t N, int32_t K, int32_t NT,
                        int32_t k_splits, int32_t flags, const float* bias, float alpha, int32_t precision, void* stream_) {
  if (!A || !B || !Cm || M <= 0 || N <= 0 || K <= 0) return fail(-62, "bfk_gemm: bad arguments");
  if (NT != 64 && NT != 128 && NT != 256) return fail(-63, "bfk_gemm: NT must be 64, 128 or 256");
  if ((N % NT) || (K % G_KC) || (ldc % 4)) return fail(-64, "bfk_gemm: need N % NT == 0, K % 32 == 0, ldc % 4 == 0");
  if (k_splits < 1) k_splits = 1;
  if (k_splits > 1 && !(flags & GEMM_ATOMIC)) return fail(-65, "bfk_gemm: k_splits > 1 needs the atomic flag");
  GemmArgs a{};
  a.A = (const unsigned char*)A; a.B = (const unsigned char*)B; a.C = Cm; a.bias = bias;
  a.M = M; a.ldc = ldc; a.N = N; a.K = K; a.NT = NT; a.flags = flags; a.alpha = alpha;
  int kps = (int)round_up64((K + k_splits - 1) / k_splits, G_KC);
  a.k_per_split = kps;
  k_splits = (K + kps - 1) / kps;
  a.idesc = umma_idesc_f16(G_TM, NT, precision == BFK_PREC_BF16 ? 1 : 0);
  static int use_persist = -1;
  if (use_persist < 0) { const char* e = getenv("BFK_GEMM_PERSIST"); use_persist = e ? atoi(e) : 1; }
  if (use_persist && k_splits == 1 && !(flags & GEMM_ATOMIC) && (K % P_KC) == 0) {
    // persistent, double-buffered-accumulator kernel: whenever there is no split-K
    static bool p_attr_dev[64] = {};
    bool& p_attr = *attr_flag(p_attr_dev);
    if (!p_attr) {
      CK(cudaFuncSetAttribute(gemm_persist_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)gemm_persist_smem_bytes(256)));
      p_attr = true;
    }
    const long long tiles = ((M + G_TM - 1) / G_TM) * (N / NT);
    int sms = bfk_num_sms();
    if (sms <= 0) sms = 148;
    const unsigned grid = (unsigned)(tiles < sms ? tiles : sms);
    gemm_persist_kernel<<<grid, P_THREADS, gemm_persist_smem_bytes(NT), (cudaStream_t)stream_>>>(a);
    CK(cudaGetLastError());
    return 0;
  }
  const size_t sh = gemm_smem_bytes(NT);
  static bool attr_set_dev[64] = {};
  bool& attr_set = *attr_flag(attr_set_dev);
  if (!attr_set) {
    CK(cudaFuncSetAttribute(gemm_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)gemm_smem_bytes(256)));
    attr_set = true;
  }
  dim3 grid((unsigned)((M + G_TM - 1) / G_TM), (unsigned)(N / NT), (unsigned)k_splits);
  gemm_kernel<false><<<grid, G_THREADS, sh, (cudaStream_t)stream_>>>(a);
  CK(cudaGetLastError());
  return 0;
}

#include "bfk_train_host.cuh"
#include "bfk_layerwise_host.cuh"
#include "bfk_edge.cuh"
#include "bfk_ot.cuh"

// ---- allocation-free entry point ---------------------------------------------------------------------------------------
// bfk_batch_moments_static: bfk_batch_moments with its scratch (the per-block partial moments) in a file-scope device
// array instead of a caller-provided workspace, for callers that cannot allocate (CUDA graph capture, serving loops,
// harnesses that forbid cudaMalloc).  No counterpart in the reference.
//  * g_moments_scratch lives in every GPU context that loads this library until the process exits
//    (BFK_STATIC_SCRATCH_FLOATS * 4 bytes = 2 MiB per device).
//  * In host code the array's name is only a host-side stand-in: the device address is taken once per device with
//    cudaGetSymbolAddress (that device current) and cached; kernels are handed that pointer.
//  * Nothing in the array has to start at zero: the partial kernel overwrites every slot (blocks * 3 * F floats) the finish
//    kernel reads, so stale contents of an earlier (larger) call are never consumed.
//  * All calls share the one array: calls that overlap on the GPU (different streams, no ordering) race.
#define BFK_STATIC_SCRATCH_FLOATS (1u << 19)
static __device__ __align__(256) float g_moments_scratch[BFK_STATIC_SCRATCH_FLOATS];

extern "C" size_t bfk_static_scratch_floats(void) { return (size_t)BFK_STATIC_SCRATCH_FLOATS; }

extern "C" int bfk_batch_moments_static(const float* x, int64_t rows, int32_t F, float* stats, void* stream) {
  if (rows < 0 || F < 1 || F > 1024) return fail(-65, "bfk_batch_moments_static: bad rows/F (F <= 1024)");
  if (!stats || (rows > 0 && !x)) return fail(-66, "bfk_batch_moments_static: null pointer");
  int dev = -1;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64)
    return fail(-62, "bfk_batch_moments_static: no CUDA device (or device ordinal >= 64)");
  const size_t need = bfk_moments_workspace_floats(rows, F);
  if (need == 0 || need > (size_t)BFK_STATIC_SCRATCH_FLOATS) {
    char msg[200];
    snprintf(msg, sizeof msg, "bfk_batch_moments_static: needs %zu scratch floats, the static array holds %u "
             "(use bfk_batch_moments with a caller workspace)", need, BFK_STATIC_SCRATCH_FLOATS);
    return fail(-69, msg);
  }
  static float* scratch_dev[64] = {};  // device address of g_moments_scratch, per device ordinal
  float* scratch = scratch_dev[dev];
  if (!scratch) {
    void* p = nullptr;
    const cudaError_t e = cudaGetSymbolAddress(&p, g_moments_scratch);
    if (e != cudaSuccess || !p) {
      char msg[200];
      snprintf(msg, sizeof msg, "bfk_batch_moments_static: cudaGetSymbolAddress failed: %s", cudaGetErrorString(e));
      return fail(-70, msg);
    }
    scratch = scratch_dev[dev] = (float*)p;
  }
  return bfk_batch_moments(x, rows, F, scratch, stats, stream);
}
