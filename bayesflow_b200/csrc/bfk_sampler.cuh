// The fused on-chip sampler: one persistent CTA per SM advances two 128-draw tiles through EVERY velocity evaluation
// of EVERY ODE step.  State, RK stages and activations never leave the SM; weights stream L2 -> SMEM through the
// TMA engine (cp.async.bulk); every Dense(W,W) is a tcgen05 UMMA (M=128, N=W, K=16) accumulating in TMEM; bias,
// activation, FiLM, residual, LayerNorm, the output projector and the RK update are the TMEM->register epilogue.
//
// Warp roles (320 threads): warps 0-3 epilogue of tile 0, warps 4-7 epilogue of tile 1 (thread == draw == TMEM lane),
// warp 8 UMMA issuer (one lane) + TMEM owner, warp 9 bulk-copy producer (one lane).
#pragma once
#include "bfk_prologue.cuh"

// Timing experiments (profiles/r01c_summary.md): compile with -DBFK_TIMING_EXPERIMENTS and set BFK_DEBUG=<bits> at run
// time to switch parts of the kernel off (results are then garbage).  Compiled out of the product build.
#ifdef BFK_TIMING_EXPERIMENTS
#define BFK_DBG(bits) ((a.dbg & (bits)) != 0)
#else
#define BFK_DBG(bits) false
#endif

namespace bfk {

struct SamplerArgs {
  const unsigned char* wg;  // packed GEMM weights [NG][W/8][W][8] 16-bit
  const float* cvec;        // constant vectors (SMEM image)
  const float* table;       // [n_evals][tstride]
  const float* step_h;      // [n_steps]
  const float* pre_add;     // [G][W] or null
  const float* post_add;    // [G][W] or null
  const float* x_in;        // [B][D]
  float* x_out;             // [B][D]
  const float* tmask;       // [D] or null
  long long B, rows_per_group, n_steps;
  int D, W, L, NG, has_merge, residual, norm, film_gamma;
  int S, mode;  // S = stages per step; mode 0 = integrate, 1 = write the velocity of the single evaluation,
                // 2 = one adaptive trial step (stage 0 = FSAL slope read from kbuf, stage 6 = evaluation at the candidate)
  int s0;       // first stage that is evaluated (1 in mode 2, else 0)
  int err_kind; // mode 2: 0 = (y5 - y4) with rk_e = 4th-order weights (rk45), 1 = h * sum e_j k_j (tsit5)
  float* ctrl;  // mode 2: device controller state (include/bfk.h BFK_CTRL_*)
  float* ybuf;  // mode 2: [2][B][D] state slots
  float* kbuf;  // mode 2: [2][B][D] FSAL slope slots
  float rk_e[BFK_MAX_STAGES];
  int dbg;      // BFK_DEBUG env (timing experiments only): bit0 = skip epilogue arithmetic, bit1 = skip pass 2
  int tstride, cvec_floats;
  int cv_wx, cv_bx, cv_blk, cv_wo, cv_bo;
  uint32_t idesc;
  SmemPlan plan;
  int* err_flag;
  float rk_a[BFK_MAX_STAGES][BFK_MAX_STAGES];  // stage i input = y + h * sum_j a[i][j] k_j
  float rk_b[BFK_MAX_STAGES];                  // y_new = y + h * sum_j b[j] k_j
};

template <bool BF16>
__device__ __forceinline__ uint32_t pack2(float a, float b) {
  if (BF16) {
    __nv_bfloat162 v = __floats2bfloat162_rn(a, b);
    return *reinterpret_cast<uint32_t*>(&v);
  } else {
    __half2 v = __floats2half2_rn(a, b);
    return *reinterpret_cast<uint32_t*>(&v);
  }
}
template <bool BF16>
__device__ __forceinline__ float2 unpack2(uint32_t u) {
  if (BF16) {
    return __bfloat1622float2(*reinterpret_cast<__nv_bfloat162*>(&u));
  } else {
    return __half22float2(*reinterpret_cast<__half2*>(&u));
  }
}

// barrier slots inside the `bars` region (8 bytes each)
struct Bars {
  uint32_t base;
  int n_stage;
  __device__ uint32_t w_full(int s) const { return base + 8 * s; }
  __device__ uint32_t w_empty(int s) const { return base + 8 * (8 + s); }
  __device__ uint32_t a_ready(int t) const { return base + 8 * (16 + t); }
  __device__ uint32_t acc_ready(int t) const { return base + 8 * (18 + t); }
  __device__ uint32_t film_full(int b) const { return base + 8 * (20 + b); }
  __device__ uint32_t film_empty(int b) const { return base + 8 * (22 + b); }
  __device__ uint32_t tmem_slot() const { return base + 8 * 24; }
};

// UMMA issuer (one warp; lane 0 issues): for every evaluation and GEMM layer, for each of the two tiles, wait for the
// tile's A operand, stream through the weight ring, commit to acc_ready.  Shared by the generic and the fast kernel.
__device__ __forceinline__ void mma_issuer_loop(const SamplerArgs& a, const Bars& bars, uint32_t smem_base,
                                                uint32_t tmem_base, long long n_pairs, long long n_evals, int lane) {
  const int W = a.W;
  const int n_chunks = W / KC;
  if (lane == 0) {
    uint32_t ph_a[2] = {0, 0};
    uint32_t ph_full = 0;  // bit s = parity of stage s
    int stage = 0;
    const uint32_t a_lbo = TILE_M * 16, b_lbo = (uint32_t)W * 16;
    for (long long pair = blockIdx.x; pair < n_pairs; pair += gridDim.x) {
      for (long long e = 0; e < n_evals; ++e) {
        for (int g = 0; g < a.NG; ++g) {
          for (int t = 0; t < TILES_PER_CTA; ++t) {
            mbar_wait(bars.a_ready(t), ph_a[t], a.err_flag, 3);
            ph_a[t] ^= 1;
            tc_fence_after();
            const uint32_t a_base = smem_base + a.plan.a_tile + t * a.plan.a_bytes;
            const uint32_t d_tmem = tmem_base + t * W;
            for (int c = 0; c < n_chunks; ++c) {
              mbar_wait(bars.w_full(stage), (ph_full >> stage) & 1, a.err_flag, 4);
              ph_full ^= (1u << stage);
              tc_fence_after();
              const uint32_t b_base = smem_base + a.plan.ring + stage * a.plan.stage_bytes;
#pragma unroll
              for (int kk = 0; kk < KC / 16; ++kk) {
                const uint64_t ad = umma_smem_desc(a_base + (uint32_t)(c * (KC / 8) + kk * 2) * a_lbo, a_lbo, 128);
                const uint64_t bd = umma_smem_desc(b_base + (uint32_t)(kk * 2) * b_lbo, b_lbo, 128);
                umma_f16(d_tmem, ad, bd, a.idesc, (uint32_t)((c | kk) != 0));
              }
              umma_commit(bars.w_empty(stage));
              stage = (stage + 1 == a.plan.n_stage) ? 0 : stage + 1;
            }
            umma_commit(bars.acc_ready(t));
          }
        }
      }
    }
  }
  __syncwarp();
}

// Bulk-copy producer (one warp; lane 0 drives the TMA engine): weight K-slabs into the ring, one table row per evaluation.
__device__ __forceinline__ void producer_loop(const SamplerArgs& a, const Bars& bars, uint32_t smem_base,
                                              long long n_pairs, long long n_evals, int lane) {
  const int W = a.W;
  const int n_chunks = W / KC;
  if (lane == 0) {
    uint32_t ph_empty = 0;
    int stage = 0;
    long long q = 0, q_film = 0;
    long long my_pairs = 0;
    for (long long pair = blockIdx.x; pair < n_pairs; pair += gridDim.x) ++my_pairs;
    const long long Q = my_pairs * n_evals;
    auto load_film = [&](long long qq) {
      const int fb = (int)(qq & 1);
      mbar_wait(bars.film_empty(fb), (uint32_t)(((qq >> 1) & 1) ^ 1), a.err_flag, 5);
      mbar_arrive_expect_tx(bars.film_full(fb), (uint32_t)a.tstride * 4);
      bulk_g2s(smem_base + a.plan.film + fb * a.plan.film_bytes, a.table + (qq % n_evals) * a.tstride,
               (uint32_t)a.tstride * 4, bars.film_full(fb));
    };
    if (Q > 0) { load_film(0); q_film = 1; }
    for (; q < Q; ++q) {
      for (int g = 0; g < a.NG; ++g) {
        for (int t = 0; t < TILES_PER_CTA; ++t) {
          for (int c = 0; c < n_chunks; ++c) {
            mbar_wait(bars.w_empty(stage), ((ph_empty >> stage) & 1) ^ 1, a.err_flag, 6);
            ph_empty ^= (1u << stage);
            if (BFK_DBG(4) && (q > 0 || g > 0 || t > 0)) {  // timing experiment: no weight streaming after the first fills
              mbar_arrive(bars.w_full(stage));
            } else {
              mbar_arrive_expect_tx(bars.w_full(stage), a.plan.stage_bytes);
              bulk_g2s(smem_base + a.plan.ring + stage * a.plan.stage_bytes,
                       a.wg + ((size_t)g * W * W + (size_t)c * KC * W) * 2, a.plan.stage_bytes, bars.w_full(stage));
            }
            stage = (stage + 1 == a.plan.n_stage) ? 0 : stage + 1;
          }
        }
        if (g == 0 && q_film < Q && q_film == q + 1) { load_film(q_film); ++q_film; }
      }
    }
  }
  __syncwarp();
}

template <int ACT, int DP, bool BF16>
__global__ void __launch_bounds__(NUM_THREADS, 1) sampler_kernel(const SamplerArgs a) {
  extern __shared__ __align__(128) unsigned char smem[];
  if (a.mode == 2 && a.ctrl[BFK_CTRL_SKIP] != 0.f) return;  // controller says: loop finished, nothing to do
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int W = a.W;
  const uint32_t smem_base = smem_u32(smem);
  Bars bars{smem_base + a.plan.bars, a.plan.n_stage};
  float* s_cvec = reinterpret_cast<float*>(smem + a.plan.cvec);
  volatile uint32_t* s_tmem = reinterpret_cast<volatile uint32_t*>(smem + a.plan.bars + 8 * 24);

  // ---------------- one-time setup
  for (int i = threadIdx.x; i < a.cvec_floats; i += NUM_THREADS) s_cvec[i] = a.cvec[i];
  if (threadIdx.x == 0) {
    for (int s = 0; s < a.plan.n_stage; ++s) { mbar_init(bars.w_full(s), 1); mbar_init(bars.w_empty(s), 1); }
    for (int t = 0; t < 2; ++t) {
      mbar_init(bars.a_ready(t), 4);     // one elected lane per epilogue warp of the tile
      mbar_init(bars.acc_ready(t), 1);   // tcgen05.commit
      mbar_init(bars.film_full(t), 1);   // producer arrive + tx bytes
      mbar_init(bars.film_empty(t), 8);  // all 8 epilogue warps
    }
    fence_barrier_init();
  }
  const uint32_t tmem_cols = (2 * W <= 32) ? 32 : (2 * W <= 64) ? 64 : (2 * W <= 128) ? 128 : (2 * W <= 256) ? 256 : 512;
  if (warp == MMA_WARP) tmem_alloc(bars.tmem_slot(), tmem_cols);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *s_tmem;

  const long long rows_per_pair = TILES_PER_CTA * TILE_M;
  const long long n_pairs = (a.B + rows_per_pair - 1) / rows_per_pair;
  const long long n_evals = a.n_steps * (a.S - a.s0);

  if (warp < EPI_WARPS) {
    // =========================================================== EPILOGUE / ODE warps
    const int tile = warp >> 2;
    const int r = (warp & 3) * 32 + lane;  // row inside the tile == TMEM lane
    unsigned char* sA = smem + a.plan.a_tile + tile * a.plan.a_bytes + r * 16;
    const uint32_t t_acc = tmem_base + ((uint32_t)((warp & 3) * 32) << 16) + tile * W;
    const float* c_wx = s_cvec + a.cv_wx;
    const float* c_bx = s_cvec + a.cv_bx;
    const float* c_blk = s_cvec + a.cv_blk;
    const float* c_wo = s_cvec + a.cv_wo;
    const float* c_bo = s_cvec + a.cv_bo;
    const float inv_w = 1.f / (float)W;
    uint32_t ph_acc = 0;
    long long q = 0;  // evaluation sequence number of this CTA (film double buffer)

    for (long long pair = blockIdx.x; pair < n_pairs; pair += gridDim.x) {
      const long long row = pair * rows_per_pair + tile * TILE_M + r;
      const bool valid = row < a.B;
      const long long grp = valid ? row / a.rows_per_group : 0;
      const float* g_pre = a.pre_add ? a.pre_add + grp * W : nullptr;
      const float* g_post = a.post_add ? a.post_add + grp * W : nullptr;
      float y[DP], xs[DP], kst[BFK_MAX_STAGES][DP], msk[DP], ynew[DP];
      const int sel = (a.mode == 2) ? (int)a.ctrl[BFK_CTRL_SEL] : 0;
      const float* y_src = (a.mode == 2) ? a.ybuf + (size_t)sel * a.B * a.D : a.x_in;
#pragma unroll
      for (int d = 0; d < DP; ++d) {
        y[d] = (valid && d < a.D) ? y_src[row * a.D + d] : 0.f;
        msk[d] = (a.tmask && d < a.D) ? a.tmask[d] : 1.f;
        ynew[d] = 0.f;
        if (a.mode == 2) kst[0][d] = (valid && d < a.D) ? a.kbuf[(size_t)sel * a.B * a.D + row * a.D + d] : 0.f;
      }

      for (long long step = 0; step < a.n_steps; ++step) {
        const float h = (a.mode == 2) ? a.ctrl[BFK_CTRL_H] : a.step_h[step];
        for (int s = a.s0; s < a.S; ++s, ++q) {
          // ---- stage input (integrate.py:47-54 add_scaled: acc = sum c_j k_j ; y + h*acc)
          if (s == 6) {  // FSAL evaluation at the candidate state (integrate.py:94-102, 184-198)
#pragma unroll
            for (int d = 0; d < DP; ++d) {
              float acc = 0.f;
              for (int j = 0; j < 6; ++j) acc = acc + a.rk_b[j] * kst[j][d];
              ynew[d] = y[d] + h * acc;
              xs[d] = ynew[d];
            }
          } else {
#pragma unroll
            for (int d = 0; d < DP; ++d) {
              float acc = 0.f;
              for (int j = 0; j < s; ++j) acc = acc + a.rk_a[s][j] * kst[j][d];
              xs[d] = (s == 0) ? y[d] : y[d] + h * acc;
            }
          }
          const int fb = (int)(q & 1);
          const float* s_film = reinterpret_cast<const float*>(smem + a.plan.film + fb * a.plan.film_bytes);
          mbar_wait(bars.film_full(fb), (uint32_t)((q >> 1) & 1), a.err_flag, 1);

          float pred[DP];
          // ---- x_proj (+ conditions for merge="add") + activation  -> A tile      time_mlp.py:131,137,140
          for (int j = 0; j < W / 8; ++j) {
            float v[8];
#pragma unroll
            for (int i = 0; i < 8; ++i) v[i] = c_bx[j * 8 + i];
            if (a.has_merge && g_pre) {
              const float4 p0 = __ldg(reinterpret_cast<const float4*>(g_pre + j * 8));
              const float4 p1 = __ldg(reinterpret_cast<const float4*>(g_pre + j * 8 + 4));
              v[0] += p0.x; v[1] += p0.y; v[2] += p0.z; v[3] += p0.w;
              v[4] += p1.x; v[5] += p1.y; v[6] += p1.z; v[7] += p1.w;
            }
#pragma unroll
            for (int d = 0; d < DP; ++d) {
#pragma unroll
              for (int i = 0; i < 8; ++i) v[i] = fmaf(xs[d], c_wx[d * W + j * 8 + i], v[i]);
            }
            uint4 pk;
            pk.x = pack2<BF16>(act_fn<ACT>(v[0]), act_fn<ACT>(v[1]));
            pk.y = pack2<BF16>(act_fn<ACT>(v[2]), act_fn<ACT>(v[3]));
            pk.z = pack2<BF16>(act_fn<ACT>(v[4]), act_fn<ACT>(v[5]));
            pk.w = pack2<BF16>(act_fn<ACT>(v[6]), act_fn<ACT>(v[7]));
            *reinterpret_cast<uint4*>(sA + j * (TILE_M * 16)) = pk;
          }
          fence_proxy_async_smem();
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(bars.a_ready(tile));

#pragma unroll
          for (int d = 0; d < DP; ++d) pred[d] = c_bo[d];

          for (int g = 0; g < a.NG; ++g) {
            mbar_wait(bars.acc_ready(tile), ph_acc, a.err_flag, 2);
            ph_acc ^= 1;
            tc_fence_after();
            const bool last = (g == a.NG - 1);
            if (a.has_merge && g == 0) {
              // ---- merge_proj epilogue: h = act(acc + [m_c + bm])                  time_mlp.py:139-140
              for (int c = 0; c < W / 32; ++c) {
                float v[32];
                tmem_ld32(t_acc + c * 32, v);
                tmem_ld_wait();
#pragma unroll
                for (int i = 0; i < 32; i += 4) {
                  const float4 p = __ldg(reinterpret_cast<const float4*>(g_post + c * 32 + i));
                  v[i] = act_fn<ACT>(v[i] + p.x); v[i + 1] = act_fn<ACT>(v[i + 1] + p.y);
                  v[i + 2] = act_fn<ACT>(v[i + 2] + p.z); v[i + 3] = act_fn<ACT>(v[i + 3] + p.w);
                }
#pragma unroll
                for (int jj = 0; jj < 4; ++jj) {
                  uint4 pk;
                  pk.x = pack2<BF16>(v[jj * 8 + 0], v[jj * 8 + 1]); pk.y = pack2<BF16>(v[jj * 8 + 2], v[jj * 8 + 3]);
                  pk.z = pack2<BF16>(v[jj * 8 + 4], v[jj * 8 + 5]); pk.w = pack2<BF16>(v[jj * 8 + 6], v[jj * 8 + 7]);
                  *reinterpret_cast<uint4*>(sA + (c * 4 + jj) * (TILE_M * 16)) = pk;
                }
              }
            } else {
              // ---- ConditionalDenseBlock epilogue                          conditional_dense_block.py:114-131
              const int l = g - a.has_merge;
              const float* bd = c_blk + (3 * l + 0) * W;
              const float* lng = c_blk + (3 * l + 1) * W;
              const float* lnb = c_blk + (3 * l + 2) * W;
              const float* fbeta = s_film + l * W;
              const float* fgam = s_film + (a.L + l) * W;
              float sum = 0.f, sumsq = 0.f;
              // pass 1: v = skip + film(act(acc + bd)); row statistics; v parked in TMEM (or finished if no norm)
              for (int c = 0; c < W / 32; ++c) {
                float v[32];
                tmem_ld32(t_acc + c * 32, v);
                tmem_ld_wait();
#pragma unroll
                for (int i = 0; i < 32; ++i) v[i] = act_fn<ACT>(v[i] + bd[c * 32 + i]);
                if (a.film_gamma) {
#pragma unroll
                  for (int i = 0; i < 32; ++i) v[i] = (1.f + fgam[c * 32 + i]) * v[i];
                }
#pragma unroll
                for (int i = 0; i < 32; ++i) v[i] += fbeta[c * 32 + i];
                if (a.residual) {
#pragma unroll
                  for (int jj = 0; jj < 4; ++jj) {
                    const uint4 sk = *reinterpret_cast<const uint4*>(sA + (c * 4 + jj) * (TILE_M * 16));
                    float2 f;
                    f = unpack2<BF16>(sk.x); v[jj * 8 + 0] += f.x; v[jj * 8 + 1] += f.y;
                    f = unpack2<BF16>(sk.y); v[jj * 8 + 2] += f.x; v[jj * 8 + 3] += f.y;
                    f = unpack2<BF16>(sk.z); v[jj * 8 + 4] += f.x; v[jj * 8 + 5] += f.y;
                    f = unpack2<BF16>(sk.w); v[jj * 8 + 6] += f.x; v[jj * 8 + 7] += f.y;
                  }
                }
                if (a.norm != BFK_NORM_NONE) {
#pragma unroll
                  for (int i = 0; i < 32; ++i) { sum += v[i]; sumsq = fmaf(v[i], v[i], sumsq); }
                  tmem_st32(t_acc + c * 32, v);
                } else if (last) {
#pragma unroll
                  for (int i = 0; i < 32; ++i) {
#pragma unroll
                    for (int d = 0; d < DP; ++d) pred[d] = fmaf(v[i], c_wo[(c * 32 + i) * DP + d], pred[d]);
                  }
                } else {
#pragma unroll
                  for (int jj = 0; jj < 4; ++jj) {
                    uint4 pk;
                    pk.x = pack2<BF16>(v[jj * 8 + 0], v[jj * 8 + 1]); pk.y = pack2<BF16>(v[jj * 8 + 2], v[jj * 8 + 3]);
                    pk.z = pack2<BF16>(v[jj * 8 + 4], v[jj * 8 + 5]); pk.w = pack2<BF16>(v[jj * 8 + 6], v[jj * 8 + 7]);
                    *reinterpret_cast<uint4*>(sA + (c * 4 + jj) * (TILE_M * 16)) = pk;
                  }
                }
              }
              if (a.norm != BFK_NORM_NONE) {
                tmem_st_wait();
                float mean = 0.f, rstd;
                if (a.norm == BFK_NORM_LAYER) {
                  mean = sum * inv_w;
                  const float var = fmaxf(sumsq * inv_w - mean * mean, 0.f);
                  rstd = rsqrtf(var + 1e-3f);  // keras LayerNormalization epsilon
                } else {
                  rstd = rsqrtf(sumsq * inv_w + 1e-6f);  // keras RMSNormalization epsilon
                }
                // pass 2: h = (v - mean) * rstd * gamma + beta  -> next A tile, or straight into the output projector
                for (int c = 0; c < W / 32; ++c) {
                  float v[32];
                  tmem_ld32(t_acc + c * 32, v);
                  tmem_ld_wait();
#pragma unroll
                  for (int i = 0; i < 32; ++i) v[i] = fmaf((v[i] - mean) * rstd, lng[c * 32 + i], lnb[c * 32 + i]);
                  if (last) {
#pragma unroll
                    for (int i = 0; i < 32; ++i) {
#pragma unroll
                      for (int d = 0; d < DP; ++d) pred[d] = fmaf(v[i], c_wo[(c * 32 + i) * DP + d], pred[d]);
                    }
                  } else {
#pragma unroll
                    for (int jj = 0; jj < 4; ++jj) {
                      uint4 pk;
                      pk.x = pack2<BF16>(v[jj * 8 + 0], v[jj * 8 + 1]); pk.y = pack2<BF16>(v[jj * 8 + 2], v[jj * 8 + 3]);
                      pk.z = pack2<BF16>(v[jj * 8 + 4], v[jj * 8 + 5]); pk.w = pack2<BF16>(v[jj * 8 + 6], v[jj * 8 + 7]);
                      *reinterpret_cast<uint4*>(sA + (c * 4 + jj) * (TILE_M * 16)) = pk;
                    }
                  }
                }
              }
            }
            if (!last) {
              fence_proxy_async_smem();
              tc_fence_before();
              __syncwarp();
              if (lane == 0) mbar_arrive(bars.a_ready(tile));
            }
          }
          // ---- velocity of this evaluation (FM: pred; DM: coefZ*z + coefP*pred), inference-time target mask
          const float coefZ = s_film[a.tstride - TABLE_SCALARS + 0], coefP = s_film[a.tstride - TABLE_SCALARS + 1];
#pragma unroll
          for (int d = 0; d < DP; ++d) kst[s][d] = (coefZ * xs[d] + coefP * pred[d]) * msk[d];
          __syncwarp();
          if (lane == 0) mbar_arrive(bars.film_empty(fb));
        }
        if (a.mode == 0) {
#pragma unroll
          for (int d = 0; d < DP; ++d) {
            float acc = 0.f;
            for (int j = 0; j < a.S; ++j) acc = acc + a.rk_b[j] * kst[j][d];
            y[d] = y[d] + h * acc;
          }
        }
      }
      if (a.mode == 2) {
        // ---- embedded error estimate, row-wise L2 norm, max over rows (integrate.py:104-118, 199-212); max|state|
        const float h = a.ctrl[BFK_CTRL_H];
        float e2 = 0.f, mabs = 0.f;
        bool nonnan = false;
#pragma unroll
        for (int d = 0; d < DP; ++d) {
          float acc = 0.f, e;
          for (int j = 0; j < 7; ++j) acc = acc + a.rk_e[j] * kst[j][d];
          if (a.err_kind == 0) e = ynew[d] - (y[d] + h * acc);
          else e = h * acc;
          if (d < a.D) {
            e2 = fmaf(e, e, e2);
            mabs = (y[d] != y[d]) ? y[d] : ((mabs != mabs) ? mabs : fmaxf(mabs, fabsf(y[d])));
            nonnan = nonnan || (ynew[d] == ynew[d]);
          }
        }
        float en = valid ? sqrtf(e2) : 0.f;
        if (!valid) { mabs = 0.f; nonnan = false; }
        bool isnan_row = (en != en) || (mabs != mabs);
        if (isnan_row) { en = 0.f; if (mabs != mabs) mabs = 0.f; }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
          en = fmaxf(en, __shfl_xor_sync(0xffffffffu, en, o));
          mabs = fmaxf(mabs, __shfl_xor_sync(0xffffffffu, mabs, o));
        }
        const bool any_nan = __any_sync(0xffffffffu, isnan_row), any_nonnan = __any_sync(0xffffffffu, nonnan);
        if (lane == 0) {
          int* acc_i = reinterpret_cast<int*>(a.ctrl + BFK_CTRL_ACC);
          atomicMax(acc_i + 0, __float_as_int(en));
          atomicMax(acc_i + 1, __float_as_int(mabs));
          if (any_nan) atomicMax(acc_i + 2, __float_as_int(1.f));
          if (any_nonnan) atomicMax(acc_i + 3, __float_as_int(1.f));
        }
        if (valid) {
          const size_t o = (size_t)(1 - sel) * a.B * a.D + row * a.D;
          for (int d = 0; d < a.D; ++d) { a.ybuf[o + d] = ynew[d]; a.kbuf[o + d] = kst[6][d]; }
        }
      } else if (valid) {
        for (int d = 0; d < a.D; ++d) a.x_out[row * a.D + d] = (a.mode == 1) ? kst[0][d] : y[d];
      }
    }
    tc_fence_before();
  } else if (warp == MMA_WARP) {
    mma_issuer_loop(a, bars, smem_base, tmem_base, n_pairs, n_evals, lane);
  } else {
    producer_loop(a, bars, smem_base, n_pairs, n_evals, lane);
  }

  __syncthreads();
  tc_fence_after();
  if (warp == MMA_WARP) tmem_dealloc(tmem_base, tmem_cols);
}

}  // namespace bfk
