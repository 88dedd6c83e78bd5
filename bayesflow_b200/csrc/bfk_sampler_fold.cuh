// The fused on-chip sampler for the reference's default block shape (LayerNorm + residual + FiLM shift,
// networks/defaults.py:13-26; time_mlp.py:125-147, conditional_dense_block.py:114-131): two 128-draw tiles per
// persistent CTA ping-pong through 2 x W TMEM columns (while one tile is in its epilogue the tensor pipe runs the other
// tile's layer), and every epilogue is ONE pass over the accumulator.
//
// Protocol as in bfk_sampler.cuh (16 epilogue warps: tile = (w>>2)&1, TMEM lane quarter = w&3, column half = w>>3; warp
// 16 UMMA issuer + TMEM owner; warp 17 bulk-copy producer).  What makes the single pass possible is that LayerNorm is
// folded into its consumer.  With v the un-normalised output of a block and (mu, r) its row statistics,
//     log2(e) (LN(v) W + b) = log2(e) r [ v (gamma o W) - mu g + cL / (log2(e) r) ]
// (g_j = sum_k fp16(gamma_k W_kj), cL = log2(e) (b + beta W): bfk_pack / fold_pack_kernel), and the whole bracket is
// accumulated by the tensor core: the A operand is the raw fp16 image of v plus XK = 16 extra K columns per row holding
// (mu, 1/(log2(e) r)) as fp16 hi/lo pairs, against XK extra weight rows holding -g and cL (one more UMMA per layer).
// Per block and element the epilogue is then
//     t  = (log2(e) r) * acc                                   pre-activation times log2(e)
//     v' = t * actw(t) + [gamma (v r - r mu) + (beta + film(t))]      act(u) = t * actw(t); skip h = LN(v) rebuilt from
//                                                                     the same fp16 image; beta rides in the FiLM table
//     row sum / sum of squares of v' in fp32; fp16(v') -> A tile
// with only two per-column vectors (gamma, beta + film) read from shared memory per element instead of four, no
// second pass over TMEM, and no multiply in front of ex2.
#pragma once
#include "bfk_fast_epilogue.cuh"
#include "bfk_sampler.cuh"

// -DBFK_PROFILE: per-warp cycle accounting into a device buffer (read with bfk_debug_read_prof); compiled out otherwise
#ifdef BFK_PROFILE
__device__ long long g_bfk_prof[148 * 18 * 8];
#define PROF_DECL long long pf_[8] = {0, 0, 0, 0, 0, 0, 0, 0}; long long pt_ = clock64();
#define PROF_ADD(slot) { const long long n_ = clock64(); pf_[slot] += n_ - pt_; pt_ = n_; }
#define PROF_FLUSH(warp) if ((threadIdx.x & 31) == 0) { for (int i_ = 0; i_ < 8; ++i_) g_bfk_prof[(blockIdx.x * 18 + (warp)) * 8 + i_] = pf_[i_]; }
#else
#define PROF_DECL
#define PROF_ADD(slot)
#define PROF_FLUSH(warp)
#endif

namespace bfk {

// Pre-activations are carried times log2(e) (bfk_pack scales the weights / biases): with t = x log2(e),
// act(x) = t * actw(t), actw(t) = ln2 * tanh(softplus(x)) = ln2 - 2 ln2 / ((1 + 2^t)^2 + 1) for mish, ln2 / (1 + 2^-t) for
// silu: the multiply by log2(e) in front of ex2 disappears.  All MUFU of the chunk issued back to back.
template <int ACT>
__device__ __forceinline__ void actw2x8(const float2* t, float2* w) {
  constexpr bool ANY_ALT = (BFK_RCP_NEWTON_MASK | BFK_EX2_POLY_MASK) != 0;
  constexpr float SGN = ACT == BFK_ACT_MISH ? 1.f : -1.f;  // mish: 2^t, silu: 2^-t
  // exponent argument, clamped where a pair leaves the MUFU (inf would poison the Newton steps / the exponent add)
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    float2 a = make_float2(SGN * t[i].x, SGN * t[i].y);
    if (ANY_ALT && (((BFK_RCP_NEWTON_MASK | BFK_EX2_POLY_MASK) >> i) & 1))
      a = make_float2(fminf(fmaxf(a.x, -100.f), 40.f), fminf(fmaxf(a.y, -100.f), 40.f));
    if ((BFK_EX2_POLY_MASK >> i) & 1) w[i] = ex2_poly2(a);
    else { w[i].x = ex2_approx(a.x); w[i].y = ex2_approx(a.y); }
  }
  if (ACT == BFK_ACT_MISH) {
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const float2 q = __fadd2_rn(w[i], f2(1.f));
      w[i] = __ffma2_rn(q, q, f2(1.f));  // d = (1 + 2^t)^2 + 1
    }
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      if ((BFK_RCP_NEWTON_MASK >> i) & 1) {
        // ln2 * (d - 2) / d: no cancellation against ln2, so the 7e-6 of the Newton reciprocal stays relative
        const float2 n = __ffma2_rn(w[i], f2(0.6931471805599453f), f2(-1.3862943611198906f));
        w[i] = __fmul2_rn(n, rcp_newton2(w[i]));
      } else {
        w[i].x = rcp_approx(w[i].x); w[i].y = rcp_approx(w[i].y);
        w[i] = __ffma2_rn(w[i], f2(-1.3862943611198906f), f2(0.6931471805599453f));
      }
    }
  } else {  // silu
#pragma unroll
    for (int i = 0; i < 8; ++i) w[i] = __fadd2_rn(w[i], f2(1.f));
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      if ((BFK_RCP_NEWTON_MASK >> i) & 1) w[i] = rcp_newton2(w[i]);
      else { w[i].x = rcp_approx(w[i].x); w[i].y = rcp_approx(w[i].y); }
    }
#pragma unroll
    for (int i = 0; i < 8; ++i) w[i] = __fmul2_rn(w[i], f2(0.6931471805599453f));
  }
}
// t (pre-activation times log2(e)) -> act(x), 16 values
template <int ACT>
__device__ __forceinline__ void act_scaled2x8(float2* t) {
  if (ACT == BFK_ACT_MISH || ACT == BFK_ACT_SILU) {
    float2 w[8];
    actw2x8<ACT>(t, w);
#pragma unroll
    for (int i = 0; i < 8; ++i) t[i] = __fmul2_rn(t[i], w[i]);
  } else {
#pragma unroll
    for (int i = 0; i < 8; ++i)
      t[i] = make_float2(act_fn<ACT>(t[i].x * 0.6931471805599453f), act_fn<ACT>(t[i].y * 0.6931471805599453f));
  }
}

__device__ __forceinline__ void ld4x2(const float* p, float2* o) {  // 8 floats -> 4 float2
#if (BFK_EXP & 32)
  o[0] = o[1] = o[2] = o[3] = make_float2(0.001f, 0.002f);  // experiment: no shared-memory vector loads
  return;
#endif
  const float4 a = *reinterpret_cast<const float4*>(p), b = *reinterpret_cast<const float4*>(p + 4);
  o[0] = make_float2(a.x, a.y); o[1] = make_float2(a.z, a.w); o[2] = make_float2(b.x, b.y); o[3] = make_float2(b.z, b.w);
}

template <bool BF16>
__device__ __forceinline__ void store_chunk16(const float2* v, int col, unsigned char* sA_row) {
#pragma unroll
  for (int jj = 0; jj < 2; ++jj)
    *reinterpret_cast<uint4*>(sA_row + (size_t)((col >> 3) + jj) * (TILE_M * 16)) =
        make_uint4(pack2f<BF16>(v[jj * 4 + 0]), pack2f<BF16>(v[jj * 4 + 1]), pack2f<BF16>(v[jj * 4 + 2]),
                   pack2f<BF16>(v[jj * 4 + 3]));
}

// One 16-column chunk of a ConditionalDenseBlock epilogue in the folded form (see the header).  v: 8 float2 accumulators
// in, block output v' out.  rs = r, nm = -r*mu of the INPUT rows (1, 0 for the first block); rsL = log2(e) r.
template <int ACT, bool BF16>
__device__ __forceinline__ void fold_block_chunk(float2* v, int col, float2 rs, float2 nm, float2 rsL,
                                                 const float* __restrict__ gam, const float* __restrict__ fbeta,
                                                 const unsigned char* sA_row, float2* sum2, float2* sq2) {
#pragma unroll
  for (int i = 0; i < 8; ++i) v[i] = __fmul2_rn(v[i], rsL);
  float2 w[8];  // v = t = u log2(e);  act(u) = t * w
  if (ACT == BFK_ACT_MISH || ACT == BFK_ACT_SILU) {
    actw2x8<ACT>(v, w);
  } else {
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      v[i] = make_float2(act_fn<ACT>(v[i].x * 0.6931471805599453f), act_fn<ACT>(v[i].y * 0.6931471805599453f));
      w[i] = f2(1.f);
    }
  }
#pragma unroll
  for (int jj = 0; jj < 2; ++jj) {  // 8 columns = one 16-byte K chunk of the A tile
    const uint4 sk = *reinterpret_cast<const uint4*>(sA_row + (size_t)((col >> 3) + jj) * (TILE_M * 16));
    const uint32_t su[4] = {sk.x, sk.y, sk.z, sk.w};
    float2 ga[4], fb[4];
    ld4x2(gam + col + jj * 8, ga);
    ld4x2(fbeta + col + jj * 8, fb);
#pragma unroll
    for (int p = 0; p < 4; ++p) {
      const int i = jj * 4 + p;
      const float2 xh = __ffma2_rn(unpack2f<BF16>(su[p]), rs, nm);
      const float2 h = __ffma2_rn(xh, ga[p], fb[p]);
      const float2 u = __ffma2_rn(v[i], w[i], h);
      sum2[p] = __fadd2_rn(sum2[p], u);
      sq2[p] = __ffma2_rn(u, u, sq2[p]);
      v[i] = u;
    }
  }
}

// per-row extra K entries of the A tile: (mu_hi, mu_lo, mu_hi, ir_hi, ir_lo, ir_hi, 0, 0), ir = 1 / (log2(e) r)
template <bool BF16>
__device__ __forceinline__ uint4 fold_extras(float mu, float ir) {
  float mh, ih, dummy;
  uint32_t m_hi, m_lo, i_hi, i_lo;
  if (BF16) {
    __nv_bfloat16 a = __float2bfloat16_rn(mu); mh = __bfloat162float(a); __nv_bfloat16 b = __float2bfloat16_rn(mu - mh);
    __nv_bfloat16 c = __float2bfloat16_rn(ir); ih = __bfloat162float(c); __nv_bfloat16 d = __float2bfloat16_rn(ir - ih);
    m_hi = *reinterpret_cast<uint16_t*>(&a); m_lo = *reinterpret_cast<uint16_t*>(&b);
    i_hi = *reinterpret_cast<uint16_t*>(&c); i_lo = *reinterpret_cast<uint16_t*>(&d);
  } else {
    __half a = __float2half_rn(mu); mh = __half2float(a); __half b = __float2half_rn(mu - mh);
    __half c = __float2half_rn(ir); ih = __half2float(c); __half d = __float2half_rn(ir - ih);
    m_hi = *reinterpret_cast<uint16_t*>(&a); m_lo = *reinterpret_cast<uint16_t*>(&b);
    i_hi = *reinterpret_cast<uint16_t*>(&c); i_lo = *reinterpret_cast<uint16_t*>(&d);
  }
  (void)dummy;
  return make_uint4(m_hi | (m_lo << 16), m_hi | (i_hi << 16), i_lo | (i_hi << 16), 0u);
}

// UMMA issuer: for every evaluation and GEMM layer, for each of the two tiles, wait for the tile's A operand, stream the
// W / FKC weight stages + the XK-row tail stage through the ring, commit to acc_ready.
__device__ __forceinline__ void fold_issuer_loop(const SamplerArgs& a, const Bars& bars, uint32_t smem_base,
                                                 uint32_t tmem_base, long long my_tiles, long long n_evals, int lane) {
  const int W = a.W;
  const int n_chunks = W / FKC;
  if (lane == 0) {
    uint32_t ph_a[2] = {0, 0};
    uint32_t ph_full = 0;  // bit s = parity of stage s
    int stage = 0;
    const uint32_t a_lbo = TILE_M * 16, b_lbo = (uint32_t)W * 16;
    const uint32_t z_base = smem_base + a.plan.a_tile + TILES_PER_CTA * a.plan.a_bytes;  // shared zero K-chunk
    for (long long j = 0; 2 * j < my_tiles; ++j) {
      const int nt = (my_tiles - 2 * j) >= 2 ? 2 : 1;  // the last round of a CTA may hold a lone tile
      for (long long e = 0; e < n_evals; ++e) {
        for (int g = 0; g < a.NG; ++g) {
          for (int t = 0; t < nt; ++t) {
            mbar_wait(bars.a_ready(t), ph_a[t], a.err_flag, 3);
            ph_a[t] ^= 1;
            tc_fence_after();
            const uint32_t a_base = smem_base + a.plan.a_tile + t * a.plan.a_bytes;
            const uint32_t d_tmem = tmem_base + t * W;
            for (int c = 0; c <= n_chunks; ++c) {
              mbar_wait(bars.w_full(stage), (ph_full >> stage) & 1, a.err_flag, 4);
              ph_full ^= (1u << stage);
              tc_fence_after();
              const uint32_t b_base = smem_base + a.plan.ring + stage * a.plan.stage_bytes;
              if (c < n_chunks) {
#pragma unroll
                for (int kk = 0; kk < FKC / 16; ++kk) {
                  const uint64_t ad = umma_smem_desc(a_base + (uint32_t)(c * (FKC / 8) + kk * 2) * a_lbo, a_lbo, 128);
                  const uint64_t bd = umma_smem_desc(b_base + (uint32_t)(kk * 2) * b_lbo, b_lbo, 128);
                  umma_f16(d_tmem, ad, bd, a.idesc, (uint32_t)((c | kk) != 0));
                }
              } else {
                // XK tail: K-chunk W/8 of the tile holds (mu, 1/r); its second K-chunk is the shared all-zero chunk
                const uint32_t x_base = a_base + (uint32_t)(W / 8) * a_lbo;
                const uint64_t ad = umma_smem_desc(x_base, z_base - x_base, 128);
                const uint64_t bd = umma_smem_desc(b_base, b_lbo, 128);
                umma_f16(d_tmem, ad, bd, a.idesc, 1u);
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

// Bulk-copy producer: per tile and layer W / FKC full stages + the XK-row tail (the layer is one contiguous byte range)
__device__ __forceinline__ void fold_producer_loop(const SamplerArgs& a, const Bars& bars, uint32_t smem_base,
                                                   long long my_tiles, long long n_evals, int lane) {
  const int W = a.W;
  const int n_chunks = W / FKC;
  if (lane == 0) {
    uint32_t ph_empty = 0;
    int stage = 0;
    long long q = 0, q_film = 0;
    const long long Q = ((my_tiles + 1) / 2) * n_evals;
    const size_t layer_bytes = (size_t)(W + XK) * W * 2;
    auto load_film = [&](long long qq) {
      const int fb = (int)(qq & 1);
      mbar_wait(bars.film_empty(fb), (uint32_t)(((qq >> 1) & 1) ^ 1), a.err_flag, 5);
      mbar_arrive_expect_tx(bars.film_full(fb), (uint32_t)a.tstride * 4);
      bulk_g2s(smem_base + a.plan.film + fb * a.plan.film_bytes, a.table + (qq % n_evals) * a.tstride,
               (uint32_t)a.tstride * 4, bars.film_full(fb));
    };
    if (Q > 0) { load_film(0); q_film = 1; }
    for (; q < Q; ++q) {
      const int nt = (my_tiles - 2 * (q / n_evals)) >= 2 ? 2 : 1;
      for (int g = 0; g < a.NG; ++g) {
        for (int t = 0; t < nt; ++t) {
          for (int c = 0; c <= n_chunks; ++c) {
            mbar_wait(bars.w_empty(stage), ((ph_empty >> stage) & 1) ^ 1, a.err_flag, 6);
            ph_empty ^= (1u << stage);
#if (BFK_EXP & 64)
            const uint32_t bytes = 128;  // experiment: (almost) no weight streaming, stale operands
#else
            const uint32_t bytes = (c < n_chunks) ? a.plan.stage_bytes : (uint32_t)(XK * W * 2);
#endif
            mbar_arrive_expect_tx(bars.w_full(stage), bytes);
            bulk_g2s(smem_base + a.plan.ring + stage * a.plan.stage_bytes,
                     a.wg + (size_t)g * layer_bytes + (size_t)c * FKC * W * 2, bytes, bars.w_full(stage));
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
__global__ void __launch_bounds__(F_NUM_THREADS, 1) sampler_fold_kernel(const SamplerArgs a) {
  extern __shared__ __align__(128) unsigned char smem[];
  if (a.mode == 2 && a.ctrl[BFK_CTRL_SKIP] != 0.f) return;  // controller says: loop finished, nothing to do
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int W = a.W;
  const uint32_t smem_base = smem_u32(smem);
  Bars bars{smem_base + a.plan.bars, a.plan.n_stage};
  float* s_cvec = reinterpret_cast<float*>(smem + a.plan.cvec);
  volatile uint32_t* s_tmem = reinterpret_cast<volatile uint32_t*>(smem + a.plan.bars + 8 * 24);

  // ---------------- one-time setup
  for (int i = threadIdx.x; i < a.cvec_floats; i += F_NUM_THREADS) s_cvec[i] = a.cvec[i];
  if (threadIdx.x < TILE_M)  // the all-zero K-chunk both tiles' XK tails point at
    *reinterpret_cast<uint4*>(smem + a.plan.a_tile + TILES_PER_CTA * a.plan.a_bytes + threadIdx.x * 16) = make_uint4(0u, 0u, 0u, 0u);
  fence_proxy_async_smem();
  if (threadIdx.x == 0) {
    for (int s = 0; s < a.plan.n_stage; ++s) { mbar_init(bars.w_full(s), 1); mbar_init(bars.w_empty(s), 1); }
    for (int t = 0; t < 2; ++t) {
      mbar_init(bars.a_ready(t), 8);      // one elected lane per epilogue warp of the tile (4 quarters x 2 halves)
      mbar_init(bars.acc_ready(t), 1);    // tcgen05.commit
      mbar_init(bars.film_full(t), 1);    // producer arrive + tx bytes
      mbar_init(bars.film_empty(t), 16);  // all 16 epilogue warps
    }
    fence_barrier_init();
  }
  const uint32_t tmem_cols = (2 * W <= 32) ? 32 : (2 * W <= 64) ? 64 : (2 * W <= 128) ? 128 : (2 * W <= 256) ? 256 : 512;
  if (warp == F_MMA_WARP) tmem_alloc(bars.tmem_slot(), tmem_cols);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *s_tmem;

  // 128-row tiles are dealt round-robin to the CTAs (tile ids c, c + G, c + 2G, ...) and taken two at a time, so CTA
  // loads differ by at most ONE tile; a lone last tile runs without its ping-pong partner
  const long long n_tiles = (a.B + TILE_M - 1) / TILE_M;
  const long long my_tiles = (long long)blockIdx.x < n_tiles ? (n_tiles - blockIdx.x + gridDim.x - 1) / gridDim.x : 0;
  const long long n_evals = a.n_steps * (a.S - a.s0);

  if (warp < F_EPI_WARPS) {
    // =========================================================== EPILOGUE / ODE warps
#if BFK_EPI_REGS
    asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;" ::"n"(BFK_EPI_REGS));
#endif
    const int tile = (warp >> 2) & 1, quarter = warp & 3, half = warp >> 3;
    const int r = quarter * 32 + lane;  // row inside the tile == TMEM lane
    const int nch = W / 32, ch0 = half * nch;  // this warp's 16-column chunks: ch0 .. ch0 + nch - 1
    unsigned char* sA = smem + a.plan.a_tile + tile * a.plan.a_bytes + r * 16;
    const uint32_t t_acc = tmem_base + ((uint32_t)(quarter * 32) << 16) + tile * W;
    const int bar_id = 1 + tile * 4 + quarter;  // the two column halves of the same 32 rows
    // exchange between the two column halves: [tile 2][half 2][TILE_M] (sum, sumsq); [tile][half][TILE_M][DP]
    float2* xst = reinterpret_cast<float2*>(smem + a.plan.xch);
    float* xpr = reinterpret_cast<float*>(smem + a.plan.xch + 2 * 2 * TILE_M * 8);
    const float* c_wx = s_cvec + a.cv_wx;
    const float* c_bx = s_cvec + a.cv_bx;
    const float* c_blk = s_cvec + a.cv_blk;  // per block: gamma_{l-1}
    const float* c_wo = s_cvec + a.cv_wo;    // [DP][W] gamma_L * Wo
    const float* c_bo = s_cvec + a.cv_bo;    // s0[DP], s1[DP]
    const float inv_w = 1.f / (float)W;
    PROF_DECL
    uint32_t ph_acc = 0;
    unsigned char* sX = sA + (size_t)(W >> 3) * (TILE_M * 16);  // my row's extra K entries (K-chunk W/8)
    long long q = 0;   // evaluation sequence number of this CTA (film double buffer)

    for (long long j = 0; 2 * j < my_tiles; ++j) {
      if (2 * j + tile >= my_tiles) {
        // lone last tile of this CTA: my tile slot is empty; keep the per-evaluation table hand-shake going
        for (long long e = 0; e < n_evals; ++e, ++q) {
          const int fb = (int)(q & 1);
          mbar_wait(bars.film_full(fb), (uint32_t)((q >> 1) & 1), a.err_flag, 1);
          __syncwarp();
          if (lane == 0) mbar_arrive(bars.film_empty(fb));
        }
        continue;
      }
      const long long row = ((long long)blockIdx.x + (2 * j + tile) * gridDim.x) * TILE_M + r;
      const bool valid = row < a.B;
      const long long grp = valid ? row / a.rows_per_group : 0;
      const float* g_pre = a.pre_add ? a.pre_add + grp * W : nullptr;
      const float* g_post = a.post_add ? a.post_add + grp * W : nullptr;
      // both column halves carry the (identical) ODE state of the row; half 0 writes results
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

      float h_next = (a.mode == 2) ? a.ctrl[BFK_CTRL_H] : a.step_h[0];
      for (long long step = 0; step < a.n_steps; ++step) {
        const float h = h_next;
        if (a.mode != 2 && step + 1 < a.n_steps) h_next = a.step_h[step + 1];  // in flight under the whole step
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
          PROF_ADD(0)
          mbar_wait(bars.film_full(fb), (uint32_t)((q >> 1) & 1), a.err_flag, 1);
          PROF_ADD(1)

          // ---- x_proj (+ conditions for merge="add") + activation -> my half of the A tile   time_mlp.py:131,137,140
          {
            float2 xb[DP];
#pragma unroll
            for (int d = 0; d < DP; ++d) xb[d] = f2(xs[d]);
            for (int ch = ch0; ch < ch0 + nch; ++ch) {
              const int col = ch * 16;
              float2 v8[8];
              ld4x2(c_bx + col, v8);
              ld4x2(c_bx + col + 8, v8 + 4);
              if (a.has_merge && g_pre) {
#pragma unroll
                for (int jj = 0; jj < 2; ++jj) {
                  const float4 p0 = __ldg(reinterpret_cast<const float4*>(g_pre + col + jj * 8));
                  const float4 p1 = __ldg(reinterpret_cast<const float4*>(g_pre + col + jj * 8 + 4));
                  float2* v4 = v8 + jj * 4;
                  v4[0] = __fadd2_rn(v4[0], make_float2(p0.x, p0.y)); v4[1] = __fadd2_rn(v4[1], make_float2(p0.z, p0.w));
                  v4[2] = __fadd2_rn(v4[2], make_float2(p1.x, p1.y)); v4[3] = __fadd2_rn(v4[3], make_float2(p1.z, p1.w));
                }
              }
#pragma unroll
              for (int d = 0; d < DP; ++d) {
                float2 w8[8];
                ld4x2(c_wx + d * W + col, w8);
                ld4x2(c_wx + d * W + col + 8, w8 + 4);
#pragma unroll
                for (int i = 0; i < 8; ++i) v8[i] = __ffma2_rn(xb[d], w8[i], v8[i]);
              }
              act_scaled2x8<ACT>(v8);
              store_chunk16<BF16>(v8, col, sA);
            }
            // the operand of the first GEMMs is not normalised: mu = 0, 1/(log2(e) r) = ln 2
            if (half == 0) *reinterpret_cast<uint4*>(sX) = fold_extras<BF16>(0.f, 0.6931471805599453f);
          }
          PROF_ADD(2)
          fence_proxy_async_smem();
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(bars.a_ready(tile));
          PROF_ADD(7)

          float pred[DP];
          // row scale / shift of the current A image (identity before block 1); rsL = log2(e) r
          float2 rs = f2(1.f), nm = f2(0.f), rsL = f2(1.4426950408889634f);
          for (int g = 0; g < a.NG; ++g) {
            mbar_wait(bars.acc_ready(tile), ph_acc, a.err_flag, 2);
            PROF_ADD(3)
            ph_acc ^= 1;
            tc_fence_after();
            PROF_ADD(6)
            const bool last = (g == a.NG - 1);
            float2 va[8], vb[8];
            if (a.has_merge && g == 0) {
              // ---- merge_proj epilogue: h = act(acc + [m_c + bm])                      time_mlp.py:139-140
              auto merge_chunk = [&](float2* v, int ch) {
                const int col = ch * 16;
#pragma unroll
                for (int jj = 0; jj < 2; ++jj) {
                  const float4 p0 = __ldg(reinterpret_cast<const float4*>(g_post + col + jj * 8));
                  const float4 p1 = __ldg(reinterpret_cast<const float4*>(g_post + col + jj * 8 + 4));
                  v[jj * 4 + 0] = __fadd2_rn(v[jj * 4 + 0], make_float2(p0.x, p0.y));
                  v[jj * 4 + 1] = __fadd2_rn(v[jj * 4 + 1], make_float2(p0.z, p0.w));
                  v[jj * 4 + 2] = __fadd2_rn(v[jj * 4 + 2], make_float2(p1.x, p1.y));
                  v[jj * 4 + 3] = __fadd2_rn(v[jj * 4 + 3], make_float2(p1.z, p1.w));
                }
                act_scaled2x8<ACT>(v);
                store_chunk16<BF16>(v, col, sA);
              };
              tmem_ld16(t_acc + ch0 * 16, reinterpret_cast<float*>(va));
              for (int c = 0; c < nch; c += 2) {
                tmem_ld_wait16(reinterpret_cast<float*>(va));
                if (c + 1 < nch) tmem_ld16(t_acc + (ch0 + c + 1) * 16, reinterpret_cast<float*>(vb));
                merge_chunk(va, ch0 + c);
                if (c + 1 < nch) {
                  tmem_ld_wait16(reinterpret_cast<float*>(vb));
                  if (c + 2 < nch) tmem_ld16(t_acc + (ch0 + c + 2) * 16, reinterpret_cast<float*>(va));
                  merge_chunk(vb, ch0 + c + 1);
                }
              }
              PROF_ADD(4)
            } else {
              // ---- ConditionalDenseBlock epilogue, folded LayerNorm, single pass     conditional_dense_block.py:114-131
              const int l = g - a.has_merge;
              const float* gam = c_blk + l * W;
              const float* fbeta = s_film + l * W;
              float2 sum2[4], sq2[4], pa[DP];
#pragma unroll
              for (int p = 0; p < 4; ++p) { sum2[p] = f2(0.f); sq2[p] = f2(0.f); }
#pragma unroll
              for (int d = 0; d < DP; ++d) pa[d] = f2(0.f);
              auto blk = [&](float2* v, int ch) {
                const int col = ch * 16;
                fold_block_chunk<ACT, BF16>(v, col, rs, nm, rsL, gam, fbeta, sA, sum2, sq2);
                if (last) {
#pragma unroll
                  for (int d = 0; d < DP; ++d) {
                    float2 w8[8];
                    ld4x2(c_wo + d * W + col, w8);
                    ld4x2(c_wo + d * W + col + 8, w8 + 4);
                    float2 t0 = __fmul2_rn(v[0], w8[0]), t1 = __fmul2_rn(v[1], w8[1]);
                    t0 = __ffma2_rn(v[2], w8[2], t0); t1 = __ffma2_rn(v[3], w8[3], t1);
                    t0 = __ffma2_rn(v[4], w8[4], t0); t1 = __ffma2_rn(v[5], w8[5], t1);
                    t0 = __ffma2_rn(v[6], w8[6], t0); t1 = __ffma2_rn(v[7], w8[7], t1);
                    pa[d] = __fadd2_rn(pa[d], __fadd2_rn(t0, t1));
                  }
                } else {
                  store_chunk16<BF16>(v, col, sA);
                }
              };
              tmem_ld16(t_acc + ch0 * 16, reinterpret_cast<float*>(va));
              for (int c = 0; c < nch; c += 2) {
                tmem_ld_wait16(reinterpret_cast<float*>(va));
                if (c + 1 < nch) tmem_ld16(t_acc + (ch0 + c + 1) * 16, reinterpret_cast<float*>(vb));
                blk(va, ch0 + c);
                if (c + 1 < nch) {
                  tmem_ld_wait16(reinterpret_cast<float*>(vb));
                  if (c + 2 < nch) tmem_ld16(t_acc + (ch0 + c + 2) * 16, reinterpret_cast<float*>(va));
                  blk(vb, ch0 + c + 1);
                }
              }
              // ---- row statistics of v': my half's partial sums + the other half's (shared memory, 64-thread barrier)
              const float2 s01 = __fadd2_rn(__fadd2_rn(sum2[0], sum2[1]), __fadd2_rn(sum2[2], sum2[3]));
              const float2 q01 = __fadd2_rn(__fadd2_rn(sq2[0], sq2[1]), __fadd2_rn(sq2[2], sq2[3]));
              // (single buffer: my partner has read these before it announced the A tile of the previous layer)
              float2* xb2 = xst + (size_t)(tile * 2) * TILE_M;
              xb2[(size_t)half * TILE_M + r] = make_float2(s01.x + s01.y, q01.x + q01.y);
              float* xp = xpr + (size_t)(tile * 2) * TILE_M * DP;
              if (last) {
#pragma unroll
                for (int d = 0; d < DP; ++d) xp[((size_t)half * TILE_M + r) * DP + d] = pa[d].x + pa[d].y;
              }
              PROF_ADD(5)
              named_bar_sync(bar_id, 64);
              const float2 p0 = xb2[r], p1 = xb2[TILE_M + r];
              const float mean = (p0.x + p1.x) * inv_w;  // identical expression in both halves
              const float var = fmaxf((p0.y + p1.y) * inv_w - mean * mean, 0.f);
              const float rstd = rsqrtf(var + 1e-3f);  // keras LayerNormalization epsilon
              rs = f2(rstd);
              nm = f2(-mean * rstd);
              rsL = f2(rstd * 1.4426950408889634f);
              if (last) {
                // output projector folded through the LayerNorm affine (flow_matching.py:229-230)
#pragma unroll
                for (int d = 0; d < DP; ++d) {
                  const float dot = xp[(size_t)r * DP + d] + xp[((size_t)TILE_M + r) * DP + d];
                  pred[d] = fmaf(rstd, dot, fmaf(-mean * rstd, c_bo[DP + d], c_bo[d]));  // rstd (dot - mean s1) + s0
                }
              } else if (half == 0) {
                // the next Dense consumes v' raw: its rank-1 and bias terms need (mu, 1 / (log2(e) r)) of this row
                *reinterpret_cast<uint4*>(sX) = fold_extras<BF16>(mean, 0.6931471805599453f * (var + 1e-3f) * rstd);
              }
              PROF_ADD(6)
            }
            if (!last) {
              fence_proxy_async_smem();
              tc_fence_before();
              __syncwarp();
              if (lane == 0) mbar_arrive(bars.a_ready(tile));
              PROF_ADD(7)
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
        if (lane == 0 && half == 0) {
          int* acc_i = reinterpret_cast<int*>(a.ctrl + BFK_CTRL_ACC);
          atomicMax(acc_i + 0, __float_as_int(en));
          atomicMax(acc_i + 1, __float_as_int(mabs));
          if (any_nan) atomicMax(acc_i + 2, __float_as_int(1.f));
          if (any_nonnan) atomicMax(acc_i + 3, __float_as_int(1.f));
        }
        if (valid && half == 0) {
          const size_t o = (size_t)(1 - sel) * a.B * a.D + row * a.D;
          for (int d = 0; d < a.D; ++d) { a.ybuf[o + d] = ynew[d]; a.kbuf[o + d] = kst[6][d]; }
        }
      } else if (valid && half == 0) {
        for (int d = 0; d < a.D; ++d) a.x_out[row * a.D + d] = (a.mode == 1) ? kst[0][d] : y[d];
      }
    }
    PROF_ADD(0)
    PROF_FLUSH(warp)
    tc_fence_before();
  } else if (warp == F_MMA_WARP) {
#if BFK_EPI_REGS
    asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(F_AUX_REGS));
#endif
    fold_issuer_loop(a, bars, smem_base, tmem_base, my_tiles, n_evals, lane);
  } else if (warp == F_LOAD_WARP) {
#if BFK_EPI_REGS
    asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(F_AUX_REGS));
#endif
    fold_producer_loop(a, bars, smem_base, my_tiles, n_evals, lane);
  } else {
#if BFK_EPI_REGS
    asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(F_AUX_REGS));  // the two idle warps of the fifth warpgroup
#endif
  }

  __syncthreads();
  tc_fence_after();
  if (warp == F_MMA_WARP) tmem_dealloc(tmem_base, tmem_cols);
}

}  // namespace bfk
