// The fused on-chip sampler, two-tile ping-pong schedule with the SINGLE-PASS folded-LayerNorm epilogue, for the
// reference's default block shape (LayerNorm + residual + FiLM shift, networks/defaults.py:13-26).
//
// Protocol, UMMA issuer and bulk-copy producer are bfk_sampler.cuh's (one persistent CTA per SM, two 128-draw tiles
// ping-ponging through 2 x W TMEM columns: while one tile is in its epilogue the tensor pipe runs the other tile's
// layer, and the two tiles' epilogues are out of phase, so their MUFU and FMA sections interleave).  The epilogue is
// bfk_sampler_chase.cuh's: LayerNorm folded into the consuming Dense kernel (bfk_pack: gamma o W, g, c), activations
// carried times log2(e), one pass over the accumulator that writes the next A operand as the raw fp16 image of v.
//
// Warp roles (576 threads): w in 0..15 epilogue (tile = (w>>2)&1, TMEM lane quarter = w&3, column half = w>>3),
// warp 16 UMMA issuer + TMEM owner, warp 17 bulk-copy producer.
#pragma once
#include "bfk_sampler_chase.cuh"

namespace bfk {

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

  const long long rows_per_pair = TILES_PER_CTA * TILE_M;
  const long long n_pairs = (a.B + rows_per_pair - 1) / rows_per_pair;
  const long long n_evals = a.n_steps * (a.S - a.s0);

  if (warp < F_EPI_WARPS) {
    // =========================================================== EPILOGUE / ODE warps
    const int tile = (warp >> 2) & 1, quarter = warp & 3, half = warp >> 3;
    const int r = quarter * 32 + lane;  // row inside the tile == TMEM lane
    const int nch = W / 32, ch0 = half * nch;  // this warp's 16-column chunks: ch0 .. ch0 + nch - 1
    unsigned char* sA = smem + a.plan.a_tile + tile * a.plan.a_bytes + r * 16;
    const uint32_t t_acc = tmem_base + ((uint32_t)(quarter * 32) << 16) + tile * W;
    const int bar_id = 1 + tile * 4 + quarter;  // the two column halves of the same 32 rows
    // exchange between the two column halves: [buf 2][tile 2][half 2][TILE_M] (sum, sumsq); [tile][half][TILE_M][DP]
    float2* xst = reinterpret_cast<float2*>(smem + a.plan.xch);
    float* xpr = reinterpret_cast<float*>(smem + a.plan.xch + 2 * 2 * 2 * TILE_M * 8);
    const float* c_wx = s_cvec + a.cv_wx;
    const float* c_bx = s_cvec + a.cv_bx;
    const float* c_blk = s_cvec + a.cv_blk;  // per block: c_l, g_l, gamma_{l-1}
    const float* c_wo = s_cvec + a.cv_wo;    // [DP][W] gamma_L * Wo
    const float* c_bo = s_cvec + a.cv_bo;    // s0[DP], s1[DP]
    const float inv_w = 1.f / (float)W;
    PROF_DECL
    uint32_t ph_acc = 0;
    long long nb = 0;  // running block-epilogue counter (stats exchange buffer)
    long long q = 0;   // evaluation sequence number of this CTA (film double buffer)

    for (long long pair = blockIdx.x; pair < n_pairs; pair += gridDim.x) {
      const long long row = pair * rows_per_pair + tile * TILE_M + r;
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
          }
          PROF_ADD(2)
          fence_proxy_async_smem();
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(bars.a_ready(tile));
          PROF_ADD(7)

          float pred[DP];
          // row scale / shift of the current A image (identity before block 1); *L = times log2(e)
          float2 rs = f2(1.f), nm = f2(0.f), rsL = f2(1.4426950408889634f), nmL = f2(0.f);
          bool stats_pending = false;
          auto read_stats = [&]() {
            named_bar_sync(bar_id, 64);  // my column-half partner wrote its partial sums before arriving here
            const float2* xb2 = xst + ((size_t)(nb & 1) * 4 + tile * 2) * TILE_M;
            const float2 p0 = xb2[r], p1 = xb2[TILE_M + r];
            const float mean = (p0.x + p1.x) * inv_w;  // identical expression in both halves
            const float var = fmaxf((p0.y + p1.y) * inv_w - mean * mean, 0.f);
            const float rstd = rsqrtf(var + 1e-3f);  // keras LayerNormalization epsilon
            rs = f2(rstd);
            nm = f2(-mean * rstd);
            rsL = f2(rstd * 1.4426950408889634f);
            nmL = f2(-mean * rstd * 1.4426950408889634f);
            ++nb;
            stats_pending = false;
          };
          for (int g = 0; g < a.NG; ++g) {
            mbar_wait(bars.acc_ready(tile), ph_acc, a.err_flag, 2);
            PROF_ADD(3)
            ph_acc ^= 1;
            tc_fence_after();
            if (stats_pending) read_stats();
            PROF_ADD(6)
#ifdef BFK_STAGGER_NS
            if (half) __nanosleep(BFK_STAGGER_NS);  // experiment: de-phase the two column halves of a tile
#endif
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
              const float* cl = c_blk + (3 * l + 0) * W;
              const float* gl = c_blk + (3 * l + 1) * W;
              const float* gam = c_blk + (3 * l + 2) * W;
              const float* fbeta = s_film + l * W;
              float2 sum2[4], sq2[4], pa[DP];
#pragma unroll
              for (int p = 0; p < 4; ++p) { sum2[p] = f2(0.f); sq2[p] = f2(0.f); }
#pragma unroll
              for (int d = 0; d < DP; ++d) pa[d] = f2(0.f);
              auto blk = [&](float2* v, int ch) {
                const int col = ch * 16;
                chase_block_chunk<ACT, BF16>(v, col, rs, nm, rsL, nmL, cl, gl, gam, fbeta, sA, sum2, sq2);
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
              // ---- row statistics of v': my half's partial sums; the exchange completes at the next epilogue's start
              const float2 s01 = __fadd2_rn(__fadd2_rn(sum2[0], sum2[1]), __fadd2_rn(sum2[2], sum2[3]));
              const float2 q01 = __fadd2_rn(__fadd2_rn(sq2[0], sq2[1]), __fadd2_rn(sq2[2], sq2[3]));
              xst[((size_t)(nb & 1) * 4 + tile * 2 + half) * TILE_M + r] = make_float2(s01.x + s01.y, q01.x + q01.y);
              stats_pending = true;
              PROF_ADD(5)
              if (last) {
                // output projector folded through the LayerNorm affine (flow_matching.py:229-230): needs every column now
                float* xp = xpr + (size_t)(tile * 2) * TILE_M * DP;
#pragma unroll
                for (int d = 0; d < DP; ++d) xp[((size_t)half * TILE_M + r) * DP + d] = pa[d].x + pa[d].y;
                read_stats();
#pragma unroll
                for (int d = 0; d < DP; ++d) {
                  const float dot = xp[(size_t)r * DP + d] + xp[((size_t)TILE_M + r) * DP + d];
                  pred[d] = fmaf(rs.x, dot, fmaf(nm.x, c_bo[DP + d], c_bo[d]));  // rstd (dot - mean s1) + s0
                }
                PROF_ADD(6)
              }
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
    mma_issuer_loop(a, bars, smem_base, tmem_base, n_pairs, n_evals, lane);
  } else {
    producer_loop(a, bars, smem_base, n_pairs, n_evals, lane);
  }

  __syncthreads();
  tc_fence_after();
  if (warp == F_MMA_WARP) tmem_dealloc(tmem_base, tmem_cols);
}

}  // namespace bfk
