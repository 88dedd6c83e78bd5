// Fast variant of the fused on-chip sampler for the reference's default block shape (LayerNorm + residual + FiLM shift,
// networks/defaults.py:13-26).  Same protocol, barriers, UMMA issuer and bulk-copy producer as bfk_sampler.cuh; what
// changes is the epilogue: 16 epilogue warps instead of 8 (4 per SM sub-partition, so MUFU / FMA / LDS latencies of one
// warp hide behind the others), each owning HALF the columns of 32 rows of one tile, packed-f32x2 arithmetic
// (bfk_fast_epilogue.cuh), 16-column double-buffered TMEM loads, row statistics and projector partials exchanged
// between the two column halves through shared memory + a 64-thread named barrier.
//
// Warp roles (576 threads): w in 0..15 epilogue (tile = (w>>2)&1, TMEM lane quarter = w&3, column half = w>>3),
// warp 16 UMMA issuer + TMEM owner, warp 17 bulk-copy producer.
#pragma once
#include "bfk_fast_epilogue.cuh"
#include "bfk_sampler.cuh"

namespace bfk {

template <int ACT, int DP, bool BF16>
__global__ void __launch_bounds__(F_NUM_THREADS, 1) sampler_fast_kernel(const SamplerArgs a) {
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
    const int WH = W >> 1, col0 = half * WH, nch = WH / 16;
    unsigned char* sA = smem + a.plan.a_tile + tile * a.plan.a_bytes + r * 16;
    const uint32_t t_acc = tmem_base + ((uint32_t)(quarter * 32) << 16) + tile * W + col0;
    const int bar_id = 1 + tile * 4 + quarter;  // the two column halves of the same 32 rows
    float* xch = reinterpret_cast<float*>(smem + a.plan.xch);
    float* x_lo = xch + (size_t)((tile * 2 + 0) * TILE_M + r) * a.plan.xch_stride;
    float* x_hi = xch + (size_t)((tile * 2 + 1) * TILE_M + r) * a.plan.xch_stride;
    float* x_me = half ? x_hi : x_lo;
    const float* c_wx = s_cvec + a.cv_wx;
    const float* c_bx = s_cvec + a.cv_bx;
    const float* c_blk = s_cvec + a.cv_blk;
    const float* c_wo = s_cvec + a.cv_wo;  // [DP][W] gamma_L * Wo
    const float* c_bo = s_cvec + a.cv_bo;  // s0[DP], s1[DP]
    const float inv_w = 1.f / (float)W;
    uint32_t ph_acc = 0;
    long long q = 0;  // evaluation sequence number of this CTA (film double buffer)

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

          // ---- x_proj (+ conditions for merge="add") + activation -> my half of the A tile   time_mlp.py:131,137,140
          {
            float2 xb[DP];
#pragma unroll
            for (int d = 0; d < DP; ++d) xb[d] = f2(xs[d]);
            for (int j = col0 / 8; j < (col0 + WH) / 8; j += 2) {
              float2 v8[8];
#pragma unroll
              for (int jj = 0; jj < 2; ++jj) {
                const float4 b0 = *reinterpret_cast<const float4*>(c_bx + (j + jj) * 8);
                const float4 b1 = *reinterpret_cast<const float4*>(c_bx + (j + jj) * 8 + 4);
                float2* v4 = v8 + jj * 4;
                v4[0] = make_float2(b0.x, b0.y); v4[1] = make_float2(b0.z, b0.w);
                v4[2] = make_float2(b1.x, b1.y); v4[3] = make_float2(b1.z, b1.w);
                if (a.has_merge && g_pre) {
                  const float4 p0 = __ldg(reinterpret_cast<const float4*>(g_pre + (j + jj) * 8));
                  const float4 p1 = __ldg(reinterpret_cast<const float4*>(g_pre + (j + jj) * 8 + 4));
                  v4[0] = __fadd2_rn(v4[0], make_float2(p0.x, p0.y)); v4[1] = __fadd2_rn(v4[1], make_float2(p0.z, p0.w));
                  v4[2] = __fadd2_rn(v4[2], make_float2(p1.x, p1.y)); v4[3] = __fadd2_rn(v4[3], make_float2(p1.z, p1.w));
                }
#pragma unroll
                for (int d = 0; d < DP; ++d) {
                  const float4 w0 = *reinterpret_cast<const float4*>(c_wx + d * W + (j + jj) * 8);
                  const float4 w1 = *reinterpret_cast<const float4*>(c_wx + d * W + (j + jj) * 8 + 4);
                  v4[0] = __ffma2_rn(xb[d], make_float2(w0.x, w0.y), v4[0]); v4[1] = __ffma2_rn(xb[d], make_float2(w0.z, w0.w), v4[1]);
                  v4[2] = __ffma2_rn(xb[d], make_float2(w1.x, w1.y), v4[2]); v4[3] = __ffma2_rn(xb[d], make_float2(w1.z, w1.w), v4[3]);
                }
              }
              act2x8<ACT>(v8);
#pragma unroll
              for (int jj = 0; jj < 2; ++jj)
                *reinterpret_cast<uint4*>(sA + (j + jj) * (TILE_M * 16)) =
                    make_uint4(pack2f<BF16>(v8[jj * 4 + 0]), pack2f<BF16>(v8[jj * 4 + 1]), pack2f<BF16>(v8[jj * 4 + 2]),
                               pack2f<BF16>(v8[jj * 4 + 3]));
            }
          }
          fence_proxy_async_smem();
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(bars.a_ready(tile));

          float pred[DP];
          for (int g = 0; g < a.NG; ++g) {
            mbar_wait(bars.acc_ready(tile), ph_acc, a.err_flag, 2);
            ph_acc ^= 1;
            tc_fence_after();
            const bool last = (g == a.NG - 1);
            float2 va[8], vb[8];
            if (a.has_merge && g == 0) {
              // ---- merge_proj epilogue: h = act(acc + [m_c + bm])                      time_mlp.py:139-140
              auto merge_chunk = [&](float2* v, int col) {
#pragma unroll
                for (int jj = 0; jj < 2; ++jj) {
                  const float4 p0 = __ldg(reinterpret_cast<const float4*>(g_post + col + jj * 8));
                  const float4 p1 = __ldg(reinterpret_cast<const float4*>(g_post + col + jj * 8 + 4));
                  v[jj * 4 + 0] = __fadd2_rn(v[jj * 4 + 0], make_float2(p0.x, p0.y));
                  v[jj * 4 + 1] = __fadd2_rn(v[jj * 4 + 1], make_float2(p0.z, p0.w));
                  v[jj * 4 + 2] = __fadd2_rn(v[jj * 4 + 2], make_float2(p1.x, p1.y));
                  v[jj * 4 + 3] = __fadd2_rn(v[jj * 4 + 3], make_float2(p1.z, p1.w));
                }
                act2x8<ACT>(v);
#pragma unroll
                for (int jj = 0; jj < 2; ++jj)
                  *reinterpret_cast<uint4*>(sA + (size_t)((col >> 3) + jj) * (TILE_M * 16)) =
                      make_uint4(pack2f<BF16>(v[jj * 4 + 0]), pack2f<BF16>(v[jj * 4 + 1]), pack2f<BF16>(v[jj * 4 + 2]),
                                 pack2f<BF16>(v[jj * 4 + 3]));
              };
              if (!BFK_DBG(1)) tmem_ld16(t_acc, reinterpret_cast<float*>(va));
              for (int c = 0; c < nch && !BFK_DBG(1); c += 2) {
                tmem_ld_wait16(reinterpret_cast<float*>(va));
                if (c + 1 < nch) tmem_ld16(t_acc + (c + 1) * 16, reinterpret_cast<float*>(vb));
                merge_chunk(va, col0 + c * 16);
                if (c + 1 < nch) {
                  tmem_ld_wait16(reinterpret_cast<float*>(vb));
                  if (c + 2 < nch) tmem_ld16(t_acc + (c + 2) * 16, reinterpret_cast<float*>(va));
                  merge_chunk(vb, col0 + (c + 1) * 16);
                }
              }
            } else {
              // ---- ConditionalDenseBlock epilogue                              conditional_dense_block.py:114-131
              const int l = g - a.has_merge;
              const float* bd = c_blk + (3 * l + 0) * W;
              const float* lng = c_blk + (3 * l + 1) * W;
              const float* lnb = c_blk + (3 * l + 2) * W;
              const float* fbeta = s_film + l * W;
              float2 sum2[4], sq2[4], pa[DP];
#pragma unroll
              for (int p = 0; p < 4; ++p) { sum2[p] = f2(0.f); sq2[p] = f2(0.f); }
#pragma unroll
              for (int d = 0; d < DP; ++d) pa[d] = f2(0.f);
              auto p1 = [&](float2* v, int c) {
                const int col = col0 + c * 16;
                fast_pass1_chunk<ACT, BF16>(v, col, bd, fbeta, sA, sum2, sq2);
                if (last) {
#pragma unroll
                  for (int d = 0; d < DP; ++d) {
                    const float* wg = c_wo + d * W + col;
                    const float4 w0 = *reinterpret_cast<const float4*>(wg), w1 = *reinterpret_cast<const float4*>(wg + 4);
                    const float4 w2 = *reinterpret_cast<const float4*>(wg + 8), w3 = *reinterpret_cast<const float4*>(wg + 12);
                    float2 t0 = __fmul2_rn(v[0], make_float2(w0.x, w0.y)), t1 = __fmul2_rn(v[1], make_float2(w0.z, w0.w));
                    t0 = __ffma2_rn(v[2], make_float2(w1.x, w1.y), t0); t1 = __ffma2_rn(v[3], make_float2(w1.z, w1.w), t1);
                    t0 = __ffma2_rn(v[4], make_float2(w2.x, w2.y), t0); t1 = __ffma2_rn(v[5], make_float2(w2.z, w2.w), t1);
                    t0 = __ffma2_rn(v[6], make_float2(w3.x, w3.y), t0); t1 = __ffma2_rn(v[7], make_float2(w3.z, w3.w), t1);
                    pa[d] = __fadd2_rn(pa[d], __fadd2_rn(t0, t1));
                  }
                } else {
                  tmem_st16(t_acc + c * 16, reinterpret_cast<const float*>(v));
                }
              };
              if (!BFK_DBG(1)) tmem_ld16(t_acc, reinterpret_cast<float*>(va));
              for (int c = 0; c < nch && !BFK_DBG(1); c += 2) {
                tmem_ld_wait16(reinterpret_cast<float*>(va));
                if (c + 1 < nch) tmem_ld16(t_acc + (c + 1) * 16, reinterpret_cast<float*>(vb));
                p1(va, c);
                if (c + 1 < nch) {
                  tmem_ld_wait16(reinterpret_cast<float*>(vb));
                  if (c + 2 < nch) tmem_ld16(t_acc + (c + 2) * 16, reinterpret_cast<float*>(va));
                  p1(vb, c + 1);
                }
              }
              // ---- row statistics: my half + the other half's (through shared memory, 64-thread named barrier)
              const float2 s01 = __fadd2_rn(__fadd2_rn(sum2[0], sum2[1]), __fadd2_rn(sum2[2], sum2[3]));
              const float2 q01 = __fadd2_rn(__fadd2_rn(sq2[0], sq2[1]), __fadd2_rn(sq2[2], sq2[3]));
              x_me[0] = s01.x + s01.y;
              x_me[1] = q01.x + q01.y;
              if (last) {
#pragma unroll
                for (int d = 0; d < DP; ++d) x_me[2 + d] = pa[d].x + pa[d].y;
              }
              named_bar_sync(bar_id, 64);
              const float mean = (x_lo[0] + x_hi[0]) * inv_w;  // identical expression in both halves
              const float var = fmaxf((x_lo[1] + x_hi[1]) * inv_w - mean * mean, 0.f);
              const float rstd = rsqrtf(var + 1e-3f);  // keras LayerNormalization epsilon
              if (last) {
                // output projector folded through the LayerNorm affine (flow_matching.py:229-230)
#pragma unroll
                for (int d = 0; d < DP; ++d)
                  pred[d] = fmaf(rstd, (x_lo[2 + d] + x_hi[2 + d]) - mean * c_bo[DP + d], c_bo[d]);
              } else {
                tmem_st_wait();
                const float nmr = -mean * rstd;
                if (!BFK_DBG(3)) tmem_ld16(t_acc, reinterpret_cast<float*>(va));
                for (int c = 0; c < nch && !BFK_DBG(3); c += 2) {
                  tmem_ld_wait16(reinterpret_cast<float*>(va));
                  if (c + 1 < nch) tmem_ld16(t_acc + (c + 1) * 16, reinterpret_cast<float*>(vb));
                  fast_pass2_chunk<BF16>(va, col0 + c * 16, rstd, nmr, lng, lnb, sA);
                  if (c + 1 < nch) {
                    tmem_ld_wait16(reinterpret_cast<float*>(vb));
                    if (c + 2 < nch) tmem_ld16(t_acc + (c + 2) * 16, reinterpret_cast<float*>(va));
                    fast_pass2_chunk<BF16>(vb, col0 + (c + 1) * 16, rstd, nmr, lng, lnb, sA);
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
