// Layer-wise sampler, fused form: ONE kernel per ConditionalDenseBlock (conditional_dense_block.py:114-131) - the Dense
// GEMM on tcgen05 and the whole block epilogue (bias, activation, FiLM shift, residual, LayerNorm statistics) in the same
// persistent kernel, so the fp32 GEMM output never leaves TMEM (r01aa: the separate block kernel was 30.4 us of the
// 53 us a 8192 x 1024 x 1024 layer cost).
//
// LayerNorm needs the whole row (W = 1024 = four 256-column tiles, more than TMEM holds), so it is not applied where it
// is computed but FOLDED into the consumer:  with v = the un-normalised block output, (mu_i, r_i) its row mean and
// 1/sqrt(var + eps),   h = gamma (v - mu) r + beta   and the next Dense is
//     h W + b = r_i * ( v (gamma o W)  -  mu_i * g )  +  c,      g_j = sum_k gamma_k W_kj,   c_j = b_j + sum_k beta_k W_kj
// i.e. the GEMM runs on the raw fp16 image of v against the pre-scaled weights (gamma o W) and the epilogue applies one
// row scale and one rank-1 correction.  Each block kernel therefore (a) reads the row statistics its producer left,
// (b) forms P, u = act(P) + film, skip = h_prev (reconstructed from the fp16 v_prev image it also multiplies),
// v = skip + u, (c) writes v as the next fp16 operand image and adds its partial (sum, sum of squares) to the row
// statistics with two fp32 atomics per row and column tile.
//
// Roles: warps 0-15 epilogue (TMEM lane quarter = warp % 4, 64-column panel = warp / 4: r01ac showed 8 epilogue warps
// latency-bound at ~20 us per tile), warp 16 UMMA issuer + TMEM owner, warp 17 bulk-copy producer; two accumulators so the epilogue of tile
// i runs under the main loop of tile i + 1.
#pragma once
#include <cuda_runtime.h>

#include "bfk_gemm.cuh"
#include "bfk_train.cuh"
#include "bfk_fast_epilogue.cuh"

namespace bfk {

struct LwFusedArgs {
  const unsigned char* A;      // v_prev (or h0) image: rows x K, K-major, row tile 128
  const unsigned char* B;      // folded weights (gamma_prev o W), row tile NT
  const float* stats_prev;     // [K/64][rows_pad][2] partial (sum, sumsq) of v_prev per 64-column panel, or null: A holds h_prev
  const float* gprev;          // LayerNorm gamma / beta of the producer (null with stats_prev)
  const float* bprev;
  const float* gcol;           // [N] g_j (read only with stats_prev)
  const float* ccol;           // [N] c_j (bias folded in)
  const float* film;           // [N] FiLM shift of this block at this evaluation
  unsigned char* out_img;      // v image of this block
  float* stats_out;            // [N/64][rows_pad][2] partials, every entry written (no atomics: bit-reproducible); null: no LayerNorm follows
  long long rows_pad;          // rows rounded up to 128 (stride of the partial-statistics planes)
  long long M;                 // valid rows
  int N, K, NT, act, residual;
  uint32_t idesc;
  // LWF_ACT: out_img = act(P + rowbias[(row0 + r) / rpg][:])  - the two activations around merge_proj (time_mlp.py:139-140)
  const float* rowbias;        // [G][N] per-dataset vector or null
  long long row0, rpg;
  // LWF_OUT: fp32 P (LayerNorm of the producer folded, bias in ccol) -> out_f32[r * ldo + col], the output projector
  float* out_f32; long long ldo;
};
enum { LWF_BLOCK = 0, LWF_ACT = 1, LWF_OUT = 2 };

constexpr int F_VEC = 5;  // g, c, film, gamma_prev, beta_prev
// BFK_LW_EPI_REGS (setmaxnreg, as in bfk_sampler_fold.cuh): 20 warps are launched, the four epilogue warpgroups raise their
// register share, the issuer / producer warpgroup (with two idle warps) drops to the rest of the 20 x 96 pool.  0 = 18 warps.
#ifndef BFK_LW_EPI_REGS
#define BFK_LW_EPI_REGS 112
#endif
constexpr int LWF_EPI_WARPS = 16, LWF_THREADS = BFK_LW_EPI_REGS ? 640 : (LWF_EPI_WARPS + 2) * 32;  // + UMMA issuer + producer
constexpr int LWF_AUX_REGS = BFK_LW_EPI_REGS ? (20 * 96 - 16 * BFK_LW_EPI_REGS) / 4 : 0;
__host__ __device__ inline size_t lw_fused_smem_bytes(int NT) {
  return (size_t)P_STAGES * (P_KC * G_TM * 2 + P_KC * NT * 2) + (size_t)2 * F_VEC * NT * 4 + 256;
}

// LWF_BLOCK arithmetic of one 16-column chunk in packed fp32 (FFMA2 / FADD2, staged MUFU: bfk_fast_epilogue.cuh), for the
// activations that have a packed form.  vc: 16 accumulator columns; rc: the 16 skip values (fp16); vb: the five staged
// per-column vectors.  Adds to the packed row statistics and returns the two 16-byte chunks of the next operand image.
template <int ACT>
__device__ __forceinline__ void lwf_block_chunk16(const float* vc, const uint4* rc, const float* vb, int NT, int c0, float rs,
                                                  float mur, bool residual, float2& s1, float2& s2, uint4* out) {
  const float2 rs2 = f2(rs), nmur2 = f2(-mur);
  float2 P[8];
#pragma unroll
  for (int jj = 0; jj < 2; ++jj) {
    const float4 g0 = *reinterpret_cast<const float4*>(vb + 0 * NT + c0 + jj * 8);
    const float4 g1 = *reinterpret_cast<const float4*>(vb + 0 * NT + c0 + jj * 8 + 4);
    const float4 k0 = *reinterpret_cast<const float4*>(vb + 1 * NT + c0 + jj * 8);
    const float4 k1 = *reinterpret_cast<const float4*>(vb + 1 * NT + c0 + jj * 8 + 4);
    const float2 gg[4] = {make_float2(g0.x, g0.y), make_float2(g0.z, g0.w), make_float2(g1.x, g1.y), make_float2(g1.z, g1.w)};
    const float2 cc[4] = {make_float2(k0.x, k0.y), make_float2(k0.z, k0.w), make_float2(k1.x, k1.y), make_float2(k1.z, k1.w)};
#pragma unroll
    for (int p = 0; p < 4; ++p) {
      const int i = jj * 4 + p;
      // P = r (acc - mu g) + c
      P[i] = __ffma2_rn(rs2, make_float2(vc[2 * i], vc[2 * i + 1]), __ffma2_rn(nmur2, gg[p], cc[p]));
    }
  }
  act2x8<ACT>(P);
#pragma unroll
  for (int jj = 0; jj < 2; ++jj) {
    const float4 f0 = *reinterpret_cast<const float4*>(vb + 2 * NT + c0 + jj * 8);
    const float4 f1 = *reinterpret_cast<const float4*>(vb + 2 * NT + c0 + jj * 8 + 4);
    const float2 ff[4] = {make_float2(f0.x, f0.y), make_float2(f0.z, f0.w), make_float2(f1.x, f1.y), make_float2(f1.z, f1.w)};
    uint32_t pk[4];
    if (residual) {
      const float4 a0 = *reinterpret_cast<const float4*>(vb + 3 * NT + c0 + jj * 8);
      const float4 a1 = *reinterpret_cast<const float4*>(vb + 3 * NT + c0 + jj * 8 + 4);
      const float4 b0 = *reinterpret_cast<const float4*>(vb + 4 * NT + c0 + jj * 8);
      const float4 b1 = *reinterpret_cast<const float4*>(vb + 4 * NT + c0 + jj * 8 + 4);
      const float2 gp[4] = {make_float2(a0.x, a0.y), make_float2(a0.z, a0.w), make_float2(a1.x, a1.y), make_float2(a1.z, a1.w)};
      const float2 bp[4] = {make_float2(b0.x, b0.y), make_float2(b0.z, b0.w), make_float2(b1.x, b1.y), make_float2(b1.z, b1.w)};
      const uint32_t su[4] = {rc[jj].x, rc[jj].y, rc[jj].z, rc[jj].w};
#pragma unroll
      for (int p = 0; p < 4; ++p) {
        // h_prev = gamma (v_prev - mu) r + beta
        const float2 hp = __ffma2_rn(gp[p], __ffma2_rn(unpack2f<false>(su[p]), rs2, nmur2), bp[p]);
        const float2 u = __fadd2_rn(__fadd2_rn(P[jj * 4 + p], ff[p]), hp);
        s1 = __fadd2_rn(s1, u);
        s2 = __ffma2_rn(u, u, s2);
        pk[p] = pack2f<false>(u);
      }
    } else {
#pragma unroll
      for (int p = 0; p < 4; ++p) {
        const float2 u = __fadd2_rn(P[jj * 4 + p], ff[p]);
        s1 = __fadd2_rn(s1, u);
        s2 = __ffma2_rn(u, u, s2);
        pk[p] = pack2f<false>(u);
      }
    }
    out[jj] = make_uint4(pk[0], pk[1], pk[2], pk[3]);
  }
}

// LWF_ACT arithmetic of one 16-column chunk in packed fp32: act(acc + c + rowbias) -> two 16-byte image chunks.
template <int ACT>
__device__ __forceinline__ void lwf_act_chunk16(const float* vc, const float* vb, int NT, int c0, const float* rb, uint4* out) {
  float2 P[8];
#pragma unroll
  for (int jj = 0; jj < 2; ++jj) {
    const float4 k0 = *reinterpret_cast<const float4*>(vb + 1 * NT + c0 + jj * 8);
    const float4 k1 = *reinterpret_cast<const float4*>(vb + 1 * NT + c0 + jj * 8 + 4);
    float2 cc[4] = {make_float2(k0.x, k0.y), make_float2(k0.z, k0.w), make_float2(k1.x, k1.y), make_float2(k1.z, k1.w)};
    if (rb) {
      const float4 r0 = *reinterpret_cast<const float4*>(rb + jj * 8);
      const float4 r1 = *reinterpret_cast<const float4*>(rb + jj * 8 + 4);
      cc[0] = __fadd2_rn(cc[0], make_float2(r0.x, r0.y)); cc[1] = __fadd2_rn(cc[1], make_float2(r0.z, r0.w));
      cc[2] = __fadd2_rn(cc[2], make_float2(r1.x, r1.y)); cc[3] = __fadd2_rn(cc[3], make_float2(r1.z, r1.w));
    }
#pragma unroll
    for (int p = 0; p < 4; ++p) {
      const int i = jj * 4 + p;
      P[i] = __fadd2_rn(make_float2(vc[2 * i], vc[2 * i + 1]), cc[p]);
    }
  }
  act2x8<ACT>(P);
#pragma unroll
  for (int jj = 0; jj < 2; ++jj)
    out[jj] = make_uint4(pack2f<false>(P[jj * 4]), pack2f<false>(P[jj * 4 + 1]), pack2f<false>(P[jj * 4 + 2]), pack2f<false>(P[jj * 4 + 3]));
}

// CL = 2: CTA pairs (a cluster of two) work on two ROW tiles of the same column tile in lockstep and share the weight
// slab - each CTA fetches half of it and multicasts it into both shared memories - which cuts the L2 -> SMEM operand
// traffic per tile from 768 KB to 512 KB (r01as: 2.0 GB of L2 traffic per launch, 11.9 TB/s, was the limiter).
template <int MODE, int ACT, int CL>
__global__ void __launch_bounds__(LWF_THREADS, 1) lw_block_gemm_kernel(const LwFusedArgs a) {
  extern __shared__ __align__(128) unsigned char smem[];
  const uint32_t crank = CL == 2 ? cluster_ctarank() : 0;
  // work items: CL == 1: (row tile, column tile); CL == 2: (row-tile PAIR, column tile), this CTA takes row tile 2*pair + rank
  // (clamped: with an odd number of row tiles the last pair computes the same tile twice and stores identical values)
  const long long w_first = CL == 2 ? blockIdx.x / 2 : blockIdx.x, w_step = CL == 2 ? gridDim.x / 2 : gridDim.x;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int NT = a.NT;
  const uint32_t a_bytes = P_KC * G_TM * 2, b_bytes = (uint32_t)P_KC * NT * 2, st_bytes = a_bytes + b_bytes;
  const uint32_t smem_base = smem_u32(smem);
  const uint32_t vec_off = P_STAGES * st_bytes;                      // [2][F_VEC][NT] floats
  const uint32_t bar_off = vec_off + 2 * F_VEC * NT * 4;
  const uint32_t bar0 = smem_base + bar_off;
  // full[s] +8s, empty[s] +8(3+s), acc_full[b] +48+8b, acc_empty[b] +64+8b, tmem slot +80
  volatile uint32_t* s_tmem = reinterpret_cast<volatile uint32_t*>(smem + bar_off + 80);
  const int n_tiles_n = a.N / NT;
  const long long m_tiles = (a.M + G_TM - 1) / G_TM;
  const long long total = (CL == 2 ? (m_tiles + 1) / 2 : m_tiles) * n_tiles_n;
  const int n_slabs = a.K / P_KC;
  const size_t K8 = (size_t)a.K / 8;
  auto m_tile_of = [&](long long t) -> long long {
    const long long mt = CL == 2 ? 2 * (t / n_tiles_n) + crank : t / n_tiles_n;
    return mt < m_tiles ? mt : m_tiles - 1;
  };

  if (threadIdx.x == 0) {
    // empty[s] collects the UMMA commits of BOTH CTAs of a pair: a stage is refilled (partly by the peer) only when both are done
    for (int s = 0; s < P_STAGES; ++s) { mbar_init(bar0 + 8 * s, 1); mbar_init(bar0 + 8 * (3 + s), CL); }
    for (int b = 0; b < 2; ++b) { mbar_init(bar0 + 48 + 8 * b, 1); mbar_init(bar0 + 64 + 8 * b, LWF_EPI_WARPS); }
    fence_barrier_init();
  }
  const uint32_t tmem_cols = NT <= 64 ? 128 : (NT <= 128 ? 256 : 512);
  const uint32_t acc_stride = tmem_cols / 2;
  if (warp == LWF_EPI_WARPS) tmem_alloc(bar0 + 80, tmem_cols);
  tc_fence_before();
  __syncthreads();
  if (CL == 2) cluster_sync_all();  // the peer's barriers are initialised before anything is multicast into them
  tc_fence_after();
  const uint32_t tmem_base = *s_tmem;

  if (warp == LWF_EPI_WARPS + 1) {
#if BFK_LW_EPI_REGS
    asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(LWF_AUX_REGS));
#endif
    if (lane == 0) {
      uint32_t g = 0;
      for (long long t = w_first; t < total; t += w_step) {
        const long long m_tile = m_tile_of(t);
        const int n_tile = (int)(t % n_tiles_n);
        for (int i = 0; i < n_slabs; ++i, ++g) {
          const uint32_t s = g % P_STAGES;
          mbar_wait(bar0 + 8 * (3 + s), ((g / P_STAGES) & 1) ^ 1, nullptr, 0);
          mbar_arrive_expect_tx(bar0 + 8 * s, st_bytes);
          const size_t k8 = (size_t)i * (P_KC / 8);
          bulk_g2s(smem_base + s * st_bytes, a.A + ((size_t)m_tile * K8 + k8) * G_TM * 16, a_bytes, bar0 + 8 * s);
          const unsigned char* bsrc = a.B + ((size_t)n_tile * K8 + k8) * NT * 16;
          if (CL == 2) {  // my half of the weight slab (4 of its 8 K octets) into both CTAs
            const uint32_t hb = b_bytes / 2;
            bulk_g2s_multicast(smem_base + s * st_bytes + a_bytes + crank * hb, bsrc + (size_t)crank * hb, hb, bar0 + 8 * s, 0x3);
          } else {
            bulk_g2s(smem_base + s * st_bytes + a_bytes, bsrc, b_bytes, bar0 + 8 * s);
          }
        }
      }
    }
    __syncwarp();
  } else if (warp == LWF_EPI_WARPS) {
#if BFK_LW_EPI_REGS
    asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(LWF_AUX_REGS));
#endif
    if (lane == 0) {
      uint32_t g = 0, j = 0;
      for (long long t = w_first; t < total; t += w_step, ++j) {
        const uint32_t acc = j & 1;
        mbar_wait(bar0 + 64 + 8 * acc, ((j >> 1) & 1) ^ 1, nullptr, 0);  // epilogue has drained this accumulator
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + acc * acc_stride;
        for (int i = 0; i < n_slabs; ++i, ++g) {
          const uint32_t s = g % P_STAGES;
          mbar_wait(bar0 + 8 * s, (g / P_STAGES) & 1, nullptr, 0);
          tc_fence_after();
          const uint32_t sa = smem_base + s * st_bytes, sb = sa + a_bytes;
#pragma unroll
          for (int kk = 0; kk < P_KC / 16; ++kk) {
            const uint64_t ad = umma_smem_desc(sa + (uint32_t)(kk * 2) * (G_TM * 16), G_TM * 16, 128);
            const uint64_t bd = umma_smem_desc(sb + (uint32_t)(kk * 2) * (NT * 16), (uint32_t)NT * 16, 128);
            umma_f16(d_tmem, ad, bd, a.idesc, (uint32_t)((i | kk) != 0));
          }
          if (CL == 2) umma_commit_multicast(bar0 + 8 * (3 + s), 0x3);
          else umma_commit(bar0 + 8 * (3 + s));
        }
        umma_commit(bar0 + 48 + 8 * acc);
      }
    }
    __syncwarp();
  } else if (warp > LWF_EPI_WARPS + 1) {
#if BFK_LW_EPI_REGS
    asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(LWF_AUX_REGS));  // the two idle warps of the fifth warpgroup
#endif
  } else {
#if BFK_LW_EPI_REGS
    asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;" ::"n"(BFK_LW_EPI_REGS));
#endif
    // ---- epilogue warps 0-15: TMEM lane quarter = warp % 4, 64-column panel = warp / 4; 16-column chunks, the TMEM and
    // skip loads of chunk i+1 in flight under the arithmetic of chunk i ------------------------------------------------
    const int q = warp & 3, pq = warp >> 2;
    const bool folded = a.stats_prev != nullptr;
    const float inv_w = 1.f / (float)a.K;
    float* vecs = reinterpret_cast<float*>(smem + vec_off);
    const int N8 = a.N / 8;
    const int n_chunks = (NT / 64 > pq) ? ((NT / 64 - pq + 3) / 4) * 4 : 0;  // 16-column chunks this warp owns per tile
    uint32_t j = 0;
    for (long long t = w_first; t < total; t += w_step, ++j) {
      const long long m_tile = m_tile_of(t);
      const int n_tile = (int)(t % n_tiles_n);
      const uint32_t acc = j & 1;
      // per-column vectors of this column tile -> shared (double buffered by tile parity, one barrier per tile)
      float* vb = vecs + (size_t)acc * F_VEC * NT;
      for (int cidx = threadIdx.x; cidx < NT; cidx += LWF_EPI_WARPS * 32) {
        const int gc = n_tile * NT + cidx;
        vb[0 * NT + cidx] = folded ? a.gcol[gc] : 0.f;
        vb[1 * NT + cidx] = a.ccol ? a.ccol[gc] : 0.f;
        vb[2 * NT + cidx] = (MODE == LWF_BLOCK && a.film) ? a.film[gc] : 0.f;
        vb[3 * NT + cidx] = (folded && a.gprev) ? a.gprev[gc] : 1.f;
        vb[4 * NT + cidx] = (folded && a.bprev) ? a.bprev[gc] : 0.f;
      }
      const long long row = m_tile * G_TM + q * 32 + lane;
      float mu = 0.f, rs = 1.f;
      if (folded && row < a.M) {
        float sx = 0.f, sy = 0.f;
        for (int pl = 0; pl < a.K / 64; ++pl) {  // fixed order: the row statistics are bit-reproducible
          const float2 st = *reinterpret_cast<const float2*>(a.stats_prev + ((size_t)pl * a.rows_pad + row) * 2);
          sx += st.x;
          sy += st.y;
        }
        mu = sx * inv_w;
        const float var = fmaxf(fmaf(sy, inv_w, -mu * mu), 0.f);
        rs = rsqrtf(var + LN_EPS);
      }
      const float mur = mu * rs;
      // chunk ci -> first column inside the tile (panels pq, pq+4, ... of 64 columns, 4 chunks each)
      auto col_of = [&](int ci) { return ((ci >> 2) * 4 + pq) * 64 + (ci & 3) * 16; };
      uint4 raw0[2], raw1[2];
      float v0[16], v1[16];
      auto fetch_skip = [&](int ci, uint4* raw) {
        if (MODE == LWF_BLOCK && a.residual) {
          const int o0 = (n_tile * NT + col_of(ci)) / 8;
          raw[0] = *reinterpret_cast<const uint4*>(a.A + img_off(row, o0, (int)K8));
          raw[1] = *reinterpret_cast<const uint4*>(a.A + img_off(row, o0 + 1, (int)K8));
        }
      };
      if (n_chunks) fetch_skip(0, raw0);  // does not depend on the accumulator
      named_bar_sync(1, LWF_EPI_WARPS * 32);
      mbar_wait(bar0 + 48 + 8 * acc, (j >> 1) & 1, nullptr, 0);
      tc_fence_after();
      const uint32_t t_acc = tmem_base + acc * acc_stride + ((uint32_t)(q * 32) << 16);
      float s1 = 0.f, s2 = 0.f;
      float2 s1p = make_float2(0.f, 0.f), s2p = make_float2(0.f, 0.f);  // packed path
      // one chunk: wait for its accumulator columns, put the next chunk's loads in flight, then the block arithmetic
      auto step = [&](int ci, float* vc, uint4* rc, float* vn, uint4* rn) {
        tmem_ld_wait16(vc);
        if (ci + 1 < n_chunks) {
          tmem_ld16(t_acc + col_of(ci + 1), vn);
          fetch_skip(ci + 1, rn);
        } else {
          // every TMEM read of this accumulator has landed: hand it back to the UMMA issuer before the arithmetic
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(bar0 + 64 + 8 * acc);
        }
        const int c0 = col_of(ci);
        const int o0 = (n_tile * NT + c0) / 8;
        if (MODE == LWF_BLOCK && ACT >= 0) {  // packed-fp32 arithmetic for mish / silu (compile-time activation)
          uint4 o16[2];
          lwf_block_chunk16<(ACT >= 0 ? ACT : 0)>(vc, rc, vb, NT, c0, rs, mur, a.residual != 0, s1p, s2p, o16);
          *reinterpret_cast<uint4*>(a.out_img + img_off(row, o0, N8)) = o16[0];
          *reinterpret_cast<uint4*>(a.out_img + img_off(row, o0 + 1, N8)) = o16[1];
          return;
        }
        const float* rb = nullptr;
        if (MODE == LWF_ACT && a.rowbias) rb = a.rowbias + ((a.row0 + (row < a.M ? row : a.M - 1)) / a.rpg) * a.N + n_tile * NT + c0;
        if (MODE == LWF_ACT && ACT >= 0) {
          uint4 o16[2];
          lwf_act_chunk16<(ACT >= 0 ? ACT : 0)>(vc, vb, NT, c0, rb, o16);
          *reinterpret_cast<uint4*>(a.out_img + img_off(row, o0, N8)) = o16[0];
          *reinterpret_cast<uint4*>(a.out_img + img_off(row, o0 + 1, N8)) = o16[1];
          return;
        }
#pragma unroll
        for (int o = 0; o < 2; ++o) {
          float sk[8];
          if (MODE == LWF_BLOCK && a.residual) {
            const __half2* hp = reinterpret_cast<const __half2*>(&rc[o]);
#pragma unroll
            for (int i = 0; i < 4; ++i) { const float2 f = __half22float2(hp[i]); sk[2 * i] = f.x; sk[2 * i + 1] = f.y; }
          }
          float out8[8];
#pragma unroll
          for (int i4 = 0; i4 < 8; i4 += 4) {
            const int cc = c0 + o * 8 + i4;
            const float4 g4 = *reinterpret_cast<const float4*>(vb + 0 * NT + cc);
            const float4 c4 = *reinterpret_cast<const float4*>(vb + 1 * NT + cc);
            const float gg[4] = {g4.x, g4.y, g4.z, g4.w}, cv[4] = {c4.x, c4.y, c4.z, c4.w};
            float ff[4] = {0.f, 0.f, 0.f, 0.f}, gp[4] = {1.f, 1.f, 1.f, 1.f}, bp[4] = {0.f, 0.f, 0.f, 0.f};
            if (MODE == LWF_BLOCK) {
              const float4 f4 = *reinterpret_cast<const float4*>(vb + 2 * NT + cc);
              const float4 gp4 = *reinterpret_cast<const float4*>(vb + 3 * NT + cc);
              const float4 bp4 = *reinterpret_cast<const float4*>(vb + 4 * NT + cc);
              ff[0] = f4.x; ff[1] = f4.y; ff[2] = f4.z; ff[3] = f4.w;
              gp[0] = gp4.x; gp[1] = gp4.y; gp[2] = gp4.z; gp[3] = gp4.w;
              bp[0] = bp4.x; bp[1] = bp4.y; bp[2] = bp4.z; bp[3] = bp4.w;
            }
            if (MODE == LWF_ACT && rb) {
              const float4 r4 = *reinterpret_cast<const float4*>(rb + o * 8 + i4);
              ff[0] = r4.x; ff[1] = r4.y; ff[2] = r4.z; ff[3] = r4.w;  // added BEFORE the activation in this mode
            }
#pragma unroll
            for (int i = 0; i < 4; ++i) {
              const float accv = vc[o * 8 + i4 + i];
              // P = r (acc - mu g) + c   (r = 1, mu = 0 when the operand image already holds h_prev)
              float P = fmaf(rs, accv, fmaf(-mur, gg[i], cv[i]));
              float u;
              if (MODE == LWF_OUT) {
                u = P;
              } else if (MODE == LWF_ACT) {
                float d;
                u = t_act(a.act, P + ff[i], d);
              } else {
                float d;
                u = t_act(a.act, P, d) + ff[i];
                if (a.residual) u += fmaf(gp[i], fmaf(sk[i4 + i], rs, -mur), bp[i]);  // h_prev = gamma (v - mu) r + beta
                s1 += u;
                s2 = fmaf(u, u, s2);
              }
              out8[i4 + i] = u;
            }
          }
          if (MODE == LWF_OUT) {
            if (row < a.M) {
              float* dst = a.out_f32 + row * a.ldo + n_tile * NT + c0 + o * 8;
              *reinterpret_cast<float4*>(dst) = make_float4(out8[0], out8[1], out8[2], out8[3]);
              *reinterpret_cast<float4*>(dst + 4) = make_float4(out8[4], out8[5], out8[6], out8[7]);
            }
          } else {
            *reinterpret_cast<uint4*>(a.out_img + img_off(row, o0 + o, N8)) = pack8(out8);
          }
        }
      };
      if (n_chunks) tmem_ld16(t_acc + col_of(0), v0);
      for (int ci = 0; ci < n_chunks; ci += 2) {  // n_chunks is a multiple of 4
        step(ci, v0, raw0, v1, raw1);
        step(ci + 1, v1, raw1, v0, raw0);
      }
      if (!n_chunks) {
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(bar0 + 64 + 8 * acc);
      }
      if (MODE == LWF_BLOCK && a.stats_out && n_chunks) {
        // one 64-column panel per warp and tile (NT <= 256): its partial sums go to plane (n_tile * NT / 64 + pq)
        const int plane = n_tile * (NT / 64) + pq;
        *reinterpret_cast<float2*>(a.stats_out + ((size_t)plane * a.rows_pad + row) * 2) =
            make_float2(s1 + (s1p.x + s1p.y), s2 + (s2p.x + s2p.y));
      }
    }
  }
  __syncthreads();
  if (CL == 2) cluster_sync_all();  // no CTA leaves while its peer can still multicast into it or signal its barriers
  tc_fence_after();
  if (warp == LWF_EPI_WARPS) tmem_dealloc(tmem_base, tmem_cols);
}

// h = gamma (v - mu) r + beta as an fp16 operand image: the LayerNorm of the LAST block, whose consumer is the output
// projector (flow_matching.py:230).  grid (K8, tile chunks), 128 threads = rows of one tile.
__global__ void lw_norm_img_kernel(const unsigned char* __restrict__ in_img, const float* __restrict__ stats,
                                   const float* __restrict__ gamma, const float* __restrict__ beta, long long rows,
                                   long long n_tiles, int K8, unsigned char* __restrict__ out_img) {
  const int k8 = blockIdx.x;
  float g[8], b[8];
  ld8(gamma + k8 * 8, g);
  ld8(beta + k8 * 8, b);
  const float inv_w = 1.f / (float)(K8 * 8);
  for (long long tile = blockIdx.y; tile < n_tiles; tile += gridDim.y) {
    const long long r = tile * 128 + threadIdx.x;
    float v[8];
    if (r < rows) {
      float sx = 0.f, sy = 0.f;
      for (int pl = 0; pl < K8 / 8; ++pl) {
        const float2 st = *reinterpret_cast<const float2*>(stats + ((size_t)pl * n_tiles * 128 + r) * 2);
        sx += st.x;
        sy += st.y;
      }
      const float mu = sx * inv_w;
      const float rs = rsqrtf(fmaxf(fmaf(sy, inv_w, -mu * mu), 0.f) + LN_EPS);
      const uint4 raw = *reinterpret_cast<const uint4*>(in_img + img_off(r, k8, K8));
      const __half2* hp = reinterpret_cast<const __half2*>(&raw);
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const float2 f = __half22float2(hp[i]);
        v[2 * i] = fmaf(g[2 * i], (f.x - mu) * rs, b[2 * i]);
        v[2 * i + 1] = fmaf(g[2 * i + 1], (f.y - mu) * rs, b[2 * i + 1]);
      }
    } else {
#pragma unroll
      for (int i = 0; i < 8; ++i) v[i] = 0.f;
    }
    *reinterpret_cast<uint4*>(out_img + img_off(r, k8, K8)) = pack8(v);
  }
}

// Once per call and block: Wf = fp16(gamma_prev o W) in the B-operand image P[n/NT][k/8][n%NT][k%8],
// g_j = sum_k float(Wf_kj) (the very values the tensor cores multiply, so that r*(acc - mu*g) cancels exactly for a
// constant row), c_j = b_j + sum_k beta_prev_k W_kj.  One thread per output column n; W is (K, N) row-major (in, out).
__global__ void __launch_bounds__(1024) lw_fold_pack_kernel(const float* __restrict__ Wk, int K, int N, int N_valid,
                                                            const float* __restrict__ gprev, const float* __restrict__ bprev,
                                                            const float* __restrict__ bias, int NT, unsigned char* __restrict__ img,
                                                            float* __restrict__ gcol, float* __restrict__ ccol) {
  // block = 32 columns (threadIdx.x) x 32 slices of K (threadIdx.y); the slice partials meet in shared memory and are
  // summed in slice order by the y = 0 threads, so g and c are bit-reproducible
  __shared__ float sg[32][33], sc[32][33];
  const int n = blockIdx.x * 32 + threadIdx.x;
  const bool on = n < N_valid;  // W is (K, N_valid) row-major (in, out); columns up to N are zero padding
  const int K8 = K / 8;
  float gs = 0.f, cs = 0.f;
  if (n < N) {
    unsigned char* base = img + ((size_t)(n / NT) * K8 * NT + (size_t)(n % NT)) * 16;
    for (int k8 = threadIdx.y; k8 < K8; k8 += 32) {
      float v[8];
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const int k = k8 * 8 + i;
        const float w = on ? Wk[(size_t)k * N_valid + n] : 0.f;
        v[i] = gprev ? gprev[k] * w : w;
        if (bprev) cs = fmaf(bprev[k], w, cs);
      }
      const uint4 pk = pack8(v);
      const __half2* hp = reinterpret_cast<const __half2*>(&pk);
#pragma unroll
      for (int i = 0; i < 4; ++i) { const float2 f = __half22float2(hp[i]); gs += f.x + f.y; }
      *reinterpret_cast<uint4*>(base + (size_t)k8 * NT * 16) = pk;
    }
  }
  sg[threadIdx.y][threadIdx.x] = gs;
  sc[threadIdx.y][threadIdx.x] = cs;
  __syncthreads();
  if (threadIdx.y == 0 && n < N) {
    float g = 0.f, c = (bias && on) ? bias[n] : 0.f;
    for (int y = 0; y < 32; ++y) { g += sg[y][threadIdx.x]; c += sc[y][threadIdx.x]; }
    gcol[n] = g;
    ccol[n] = c;
  }
}

}  // namespace bfk
