// Host/device-shared description of the packed weight blob and of the sampler's shared-memory carve-up.
#pragma once
#include <stddef.h>
#include <stdint.h>

#include "../../include/bfk.h"

namespace bfk {

constexpr int TILE_M = 128;       // rows (draws) per tile = UMMA M = TMEM lanes
constexpr int TILES_PER_CTA = 2;  // two tiles ping-pong through TMEM
constexpr int KC = 32;            // K rows of a weight stage (2 UMMAs of K=16)
constexpr int XK = 16;            // folded layout: extra K rows per layer (one more UMMA): -g, c as fp16 hi/lo pairs
constexpr int FKC = 32;           // folded kernel: K rows of a full weight stage (16-row stages, 6 of them, measured slower:
                                  // the issuer pays one mbarrier wait per stage); the XK tail is one more, half-filled stage
constexpr int EPI_WARPS = 8, MMA_WARP = 8, LOAD_WARP = 9, NUM_THREADS = 320;
// fast kernel: 16 epilogue warps (tile = (w>>2)&1, TMEM lane quarter = w&3, column half = w>>3) + issuer + producer
// BFK_EPI_REGS (setmaxnreg): 20 warps are launched (two idle ones complete the fifth warpgroup); after set-up the four
// epilogue warpgroups raise their register file share to BFK_EPI_REGS and the issuer / producer warpgroup drops to what is
// left of the 20 x 96 pool.  0 = 18 warps, 96 registers each (ptxas rounds 576 threads up to 640 for the register cap).
#ifndef BFK_EPI_REGS
#define BFK_EPI_REGS 104
#endif
constexpr int F_EPI_WARPS = 16, F_MMA_WARP = 16, F_LOAD_WARP = 17, F_NUM_THREADS = BFK_EPI_REGS ? 640 : 576;
constexpr int F_AUX_REGS = BFK_EPI_REGS ? (20 * 96 - 16 * BFK_EPI_REGS) / 4 : 0;
constexpr int TABLE_SCALARS = 8;  // per-evaluation scalars after the FiLM vectors: coefZ, coefP, time, norm-time, ...
constexpr size_t SMEM_LIMIT = 232448;

__host__ __device__ inline int dp_of(int D) { return D <= 2 ? 2 : (D <= 4 ? 4 : 8); }
// the packed-f32x2 epilogue covers the reference's default block shape (defaults.py:13-26); anything else -> generic
__host__ __device__ inline int is_fast(const bfk_net_config& c) {
  return (c.norm == BFK_NORM_LAYER && c.residual && !c.film_use_gamma) ? 1 : 0;
}
__host__ __device__ inline size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

// byte offsets inside the packed blob
struct PackedLayout {
  int NG, DP, has_merge, Th;  // Th = (T-1)/2 fourier frequencies
  size_t gemm;                // [NG][W/8][W][8] 16-bit
  size_t cvec;                // SMEM image of the constant vectors (floats), see below
  int cvec_floats;
  int cv_wx, cv_bx, cv_blk, cv_wo, cv_bo;  // float offsets inside cvec: Wx[DP][W], bx[W], per block {bd,lng,lnb}[W], Wo[W][DP], bo[DP]
                                           // fast layout: cv_wo = [DP][W] gamma_L*Wo (projector folded through the last
                                           // LayerNorm), cv_bo = s0[DP] (= beta_L.Wo + bo) then s1[DP] (= sum_c gamma_L*Wo)
  size_t freq;                // [Th]
  size_t film_b_k, film_b_b;  // [L][T][W], [L][W]
  size_t film_g_k, film_g_b;  // same, only if film_use_gamma
  size_t c_k, c_b;            // [C][W], [W]
  size_t m_c, m_b;            // conditions half of merge_proj [W][W] (concat only), [W]
  size_t ln_beta;             // [L][W] LayerNorm beta of every block (folded layout: goes into the per-evaluation table)
  size_t gemm_x;              // folded layout: [NG][(W+XK)/8][W][8] 16-bit, XK extra K rows per layer (rank-1 + bias terms)
  size_t total;
};

__host__ __device__ inline PackedLayout make_layout(const bfk_net_config& c) {
  PackedLayout L{};
  const size_t W = c.W;
  L.has_merge = (c.C > 0 && c.merge != BFK_MERGE_NONE) ? 1 : 0;
  L.NG = c.L + L.has_merge;
  L.DP = dp_of(c.D);
  L.Th = (c.T - 1) / 2;
  size_t off = 0;
  L.gemm = off; off += (size_t)L.NG * W * W * 2; off = align_up(off, 256);
  L.cv_wx = 0;
  L.cv_bx = L.cv_wx + L.DP * c.W;
  L.cv_blk = L.cv_bx + c.W;
  L.cv_wo = L.cv_blk + (is_fast(c) ? 1 : 3) * c.L * c.W;  // folded layout: only gamma_{l-1} per block
  L.cv_bo = L.cv_wo + c.W * L.DP;
  L.cvec_floats = L.cv_bo + 2 * L.DP;
  L.cvec = off; off += (size_t)L.cvec_floats * 4; off = align_up(off, 256);
  L.freq = off; off += align_up((size_t)(L.Th > 0 ? L.Th : 1) * 4, 256);
  L.film_b_k = off; off += (size_t)c.L * c.T * W * 4;
  L.film_b_b = off; off += (size_t)c.L * W * 4;
  L.film_g_k = off; if (c.film_use_gamma) off += (size_t)c.L * c.T * W * 4;
  L.film_g_b = off; if (c.film_use_gamma) off += (size_t)c.L * W * 4;
  L.c_k = off; off += (size_t)(c.C > 0 ? c.C : 0) * W * 4;
  L.c_b = off; off += W * 4;
  L.m_c = off; off += W * W * 4;
  L.m_b = off; off += W * 4;
  L.ln_beta = off; off += (size_t)c.L * W * 4;
  off = align_up(off, 256);
  L.gemm_x = off; off += (size_t)L.NG * (W + XK) * W * 2;
  L.total = align_up(off, 256);
  return L;
}

__host__ __device__ inline int table_stride(const bfk_net_config& c) {
  return (c.film_use_gamma ? 2 : 1) * c.L * c.W + TABLE_SCALARS;
}

struct SmemPlan {
  int n_stage;
  uint32_t a_tile, ring, film, cvec, bars, xch, total;  // byte offsets
  uint32_t xch_stride;                                  // floats per (tile, half, row) exchange slot (fast kernel)
  uint32_t a_bytes, stage_bytes, film_bytes;
};

__host__ __device__ inline SmemPlan make_smem_plan(const bfk_net_config& c) {
  SmemPlan p{};
  PackedLayout L = make_layout(c);
  p.a_bytes = TILE_M * c.W * 2;
  p.stage_bytes = KC * c.W * 2;
  p.film_bytes = (uint32_t)align_up((size_t)table_stride(c) * 4, 16);
  uint32_t off = 0;
  p.a_tile = off; off += TILES_PER_CTA * p.a_bytes;
  p.film = off; off += 2 * p.film_bytes;
  p.cvec = off; off += (uint32_t)align_up((size_t)L.cvec_floats * 4, 16);
  p.bars = off; off += 256;
  p.xch = off;
  p.xch_stride = (uint32_t)((2 + L.DP + 3) / 4 * 4);  // row sum, sum of squares, DP projector partials
  if (is_fast(c)) off += 2 * 2 * TILE_M * p.xch_stride * 4;
  p.ring = (uint32_t)align_up(off, 128);
  int n = (int)((SMEM_LIMIT - p.ring) / p.stage_bytes);
  if (n > 8) n = 8;
  p.n_stage = n;
  p.total = p.ring + (n > 0 ? n : 0) * p.stage_bytes;
  return p;
}

// two-tile ping-pong schedule with the folded single-pass epilogue (bfk_sampler_fold.cuh)
__host__ __device__ inline SmemPlan make_fold_plan(const bfk_net_config& c) {
  SmemPlan p{};
  PackedLayout L = make_layout(c);
  p.a_bytes = TILE_M * (c.W + 8) * 2;  // K-chunk W/8 of each tile: per-row (mu, 1/r) as fp16 hi/lo pairs; the second
                                       // K-chunk of the XK tail is all zero and shared by both tiles (after the tiles)
  p.stage_bytes = FKC * c.W * 2;
  p.film_bytes = (uint32_t)align_up((size_t)table_stride(c) * 4, 16);
  uint32_t off = 0;
  p.a_tile = off; off += TILES_PER_CTA * p.a_bytes + TILE_M * 16;
  p.film = off; off += 2 * p.film_bytes;
  p.cvec = off; off += (uint32_t)align_up((size_t)L.cvec_floats * 4, 16);
  p.bars = off; off += 256;
  p.xch = off;
  p.xch_stride = 2;
  off += 2 * 2 * TILE_M * 8;           // [2 tiles][2 column halves][TILE_M] (row sum, sum of squares)
  off += 2 * 2 * TILE_M * L.DP * 4;    // [2 tiles][2 halves][TILE_M][DP] output-projector partials
  p.ring = (uint32_t)align_up(off, 128);
  int n = (int)((SMEM_LIMIT - p.ring) / p.stage_bytes);
  if (n > 8) n = 8;
  p.n_stage = n;
  p.total = p.ring + (n > 0 ? n : 0) * p.stage_bytes;
  return p;
}

}  // namespace bfk
