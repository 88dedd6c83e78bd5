// Training forward of one ConditionalDenseBlock (conditional_dense_block.py:114-131) as ONE tcgen05 kernel + one light
// normalisation pass, instead of Dense GEMM -> FiLM GEMM -> block kernel (r01ah: 25.6 + 25.6 + 71.5 us per block at
// 65,536 x 256, all three bound by fp32 round trips of P and of the FiLM projection through HBM).
//
// Per 128 x 128 output tile the kernel runs TWO contractions into two TMEM accumulators - P = h_prev . Wd over K = W and
// F = emb(t) . Wf over K = 64 (per-row time: the FiLM shift is a matrix, it cannot be hoisted as in sampling) - and the
// epilogue forms  u = dropout(act(P + bd)),  v = h_prev + u + F + bf  straight from TMEM.  It writes what backward needs
// (P in fp32) and the un-normalised v as an fp16 image, and adds each row's partial (sum, sum of squares) to the row
// statistics; `train_norm_kernel` then turns (v, statistics) into xhat / rstd / the normalised fp16 operand image of the
// next block (LayerNorm needs the whole row; W = 2..8 column tiles).
// Roles: warps 0-15 epilogue (TMEM lane quarter = warp % 4, 32-column group = warp / 4), warp 16 UMMA issuer + TMEM
// owner, warp 17 bulk-copy producer; two accumulator PAIRS so the epilogue of tile i runs under the MMAs of tile i+1.
#pragma once
#include <cuda_runtime.h>

#include "bfk_gemm.cuh"
#include "bfk_train.cuh"

namespace bfk {

constexpr int TF_NT = 128, TF_KC = 64, TF_STAGES = 4, TF_EPI_WARPS = 16, TF_THREADS = (TF_EPI_WARPS + 2) * 32;
constexpr uint32_t TF_SLAB = TF_KC * 128 * 2;  // bytes of one operand slab (128 rows x 64 K x fp16), A and B alike

struct TrainFusedArgs {
  const unsigned char* A;    // h_prev image (rows x W), also the skip source
  const unsigned char* Bd;   // dense weights, K-major, row tile 128
  const unsigned char* T;    // time-embedding image (rows x 64)
  const unsigned char* Bf;   // FiLM weights, K-major, row tile 128, K = 64
  const float* bias_d; const float* bias_f;
  float* P;                  // (rows, W) fp32 pre-activation incl. bias (saved for backward)
  unsigned char* v_img;      // un-normalised block output, fp16 image
  float* stats;              // [rows][2], zeroed by the host
  long long M;
  int W, act, residual, layer;
  float drop_rate, inv_keep; unsigned long long seed;
  uint32_t idesc;
};

__host__ __device__ inline size_t train_fused_smem_bytes() { return (size_t)TF_STAGES * 2 * TF_SLAB + 2 * 2 * TF_NT * 4 + 256; }

__global__ void __launch_bounds__(TF_THREADS, 1) train_block_gemm_kernel(const TrainFusedArgs a) {
  extern __shared__ __align__(128) unsigned char smem[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t st_bytes = 2 * TF_SLAB;
  const uint32_t smem_base = smem_u32(smem);
  const uint32_t vec_off = TF_STAGES * st_bytes;  // [2][2][128] floats: bias_d, bias_f per tile parity
  const uint32_t bar_off = vec_off + 2 * 2 * TF_NT * 4;
  const uint32_t bar0 = smem_base + bar_off;
  // full[s] +8s, empty[s] +8(4+s), acc_full[b] +64+8b, acc_empty[b] +80+8b, tmem slot +96
  volatile uint32_t* s_tmem = reinterpret_cast<volatile uint32_t*>(smem + bar_off + 96);
  const int n_tiles_n = a.W / TF_NT;
  const long long m_tiles = (a.M + 127) / 128;
  const long long total = m_tiles * n_tiles_n;
  const int n_slabs = a.W / TF_KC;  // dense slabs; slab index n_slabs is the FiLM contraction
  const size_t K8 = (size_t)a.W / 8, K8T = TF_KC / 8;

  if (threadIdx.x == 0) {
    for (int s = 0; s < TF_STAGES; ++s) { mbar_init(bar0 + 8 * s, 1); mbar_init(bar0 + 8 * (4 + s), 1); }
    for (int b = 0; b < 2; ++b) { mbar_init(bar0 + 64 + 8 * b, 1); mbar_init(bar0 + 80 + 8 * b, TF_EPI_WARPS); }
    fence_barrier_init();
  }
  if (warp == TF_EPI_WARPS) tmem_alloc(bar0 + 96, 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *s_tmem;

  if (warp == TF_EPI_WARPS + 1) {
    if (lane == 0) {
      uint32_t g = 0;
      for (long long t = blockIdx.x; t < total; t += gridDim.x) {
        const long long m_tile = t / n_tiles_n;
        const int n_tile = (int)(t % n_tiles_n);
        for (int i = 0; i <= n_slabs; ++i, ++g) {
          const uint32_t s = g % TF_STAGES;
          mbar_wait(bar0 + 8 * (4 + s), ((g / TF_STAGES) & 1) ^ 1, nullptr, 0);
          mbar_arrive_expect_tx(bar0 + 8 * s, st_bytes);
          const unsigned char *pa, *pb;
          if (i < n_slabs) {
            const size_t k8 = (size_t)i * (TF_KC / 8);
            pa = a.A + ((size_t)m_tile * K8 + k8) * 128 * 16;
            pb = a.Bd + ((size_t)n_tile * K8 + k8) * TF_NT * 16;
          } else {
            pa = a.T + (size_t)m_tile * K8T * 128 * 16;
            pb = a.Bf + (size_t)n_tile * K8T * TF_NT * 16;
          }
          bulk_g2s(smem_base + s * st_bytes, pa, TF_SLAB, bar0 + 8 * s);
          bulk_g2s(smem_base + s * st_bytes + TF_SLAB, pb, TF_SLAB, bar0 + 8 * s);
        }
      }
    }
    __syncwarp();
  } else if (warp == TF_EPI_WARPS) {
    if (lane == 0) {
      uint32_t g = 0, j = 0;
      for (long long t = blockIdx.x; t < total; t += gridDim.x, ++j) {
        const uint32_t acc = j & 1;
        mbar_wait(bar0 + 80 + 8 * acc, ((j >> 1) & 1) ^ 1, nullptr, 0);  // epilogue has drained this accumulator pair
        tc_fence_after();
        const uint32_t d_p = tmem_base + acc * 256, d_f = d_p + 128;
        for (int i = 0; i <= n_slabs; ++i, ++g) {
          const uint32_t s = g % TF_STAGES;
          mbar_wait(bar0 + 8 * s, (g / TF_STAGES) & 1, nullptr, 0);
          tc_fence_after();
          const uint32_t sa = smem_base + s * st_bytes, sb = sa + TF_SLAB;
          const bool film = i == n_slabs;
#pragma unroll
          for (int kk = 0; kk < TF_KC / 16; ++kk) {
            const uint64_t ad = umma_smem_desc(sa + (uint32_t)(kk * 2) * (128 * 16), 128 * 16, 128);
            const uint64_t bd = umma_smem_desc(sb + (uint32_t)(kk * 2) * (TF_NT * 16), TF_NT * 16, 128);
            umma_f16(film ? d_f : d_p, ad, bd, a.idesc, (uint32_t)(film ? (kk != 0) : ((i | kk) != 0)));
          }
          umma_commit(bar0 + 8 * (4 + s));
        }
        umma_commit(bar0 + 64 + 8 * acc);
      }
    }
    __syncwarp();
  } else {
    // ---- epilogue warps 0-15 ----------------------------------------------------------------------------------------
    const int q = warp & 3, cg = warp >> 2;  // 32-column group inside the 128-column tile
    float* vecs = reinterpret_cast<float*>(smem + vec_off);
    const int N8 = a.W / 8;
    uint32_t j = 0;
    for (long long t = blockIdx.x; t < total; t += gridDim.x, ++j) {
      const long long m_tile = t / n_tiles_n;
      const int n_tile = (int)(t % n_tiles_n);
      const uint32_t acc = j & 1;
      float* vb = vecs + (size_t)acc * 2 * TF_NT;
      for (int cidx = threadIdx.x; cidx < TF_NT; cidx += TF_EPI_WARPS * 32) {
        vb[cidx] = a.bias_d[n_tile * TF_NT + cidx];
        vb[TF_NT + cidx] = a.bias_f[n_tile * TF_NT + cidx];
      }
      const long long row = m_tile * 128 + q * 32 + lane;
      const bool valid = row < a.M;
      const int cbase = n_tile * TF_NT + cg * 32;  // first global column of this thread's 32
      uint4 raw[4];
      if (a.residual) {
#pragma unroll
        for (int o = 0; o < 4; ++o) raw[o] = *reinterpret_cast<const uint4*>(a.A + img_off(row, cbase / 8 + o, (int)K8));
      }
      named_bar_sync(1, TF_EPI_WARPS * 32);
      mbar_wait(bar0 + 64 + 8 * acc, (j >> 1) & 1, nullptr, 0);
      tc_fence_after();
      const uint32_t t_p = tmem_base + acc * 256 + cg * 32 + ((uint32_t)(q * 32) << 16), t_f = t_p + 128;
      float s1 = 0.f, s2 = 0.f;
#pragma unroll
      for (int hf = 0; hf < 2; ++hf) {
        float vp[16], vf[16];
        tmem_ld16(t_p + hf * 16, vp);
        tmem_ld16(t_f + hf * 16, vf);
        tmem_ld_wait16(vp);
        tmem_ld_wait16(vf);
        if (hf == 1) {
          // every TMEM read of this accumulator pair has landed: hand it back to the UMMA issuer before the arithmetic
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(bar0 + 80 + 8 * acc);
        }
#pragma unroll
        for (int oo = 0; oo < 2; ++oo) {
          const int o = hf * 2 + oo;
          float sk[8];
          if (a.residual) {
            const __half2* hp = reinterpret_cast<const __half2*>(&raw[o]);
#pragma unroll
            for (int i = 0; i < 4; ++i) { const float2 f = __half22float2(hp[i]); sk[2 * i] = f.x; sk[2 * i + 1] = f.y; }
          }
          float p8[8], out8[8];
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            const int ct = cg * 32 + o * 8 + i;  // column inside the tile
            const float P = vp[oo * 8 + i] + vb[ct];
            p8[i] = P;
            float d;
            float u = t_act(a.act, P, d);
            if (a.drop_rate > 0.f) u *= drop_scale(a.seed, a.layer, row, cbase + o * 8 + i, a.drop_rate, a.inv_keep);
            u += vf[oo * 8 + i] + vb[TF_NT + ct];
            if (a.residual) u += sk[i];
            out8[i] = u;
            s1 += u;
            s2 = fmaf(u, u, s2);
          }
          if (valid) st8(a.P + row * a.W + cbase + o * 8, p8);
          *reinterpret_cast<uint4*>(a.v_img + img_off(row, cbase / 8 + o, N8)) = pack8(out8);
        }
      }
      if (valid) {
        atomicAdd(a.stats + row * 2, s1);
        atomicAdd(a.stats + row * 2 + 1, s2);
      }
    }
  }
  __syncthreads();
  tc_fence_after();
  if (warp == TF_EPI_WARPS) tmem_dealloc(tmem_base, 512);
}

// (v image, row statistics) -> xhat (fp32, saved for backward and the LayerNorm gradients), rstd, and the normalised
// fp16 operand image h = gamma * xhat + beta of the next block.  norm none: xhat = h = v, rstd = 1.
// grid (K8, tile chunks), 128 threads = rows of one tile; rows >= B get a zero image (K padding of the dW contraction).
__global__ void train_norm_kernel(const unsigned char* __restrict__ v_img, const float* __restrict__ stats,
                                  const float* __restrict__ gamma, const float* __restrict__ beta, long long B,
                                  long long n_tiles, int K8, int norm, float* __restrict__ xhat, float* __restrict__ rstd,
                                  unsigned char* __restrict__ h_img) {
  const int k8 = blockIdx.x;
  float g[8], b[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) { g[i] = 1.f; b[i] = 0.f; }
  if (gamma) { ld8(gamma + k8 * 8, g); ld8(beta + k8 * 8, b); }
  const int W = K8 * 8;
  const float inv_w = 1.f / (float)W;
  for (long long tile = blockIdx.y; tile < n_tiles; tile += gridDim.y) {
    const long long r = tile * 128 + threadIdx.x;
    float h[8];
    if (r < B) {
      float mu = 0.f, rs = 1.f;
      if (norm == BFK_NORM_LAYER) {
        const float2 st = *reinterpret_cast<const float2*>(stats + r * 2);
        mu = st.x * inv_w;
        rs = rsqrtf(fmaxf(fmaf(st.y, inv_w, -mu * mu), 0.f) + LN_EPS);
      }
      if (k8 == 0) rstd[r] = rs;
      const uint4 raw = *reinterpret_cast<const uint4*>(v_img + img_off(r, k8, K8));
      const __half2* hp = reinterpret_cast<const __half2*>(&raw);
      float xh[8];
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const float2 f = __half22float2(hp[i]);
        xh[2 * i] = (f.x - mu) * rs;
        xh[2 * i + 1] = (f.y - mu) * rs;
      }
      st8(xhat + r * W + k8 * 8, xh);
#pragma unroll
      for (int i = 0; i < 8; ++i) h[i] = fmaf(xh[i], g[i], b[i]);
    } else {
#pragma unroll
      for (int i = 0; i < 8; ++i) h[i] = 0.f;
    }
    *reinterpret_cast<uint4*>(h_img + img_off(r, k8, K8)) = pack8(h);
  }
}

}  // namespace bfk
