// Layer-wise tensor-core GEMM for the paths that cannot live in one SM (training forward/backward with per-row time,
// widths > 256):  C[M,N] (fp32, row-major)  (+)=  alpha * A[M,K] * B[K,N]  (+ bias[N]),  fp16/bf16 operands, fp32
// accumulate in TMEM.  Both operands arrive PRE-PACKED in the K-major no-swizzle UMMA image
//     P[r / RT][k / 8][r % RT][k % 8]            (RT = 128 for the A role, RT = NT for the B role)
// so that any (row tile, K slab) is one contiguous byte range: one cp.async.bulk per operand per stage, LBO = RT*16,
// SBO = 128.  The producers of activations write this layout directly (bfk_train.cuh), weights are packed once.
//
// One CTA per (M tile of 128, N tile of NT, K split); 192 threads: warps 0-3 epilogue (thread == row == TMEM lane),
// warp 4 UMMA issuer + TMEM owner, warp 5 bulk-copy producer.  4-stage ring of K=32 slabs (24 KB per stage at NT=256)
// so that two CTAs share an SM and one CTA's epilogue overlaps the other's main loop.
#pragma once
#include <cuda_runtime.h>

#include "bfk_ptx.cuh"

namespace bfk {

constexpr int G_KC = 32, G_STAGES = 4, G_THREADS = 192, G_TM = 128;
enum { GEMM_ATOMIC = 1, GEMM_ADD_EXISTING = 2 };

struct GemmArgs {
  const unsigned char* A;  // packed, row tile 128
  const unsigned char* B;  // packed, row tile NT
  float* C;
  const float* bias;  // [N] or null (added once: by split 0)
  long long M, ldc;   // valid rows of C, leading dimension
  int N, K, NT, k_per_split, flags;
  float alpha;
  uint32_t idesc;
};

__host__ __device__ inline size_t gemm_smem_bytes(int NT) { return (size_t)G_STAGES * (G_KC * G_TM * 2 + G_KC * NT * 2) + 256; }

template <bool BF16_UNUSED = false>
__global__ void __launch_bounds__(G_THREADS, 2) gemm_kernel(const GemmArgs a) {
  extern __shared__ __align__(128) unsigned char smem[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int NT = a.NT;
  const uint32_t a_bytes = G_KC * G_TM * 2, b_bytes = (uint32_t)G_KC * NT * 2, st_bytes = a_bytes + b_bytes;
  const uint32_t smem_base = smem_u32(smem);
  const uint32_t bar0 = smem_base + G_STAGES * st_bytes;  // full[s] at +8s, empty[s] at +8(4+s), acc at +64, tmem slot +72
  volatile uint32_t* s_tmem = reinterpret_cast<volatile uint32_t*>(smem + G_STAGES * st_bytes + 72);
  const int m_tile = blockIdx.x, n_tile = blockIdx.y, split = blockIdx.z;
  const int k0 = split * a.k_per_split;
  const int k1 = min(a.K, k0 + a.k_per_split);
  const int n_slabs = (k1 - k0) / G_KC;

  if (threadIdx.x == 0) {
    for (int s = 0; s < G_STAGES; ++s) { mbar_init(bar0 + 8 * s, 1); mbar_init(bar0 + 8 * (4 + s), 1); }
    mbar_init(bar0 + 64, 1);
    fence_barrier_init();
  }
  const uint32_t tmem_cols = NT <= 32 ? 32 : NT <= 64 ? 64 : NT <= 128 ? 128 : 256;
  if (warp == 4) tmem_alloc(bar0 + 72, tmem_cols);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *s_tmem;

  if (warp == 5) {
    if (lane == 0) {
      const size_t K8 = (size_t)a.K / 8;
      for (int i = 0; i < n_slabs; ++i) {
        const int s = i % G_STAGES;
        mbar_wait(bar0 + 8 * (4 + s), ((i / G_STAGES) & 1) ^ 1, nullptr, 0);
        mbar_arrive_expect_tx(bar0 + 8 * s, st_bytes);
        const size_t k8 = (size_t)(k0 + i * G_KC) / 8;
        bulk_g2s(smem_base + s * st_bytes, a.A + ((size_t)m_tile * K8 + k8) * G_TM * 16, a_bytes, bar0 + 8 * s);
        bulk_g2s(smem_base + s * st_bytes + a_bytes, a.B + ((size_t)n_tile * K8 + k8) * NT * 16, b_bytes, bar0 + 8 * s);
      }
    }
    __syncwarp();
  } else if (warp == 4) {
    if (lane == 0) {
      for (int i = 0; i < n_slabs; ++i) {
        const int s = i % G_STAGES;
        mbar_wait(bar0 + 8 * s, (i / G_STAGES) & 1, nullptr, 0);
        tc_fence_after();
        const uint32_t sa = smem_base + s * st_bytes, sb = sa + a_bytes;
#pragma unroll
        for (int kk = 0; kk < G_KC / 16; ++kk) {
          const uint64_t ad = umma_smem_desc(sa + (uint32_t)(kk * 2) * (G_TM * 16), G_TM * 16, 128);
          const uint64_t bd = umma_smem_desc(sb + (uint32_t)(kk * 2) * (NT * 16), (uint32_t)NT * 16, 128);
          umma_f16(tmem_base, ad, bd, a.idesc, (uint32_t)((i | kk) != 0));
        }
        umma_commit(bar0 + 8 * (4 + s));
      }
      umma_commit(bar0 + 64);
    }
    __syncwarp();
  } else {
    // epilogue.  TMEM hands each thread one ROW; written straight to global that is 32 rows x 16 B per store
    // instruction (32 half-used sectors).  So 64-column panels are transposed through the (now idle) ring: thread == row
    // on the way in, 16 lanes x float4 == 256 contiguous bytes of one row on the way out.
    const uint32_t t_acc = tmem_base + ((uint32_t)(warp * 32) << 16);
    if (n_slabs > 0) {
      mbar_wait(bar0 + 64, 0, nullptr, 0);
      tc_fence_after();
    }
    float* stg = reinterpret_cast<float*>(smem) + warp * (32 * 68);
    const long long row0 = (long long)m_tile * G_TM + warp * 32;
    const int sub = lane >> 4, c4 = (lane & 15) * 4;
    const float* bias = (a.bias && split == 0) ? a.bias + (size_t)n_tile * NT : nullptr;
    for (int p = 0; p < NT / 64; ++p) {
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        float v[32];
        if (n_slabs > 0) {
          tmem_ld32(t_acc + p * 64 + h * 32, v);
          tmem_ld_wait32(v);
        } else {
#pragma unroll
          for (int i = 0; i < 32; ++i) v[i] = 0.f;
        }
#pragma unroll
        for (int i = 0; i < 32; i += 4) *reinterpret_cast<float4*>(stg + lane * 68 + h * 32 + i) = make_float4(v[i], v[i + 1], v[i + 2], v[i + 3]);
      }
      __syncwarp();
      float4 b4 = make_float4(0.f, 0.f, 0.f, 0.f);
      if (bias) b4 = *reinterpret_cast<const float4*>(bias + p * 64 + c4);
      float4* dst[16];
      float4 o[16];
#pragma unroll
      for (int it = 0; it < 16; ++it) {
        const int rr = it * 2 + sub;
        dst[it] = reinterpret_cast<float4*>(a.C + (row0 + rr) * a.ldc + (size_t)n_tile * NT + p * 64 + c4);
        const float4 t = *reinterpret_cast<const float4*>(stg + rr * 68 + c4);
        o[it] = make_float4(fmaf(t.x, a.alpha, b4.x), fmaf(t.y, a.alpha, b4.y), fmaf(t.z, a.alpha, b4.z), fmaf(t.w, a.alpha, b4.w));
      }
      if (a.flags & GEMM_ATOMIC) {
#pragma unroll
        for (int it = 0; it < 16; ++it)
          if (row0 + it * 2 + sub < a.M) atomicAdd(dst[it], o[it]);
      } else {
        if (a.flags & GEMM_ADD_EXISTING) {  // all 16 loads of the existing tile in flight before the first add
          float4 e[16];
#pragma unroll
          for (int it = 0; it < 16; ++it) e[it] = (row0 + it * 2 + sub < a.M) ? *dst[it] : make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
          for (int it = 0; it < 16; ++it) { o[it].x += e[it].x; o[it].y += e[it].y; o[it].z += e[it].z; o[it].w += e[it].w; }
        }
#pragma unroll
        for (int it = 0; it < 16; ++it)
          if (row0 + it * 2 + sub < a.M) *dst[it] = o[it];
      }
      __syncwarp();
    }
    tc_fence_before();
  }
  __syncthreads();
  tc_fence_after();
  if (warp == 4) tmem_dealloc(tmem_base, tmem_cols);
}

// ---------------------------------------------------------------------------------------------------------------
// Persistent form of the same GEMM (no split-K): one CTA per SM walks the output tiles (n fastest, so the CTAs that run
// together share an A tile in L2), TWO accumulators in TMEM so that the epilogue of tile i runs under the main loop of
// tile i+1, K = 64 per ring stage (4 UMMAs per hand-shake), 8 epilogue warps (TMEM lane quarter = warp % 4, odd / even
// 64-column panels = warp / 4).  r01w: the one-tile-per-CTA kernel reached 34 % tensor-pipe at 8192 x 1024 x 1024
// because prologue and epilogue of co-resident CTAs coincide instead of overlapping.
constexpr int P_KC = 64, P_STAGES = 3, P_THREADS = 320, P_EPI_WARPS = 8;
__host__ __device__ inline size_t gemm_persist_smem_bytes(int NT) {
  return (size_t)P_STAGES * (P_KC * G_TM * 2 + P_KC * NT * 2) + (size_t)P_EPI_WARPS * 32 * 68 * 4 + 256;
}

__global__ void __launch_bounds__(P_THREADS, 1) gemm_persist_kernel(const GemmArgs a) {
  extern __shared__ __align__(128) unsigned char smem[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int NT = a.NT;
  const uint32_t a_bytes = P_KC * G_TM * 2, b_bytes = (uint32_t)P_KC * NT * 2, st_bytes = a_bytes + b_bytes;
  const uint32_t smem_base = smem_u32(smem);
  const uint32_t stg_off = P_STAGES * st_bytes;
  const uint32_t bar0 = smem_base + stg_off + P_EPI_WARPS * 32 * 68 * 4;
  // full[s] +8s, empty[s] +8(3+s), acc_full[b] +48+8b, acc_empty[b] +64+8b, tmem slot +80
  volatile uint32_t* s_tmem = reinterpret_cast<volatile uint32_t*>(smem + stg_off + P_EPI_WARPS * 32 * 68 * 4 + 80);
  const int n_tiles_n = a.N / NT;
  const long long m_tiles = (a.M + G_TM - 1) / G_TM;
  const long long total = m_tiles * n_tiles_n;
  const int n_slabs = a.K / P_KC;
  const size_t K8 = (size_t)a.K / 8;

  if (threadIdx.x == 0) {
    for (int s = 0; s < P_STAGES; ++s) { mbar_init(bar0 + 8 * s, 1); mbar_init(bar0 + 8 * (3 + s), 1); }
    for (int b = 0; b < 2; ++b) { mbar_init(bar0 + 48 + 8 * b, 1); mbar_init(bar0 + 64 + 8 * b, P_EPI_WARPS); }
    fence_barrier_init();
  }
  const uint32_t tmem_cols = NT <= 64 ? 128 : (NT <= 128 ? 256 : 512);
  const uint32_t acc_stride = tmem_cols / 2;
  if (warp == 8) tmem_alloc(bar0 + 80, tmem_cols);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *s_tmem;

  if (warp == 9) {
    if (lane == 0) {
      uint32_t g = 0;
      for (long long t = blockIdx.x; t < total; t += gridDim.x) {
        const long long m_tile = t / n_tiles_n;
        const int n_tile = (int)(t % n_tiles_n);
        for (int i = 0; i < n_slabs; ++i, ++g) {
          const uint32_t s = g % P_STAGES;
          mbar_wait(bar0 + 8 * (3 + s), ((g / P_STAGES) & 1) ^ 1, nullptr, 0);
          mbar_arrive_expect_tx(bar0 + 8 * s, st_bytes);
          const size_t k8 = (size_t)i * (P_KC / 8);
          bulk_g2s(smem_base + s * st_bytes, a.A + ((size_t)m_tile * K8 + k8) * G_TM * 16, a_bytes, bar0 + 8 * s);
          bulk_g2s(smem_base + s * st_bytes + a_bytes, a.B + ((size_t)n_tile * K8 + k8) * NT * 16, b_bytes, bar0 + 8 * s);
        }
      }
    }
    __syncwarp();
  } else if (warp == 8) {
    if (lane == 0) {
      uint32_t g = 0, j = 0;
      for (long long t = blockIdx.x; t < total; t += gridDim.x, ++j) {
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
          umma_commit(bar0 + 8 * (3 + s));
        }
        umma_commit(bar0 + 48 + 8 * acc);
      }
    }
    __syncwarp();
  } else {
    const int q = warp & 3, half = warp >> 2;
    float* stg = reinterpret_cast<float*>(smem + stg_off) + warp * (32 * 68);
    const int sub = lane >> 4, c4 = (lane & 15) * 4;
    uint32_t j = 0;
    for (long long t = blockIdx.x; t < total; t += gridDim.x, ++j) {
      const long long m_tile = t / n_tiles_n;
      const int n_tile = (int)(t % n_tiles_n);
      const uint32_t acc = j & 1;
      mbar_wait(bar0 + 48 + 8 * acc, (j >> 1) & 1, nullptr, 0);
      tc_fence_after();
      const uint32_t t_acc = tmem_base + acc * acc_stride + ((uint32_t)(q * 32) << 16);
      const long long row0 = m_tile * G_TM + q * 32;
      const float* bias = a.bias ? a.bias + (size_t)n_tile * NT : nullptr;
      for (int p = half; p < NT / 64; p += 2) {
#pragma unroll
        for (int h = 0; h < 2; ++h) {
          float v[32];
          tmem_ld32(t_acc + p * 64 + h * 32, v);
          tmem_ld_wait32(v);
#pragma unroll
          for (int i = 0; i < 32; i += 4) *reinterpret_cast<float4*>(stg + lane * 68 + h * 32 + i) = make_float4(v[i], v[i + 1], v[i + 2], v[i + 3]);
        }
        __syncwarp();
        float4 b4 = make_float4(0.f, 0.f, 0.f, 0.f);
        if (bias) b4 = *reinterpret_cast<const float4*>(bias + p * 64 + c4);
        float4* dst[16];
        float4 o[16];
#pragma unroll
        for (int it = 0; it < 16; ++it) {
          const int rr = it * 2 + sub;
          dst[it] = reinterpret_cast<float4*>(a.C + (row0 + rr) * a.ldc + (size_t)n_tile * NT + p * 64 + c4);
          const float4 tt = *reinterpret_cast<const float4*>(stg + rr * 68 + c4);
          o[it] = make_float4(fmaf(tt.x, a.alpha, b4.x), fmaf(tt.y, a.alpha, b4.y), fmaf(tt.z, a.alpha, b4.z), fmaf(tt.w, a.alpha, b4.w));
        }
        if (a.flags & GEMM_ADD_EXISTING) {
          float4 e[16];
#pragma unroll
          for (int it = 0; it < 16; ++it) e[it] = (row0 + it * 2 + sub < a.M) ? *dst[it] : make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
          for (int it = 0; it < 16; ++it) { o[it].x += e[it].x; o[it].y += e[it].y; o[it].z += e[it].z; o[it].w += e[it].w; }
        }
#pragma unroll
        for (int it = 0; it < 16; ++it)
          if (row0 + it * 2 + sub < a.M) *dst[it] = o[it];
        __syncwarp();
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(bar0 + 64 + 8 * acc);
    }
  }
  __syncthreads();
  tc_fence_after();
  if (warp == 8) tmem_dealloc(tmem_base, tmem_cols);
}

// ---------------------------------------------------------------------------------------------------------------
// dW GEMM: C[Mf, N] += alpha * X^T * Y with the contraction over the BATCH rows.  X and Y are activation images
// IMG[r/128][k/8][r%128][k%8] (r = batch row, k = feature).  Read as an MN-major UMMA operand (MN = feature, K = batch
// row) the very same bytes are the canonical no-swizzle layout: core matrix = 8 batch rows x 8 features (128 B), next
// 8 batch rows +128 B (LBO), next feature octet +2048 B (SBO) - so the images the forward pass wrote are consumed
// without a transposed re-pack.  One stage = one 128-row batch tile: X part 16 octets (32 KB), Y part NT/8 octets,
// each ONE contiguous bulk copy.  Split over batch tiles (blockIdx.z), fp32 atomics into a zeroed C.
constexpr int DW_STAGES = 3;
struct GemmDwArgs {
  const unsigned char* X; const unsigned char* Y;
  float* C; long long ldc;
  int XK8, YK8;  // feature octets per row in the images
  int Mv, Nv;    // valid rows / columns of C
  int NT, n_btiles, tiles_per_split, swap;
  float alpha;
  uint32_t idesc;
};
__host__ __device__ inline size_t gemm_dw_smem_bytes(int NT) { return (size_t)DW_STAGES * (32768 + (size_t)NT * 256) + 256; }
__host__ __device__ constexpr uint32_t umma_idesc_f16_mn(int M, int N, int ab_format) {
  return umma_idesc_f16(M, N, ab_format) | (1u << 15) | (1u << 16);  // a_major = b_major = MN
}

__global__ void __launch_bounds__(G_THREADS, 1) gemm_dw_kernel(const GemmDwArgs a) {
  extern __shared__ __align__(128) unsigned char smem[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int NT = a.NT;
  const uint32_t a_bytes = 32768, b_bytes = (uint32_t)NT * 256, st_bytes = a_bytes + b_bytes;
  const uint32_t smem_base = smem_u32(smem);
  const uint32_t bar0 = smem_base + DW_STAGES * st_bytes;
  volatile uint32_t* s_tmem = reinterpret_cast<volatile uint32_t*>(smem + DW_STAGES * st_bytes + 72);
  const int m_tile = blockIdx.x, n_tile = blockIdx.y, split = blockIdx.z;
  const int t0 = split * a.tiles_per_split;
  const int n_it = min(a.n_btiles, t0 + a.tiles_per_split) - t0;
  if (n_it <= 0) return;
  const int a_oct = min(16, a.XK8 - m_tile * 16), b_oct = min(NT / 8, a.YK8 - n_tile * (NT / 8));
  if (a_oct < 16 || b_oct < NT / 8) {  // partial tiles: the never-written octets must read as zero
    for (uint32_t i = threadIdx.x; i < DW_STAGES * st_bytes / 16; i += blockDim.x) reinterpret_cast<uint4*>(smem)[i] = make_uint4(0, 0, 0, 0);
    fence_proxy_async_smem();
  }
  if (threadIdx.x == 0) {
    for (int s = 0; s < DW_STAGES; ++s) { mbar_init(bar0 + 8 * s, 1); mbar_init(bar0 + 8 * (4 + s), 1); }
    mbar_init(bar0 + 64, 1);
    fence_barrier_init();
  }
  const uint32_t tmem_cols = NT <= 32 ? 32 : NT <= 64 ? 64 : NT <= 128 ? 128 : 256;
  if (warp == 4) tmem_alloc(bar0 + 72, tmem_cols);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *s_tmem;

  if (warp == 5) {
    if (lane == 0) {
      for (int i = 0; i < n_it; ++i) {
        const int s = i % DW_STAGES;
        mbar_wait(bar0 + 8 * (4 + s), ((i / DW_STAGES) & 1) ^ 1, nullptr, 0);
        mbar_arrive_expect_tx(bar0 + 8 * s, (uint32_t)(a_oct + b_oct) * 2048);
        const size_t t = (size_t)(t0 + i);
        bulk_g2s(smem_base + s * st_bytes, a.X + (t * a.XK8 + (size_t)m_tile * 16) * 2048, (uint32_t)a_oct * 2048, bar0 + 8 * s);
        bulk_g2s(smem_base + s * st_bytes + a_bytes, a.Y + (t * a.YK8 + (size_t)n_tile * (NT / 8)) * 2048, (uint32_t)b_oct * 2048,
                 bar0 + 8 * s);
      }
    }
    __syncwarp();
  } else if (warp == 4) {
    if (lane == 0) {
      const uint32_t lbo = a.swap ? 2048 : 128, sbo = a.swap ? 128 : 2048;
      for (int i = 0; i < n_it; ++i) {
        const int s = i % DW_STAGES;
        mbar_wait(bar0 + 8 * s, (i / DW_STAGES) & 1, nullptr, 0);
        tc_fence_after();
        const uint32_t sa = smem_base + s * st_bytes, sb = sa + a_bytes;
#pragma unroll
        for (int kk = 0; kk < 8; ++kk) {  // 16 batch rows per UMMA = two 128-byte core-matrix rows
          const uint64_t ad = umma_smem_desc(sa + (uint32_t)kk * 256, lbo, sbo);
          const uint64_t bd = umma_smem_desc(sb + (uint32_t)kk * 256, lbo, sbo);
          umma_f16(tmem_base, ad, bd, a.idesc, (uint32_t)((i | kk) != 0));
        }
        umma_commit(bar0 + 8 * (4 + s));
      }
      umma_commit(bar0 + 64);
    }
    __syncwarp();
  } else {
    const int m = m_tile * G_TM + warp * 32 + lane;
    const uint32_t t_acc = tmem_base + ((uint32_t)(warp * 32) << 16);
    mbar_wait(bar0 + 64, 0, nullptr, 0);
    tc_fence_after();
    const bool vec = (a.ldc % 4 == 0) && (a.Nv % 4 == 0);
    for (int c = 0; c < NT / 32; ++c) {
      float v[32];
      tmem_ld32(t_acc + c * 32, v);
      tmem_ld_wait32(v);
      if (m >= a.Mv) continue;
      const int n0 = n_tile * NT + c * 32;
      float* crow = a.C + (long long)m * a.ldc + n0;
      if (vec) {
#pragma unroll
        for (int i = 0; i < 32; i += 4)
          if (n0 + i < a.Nv) atomicAdd(reinterpret_cast<float4*>(crow + i), make_float4(v[i] * a.alpha, v[i + 1] * a.alpha, v[i + 2] * a.alpha, v[i + 3] * a.alpha));
      } else {
#pragma unroll
        for (int i = 0; i < 32; ++i)
          if (n0 + i < a.Nv) atomicAdd(crow + i, v[i] * a.alpha);
      }
    }
    tc_fence_before();
  }
  __syncthreads();
  tc_fence_after();
  if (warp == 4) tmem_dealloc(tmem_base, tmem_cols);
}

// src[R][ld] fp32 row-major, contraction along columns (Kc) -> P[r/RT][k/8][r%RT][k%8]; rows >= R are zero filled up to
// R_pad (multiple of RT); columns >= Kc_valid zero filled up to Kc (multiple of 8).
__global__ void pack_kmajor_kernel(const float* __restrict__ src, long long ld, long long R, long long R_pad, int Kc_valid,
                                   int Kc, int RT, int bf16, uint16_t* __restrict__ dst) {
  const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;  // one thread per (r, k8): 16 bytes out
  const int K8 = Kc / 8;
  if (idx >= R_pad * K8) return;
  // consecutive threads -> consecutive rows of the same k8 (coalesced 16-byte stores)
  const long long tile = idx / ((long long)K8 * RT);
  const long long rem = idx % ((long long)K8 * RT);
  const int k8 = (int)(rem / RT), rin = (int)(rem % RT);
  const long long r = tile * RT + rin;
  uint16_t out[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const int k = k8 * 8 + i;
    const float v = (r < R && k < Kc_valid) ? src[r * ld + k] : 0.f;
    if (bf16) { __nv_bfloat16 b = __float2bfloat16_rn(v); out[i] = *reinterpret_cast<uint16_t*>(&b); }
    else { __half h = __float2half_rn(v); out[i] = *reinterpret_cast<uint16_t*>(&h); }
  }
  *reinterpret_cast<uint4*>(dst + (size_t)idx * 8) = *reinterpret_cast<uint4*>(out);
}

// src[Kc][ld] fp32 row-major, contraction along ROWS (Kc), operand rows = columns f of src (F of them)
//   -> P[f/RT][k/8][f%RT][k%8];  f >= F zero filled up to F_pad, k >= Kc_valid zero filled up to Kc.
__global__ void pack_kmajor_t_kernel(const float* __restrict__ src, long long ld, long long Kc_valid, long long Kc, int F,
                                     int F_pad, int RT, int bf16, uint16_t* __restrict__ dst) {
  const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;  // one thread per (f, k8)
  const long long K8 = Kc / 8;
  if (idx >= (long long)F_pad * K8) return;
  const long long tile = idx / (K8 * RT);
  const long long rem = idx % (K8 * RT);
  const long long k8 = rem / RT;
  const int fin = (int)(rem % RT);
  const int f = (int)(tile * RT + fin);  // consecutive threads -> consecutive f: coalesced reads of src rows
  uint16_t out[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const long long k = k8 * 8 + i;
    const float v = (f < F && k < Kc_valid) ? src[k * ld + f] : 0.f;
    if (bf16) { __nv_bfloat16 b = __float2bfloat16_rn(v); out[i] = *reinterpret_cast<uint16_t*>(&b); }
    else { __half h = __float2half_rn(v); out[i] = *reinterpret_cast<uint16_t*>(&h); }
  }
  *reinterpret_cast<uint4*>(dst + (size_t)idx * 8) = *reinterpret_cast<uint4*>(out);
}

// Several weight images in ONE launch (the training step re-packs ~25 small matrices every step; as separate launches they
// are a quarter of the step's launch count).  blockIdx.y = job; the bodies are the two kernels above.
struct PackJob {
  const float* src; uint16_t* dst;
  long long ld, a, b, n;   // t == 0: a = R, b = R_pad;  t == 1: a = Kc_valid, b = Kc;  n = 16-byte outputs of the job
  int c, d, RT, t;         // t == 0: c = Kc_valid, d = Kc;  t == 1: c = F, d = F_pad
};
constexpr int PACK_MAX_JOBS = 68;  // 4 * BFK_MAX_BLOCKS + 4 (needs the > 4 KB kernel-parameter space of CUDA 12.1+)
struct PackBatch { PackJob j[PACK_MAX_JOBS]; int n; };
__global__ void pack_batch_kernel(const PackBatch pb) {
  const PackJob& q = pb.j[blockIdx.y];
  const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= q.n) return;
  uint16_t out[8];
  if (q.t == 0) {
    const int K8 = q.d / 8;
    const long long tile = idx / ((long long)K8 * q.RT), rem = idx % ((long long)K8 * q.RT);
    const int k8 = (int)(rem / q.RT), rin = (int)(rem % q.RT);
    const long long r = tile * q.RT + rin;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const int k = k8 * 8 + i;
      const float v = (r < q.a && k < q.c) ? q.src[r * q.ld + k] : 0.f;
      __half h = __float2half_rn(v);
      out[i] = *reinterpret_cast<uint16_t*>(&h);
    }
  } else {
    const long long K8 = q.b / 8;
    const long long tile = idx / (K8 * q.RT), rem = idx % (K8 * q.RT);
    const long long k8 = rem / q.RT;
    const int f = (int)(tile * q.RT + (int)(rem % q.RT));
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const long long k = k8 * 8 + i;
      const float v = (f < q.c && k < q.a) ? q.src[k * q.ld + f] : 0.f;
      __half h = __float2half_rn(v);
      out[i] = *reinterpret_cast<uint16_t*>(&h);
    }
  }
  *reinterpret_cast<uint4*>(q.dst + (size_t)idx * 8) = *reinterpret_cast<uint4*>(out);
}

}  // namespace bfk
