// Layer-wise training path (FlowMatching.compute_metrics / DiffusionModel.compute_metrics, flow_matching.py:132-178,
// diffusion_model.py:142-232, and the torch autograd transposes the reference gets from torch_approximator.py:88-99).
// The dense contractions run in bfk_gemm.cuh (tcgen05); the kernels here are everything BETWEEN two GEMMs, each one a
// single pass over a (rows, width) activation: they read the fp32 GEMM output and write (a) what backward needs in fp32
// and (b) the next GEMM's operand directly in the packed fp16 K-major image
//     IMG[r/128][k/8][r%128][k%8]
// which serves three roles without re-packing: A operand (K-major) of the next forward GEMM / of the dX GEMM, and BOTH
// operands (MN-major) of the dW GEMM whose contraction runs over the batch rows (gemm_dw_kernel).
// Gradients are propagated UN-normalised (the 1/(B*D) of the mean is applied where a gradient leaves the chain), which
// keeps the fp16 gradient operands at O(1) magnitudes.
#pragma once
#include <cuda_runtime.h>

#include "bfk_ptx.cuh"
#include "../../include/bfk.h"

namespace bfk {

#ifndef BFK_BWD_MINB
#define BFK_BWD_MINB 2  // resident CTAs per SM asked of ptxas for the W <= 256 block kernels: 3 / 4 spill and measured
                        // slower (profiles/r02ab_train_variants.md)
#endif
#ifndef BFK_BLOCK_PREFETCH
#define BFK_BLOCK_PREFETCH 1  // W <= 256 block kernels: next row's streams fetched under this row's reductions
#endif
#ifndef BFK_FWD_MINB
#define BFK_FWD_MINB 3
#endif
constexpr float LN_EPS = 1e-3f;  // keras.layers.LayerNormalization default epsilon
constexpr float RMS_EPS = 1e-6f; // keras.layers.RMSNormalization (ops.rms_normalization): x * rsqrt(mean(x^2) + eps) * scale

// value and derivative of the Keras activation (time_mlp.py:106-109), fp32
__device__ __forceinline__ float t_act(int act, float x, float& d) {
  switch (act) {
    case BFK_ACT_MISH: {
      // x*tanh(softplus(x)) with tanh(ln(1+e)) = w/(w+2), w = e*(e+2), e = exp(x): one ex2 + two rcp (fp32 accurate to a
      // few ulp; libm tanhf/log1pf made the block kernels issue-bound)
      if (x > 15.f) { d = 1.f; return x; }
      const float e = __expf(x);
      const float w = e * (e + 2.f);
      const float t = __fdividef(w, w + 2.f);
      const float sg = __fdividef(e, 1.f + e);
      d = fmaf(x * (1.f - t * t), sg, t);
      return x * t;
    }
    case BFK_ACT_SILU: {
      const float sg = __fdividef(1.f, 1.f + __expf(-x));
      d = sg * fmaf(x, 1.f - sg, 1.f);
      return x * sg;
    }
    case BFK_ACT_RELU:
      d = x > 0.f ? 1.f : 0.f;
      return x > 0.f ? x : 0.f;
    case BFK_ACT_GELU: {
      // keras.activations.gelu defaults to approximate=True (tanh form)
      const float u = 0.7978845608028654f * (x + 0.044715f * x * x * x);
      const float t = tanhf(u);
      const float cdf = 0.5f * (1.f + t);
      d = cdf + 0.5f * x * (1.f - t * t) * 0.7978845608028654f * (1.f + 3.f * 0.044715f * x * x);
      return x * cdf;
    }
    default: {  // tanh
      const float t = tanhf(x);
      d = 1.f - t * t;
      return t;
    }
  }
}

// 8 values at once with the (warp-uniform, run-time) activation switch taken ONCE: inside the unrolled element loop the
// switch was re-evaluated per element (7 % of block_fwd_kernel's instructions)
__device__ __forceinline__ void t_act8(int act, const float* x, float* u, float* d) {
#define BFK_ACT8(A)                                              \
  _Pragma("unroll") for (int i = 0; i < 8; ++i) u[i] = t_act(A, x[i], d[i]); \
  break;
  switch (act) {
    case BFK_ACT_MISH: BFK_ACT8(BFK_ACT_MISH)
    case BFK_ACT_SILU: BFK_ACT8(BFK_ACT_SILU)
    case BFK_ACT_RELU: BFK_ACT8(BFK_ACT_RELU)
    case BFK_ACT_GELU: BFK_ACT8(BFK_ACT_GELU)
    default: BFK_ACT8(BFK_ACT_TANH)
  }
#undef BFK_ACT8
}

// Counter-based keep mask for Dropout (conditional_dense_block.py:120-121): identical in forward and backward, a function
// of (seed, layer, GLOBAL row, column) only, so a data-parallel shard draws the mask of the same rows of the whole batch.
// The seed and layer are avalanched into a 64-bit key first (splitmix64), so consecutive seeds give unrelated masks; the
// index of a column pair then goes through two multiply-xorshift rounds.
__device__ __forceinline__ unsigned long long drop_key(unsigned long long seed, const uint32_t* seed_dev, int layer) {
  if (seed_dev) seed = ((unsigned long long)seed_dev[1] << 32) | seed_dev[0];  // graph replay: the seed lives in HBM
  unsigned long long z = seed + 0x9E3779B97F4A7C15ull * (unsigned long long)(layer + 1);
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  return z ^ (z >> 31);
}
// Keep scales of the 8 consecutive columns c0 .. c0+7 (c0 a multiple of 8).  One hash serves a PAIR of columns (the top
// 24 bits of each 32-bit half), and the uniform is compared as an integer: k / 2^24 >= rate  <=>  k >= ceil(rate * 2^24),
// exact because both sides are exact in fp32.  Per-element hashing with an int->float convert was a third of the
// instructions of block_fwd_kernel / block_bwd_kernel (profiles/r02aa_train_blocks.md).
__device__ __forceinline__ void drop_scale8(unsigned long long key, long long row, int c0, float rate, float inv_keep,
                                            float* ds) {
  const uint32_t thr = (uint32_t)ceilf(rate * 16777216.f);
  const unsigned long long base = (unsigned long long)row * 512ull + (unsigned long long)(c0 >> 1);
#pragma unroll
  for (int q = 0; q < 4; ++q) {
    unsigned long long x = (base + (unsigned long long)q) ^ key;
    x *= 0xD6E8FEB86659FD93ull; x ^= x >> 32;
    x *= 0xD6E8FEB86659FD93ull; x ^= x >> 32;
    ds[2 * q] = ((uint32_t)x >> 8) >= thr ? inv_keep : 0.f;
    ds[2 * q + 1] = (uint32_t)(x >> 40) >= thr ? inv_keep : 0.f;
  }
}

__device__ __forceinline__ uint32_t h2pack(float a, float b) {
  __half2 h = __floats2half2_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&h);
}
__device__ __forceinline__ uint4 pack8(const float* v) {
  return make_uint4(h2pack(v[0], v[1]), h2pack(v[2], v[3]), h2pack(v[4], v[5]), h2pack(v[6], v[7]));
}
// byte offset of the 16-byte chunk (row r, octet k8) in an image with K8 octets per row
__device__ __forceinline__ size_t img_off(long long r, int k8, int K8) {
  return (((size_t)(r >> 7) * K8 + k8) * 128 + (size_t)(r & 127)) * 16;
}
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ void ld8(const float* p, float* v) {
  const float4 a = *reinterpret_cast<const float4*>(p), b = *reinterpret_cast<const float4*>(p + 4);
  v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
}
__device__ __forceinline__ void st8(float* p, const float* v) {
  *reinterpret_cast<float4*>(p) = make_float4(v[0], v[1], v[2], v[3]);
  *reinterpret_cast<float4*>(p + 4) = make_float4(v[4], v[5], v[6], v[7]);
}

// ---------------------------------------------------------------------------------------------------------------
// Inputs: u = [x | conditions] image (KU8 octets) and the Fourier time embedding image (KT8 octets):
//   emb = concat[t, sin(t*w*2pi*scale), cos(..)]   fourier_embedding.py:71-73 (same left-to-right product)
// One thread per row (rows >= B are zero filled: they are the K padding of the dW contraction).
__global__ void train_input_kernel(const float* __restrict__ x, const float* __restrict__ cond, const float* __restrict__ t,
                                   const float* __restrict__ freq, float fscale, long long B, long long R_pad, int D, int C,
                                   int T, int KU8, int KT8, unsigned char* __restrict__ u_img, unsigned char* __restrict__ te_img) {
  // grid (row blocks, KU8 + KT8): one 8-column group of one row per thread (one thread per ROW left a 8192-row shard
  // with 64 CTAs each walking ~70 columns serially)
  const long long r = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= R_pad) return;
  const bool valid = r < B;
  int k8 = blockIdx.y;
  float v[8];
  if (k8 < KU8) {
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const int k = k8 * 8 + i;
      float e = 0.f;
      if (valid) {
        if (k < D) e = x[r * D + k];
        else if (k < D + C) e = cond[r * C + (k - D)];
      }
      v[i] = e;
    }
    *reinterpret_cast<uint4*>(u_img + img_off(r, k8, KU8)) = pack8(v);
    return;
  }
  k8 -= KU8;
  const float tt = valid ? t[r] : 0.f;
  const int Th = (T - 1) / 2;
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const int k = k8 * 8 + i;
    float e = 0.f;
    if (valid) {
      if (T == 1) e = k == 0 ? tt : 0.f;
      else if (k == 0) e = tt;
      else if (k <= Th) e = sinf(tt * freq[k - 1] * 2.f * 3.14159265358979323846f * fscale);
      else if (k < T) e = cosf(tt * freq[k - 1 - Th] * 2.f * 3.14159265358979323846f * fscale);
    }
    v[i] = e;
  }
  *reinterpret_cast<uint4*>(te_img + img_off(r, k8, KT8)) = pack8(v);
}

// act(P) -> packed fp16 image (+ optional fp32 copy).  One WARP per row, a lane per 8 columns: the row is read as one
// contiguous kilobyte (thread == row read 32 lanes x 32 B spread over 32 KB and ran at a third of the HBM rate).
__global__ void __launch_bounds__(256) act_pack_kernel(const float* __restrict__ P, long long ld, long long B, long long R_pad,
                                                       int K8, int act, unsigned char* __restrict__ img,
                                                       float* __restrict__ out_f32) {
  const int lane = threadIdx.x & 31;
  const long long warp0 = (long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const long long nwarps = (long long)gridDim.x * (blockDim.x >> 5);
  for (long long r = warp0; r < R_pad; r += nwarps) {
    for (int ob = 0; ob < K8; ob += 128) {  // up to 4 octets per lane with all their loads in flight first
      float pv[4][8];
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const int o = ob + q * 32 + lane;
        if (o < K8 && r < B) ld8(P + r * ld + o * 8, pv[q]);
      }
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const int o = ob + q * 32 + lane;
        if (o < K8) {
          float v[8];
          if (r < B) {
            float d[8];
            t_act8(act, pv[q], v, d);
            if (out_f32) st8(out_f32 + r * ld + o * 8, v);
          } else {
#pragma unroll
            for (int i = 0; i < 8; ++i) v[i] = 0.f;
          }
          *reinterpret_cast<uint4*>(img + img_off(r, o, K8)) = pack8(v);
        }
      }
    }
  }
}

// dP = dY * act'(P): packed image, column sums of dP -> bias gradient (scaled).  Same row-per-warp form; NJ = columns / 256.
template <int NJ>
__global__ void __launch_bounds__(256) act_bwd_kernel(const float* __restrict__ dY, const float* __restrict__ P, long long ld,
                                                      long long B, long long R_pad, int K8, int act,
                                                      unsigned char* __restrict__ img, float* __restrict__ dbias, float gscale) {
  extern __shared__ float s_cs[];  // [K8 * 8], column c = o*8 + i at [i*K8 + o]
  const int lane = threadIdx.x & 31;
  const long long warp0 = (long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const long long nwarps = (long long)gridDim.x * (blockDim.x >> 5);
  float cs[NJ][8];
#pragma unroll
  for (int j = 0; j < NJ; ++j)
#pragma unroll
    for (int i = 0; i < 8; ++i) cs[j][i] = 0.f;
  for (long long r = warp0; r < R_pad; r += nwarps) {
#pragma unroll
    for (int j = 0; j < NJ; ++j) {
      const int o = j * 32 + lane;
      if (o < K8) {
        float v[8];
        if (r < B) {
          float p[8], uu[8], d[8];
          ld8(P + r * ld + o * 8, p);
          ld8(dY + r * ld + o * 8, v);
          t_act8(act, p, uu, d);
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            v[i] *= d[i];
            cs[j][i] += v[i];
          }
        } else {
#pragma unroll
          for (int i = 0; i < 8; ++i) v[i] = 0.f;
        }
        *reinterpret_cast<uint4*>(img + img_off(r, o, K8)) = pack8(v);
      }
    }
  }
  if (!dbias) return;
  for (int i = threadIdx.x; i < K8 * 8; i += blockDim.x) s_cs[i] = 0.f;
  __syncthreads();
  for (int wv = 0; wv < (int)(blockDim.x >> 5); ++wv) {  // the warps take turns (no shared float atomics)
    if ((int)(threadIdx.x >> 5) == wv) {
#pragma unroll
      for (int j = 0; j < NJ; ++j) {
        const int o = j * 32 + lane;
        if (o < K8) {
#pragma unroll
          for (int i = 0; i < 8; ++i) s_cs[i * K8 + o] += cs[j][i];
        }
      }
    }
    __syncthreads();
  }
  for (int c = threadIdx.x; c < K8 * 8; c += blockDim.x) atomicAdd(dbias + c, gscale * s_cs[(c & 7) * K8 + (c >> 3)]);
}

// ---------------------------------------------------------------------------------------------------------------
// ConditionalDenseBlock forward between its GEMMs (conditional_dense_block.py:117-130):
//   u = dropout(act(P));  v = skip + film(u);  out = LayerNorm(v)          one warp per row, lane owns octets lane+32j
struct BlockFwdArgs {
  const float* P;       // (B,W) dense output incl. bias
  const float* FB;      // FiLM shift: (B,W) per row, or one vector (fb_stride 0)
  const float* FG;      // FiLM scale (film_use_gamma) or null
  long long fb_stride;  // row stride of FB (the training step keeps the shifts of ALL blocks side by side: L * W)
  long long fg_stride;  // row stride of FG
  const float* prev;    // (B,W) skip source: previous block's normalised activations (or h0), or null (residual off)
  const float* pg; const float* pb;  // affine of the previous LayerNorm applied to prev (null: prev is used as is)
  const unsigned char* prev_img;     // inference: skip source = the fp16 operand image itself (may equal img: in place)
  const float* gamma; const float* beta;  // this block's LayerNorm parameters (null with norm none)
  float* xhat;          // (B,W) normalised activations before the affine (saved for backward / next skip), or null
  float* rstd;          // (B) or null
  unsigned char* img;   // packed fp16 image of the block output
  float* out_f32;       // optional fp32 copy of the block output
  long long B, R_pad;
  int W, act, norm, layer;
  float drop_rate, inv_keep; unsigned long long seed; const uint32_t* seed_dev; long long row0;
  int Weff;             // true width of this block when W is a padded common width (0: W); the padded columns of v are
                        // exactly zero (zero-padded weights, act(0) = 0), so only the norm's divisor and variance change
};

template <int NJ>
__global__ void __launch_bounds__(256, NJ == 1 ? BFK_FWD_MINB : 1) block_fwd_kernel(const BlockFwdArgs a) {
  const unsigned long long dkey = a.drop_rate > 0.f ? drop_key(a.seed, a.seed_dev, a.layer) : 0ull;
  const int lane = threadIdx.x & 31;
  const long long warp0 = (long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const long long nwarps = (long long)gridDim.x * (blockDim.x >> 5);
  const int K8 = a.W / 8;
  // NJ == 1 (W <= 256, the only instantiation launched): the three row streams of the NEXT row are in flight under this
  // row's two warp reductions and its stores - one row per warp at a time left the kernel at half the HBM rate
  constexpr bool PF = NJ == 1 && BFK_BLOCK_PREFETCH;
  float pn[8], fn[8], sn[8];
  auto fetch = [&](long long rr) {
    if (PF && lane < K8 && rr < a.B) {
      ld8(a.P + rr * a.W + lane * 8, pn);
      ld8(a.FB + rr * a.fb_stride + lane * 8, fn);
      if (a.prev) ld8(a.prev + rr * a.W + lane * 8, sn);
    }
  };
  fetch(warp0);
  for (long long r = warp0; r < a.R_pad; r += nwarps) {
    if (r >= a.B) {
#pragma unroll
      for (int j = 0; j < NJ; ++j) {
        const int o = j * 32 + lane;
        if (o < K8) *reinterpret_cast<uint4*>(a.img + img_off(r, o, K8)) = make_uint4(0, 0, 0, 0);
      }
      continue;
    }
    float v[NJ][8];
    float sum = 0.f;
#pragma unroll
    for (int j = 0; j < NJ; ++j) {
      const int o = j * 32 + lane;
      if (o < K8) {
        const int c0 = o * 8;
        float p[8], f[8];
        if (PF) {
#pragma unroll
          for (int i = 0; i < 8; ++i) { p[i] = pn[i]; f[i] = fn[i]; }
        } else {
          ld8(a.P + r * a.W + c0, p);
          ld8(a.FB + r * a.fb_stride + c0, f);
        }
        float ds[8], ud[8];
        if (a.drop_rate > 0.f) drop_scale8(dkey, a.row0 + r, c0, a.drop_rate, a.inv_keep, ds);
        t_act8(a.act, p, v[j], ud);
        if (a.drop_rate > 0.f) {
#pragma unroll
          for (int i = 0; i < 8; ++i) v[j][i] *= ds[i];
        }
        if (a.FG) {
          float g[8];
          ld8(a.FG + r * a.fg_stride + c0, g);
#pragma unroll
          for (int i = 0; i < 8; ++i) v[j][i] *= 1.f + g[i];
        }
#pragma unroll
        for (int i = 0; i < 8; ++i) v[j][i] += f[i];
        if (a.prev) {
          float s[8];
          if (PF) {
#pragma unroll
            for (int i = 0; i < 8; ++i) s[i] = sn[i];
          } else {
            ld8(a.prev + r * a.W + c0, s);
          }
          if (a.pg) {
            float g[8], b[8];
            ld8(a.pg + c0, g);
            if (a.pb) ld8(a.pb + c0, b);
            else {
#pragma unroll
              for (int i = 0; i < 8; ++i) b[i] = 0.f;
            }
#pragma unroll
            for (int i = 0; i < 8; ++i) s[i] = fmaf(s[i], g[i], b[i]);
          }
#pragma unroll
          for (int i = 0; i < 8; ++i) v[j][i] += s[i];
        }
        if (a.prev_img) {
          const uint4 raw = *reinterpret_cast<const uint4*>(a.prev_img + img_off(r, o, K8));
          const __half2* hp = reinterpret_cast<const __half2*>(&raw);
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            const float2 f = __half22float2(hp[i]);
            v[j][2 * i] += f.x; v[j][2 * i + 1] += f.y;
          }
        }
#pragma unroll
        for (int i = 0; i < 8; ++i) sum += v[j][i];
      } else {
#pragma unroll
        for (int i = 0; i < 8; ++i) v[j][i] = 0.f;
      }
    }
    fetch(r + nwarps);
    float mean = 0.f, rs = 1.f;
    if (a.norm != BFK_NORM_NONE) {  // layer: centred; rms: mean stays 0
      const float we = (float)(a.Weff > 0 ? a.Weff : a.W);
      if (a.norm == BFK_NORM_LAYER) mean = warp_sum(sum) / we;
      float sq = 0.f;
#pragma unroll
      for (int j = 0; j < NJ; ++j)
        if (j * 32 + lane < K8) {
#pragma unroll
          for (int i = 0; i < 8; ++i) { const float d = v[j][i] - mean; sq = fmaf(d, d, sq); }
        }
      // the W - Weff padded columns hold v = 0 and each added mean^2 to the sum of squares
      rs = rsqrtf((warp_sum(sq) - ((float)a.W - we) * mean * mean) / we + (a.norm == BFK_NORM_LAYER ? LN_EPS : RMS_EPS));
    }
    if (a.rstd && lane == 0) a.rstd[r] = rs;
#pragma unroll
    for (int j = 0; j < NJ; ++j) {
      const int o = j * 32 + lane;
      if (o < K8) {
        const int c0 = o * 8;
        float xh[8], h[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) xh[i] = (v[j][i] - mean) * rs;
        if (a.xhat) st8(a.xhat + r * a.W + c0, xh);
        if (a.gamma) {
          float g[8], b[8];
          ld8(a.gamma + c0, g);
          if (a.beta) ld8(a.beta + c0, b);
          else {
#pragma unroll
            for (int i = 0; i < 8; ++i) b[i] = 0.f;  // rms: scale only
          }
#pragma unroll
          for (int i = 0; i < 8; ++i) h[i] = fmaf(xh[i], g[i], b[i]);
        } else {
#pragma unroll
          for (int i = 0; i < 8; ++i) h[i] = xh[i];
        }
        if (a.out_f32) st8(a.out_f32 + r * a.W + c0, h);
        *reinterpret_cast<uint4*>(a.img + img_off(r, o, K8)) = pack8(h);
      }
    }
  }
}

// Same block epilogue for WIDE rows (W > 256): WPR warps share one row (256 columns each, 8 per lane) and exchange the two
// LayerNorm partials through shared memory behind a named barrier.  The one-warp-per-row form needs 32 values per lane
// at W = 1024 (108 registers, 2.3 rows per warp at 8192 rows): measured 46 us per 8192 x 1024 panel, 1.4 TB/s (r01w).
template <int WPR>
__global__ void __launch_bounds__(256) block_fwd_wide_kernel(const BlockFwdArgs a) {
  const unsigned long long dkey = a.drop_rate > 0.f ? drop_key(a.seed, a.seed_dev, a.layer) : 0ull;
  __shared__ float s_sum[8], s_sq[8];
  constexpr int GPB = 8 / WPR;  // rows in flight per block
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int grp = warp / WPR, wig = warp % WPR;
  const int K8 = a.W / 8, o = wig * 32 + lane, c0 = o * 8;
  const bool on = o < K8;
  const long long stride = (long long)gridDim.x * GPB;
  // software pipeline: the loads of row r+stride are issued before the reductions / barriers of row r, otherwise every
  // row costs two exposed memory round trips (measured 2.3 TB/s on L2-resident panels, r01x)
  float p[8], f[8], sk[8], fg[8];
  uint4 raw = make_uint4(0, 0, 0, 0);
  auto fetch = [&](long long r) {
    if (on && r < a.B) {
      ld8(a.P + r * a.W + c0, p);
      ld8(a.FB + r * a.fb_stride + c0, f);
      if (a.FG) ld8(a.FG + r * a.fg_stride + c0, fg);
      if (a.prev) ld8(a.prev + r * a.W + c0, sk);
      if (a.prev_img) raw = *reinterpret_cast<const uint4*>(a.prev_img + img_off(r, o, K8));
    }
  };
  long long r = (long long)blockIdx.x * GPB + grp;
  fetch(r);
  for (; r < a.R_pad; r += stride) {
    if (r >= a.B) {
      if (on) *reinterpret_cast<uint4*>(a.img + img_off(r, o, K8)) = make_uint4(0, 0, 0, 0);
      continue;
    }
    float v[8];
    float sum = 0.f;
    if (on) {
      float ds[8], uu[8], ud[8];
      if (a.drop_rate > 0.f) drop_scale8(dkey, a.row0 + r, c0, a.drop_rate, a.inv_keep, ds);
      t_act8(a.act, p, uu, ud);
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        float u = uu[i];
        if (a.drop_rate > 0.f) u *= ds[i];
        v[i] = a.FG ? fmaf(1.f + fg[i], u, f[i]) : u + f[i];  // (1 + gamma) * u + beta, featurewise_linear_modulation.py:25-33
      }
      if (a.prev) {
        if (a.pg) {
          float g[8], b[8];
          ld8(a.pg + c0, g);
          if (a.pb) ld8(a.pb + c0, b);
          else {
#pragma unroll
            for (int i = 0; i < 8; ++i) b[i] = 0.f;
          }
#pragma unroll
          for (int i = 0; i < 8; ++i) sk[i] = fmaf(sk[i], g[i], b[i]);
        }
#pragma unroll
        for (int i = 0; i < 8; ++i) v[i] += sk[i];
      }
      if (a.prev_img) {
        const __half2* hp = reinterpret_cast<const __half2*>(&raw);
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const float2 q = __half22float2(hp[i]);
          v[2 * i] += q.x; v[2 * i + 1] += q.y;
        }
      }
#pragma unroll
      for (int i = 0; i < 8; ++i) sum += v[i];
    } else {
#pragma unroll
      for (int i = 0; i < 8; ++i) v[i] = 0.f;
    }
    fetch(r + stride);  // in place (img == prev_img) is safe: another row
    float mean = 0.f, rs = 1.f;
    if (a.norm != BFK_NORM_NONE) {  // layer: centred; rms: mean stays 0 (the first exchange is kept: it also orders s_sq)
      sum = warp_sum(sum);
      if (lane == 0) s_sum[warp] = sum;
      named_bar_sync(1 + grp, WPR * 32);
      float ts = 0.f;
#pragma unroll
      for (int w = 0; w < WPR; ++w) ts += s_sum[grp * WPR + w];
      const float we = (float)(a.Weff > 0 ? a.Weff : a.W);
      mean = a.norm == BFK_NORM_LAYER ? ts / we : 0.f;
      float sq = 0.f;
      if (on) {
#pragma unroll
        for (int i = 0; i < 8; ++i) { const float d = v[i] - mean; sq = fmaf(d, d, sq); }
      }
      sq = warp_sum(sq);
      if (lane == 0) s_sq[warp] = sq;
      named_bar_sync(1 + grp, WPR * 32);  // also orders this row's s_sum reads before the next row's s_sum writes
      float tq = 0.f;
#pragma unroll
      for (int w = 0; w < WPR; ++w) tq += s_sq[grp * WPR + w];
      rs = rsqrtf((tq - ((float)a.W - we) * mean * mean) / we + (a.norm == BFK_NORM_LAYER ? LN_EPS : RMS_EPS));
    }
    if (a.rstd && wig == 0 && lane == 0) a.rstd[r] = rs;
    if (on) {
      float xh[8], h[8];
#pragma unroll
      for (int i = 0; i < 8; ++i) xh[i] = (v[i] - mean) * rs;
      if (a.xhat) st8(a.xhat + r * a.W + c0, xh);
      if (a.gamma) {
        float g[8], b[8];
        ld8(a.gamma + c0, g);
        if (a.beta) ld8(a.beta + c0, b);
        else {
#pragma unroll
          for (int i = 0; i < 8; ++i) b[i] = 0.f;  // rms: scale only
        }
#pragma unroll
        for (int i = 0; i < 8; ++i) h[i] = fmaf(xh[i], g[i], b[i]);
      } else {
#pragma unroll
        for (int i = 0; i < 8; ++i) h[i] = xh[i];
      }
      if (a.out_f32) st8(a.out_f32 + r * a.W + c0, h);
      *reinterpret_cast<uint4*>(a.img + img_off(r, o, K8)) = pack8(h);
    }
  }
}

// ConditionalDenseBlock backward between its GEMMs.  In: dH (gradient wrt the block output, fp32, overwritten by dV =
// gradient wrt the pre-norm sum, which is also the skip gradient the dX GEMM adds onto).  Out: dP and dV images, column
// sums -> d(norm.gamma), d(norm.beta), d(dense.bias), d(film.to_beta.bias).
struct BlockBwdArgs {
  float* dH;
  const float* xhat; const float* rstd; const float* P;
  const float* gamma;
  unsigned char* dP_img; unsigned char* dV_img;
  float* d_gamma; float* d_beta; float* d_bd; float* d_bfilm;
  // film_use_gamma: FG (B,W) = the FiLM scale to_gamma(t_emb) of this block (the forward applied 1 + FG); dG_img gets the
  // cotangent of FG (= dV * u), d_bgfilm the gradient of to_gamma's bias
  const float* FG; unsigned char* dG_img; float* d_bgfilm;
  int Weff;        // true width when W is a padded common width (0: W): the divisor of the norm backward
  float* d_bskip;  // bias gradient of the residual projector (= column sums of dV, like d_bfilm) or null
  long long B, R_pad;
  int W, act, norm, layer;
  float drop_rate, inv_keep; unsigned long long seed; const uint32_t* seed_dev; long long row0;
  float gscale;
};

template <int NJ, bool FGAM = false>
__global__ void __launch_bounds__(256, (NJ == 1 && !FGAM) ? BFK_BWD_MINB : 1) block_bwd_kernel(const BlockBwdArgs a) {
  const unsigned long long dkey = a.drop_rate > 0.f ? drop_key(a.seed, a.seed_dev, a.layer) : 0ull;
  extern __shared__ float s_red[];  // [4 (+1 with FGAM)][W]
  const int lane = threadIdx.x & 31;
  const long long warp0 = (long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const long long nwarps = (long long)gridDim.x * (blockDim.x >> 5);
  const int K8 = a.W / 8, W = a.W;
  for (int i = threadIdx.x; i < (FGAM ? 5 : 4) * W; i += blockDim.x) s_red[i] = 0.f;
  __syncthreads();
  float ag[NJ][8], ab[NJ][8], ad[NJ][8], af[NJ][8], aq[FGAM ? NJ : 1][8];
#pragma unroll
  for (int j = 0; j < NJ; ++j)
#pragma unroll
    for (int i = 0; i < 8; ++i) ag[j][i] = ab[j][i] = ad[j][i] = af[j][i] = 0.f;
#pragma unroll
  for (int j = 0; j < (FGAM ? NJ : 1); ++j)
#pragma unroll
    for (int i = 0; i < 8; ++i) aq[j][i] = 0.f;

  // NJ == 1: the next row's three streams are fetched under this row's reductions and stores (see block_fwd_kernel);
  // dH is rewritten in place for row r only, the prefetch reads another row
  constexpr bool PF = NJ == 1 && BFK_BLOCK_PREFETCH;
  float ppn[8], dhn[8], xhn[8];
  auto fetch = [&](long long rr) {
    if (PF && lane < K8 && rr < a.B) {
      ld8(a.P + rr * W + lane * 8, ppn);
      ld8(a.dH + rr * W + lane * 8, dhn);
      ld8(a.xhat + rr * W + lane * 8, xhn);
    }
  };
  fetch(warp0);
  for (long long r = warp0; r < a.R_pad; r += nwarps) {
    if (r >= a.B) {
#pragma unroll
      for (int j = 0; j < NJ; ++j) {
        const int o = j * 32 + lane;
        if (o < K8) {
          *reinterpret_cast<uint4*>(a.dP_img + img_off(r, o, K8)) = make_uint4(0, 0, 0, 0);
          *reinterpret_cast<uint4*>(a.dV_img + img_off(r, o, K8)) = make_uint4(0, 0, 0, 0);
          if (FGAM && a.dG_img) *reinterpret_cast<uint4*>(a.dG_img + img_off(r, o, K8)) = make_uint4(0, 0, 0, 0);
        }
      }
      continue;
    }
    float g[NJ][8], xh[NJ][8], pp[NJ][8];
    float s1 = 0.f, s2 = 0.f;
#pragma unroll
    for (int j = 0; j < NJ; ++j) {  // all three streams of this row in flight before the first reduction
      const int o = j * 32 + lane;
      if (o < K8) {
        if (PF) {
#pragma unroll
          for (int i = 0; i < 8; ++i) pp[j][i] = ppn[i];
        } else {
          ld8(a.P + r * W + o * 8, pp[j]);
        }
      }
    }
#pragma unroll
    for (int j = 0; j < NJ; ++j) {
      const int o = j * 32 + lane;
      if (o < K8) {
        const int c0 = o * 8;
        float dh[8];
        if (PF) {
#pragma unroll
          for (int i = 0; i < 8; ++i) { dh[i] = dhn[i]; xh[j][i] = xhn[i]; }
        } else {
          ld8(a.dH + r * W + c0, dh);
          ld8(a.xhat + r * W + c0, xh[j]);
        }
        if (a.gamma) {
          float gm[8];
          ld8(a.gamma + c0, gm);
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            ag[j][i] = fmaf(dh[i], xh[j][i], ag[j][i]);
            ab[j][i] += dh[i];
            g[j][i] = dh[i] * gm[i];
          }
        } else {
#pragma unroll
          for (int i = 0; i < 8; ++i) g[j][i] = dh[i];
        }
#pragma unroll
        for (int i = 0; i < 8; ++i) { s1 += g[j][i]; s2 = fmaf(g[j][i], xh[j][i], s2); }
      } else {
#pragma unroll
        for (int i = 0; i < 8; ++i) g[j][i] = xh[j][i] = 0.f;
      }
    }
    fetch(r + nwarps);
    float m1 = 0.f, m2 = 0.f, rs = 1.f;
    if (a.norm != BFK_NORM_NONE) {  // rms: no centring, so the m1 term is absent
      const float we = (float)(a.Weff > 0 ? a.Weff : W);
      if (a.norm == BFK_NORM_LAYER) m1 = warp_sum(s1) / we;
      m2 = warp_sum(s2) / we;
      rs = a.rstd[r];
    }
#pragma unroll
    for (int j = 0; j < NJ; ++j) {
      const int o = j * 32 + lane;
      if (o < K8) {
        const int c0 = o * 8;
        float dv[8], dp[8], dq[8], fg[8], ds[8];
        if (FGAM) ld8(a.FG + r * W + c0, fg);
        if (a.drop_rate > 0.f) drop_scale8(dkey, a.row0 + r, c0, a.drop_rate, a.inv_keep, ds);
        float uu[8], ud[8];
        t_act8(a.act, pp[j], uu, ud);
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          dv[i] = a.norm != BFK_NORM_NONE ? rs * (g[j][i] - m1 - xh[j][i] * m2) : g[j][i];
          float d = ud[i], u = uu[i];
          if (a.drop_rate > 0.f) {
            d *= ds[i];
            u *= ds[i];
          }
          if (FGAM) {  // v = skip + (1 + FG) u + FB:  dU = dV (1 + FG),  dFG = dV u
            dq[i] = dv[i] * u;
            aq[j][i] += dq[i];
            d *= 1.f + fg[i];
          }
          dp[i] = dv[i] * d;
          ad[j][i] += dp[i];
          af[j][i] += dv[i];
        }
        st8(a.dH + r * W + c0, dv);
        *reinterpret_cast<uint4*>(a.dP_img + img_off(r, o, K8)) = pack8(dp);
        *reinterpret_cast<uint4*>(a.dV_img + img_off(r, o, K8)) = pack8(dv);
        if (FGAM && a.dG_img) *reinterpret_cast<uint4*>(a.dG_img + img_off(r, o, K8)) = pack8(dq);
      }
    }
  }
  if (!a.d_bd) return;  // cotangent-only pass (bfk_train_vjp): no parameter gradients
  // column sums across the CTA's warps: the warps take turns on s_red (no shared float atomics: those are CAS loops, and
  // 8 warps on one address cost ~6 us per CTA - a third of the kernel on a 8192-row shard).  Column c = o*8 + i lives at
  // [i*K8 + o], so a warp's 32 lanes touch consecutive words.
  for (int wv = 0; wv < (int)(blockDim.x >> 5); ++wv) {
    if ((int)(threadIdx.x >> 5) == wv) {
#pragma unroll
      for (int j = 0; j < NJ; ++j) {
        const int o = j * 32 + lane;
        if (o < K8) {
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            const int c = i * K8 + o;
            if (a.gamma) { s_red[c] += ag[j][i]; s_red[W + c] += ab[j][i]; }
            s_red[2 * W + c] += ad[j][i];
            s_red[3 * W + c] += af[j][i];
            if (FGAM) s_red[4 * W + c] += aq[j][i];
          }
        }
      }
    }
    __syncthreads();
  }
  for (int c = threadIdx.x; c < W; c += blockDim.x) {
    const int sc = (c & 7) * K8 + (c >> 3);
    if (a.gamma) {
      atomicAdd(a.d_gamma + c, a.gscale * s_red[sc]);
      if (a.d_beta) atomicAdd(a.d_beta + c, a.gscale * s_red[W + sc]);
    }
    atomicAdd(a.d_bd + c, a.gscale * s_red[2 * W + sc]);
    atomicAdd(a.d_bfilm + c, a.gscale * s_red[3 * W + sc]);
    if (a.d_bskip) atomicAdd(a.d_bskip + c, a.gscale * s_red[3 * W + sc]);
    if (FGAM && a.d_bgfilm) atomicAdd(a.d_bgfilm + c, a.gscale * s_red[4 * W + sc]);
  }
}

// Wide rows (W > 256) without film_use_gamma: block_bwd_kernel<4> keeps 4 x 32 column accumulators per lane (255
// registers, ONE 8-warp CTA per SM) and ran at 44 % of the HBM rate.  Here a row is shared by CW warps (CW * 256 columns),
// each lane owns ONE octet like the W <= 256 kernel (two CTAs per SM), and the two row sums of the norm backward cross the
// warps through shared memory and one named barrier per row (double-buffered slots).  Same arithmetic as
// block_bwd_kernel; rows of a CTA: 8 / CW at a time.
template <int CW>
__global__ void __launch_bounds__(256, 2) block_bwd_split_kernel(const BlockBwdArgs a) {
  const unsigned long long dkey = a.drop_rate > 0.f ? drop_key(a.seed, a.seed_dev, a.layer) : 0ull;
  extern __shared__ float s_red[];  // [4][W]
  constexpr int RG = 8 / CW;        // row groups (rows in flight) per CTA
  __shared__ float2 s_part[2][RG][CW];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int rg = warp / CW, cw = warp % CW;
  const int K8 = a.W / 8, W = a.W;
  const int o = cw * 32 + lane;
  const bool on = o < K8;
  const int c0 = o * 8;
  for (int i = threadIdx.x; i < 4 * W; i += blockDim.x) s_red[i] = 0.f;
  __syncthreads();
  float ag[8], ab[8], ad[8], af[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) ag[i] = ab[i] = ad[i] = af[i] = 0.f;
  const long long grp0 = (long long)blockIdx.x * RG + rg, ngrp = (long long)gridDim.x * RG;
  float ppn[8], dhn[8], xhn[8];
  auto fetch = [&](long long rr) {
    if (on && rr < a.B) {
      ld8(a.P + rr * W + c0, ppn);
      ld8(a.dH + rr * W + c0, dhn);
      ld8(a.xhat + rr * W + c0, xhn);
    }
  };
  fetch(grp0);
  float gm[8];
  if (a.gamma && on) ld8(a.gamma + c0, gm);
  uint32_t it = 0;
  for (long long r = grp0; r < a.R_pad; r += ngrp) {
    if (r >= a.B) {  // (uniform over the row's warps)
      if (on) {
        *reinterpret_cast<uint4*>(a.dP_img + img_off(r, o, K8)) = make_uint4(0, 0, 0, 0);
        *reinterpret_cast<uint4*>(a.dV_img + img_off(r, o, K8)) = make_uint4(0, 0, 0, 0);
      }
      continue;
    }
    float g[8], xh[8], pp[8];
    float s1 = 0.f, s2 = 0.f;
    if (on) {
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        pp[i] = ppn[i];
        xh[i] = xhn[i];
        const float dh = dhn[i];
        if (a.gamma) {
          ag[i] = fmaf(dh, xh[i], ag[i]);
          ab[i] += dh;
          g[i] = dh * gm[i];
        } else {
          g[i] = dh;
        }
        s1 += g[i];
        s2 = fmaf(g[i], xh[i], s2);
      }
    } else {
#pragma unroll
      for (int i = 0; i < 8; ++i) g[i] = xh[i] = pp[i] = 0.f;
    }
    fetch(r + ngrp);
    float m1 = 0.f, m2 = 0.f, rs = 1.f;
    if (a.norm != BFK_NORM_NONE) {
      s1 = warp_sum(s1);
      s2 = warp_sum(s2);
      const uint32_t slot = it & 1;
      if (lane == 0) s_part[slot][rg][cw] = make_float2(s1, s2);
      asm volatile("bar.sync %0, %1;" ::"r"(1 + rg), "r"(CW * 32) : "memory");
      float t1 = 0.f, t2 = 0.f;
#pragma unroll
      for (int q = 0; q < CW; ++q) { const float2 pq = s_part[slot][rg][q]; t1 += pq.x; t2 += pq.y; }
      const float we = (float)(a.Weff > 0 ? a.Weff : W);
      if (a.norm == BFK_NORM_LAYER) m1 = t1 / we;
      m2 = t2 / we;
      rs = a.rstd[r];
      ++it;
    }
    if (on) {
      float dv[8], dp[8], ds[8], uu[8], ud[8];
      if (a.drop_rate > 0.f) drop_scale8(dkey, a.row0 + r, c0, a.drop_rate, a.inv_keep, ds);
      t_act8(a.act, pp, uu, ud);
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        dv[i] = a.norm != BFK_NORM_NONE ? rs * (g[i] - m1 - xh[i] * m2) : g[i];
        float d = ud[i];
        if (a.drop_rate > 0.f) d *= ds[i];
        dp[i] = dv[i] * d;
        ad[i] += dp[i];
        af[i] += dv[i];
      }
      st8(a.dH + r * W + c0, dv);
      *reinterpret_cast<uint4*>(a.dP_img + img_off(r, o, K8)) = pack8(dp);
      *reinterpret_cast<uint4*>(a.dV_img + img_off(r, o, K8)) = pack8(dv);
    }
  }
  if (!a.d_bd) return;  // cotangent-only pass (bfk_train_vjp)
  for (int wv = 0; wv < RG; ++wv) {  // the row groups take turns on s_red; the CW warps of a group own disjoint columns
    if (rg == wv && on) {
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const int c = i * K8 + o;
        if (a.gamma) { s_red[c] += ag[i]; s_red[W + c] += ab[i]; }
        s_red[2 * W + c] += ad[i];
        s_red[3 * W + c] += af[i];
      }
    }
    __syncthreads();
  }
  for (int c = threadIdx.x; c < W; c += blockDim.x) {
    const int sc = (c & 7) * K8 + (c >> 3);
    if (a.gamma) {
      atomicAdd(a.d_gamma + c, a.gscale * s_red[sc]);
      if (a.d_beta) atomicAdd(a.d_beta + c, a.gscale * s_red[W + sc]);
    }
    atomicAdd(a.d_bd + c, a.gscale * s_red[2 * W + sc]);
    atomicAdd(a.d_bfilm + c, a.gscale * s_red[3 * W + sc]);
    if (a.d_bskip) atomicAdd(a.d_bskip + c, a.gscale * s_red[3 * W + sc]);
  }
}

// ---------------------------------------------------------------------------------------------------------------
// Loss head.  resid = a_b*pred - target;  loss_b = w_b * mean_i resid^2 (mse) | mean_i |resid| (mae)
//   | mean_i (sqrt(resid^2 + c^2) - c) (kind 2, pseudo-Huber: ConsistencyModel, consistency_model.py:413-414 with
//   a = c_out(t2), target = teacher - c_skip(t2) * input, w = 1 / (t2 - t1))
//   FlowMatching: a = 1, target = x1 - x0 (flow_matching.py:170-176);  DiffusionModel: a_b and target from the schedule
//   (diffusion_model.py:197-230).  Writes the un-normalised dLoss/dpred image (64 features, zeros beyond D), adds
//   loss_scale * sum_b loss_b to *loss and gscale * column sums to d_bo.  One thread per row.
__global__ void loss_kernel(const float* __restrict__ pred, long long ldp, const float* __restrict__ target,
                            const float* __restrict__ a_row, const float* __restrict__ w_row, long long B, long long R_pad,
                            int D, int kind, unsigned char* __restrict__ dout_img, int KO8, float* __restrict__ loss,
                            float loss_scale, float* __restrict__ d_bo, float gscale, float huber_c,
                            const float* __restrict__ elem_mask = nullptr) {
  __shared__ float s_loss;
  __shared__ float s_db[64];
  if (threadIdx.x == 0) s_loss = 0.f;
  if (threadIdx.x < 64) s_db[threadIdx.x] = 0.f;
  __syncthreads();
  const long long r = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  float lrow = 0.f;
  if (r < R_pad) {
    const bool valid = r < B;
    const float av = (valid && a_row) ? a_row[r] : 1.f;
    const float wv = (valid && w_row) ? w_row[r] : 1.f;
    for (int k8 = 0; k8 < KO8; ++k8) {
      float v[8];
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const int k = k8 * 8 + i;
        float g = 0.f;
        if (valid && k < D) {
          // elem_mask: the 0 / 1 target-dropout mask of flow_matching.py:170-175 / diffusion_model.py:177-216 (loss of
          // mask * target against mask * prediction)
          const float res = (av * pred[r * ldp + k] - target[r * D + k]) * (elem_mask ? elem_mask[r * D + k] : 1.f);
          if (kind == 1) { lrow += fabsf(res); g = wv * av * (res > 0.f ? 1.f : (res < 0.f ? -1.f : 0.f)); }
          else if (kind == 2) {
            const float rt = sqrtf(fmaf(res, res, huber_c * huber_c));
            lrow += rt - huber_c;
            g = wv * av * res / rt;
          }
          else { lrow = fmaf(res, res, lrow); g = 2.f * wv * av * res; }
        }
        if (d_bo && k < D) {  // (uniform) one shared atomic per warp and column: 256 threads on one address is a CAS storm
          const float gs = warp_sum(g);
          if ((threadIdx.x & 31) == 0) atomicAdd(&s_db[k], gs);
        }
        v[i] = g;
      }
      *reinterpret_cast<uint4*>(dout_img + img_off(r, k8, KO8)) = pack8(v);
    }
    lrow *= wv;
  }
  lrow = warp_sum(lrow);
  if ((threadIdx.x & 31) == 0) atomicAdd(&s_loss, lrow);
  __syncthreads();
  if (threadIdx.x == 0) atomicAdd(loss, s_loss * loss_scale);
  if (d_bo && threadIdx.x < D) atomicAdd(d_bo + threadIdx.x, gscale * s_db[threadIdx.x]);
}

// d(frequencies) from d(time embedding) (fourier_embedding.py:71-73): one thread per row, block reduction.
__global__ void freq_bwd_kernel(const float* __restrict__ dte, long long ld, const float* __restrict__ t,
                                const float* __restrict__ freq, float fscale, long long B, int Th, float* __restrict__ dfreq,
                                float gscale) {
  // grid (row blocks, Th): one frequency per blockIdx.y
  __shared__ float s;
  if (threadIdx.x == 0) s = 0.f;
  __syncthreads();
  const long long r = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const float c = 2.f * 3.14159265358979323846f * fscale;
  const int j = blockIdx.y;
  float g = 0.f;
  if (r < B) {
    const float tt = t[r];
    const float p = tt * freq[j] * 2.f * 3.14159265358979323846f * fscale;
    g = tt * c * (dte[r * ld + 1 + j] * cosf(p) - dte[r * ld + 1 + Th + j] * sinf(p));
  }
  g = warp_sum(g);
  if ((threadIdx.x & 31) == 0) atomicAdd(&s, g);
  __syncthreads();
  if (threadIdx.x == 0) atomicAdd(dfreq + j, gscale * s);
}

// First layer as ONE contraction over u = [x | c]:  concat: W1 = blockdiag(Wx, Wc) (KU x 2W), b1 = [bx, bc];
// add: W1 = [Wx; Wc] (KU x W), b1 = bx + bc; unconditional: W1 = Wx.      time_mlp.py:131-137
__global__ void build_w1_kernel(const float* __restrict__ Wx, const float* __restrict__ bx, const float* __restrict__ Wc,
                                const float* __restrict__ bc, int D, int C, int W, int merge, int KU, int N1,
                                float* __restrict__ W1, float* __restrict__ b1) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx < KU * N1) {
    const int k = idx / N1, n = idx % N1;
    float v = 0.f;
    if (k < D) { if (n < W) v = Wx[k * W + n]; }
    else if (k < D + C) {
      if (merge == BFK_MERGE_CONCAT) { if (n >= W) v = Wc[(k - D) * W + (n - W)]; }
      else v = Wc[(k - D) * W + n];
    }
    W1[idx] = v;
  }
  if (idx < N1) {
    float b = 0.f;
    if (merge == BFK_MERGE_CONCAT) b = idx < W ? bx[idx] : bc[idx - W];
    else b = bx[idx] + (C > 0 ? bc[idx] : 0.f);
    b1[idx] = b;
  }
}
__global__ void scatter_w1_grad_kernel(const float* __restrict__ dW1, const float* __restrict__ db1, int D, int C, int W,
                                       int merge, int N1, float gscale, float* __restrict__ dWx, float* __restrict__ dbx,
                                       float* __restrict__ dWc, float* __restrict__ dbc) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx < (D + C) * W) {
    const int k = idx / W, n = idx % W;
    if (k < D) dWx[k * W + n] += gscale * dW1[k * N1 + n];
    else dWc[(k - D) * W + n] += gscale * dW1[k * N1 + (merge == BFK_MERGE_CONCAT ? W + n : n)];
  }
  if (idx < W) {
    dbx[idx] += gscale * db1[idx];
    if (C > 0) dbc[idx] += gscale * db1[merge == BFK_MERGE_CONCAT ? W + idx : idx];
  }
}

// FlowMatching.compute_metrics inputs (flow_matching.py:159-168): x = t*x1 + (1-t)*x0, target = x1 - x0
__global__ void fm_interp_kernel(const float* __restrict__ x0, const float* __restrict__ x1, const float* __restrict__ t,
                                 long long B, int D, float* __restrict__ x, float* __restrict__ target) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= B * D) return;
  const float tt = t[i / D];
  x[i] = tt * x1[i] + (1.f - tt) * x0[i];
  target[i] = x1[i] - x0[i];
}

// ---------------------------------------------------------------------------------------------------------------
// Vector-Jacobian products of the velocity w.r.t. its input (utils/jacobian/vjp.py:44-68, jacobian_trace.py:44-55):
// the cotangent enters where the loss head would (dout image), the dX chain of the training backward carries it to the
// first layer's pre-activation, and dx = d(h0pre)[:, :W] . Wx^T closes it.
// cotangent image: row r gets dout[r,:] (or the one-hot e_col when dout == nullptr); K padding and rows >= B are zero.
__global__ void cotangent_pack_kernel(const float* __restrict__ dout, int col, long long B, long long R_pad, int D, int KO8,
                                      unsigned char* __restrict__ img) {
  const long long r = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= R_pad) return;
  for (int k8 = 0; k8 < KO8; ++k8) {
    float v[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const int k = k8 * 8 + i;
      v[i] = (r < B && k < D) ? (dout ? dout[r * D + k] : (k == col ? 1.f : 0.f)) : 0.f;
    }
    *reinterpret_cast<uint4*>(img + img_off(r, k8, KO8)) = pack8(v);
  }
}

// dx[r,d] = sum_k g[r,k] * Wx[d,k] for the first W columns of the d(h0pre) image; one thread per row, DC outputs per pass.
// trace[r] += dx[r,col] when asked (exact Jacobian trace accumulates one diagonal entry per reverse pass).
template <int DC>
__global__ void __launch_bounds__(128) vjp_input_kernel(const unsigned char* __restrict__ g_img, int K8, const float* __restrict__ Wx,
                                                        long long B, int D, int W, float* __restrict__ dx, float* __restrict__ trace,
                                                        int col) {
  const long long r = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= B) return;
  for (int d0 = 0; d0 < D; d0 += DC) {
    if (!dx && (col < d0 || col >= d0 + DC)) continue;  // trace only: one diagonal entry is needed
    float acc[DC];
#pragma unroll
    for (int d = 0; d < DC; ++d) acc[d] = 0.f;
    for (int k8 = 0; k8 < W / 8; ++k8) {
      const uint4 q = *reinterpret_cast<const uint4*>(g_img + img_off(r, k8, K8));
      const __half2* h = reinterpret_cast<const __half2*>(&q);
      float v[8];
#pragma unroll
      for (int i = 0; i < 4; ++i) { const float2 f = __half22float2(h[i]); v[2 * i] = f.x; v[2 * i + 1] = f.y; }
#pragma unroll
      for (int d = 0; d < DC; ++d) {
        if (d0 + d < D) {
          const float* wrow = Wx + (size_t)(d0 + d) * W + k8 * 8;
#pragma unroll
          for (int i = 0; i < 8; ++i) acc[d] = fmaf(v[i], __ldg(wrow + i), acc[d]);
        }
      }
    }
#pragma unroll
    for (int d = 0; d < DC; ++d) {
      if (d0 + d < D) {
        if (dx) dx[r * D + d0 + d] = acc[d];
        if (trace && d0 + d == col) trace[r] += acc[d];
      }
    }
  }
}

// out = y + h * sum_j c_j k_j (integrate.py:47-54): products and sums rounded separately, left to right, as torch does.
struct RkCombineArgs { const float* k[8]; float c[8]; int n; };
__global__ void rk_combine_kernel(const float* __restrict__ y, const RkCombineArgs a, float h, long long numel,
                                  float* __restrict__ out) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= numel) return;
  float acc = 0.f;
  for (int j = 0; j < a.n; ++j) acc = __fadd_rn(acc, __fmul_rn(a.c[j], a.k[j][i]));
  out[i] = __fadd_rn(y[i], __fmul_rn(h, acc));
}

// ---------------------------------------------------------------------------------------------------------------
// Optimizer (keras.optimizers.Adam as BasicWorkflow builds it, basic_workflow.py:514-526: per-variable clipnorm).
// Segments = parameter tensors inside one flat buffer.
__global__ void seg_sqnorm_kernel(const float* __restrict__ g, const long long* __restrict__ seg_off, int n_seg,
                                  float* __restrict__ sq) {
  const int s = blockIdx.x;
  const long long a = seg_off[s], b = seg_off[s + 1];
  float acc = 0.f;
  for (long long i = a + blockIdx.y * blockDim.x + threadIdx.x; i < b; i += (long long)gridDim.y * blockDim.x) acc = fmaf(g[i], g[i], acc);
  acc = warp_sum(acc);
  __shared__ float red[32];
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x < 32) {
    float v = threadIdx.x < (blockDim.x >> 5) ? red[threadIdx.x] : 0.f;
    v = warp_sum(v);
    if (threadIdx.x == 0) atomicAdd(sq + s, v);
  }
}
__global__ void adam_kernel(float* __restrict__ p, const float* __restrict__ g, float* __restrict__ m, float* __restrict__ v,
                            const long long* __restrict__ seg_off, const float* __restrict__ sq, float clipnorm, float lr_t,
                            float b1, float b2, float eps, float wd_lr, const float* __restrict__ hyper_dev) {
  if (hyper_dev) { lr_t = hyper_dev[0]; wd_lr = hyper_dev[1]; }  // graph replay: step-dependent scalars live in HBM
  const int s = blockIdx.x;
  const long long a = seg_off[s], b = seg_off[s + 1];
  float clip = 1.f;
  if (clipnorm > 0.f) {
    const float nrm = sqrtf(sq[s]);
    if (nrm > clipnorm) clip = clipnorm / nrm;
  }
#pragma unroll 4
  for (long long i = a + blockIdx.y * blockDim.x + threadIdx.x; i < b; i += (long long)gridDim.y * blockDim.x) {
    const float gi = g[i] * clip;
    const float mi = m[i] + (gi - m[i]) * (1.f - b1);
    const float vi = v[i] + (gi * gi - v[i]) * (1.f - b2);
    m[i] = mi; v[i] = vi;
    float pi = p[i];
    if (wd_lr > 0.f) pi -= wd_lr * pi;
    p[i] = pi - lr_t * mi / (sqrtf(vi) + eps);
  }
}


}  // namespace bfk
