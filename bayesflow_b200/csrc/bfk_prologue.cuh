// Weight packing and the two per-call prologue kernels (conditions hoist, per-evaluation FiLM / schedule table).
// All fp32 CUDA-core work: O(params + datasets + evaluations), not O(draws x steps).
#pragma once
#include <cuda_runtime.h>
#include <math_constants.h>

#include "bfk_layout.h"
#include "bfk_ptx.cuh"

namespace bfk {

template <int ACT>
__device__ __forceinline__ float act_fn(float x) {
  if (ACT == BFK_ACT_MISH) {  // x*tanh(softplus(x)) = x*p/(p+2), p = e^x (e^x + 2)
    float e = ex2_approx(fminf(x, 20.f) * 1.4426950408889634f);
    float p = e * (e + 2.f);
    return x * p * rcp_approx(p + 2.f);
  } else if (ACT == BFK_ACT_SILU) {
    float e = ex2_approx(fminf(-x, 80.f) * 1.4426950408889634f);
    return x * rcp_approx(1.f + e);
  } else if (ACT == BFK_ACT_RELU) {
    return fmaxf(x, 0.f);
  } else if (ACT == BFK_ACT_GELU) {
    // keras.activations.gelu defaults to approximate=True: 0.5 x (1 + tanh(sqrt(2/pi) (x + 0.044715 x^3)))
    return 0.5f * x * (1.f + tanhf(0.7978845608028654f * (x + 0.044715f * x * x * x)));
  } else {
    return tanhf(x);
  }
}
__device__ __forceinline__ float act_rt(int act, float x) {
  switch (act) {
    case BFK_ACT_MISH: return act_fn<BFK_ACT_MISH>(x);
    case BFK_ACT_SILU: return act_fn<BFK_ACT_SILU>(x);
    case BFK_ACT_RELU: return act_fn<BFK_ACT_RELU>(x);
    case BFK_ACT_GELU: return act_fn<BFK_ACT_GELU>(x);
    default: return act_fn<BFK_ACT_TANH>(x);
  }
}

// (K,N) fp32 row-major (Keras (in,out))  ->  [K/8][N][8] 16-bit: the K-major no-swizzle UMMA operand image, so that any
// slab of K rows is one contiguous byte range (bulk-copyable) and LBO = N*16, SBO = 128.
__global__ void pack_gemm_kernel(const float* __restrict__ src, int K, int N, int bf16, uint16_t* __restrict__ dst,
                                 float scale = 1.f) {
  int idx = blockIdx.x * blockDim.x + threadIdx.x;  // one thread per (k8, n)
  if (idx >= (K / 8) * N) return;
  int n = idx % N, k8 = idx / N;
  uint16_t out[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    float v = src[(size_t)(k8 * 8 + i) * N + n] * scale;
    if (bf16) {
      __nv_bfloat16 b = __float2bfloat16_rn(v);
      out[i] = *reinterpret_cast<uint16_t*>(&b);
    } else {
      __half h = __float2half_rn(v);
      out[i] = *reinterpret_cast<uint16_t*>(&h);
    }
  }
  *reinterpret_cast<uint4*>(dst + (size_t)idx * 8) = *reinterpret_cast<uint4*>(out);
}

__global__ void copy_f32_kernel(const float* __restrict__ src, float* __restrict__ dst, size_t n) {
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) dst[i] = src ? src[i] : 0.f;
}
__global__ void scale_f32_kernel(float* __restrict__ dst, float s, size_t n) {
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) dst[i] *= s;
}
__global__ void fill_f32_kernel(float* __restrict__ dst, float v, size_t n) {
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) dst[i] = v;
}
// src (rows, cols) -> dst (rows_pad, cols_pad) zero padded;  transpose=1: dst[c][r] = src[r][c] with dst (cols_pad? no: see use)
__global__ void pad2d_kernel(const float* __restrict__ src, int rows, int cols, float* __restrict__ dst, int drows, int dcols) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= drows * dcols) return;
  int r = i / dcols, c = i % dcols;
  dst[i] = (r < rows && c < cols) ? src[(size_t)r * cols + c] : 0.f;
}

// Output projector folded through the last block's LayerNorm affine (fast epilogue):
//   wog[d][c] = gamma[c] * Wo[c][d];  s0[d] = sum_c beta[c] * Wo[c][d] + bo[d];  s1[d] = sum_c wog[d][c]
__global__ void fold_proj_kernel(const float* __restrict__ out_k, const float* __restrict__ out_b,
                                 const float* __restrict__ gamma, const float* __restrict__ beta, int W, int D, int DP,
                                 float* __restrict__ wog, float* __restrict__ s0s1) {
  const int d = threadIdx.x;
  if (d >= DP) return;
  double a0 = 0.0, a1 = 0.0;
  for (int c = 0; c < W; ++c) {
    const float w = d < D ? out_k[(size_t)c * D + d] : 0.f;
    const float g = (gamma ? gamma[c] : 1.f) * w;
    wog[(size_t)d * W + c] = g;
    a1 += (double)g;
    a0 += (double)(beta ? beta[c] : 0.f) * (double)w;
  }
  s0s1[d] = (float)(a0 + (d < D ? (double)out_b[d] : 0.0));
  s0s1[DP + d] = (float)a1;
}

// Folded layout (bfk_sampler_fold.cuh): the Dense kernel of block l >= 1 consumes the UN-normalised output v of block
// l-1, so LayerNorm's affine moves into the weights: W' = gamma_prev o W (rows scaled, then rounded to the operand type).
// With (mu, r) the row statistics of v:  log2(e) (LN(v) W + b) = log2(e) r [ v W' - mu g + (1 / (log2(e) r)) cL ],
// g_j = sum_k W'_kj (of the ROUNDED values: exactly what the tensor core multiplies v by), cL_j = log2(e) (b_j +
// sum_k beta_prev_k W_kj).  The bracket is ONE accumulation: XK extra K rows carry -g and cL as 16-bit hi/lo pairs
// against per-row A entries (mu_hi, mu_lo, mu_hi, ir_hi, ir_lo, ir_hi), ir = 1 / (log2(e) r), written by the epilogue:
//   rows K+0..2 = (-g_hi, -g_hi, -g_lo)   rows K+3..5 = (cL_hi, cL_hi, cL_lo)   rest 0
// One thread per output column.  gamma_prev == nullptr (block 0): plain pack, g = 0, gamma image = 1.
__device__ __forceinline__ uint16_t round16(float x, int bf16, float* back) {
  if (bf16) {
    __nv_bfloat16 b = __float2bfloat16_rn(x);
    *back = __bfloat162float(b);
    return *reinterpret_cast<uint16_t*>(&b);
  }
  __half b = __float2half_rn(x);
  *back = __half2float(b);
  return *reinterpret_cast<uint16_t*>(&b);
}
__global__ void fold_pack_kernel(const float* __restrict__ Wk, const float* __restrict__ bd, const float* __restrict__ gamma_prev,
                                 const float* __restrict__ beta_prev, int K, int N, int bf16, uint16_t* __restrict__ dst,
                                 float* __restrict__ gam_out) {
  const int n = blockIdx.x * blockDim.x + threadIdx.x;
  if (n >= N) return;
  float g = 0.f;
  double cacc = 0.0;
  for (int k = 0; k < K; ++k) {
    const float w = Wk[(size_t)k * N + n];
    float back;
    const uint16_t h = round16(gamma_prev ? gamma_prev[k] * w : w, bf16, &back);
    dst[((size_t)(k >> 3) * N + n) * 8 + (k & 7)] = h;
    g += back;
    if (beta_prev) cacc += (double)beta_prev[k] * (double)w;
  }
  const float cL = (bd[n] + (float)cacc) * 1.4426950408889634f;
  const float gg = gamma_prev ? -g : 0.f;
  float g_hi, g_lo, c_hi, c_lo;
  const uint16_t gh = round16(gg, bf16, &g_hi), gl = round16(gg - g_hi, bf16, &g_lo);
  const uint16_t ch = round16(cL, bf16, &c_hi), cl = round16(cL - c_hi, bf16, &c_lo);
  const uint16_t x[6] = {gh, gh, gl, ch, ch, cl};
  uint16_t* e = dst + ((size_t)(K >> 3) * N + n) * 8;  // K-chunk K/8 of this column (K-chunk K/8+1 stays zero)
  for (int i = 0; i < 6; ++i) e[i] = x[i];
  gam_out[n] = gamma_prev ? gamma_prev[n] : 1.f;
}

// out_scale: 1, or log2(e) for the folded layout, whose pre-activations are carried times log2(e)
__global__ void cond_prologue_kernel(bfk_net_config cfg, const unsigned char* __restrict__ packed, PackedLayout L,
                                     const float* __restrict__ cond, long long G, float* __restrict__ pre_add,
                                     float* __restrict__ post_add, float out_scale) {
  extern __shared__ float sh[];  // [C] conditions, [W] act(hc)
  const int W = cfg.W, C = cfg.C;
  float* sc = sh;
  float* sa = sh + C;
  const float* ck = reinterpret_cast<const float*>(packed + L.c_k);
  const float* cb = reinterpret_cast<const float*>(packed + L.c_b);
  const float* mc = reinterpret_cast<const float*>(packed + L.m_c);
  const float* mb = reinterpret_cast<const float*>(packed + L.m_b);
  for (long long g = blockIdx.x; g < G; g += gridDim.x) {
    __syncthreads();
    for (int i = threadIdx.x; i < C; i += blockDim.x) sc[i] = cond[g * C + i];
    __syncthreads();
    for (int k = threadIdx.x; k < W; k += blockDim.x) {
      float hc = cb[k];
      for (int i = 0; i < C; ++i) hc = fmaf(sc[i], ck[(size_t)i * W + k], hc);
      if (cfg.merge == BFK_MERGE_ADD) {
        pre_add[g * W + k] = hc * out_scale;
        post_add[g * W + k] = mb[k] * out_scale;
      } else {
        sa[k] = act_rt(cfg.activation, hc);
      }
    }
    if (cfg.merge == BFK_MERGE_CONCAT) {
      __syncthreads();
      for (int k = threadIdx.x; k < W; k += blockDim.x) {
        float acc = mb[k];
        for (int j = 0; j < W; ++j) acc = fmaf(sa[j], mc[(size_t)j * W + k], acc);
        post_add[g * W + k] = acc * out_scale;
      }
    }
  }
}

// log-SNR lambda(t) and beta(t) = g(t)^2 of the inference-time schedule (noise_schedule.py:70-99;
// edm_noise_schedule.py:61-108; cosine_noise_schedule.py:60-82), in double.
__host__ __device__ inline void schedule_lam_beta(const bfk_schedule& s, double t, double* lam_out, double* beta_out) {
  double lam, beta;
  if (s.schedule == BFK_SCHED_EDM) {
    double smin = pow(s.sigma_min, 1.0 / s.rho), smax = pow(s.sigma_max, 1.0 / s.rho);
    double u = smax + (1.0 - t) * (smin - smax);
    lam = -2.0 * s.rho * log(u);
    double dsnr = 2.0 * s.rho * (smin - smax) / u;
    double factor = exp(-lam) / (1.0 + exp(-lam));
    beta = -factor * dsnr;
  } else {
    double tt = s.t_min + (s.t_max - s.t_min) * t;
    lam = -2.0 * log(tan(CUDART_PI * tt * 0.5)) + 2.0 * s.shift;
    // reference quirk kept: derivative_log_snr truncates the already truncated time again (cosine_noise_schedule.py:70-75)
    double t_back = 2.0 / CUDART_PI * atan(exp((2.0 * s.shift - lam) * 0.5));
    double t2 = s.t_min + (s.t_max - s.t_min) * t_back;
    double dsnr_dt = -(2.0 * CUDART_PI) / sin(CUDART_PI * t2) * (s.t_max - s.t_min);
    double factor = exp(-lam) / (1.0 + exp(-lam));
    beta = -factor * dsnr_dt;
  }
  *lam_out = lam;
  *beta_out = beta;
}

// DiffusionModel probability-flow velocity as  v = coefZ * z + coefP * pred  (diffusion_model.py:351-405, 466-489,
// 539-552; noise_schedule.py:70-119; edm_noise_schedule.py:61-108; cosine_noise_schedule.py:60-82), evaluated in double.
__host__ __device__ inline void schedule_coefs(const bfk_schedule& s, double t, double* coefZ, double* coefP, double* net_time) {
  if (s.model == BFK_MODEL_FM) {
    *coefZ = 0.0; *coefP = 1.0; *net_time = t;
    return;
  }
  double lam, beta;
  schedule_lam_beta(s, t, &lam, &beta);
  double alpha, sigma;
  if (s.exploding) { alpha = 1.0; sigma = sqrt(exp(-lam)); }
  else { alpha = sqrt(1.0 / (1.0 + exp(-lam))); sigma = sqrt(1.0 / (1.0 + exp(lam))); }
  double pz, pp;  // x_pred = pz*z + pp*pred
  switch (s.prediction) {
    case BFK_PRED_VELOCITY: pz = alpha; pp = -sigma; break;
    case BFK_PRED_NOISE: pz = 1.0 / alpha; pp = -sigma / alpha; break;
    case BFK_PRED_F: {
      double sd = s.sigma_data, e = exp(-lam);
      pz = sd * sd * alpha / (e + sd * sd);
      pp = exp(-lam * 0.5) * sd / sqrt(e + sd * sd);
    } break;
    case BFK_PRED_X: pz = 0.0; pp = 1.0; break;
    default: pz = 1.0 / alpha; pp = sigma * sigma / alpha; break;
  }
  // score = ((alpha*pz - 1) z + alpha*pp pred)/sigma^2 ; v = f - 0.5 beta score, f = -0.5 beta z (preserving) or 0
  double s2 = sigma * sigma;
  double fz = s.exploding ? 0.0 : -0.5 * beta;
  *coefZ = fz - 0.5 * beta * (alpha * pz - 1.0) / s2;
  *coefP = -0.5 * beta * alpha * pp / s2;
  *net_time = 2.0 * ((lam - s.log_snr_min) / (s.log_snr_max - s.log_snr_min)) - 1.0;  // diffusion_model.py:711-717
}

// times (n) -> table (n, stride): [L][W] FiLM beta, ([L][W] FiLM gamma), TABLE_SCALARS scalars
// chase != 0 (folded layout): the FiLM shift of block l >= 1 also carries LayerNorm beta of block l-1 (bfk_sampler_fold.cuh)
__global__ void eval_table_kernel(bfk_net_config cfg, const unsigned char* __restrict__ packed, PackedLayout L,
                                  bfk_schedule sched, const float* __restrict__ times, long long n, int stride,
                                  float* __restrict__ table, int chase) {
  __shared__ float emb[80];
  __shared__ float sc[4];
  const int W = cfg.W, T = cfg.T;
  const float* freq = reinterpret_cast<const float*>(packed + L.freq);
  for (long long e = blockIdx.x; e < n; e += gridDim.x) {
    __syncthreads();
    if (threadIdx.x == 0) {
      double cz, cp, nt;
      schedule_coefs(sched, (double)times[e], &cz, &cp, &nt);
      sc[0] = (float)cz; sc[1] = (float)cp; sc[2] = times[e]; sc[3] = (float)nt;
    }
    __syncthreads();
    const float t = sc[3];
    // fourier_embedding.py:71-73: proj = t * w * 2 * pi * scale (left to right, fp32); emb = [t, sin, cos]
    if (T == 1) {
      if (threadIdx.x == 0) emb[0] = t;
    } else {
      for (int i = threadIdx.x; i < L.Th; i += blockDim.x) {
        float proj = t * freq[i];
        proj = proj * 2.f;
        proj = proj * 3.14159265358979323846f;
        proj = proj * cfg.fourier_scale;
        emb[1 + i] = sinf(proj);
        emb[1 + L.Th + i] = cosf(proj);
      }
      if (threadIdx.x == 0) emb[0] = t;
    }
    __syncthreads();
    float* row = table + e * stride;
    const int nv = cfg.film_use_gamma ? 2 : 1;
    for (int idx = threadIdx.x; idx < nv * cfg.L * W; idx += blockDim.x) {
      int which = idx / (cfg.L * W), rem = idx % (cfg.L * W), l = rem / W, k = rem % W;
      const float* fk = reinterpret_cast<const float*>(packed + (which ? L.film_g_k : L.film_b_k)) + (size_t)l * T * W;
      const float* fb = reinterpret_cast<const float*>(packed + (which ? L.film_g_b : L.film_b_b)) + (size_t)l * W;
      float acc = 0.f;
      for (int j = 0; j < T; ++j) acc = fmaf(emb[j], fk[(size_t)j * W + k], acc);
      acc = acc + fb[k];
      if (chase && which == 0 && l > 0) acc += reinterpret_cast<const float*>(packed + L.ln_beta)[(size_t)(l - 1) * W + k];
      row[idx] = acc;
    }
    if (threadIdx.x < TABLE_SCALARS) row[nv * cfg.L * W + threadIdx.x] = threadIdx.x < 4 ? sc[threadIdx.x] : 0.f;
  }
}

}  // namespace bfk
