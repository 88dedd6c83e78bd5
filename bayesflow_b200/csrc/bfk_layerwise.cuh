// Layer-wise sampler for networks that do not fit the fused on-chip kernel (W > 256 or D > 8: BASELINE.json configs[3],
// widths (1024,)*8, D = 64).  Same math as bfk_sampler.cuh (time_mlp.py:125-147, conditional_dense_block.py:114-131,
// integrate.py:31-288), but one tcgen05 GEMM launch per Dense layer (bfk_gemm.cuh) and one elementwise kernel between
// two GEMMs.  Rows are processed in CHUNKS sized so that a chunk's activations (fp32 GEMM output + fp16 operand image)
// stay in the 126 MB L2 between producer and consumer; a chunk runs its whole ODE trajectory before the next starts
// (rows are independent, fixed-step schedules are shared), so HBM sees the state once in and once out.
#pragma once
#include <cuda_runtime.h>
#include <math_constants.h>

#include "bfk_train.cuh"
#include "bfk_prologue.cuh"

namespace bfk {

constexpr int LW_SCALARS = 8;
__host__ __device__ inline int lw_table_stride(const bfk_net_config& c) { return c.L * c.W + LW_SCALARS; }

// times (n) -> table (n, L*W + 8): FiLM shift of every block at that evaluation (fourier_embedding.py:57-76,
// featurewise_linear_modulation.py:25-33) + {coefZ, coefP, time, network time}; weights by pointer (no packed blob).
__global__ void lw_eval_table_kernel(bfk_net_config cfg, bfk_weights w, bfk_schedule sched, const float* __restrict__ times,
                                     long long n, float* __restrict__ table) {
  __shared__ float emb[80];
  __shared__ float sc[4];
  const int W = cfg.W, T = cfg.T, Th = (cfg.T - 1) / 2, stride = lw_table_stride(cfg);
  for (long long e = blockIdx.x; e < n; e += gridDim.x) {
    __syncthreads();
    if (threadIdx.x == 0) {
      double cz, cp, nt;
      schedule_coefs(sched, (double)times[e], &cz, &cp, &nt);
      sc[0] = (float)cz; sc[1] = (float)cp; sc[2] = times[e]; sc[3] = (float)nt;
    }
    __syncthreads();
    const float t = sc[3];
    if (T == 1) {
      if (threadIdx.x == 0) emb[0] = t;
    } else {
      for (int i = threadIdx.x; i < Th; i += blockDim.x) {
        float proj = t * w.freq[i];
        proj = proj * 2.f;
        proj = proj * 3.14159265358979323846f;
        proj = proj * cfg.fourier_scale;
        emb[1 + i] = sinf(proj);
        emb[1 + Th + i] = cosf(proj);
      }
      if (threadIdx.x == 0) emb[0] = t;
    }
    __syncthreads();
    float* row = table + e * stride;
    for (int idx = threadIdx.x; idx < cfg.L * W; idx += blockDim.x) {
      const int l = idx / W, k = idx % W;
      const float* fk = w.film_b_k[l];
      float acc = 0.f;
      for (int j = 0; j < T; ++j) acc = fmaf(emb[j], fk[(size_t)j * W + k], acc);
      row[idx] = acc + w.film_b_b[l][k];
    }
    if (threadIdx.x < LW_SCALARS) row[cfg.L * W + threadIdx.x] = threadIdx.x < 4 ? sc[threadIdx.x] : 0.f;
  }
}

// conditions half of TimeMLP.call, once per dataset (time_mlp.py:133-139; samplers.py:51-63 repeats it per draw):
//   concat: post[g] = bm + act(c Wc + bc) . Wm[W:2W, :]      add: pre[g] = c Wc + bc, post[g] = bm
__global__ void lw_cond_prologue_kernel(bfk_net_config cfg, bfk_weights w, const float* __restrict__ cond, long long G,
                                        float* __restrict__ pre_add, float* __restrict__ post_add) {
  extern __shared__ float sh[];  // [C] conditions, [W] act(hc)
  const int W = cfg.W, C = cfg.C;
  float* sc = sh;
  float* sa = sh + C;
  const float* mc = w.m_k + (size_t)W * W;
  for (long long g = blockIdx.x; g < G; g += gridDim.x) {
    __syncthreads();
    for (int i = threadIdx.x; i < C; i += blockDim.x) sc[i] = cond[g * C + i];
    __syncthreads();
    for (int k = threadIdx.x; k < W; k += blockDim.x) {
      float hc = w.c_b[k];
      for (int i = 0; i < C; ++i) hc = fmaf(sc[i], w.c_k[(size_t)i * W + k], hc);
      if (cfg.merge == BFK_MERGE_ADD) {
        pre_add[g * W + k] = hc;
        post_add[g * W + k] = w.m_b[k];
      } else {
        float d;
        sa[k] = t_act(cfg.activation, hc, d);
      }
    }
    if (cfg.merge == BFK_MERGE_CONCAT) {
      __syncthreads();
      for (int k = threadIdx.x; k < W; k += blockDim.x) {
        float acc = w.m_b[k];
        for (int j = 0; j < W; ++j) acc = fmaf(sa[j], mc[(size_t)j * W + k], acc);
        post_add[g * W + k] = acc;
      }
    }
  }
}

// out = y + h * sum_j coef_j * k_j in the reference's operation order (integrate.py:47-54: acc = 0; acc = acc + c*k;
// y + h*acc; separate roundings).  Writes any of: fp32 state (ldo), the fp16 operand image of the next evaluation,
// and x_out (last step).  grid (KU8, tile chunks), 128 threads = rows of one tile.
struct LwCombArgs {
  const float* y; long long ldy;
  const float* k[BFK_MAX_STAGES]; float coef[BFK_MAX_STAGES]; int nk; long long ldk;
  const float* h_ptr;
  float* out; long long ldo;
  float* x_out; long long ldx;
  unsigned char* img; int KU8;
  long long rows, n_tiles; int D;
};
__global__ void lw_combine_kernel(const LwCombArgs a) {
  const int k8 = blockIdx.x;
  const float h = a.nk > 0 ? *a.h_ptr : 0.f;
  for (long long tile = blockIdx.y; tile < a.n_tiles; tile += gridDim.y) {
    const long long r = tile * 128 + threadIdx.x;
    float v[8];
    if (r < a.rows && k8 * 8 + 8 <= a.D && !((a.ldy | a.ldk | a.ldo | a.ldx) & 3)) {
      // whole octet inside the state: 128-bit loads / stores (the scalar path below costs 8 strided sectors per value)
      ld8(a.y + r * a.ldy + k8 * 8, v);
      if (a.nk > 0) {
        float acc[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) acc[i] = 0.f;
        for (int j = 0; j < a.nk; ++j) {
          float kk[8];
          ld8(a.k[j] + r * a.ldk + k8 * 8, kk);
#pragma unroll
          for (int i = 0; i < 8; ++i) acc[i] = __fadd_rn(acc[i], __fmul_rn(a.coef[j], kk[i]));
        }
#pragma unroll
        for (int i = 0; i < 8; ++i) v[i] = __fadd_rn(v[i], __fmul_rn(h, acc[i]));
      }
      if (a.out) st8(a.out + r * a.ldo + k8 * 8, v);
      if (a.x_out) st8(a.x_out + r * a.ldx + k8 * 8, v);
      if (a.img) *reinterpret_cast<uint4*>(a.img + img_off(r, k8, a.KU8)) = pack8(v);
      continue;
    }
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const int c = k8 * 8 + i;
      float o = 0.f;
      if (r < a.rows && c < a.D) {
        o = a.y[r * a.ldy + c];
        if (a.nk > 0) {
          float acc = 0.f;
          for (int j = 0; j < a.nk; ++j) acc = __fadd_rn(acc, __fmul_rn(a.coef[j], a.k[j][r * a.ldk + c]));
          o = __fadd_rn(o, __fmul_rn(h, acc));
        }
        if (a.out) a.out[r * a.ldo + c] = o;
        if (a.x_out) a.x_out[r * a.ldx + c] = o;
      }
      v[i] = o;
    }
    if (a.img) *reinterpret_cast<uint4*>(a.img + img_off(r, k8, a.KU8)) = pack8(v);
  }
}

// img = act(P + group_bias[(row0 + r) / rpg]) : the two activations around merge_proj (time_mlp.py:139-140) with the
// hoisted per-dataset vector added first.  grid (K8, tile chunks), 128 threads.
__global__ void lw_act_kernel(const float* __restrict__ P, long long ld, long long rows, long long n_tiles, int K8, int act,
                              const float* __restrict__ gbias, long long row0, long long rpg, unsigned char* __restrict__ img) {
  const int k8 = blockIdx.x;
  for (long long tile = blockIdx.y; tile < n_tiles; tile += gridDim.y) {
    const long long r = tile * 128 + threadIdx.x;
    float v[8];
    if (r < rows) {
      ld8(P + r * ld + k8 * 8, v);
      if (gbias) {
        float b[8];
        ld8(gbias + ((row0 + r) / rpg) * ld + k8 * 8, b);
#pragma unroll
        for (int i = 0; i < 8; ++i) v[i] += b[i];
      }
      float d;
#pragma unroll
      for (int i = 0; i < 8; ++i) v[i] = t_act(act, v[i], d);
    } else {
#pragma unroll
      for (int i = 0; i < 8; ++i) v[i] = 0.f;
    }
    *reinterpret_cast<uint4*>(img + img_off(r, k8, K8)) = pack8(v);
  }
}

// k = (coefZ * y_stage + coefP * pred) * target_mask   (diffusion_model.py:539-557, flow_matching.py:233-235)
__global__ void lw_finish_kernel(float* __restrict__ k, long long ldk, const float* __restrict__ ys, long long ldy,
                                 const float* __restrict__ scal, const float* __restrict__ tmask, long long rows, int D) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= rows * D) return;
  const long long r = i / D;
  const int c = (int)(i % D);
  float v = k[r * ldk + c];
  if (ys) v = fmaf(scal[0], ys[r * ldy + c], scal[1] * v);
  if (tmask) v *= tmask[c];
  k[r * ldk + c] = v;
}

}  // namespace bfk
