// bfk_edge.cuh — the approximator edge of the sampling path: running-moment standardization
// (networks/helpers/standardization/standardize.py:50-170) and base-distribution noise generated in HBM
// (distributions/diagonal_normal.py:100-105).  All three kernels are HBM-bound elementwise / reduction work: the rules
// that matter are coalescing, 16-byte accesses where alignment allows and grids sized from the SM count.
// Included at the end of bfk_host.cu (uses its fail()/CK helpers).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace bfk {

// ---------------------------------------------------------------------------------------------------------------------
// Philox4x32-10 (Salmon et al., "Parallel random numbers: as easy as 1, 2, 3", SC'11; same generator as curand /
// torch.cuda).  Counter = {q_lo, q_hi, 0, 0} with q = global element index / 4, key = seed.  The stream depends only on
// (seed, global element index), so a row shard generated on any rank equals the same rows of a single-GPU run.
// ---------------------------------------------------------------------------------------------------------------------
__host__ __device__ inline void philox4x32_10(uint32_t c[4], uint32_t k0, uint32_t k1) {
  const uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u, W0 = 0x9E3779B9u, W1 = 0xBB67AE85u;
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    const uint64_t p0 = (uint64_t)M0 * c[0], p1 = (uint64_t)M1 * c[2];
    const uint32_t n0 = (uint32_t)(p1 >> 32) ^ c[1] ^ k0, n2 = (uint32_t)(p0 >> 32) ^ c[3] ^ k1;
    c[1] = (uint32_t)p1;
    c[3] = (uint32_t)p0;
    c[0] = n0;
    c[2] = n2;
    k0 += W0;
    k1 += W1;
  }
}

// uniform in (0,1): 24 random bits + 1/2, exactly representable, never 0 or 1
__device__ __forceinline__ float u01(uint32_t r) { return ((float)(r >> 8) + 0.5f) * 5.9604644775390625e-8f; }

// 4 standard normals of quad q: Box-Muller on word pairs (0,1) and (2,3), z = rad * (cos, sin)(2*pi*(u2 - 1/2)) with
// rad = sqrt(-2 ln u1).  lg2 / sin / cos go to the MUFU pipe (the angle is centred so |theta| <= pi, where sin.approx and
// cos.approx have 2^-21.4 absolute error); with libm calls the kernel is issue-bound at 0.28 of the HBM roofline.
__device__ __forceinline__ void normal_quad(uint64_t q, uint64_t seed, float z[4]) {
  uint32_t c[4] = {(uint32_t)q, (uint32_t)(q >> 32), 0u, 0u};
  philox4x32_10(c, (uint32_t)seed, (uint32_t)(seed >> 32));
#pragma unroll
  for (int p = 0; p < 2; ++p) {
    const float rad = sqrtf(-1.3862943611198906f * __log2f(u01(c[2 * p])));  // -2 ln u = -2 ln2 * lg2 u
    const float th = 6.283185307179586f * (u01(c[2 * p + 1]) - 0.5f);
    z[2 * p] = rad * __cosf(th);
    z[2 * p + 1] = rad * __sinf(th);
  }
}

// out[e] = mean[(e0+e) % D] + std[(e0+e) % D] * z(e0 + e),  e in [0, n): e0 = global element offset of this shard.
// One thread per quad of the GLOBAL stream; float4 stores when the quad lies inside the shard and is 16-byte aligned.
__global__ void __launch_bounds__(256) normal_fill_kernel(float* __restrict__ out, long long n, long long e0, int D,
                                                          const float* __restrict__ mean, const float* __restrict__ stdv,
                                                          unsigned long long seed, uint32_t* __restrict__ raw) {
  const long long q_first = e0 >> 2, q_last = (e0 + n - 1) >> 2;
  const bool vec = ((e0 & 3) == 0) && ((reinterpret_cast<uintptr_t>(out) & 15) == 0);
  for (long long q = q_first + (long long)blockIdx.x * blockDim.x + threadIdx.x; q <= q_last;
       q += (long long)gridDim.x * blockDim.x) {
    float z[4];
    normal_quad((uint64_t)q, seed, z);
    if (raw) {  // known-answer hook: the raw Philox words of this quad
      uint32_t c[4] = {(uint32_t)q, (uint32_t)((uint64_t)q >> 32), 0u, 0u};
      philox4x32_10(c, (uint32_t)seed, (uint32_t)(seed >> 32));
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const long long e = (q << 2) + j - e0;
        if (e >= 0 && e < n) raw[e] = c[j];
      }
    }
    const long long g = q << 2;
    if (mean || stdv) {
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int d = (int)((g + j) % D);
        z[j] = (mean ? mean[d] : 0.0f) + (stdv ? stdv[d] : 1.0f) * z[j];
      }
    }
    const long long e = g - e0;
    if (vec && e >= 0 && e + 4 <= n) {
      *reinterpret_cast<float4*>(out + e) = make_float4(z[0], z[1], z[2], z[3]);
    } else {
#pragma unroll
      for (int j = 0; j < 4; ++j)
        if (e + j >= 0 && e + j < n) out[e + j] = z[j];
    }
  }
}

// ---------------------------------------------------------------------------------------------------------------------
// Standardize.call (standardize.py:50-133), transformation_type == "location_scale":
//   forward: (x - mean) / std      inverse: x * std + mean      std = m2 > 0 ? sqrt(m2 / count) : 1   (:27-37)
// x is (rows, F) row-major, flat index i -> feature i % F.  In place allowed.
// ---------------------------------------------------------------------------------------------------------------------
template <bool FWD>
__global__ void __launch_bounds__(256) standardize_kernel(const float* __restrict__ x, float* __restrict__ out,
                                                          long long n, int F, const float* __restrict__ mean,
                                                          const float* __restrict__ m2, const float* __restrict__ count) {
  extern __shared__ float sm[];  // mean[F], std[F] (forward: 1/std)
  float* s_mean = sm;
  float* s_std = sm + F;
  const float cnt = count[0];
  for (int f = threadIdx.x; f < F; f += blockDim.x) {
    const float v = m2[f];
    s_mean[f] = mean[f];
    const float sd = v > 0.0f ? sqrtf(v / cnt) : 1.0f;
    s_std[f] = FWD ? 1.0f / sd : sd;  // forward multiplies by 1/std (<= 1 ulp from the reference's division; the IEEE
                                      // division sequence holds the kernel at 0.71 of the copy bandwidth)
  }
  __syncthreads();
  const bool vec = ((reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(out)) & 15) == 0;
  const long long nq = vec ? (n >> 2) : 0;
  const long long stride = (long long)gridDim.x * blockDim.x, t0 = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  // four independent 16-byte loads in flight per thread (Little's law: ~40 KB per SM must be outstanding at 7.7 TB/s)
  for (long long q0 = t0; q0 < nq; q0 += 4 * stride) {
    float4 v[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const long long q = q0 + u * stride;
      if (q < nq) v[u] = __ldcs(reinterpret_cast<const float4*>(x) + q);
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const long long q = q0 + u * stride;
      if (q >= nq) break;
      float* p = reinterpret_cast<float*>(&v[u]);
      int f = (int)((q << 2) % F);
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        p[j] = FWD ? (p[j] - s_mean[f]) * s_std[f] : p[j] * s_std[f] + s_mean[f];
        f = (f + 1 == F) ? 0 : f + 1;
      }
      __stcs(reinterpret_cast<float4*>(out) + q, v[u]);
    }
  }
  for (long long i = (nq << 2) + t0; i < n; i += stride) {
    const int f = (int)(i % F);
    out[i] = FWD ? (x[i] - s_mean[f]) * s_std[f] : x[i] * s_std[f] + s_mean[f];
  }
}

// ---------------------------------------------------------------------------------------------------------------------
// Standardize._update_moments (standardize.py:135-170).  The reference takes the batch mean, then the batch M2 about
// it (two reads of x), then merges with the running totals.  Here x is read ONCE: each thread takes shifted sums over its
// rows of one feature, threads of a block and then the blocks are combined with the same pairwise merge the reference
// uses between batch and running totals (:159-166) — always in a fixed order, so the result is bit-reproducible.
// Thread layout: tx = tid % F is the feature, ty = tid / F the row lane, so a warp reads consecutive floats.
// ---------------------------------------------------------------------------------------------------------------------
struct Moments {
  float n, mean, m2;
};
__host__ __device__ inline Moments merge_moments(Moments a, Moments b) {
  if (b.n == 0.0f) return a;
  if (a.n == 0.0f) return b;
  const float tot = a.n + b.n, delta = b.mean - a.mean;
  Moments r;
  r.n = tot;
  r.mean = a.mean + delta * (b.n / tot);
  r.m2 = a.m2 + b.m2 + delta * delta * (a.n * b.n / tot);
  return r;
}

// partial (gridDim.x, 3, F): per-block count / mean / M2 of each feature
__global__ void moments_partial_kernel(const float* __restrict__ x, long long rows, int F, float* __restrict__ partial) {
  extern __shared__ float sm[];  // [3][TY][F]
  const int TY = blockDim.x / F, tx = threadIdx.x % F, ty = threadIdx.x / F;
  const long long rows_per_block = (rows + gridDim.x - 1) / gridDim.x;
  const long long r_lo = (long long)blockIdx.x * rows_per_block;
  const long long r_hi = r_lo + rows_per_block < rows ? r_lo + rows_per_block : rows;
  Moments m = {0.0f, 0.0f, 0.0f};
  if (ty < TY) {
    // shifted sums about the thread's first value K: no division and no dependent chain in the loop
    float K = 0.0f, s1 = 0.0f, s2 = 0.0f, cnt = 0.0f;
    if (r_lo + ty < r_hi) K = x[(r_lo + ty) * F + tx];
#pragma unroll 4
    for (long long r = r_lo + ty; r < r_hi; r += TY) {
      const float d = x[r * F + tx] - K;
      s1 += d;
      s2 = fmaf(d, d, s2);
      cnt += 1.0f;
    }
    if (cnt > 0.0f) {
      m.n = cnt;
      m.mean = K + s1 / cnt;
      m.m2 = fmaxf(s2 - s1 * s1 / cnt, 0.0f);
    }
    sm[(0 * TY + ty) * F + tx] = m.n;
    sm[(1 * TY + ty) * F + tx] = m.mean;
    sm[(2 * TY + ty) * F + tx] = m.m2;
  }
  __syncthreads();
  int span = 1;
  while (span < TY) span <<= 1;
  for (int s = span >> 1; s >= 1; s >>= 1) {
    if (ty < s && ty + s < TY) {
      Moments a = {sm[(0 * TY + ty) * F + tx], sm[(1 * TY + ty) * F + tx], sm[(2 * TY + ty) * F + tx]};
      Moments b = {sm[(0 * TY + ty + s) * F + tx], sm[(1 * TY + ty + s) * F + tx], sm[(2 * TY + ty + s) * F + tx]};
      a = merge_moments(a, b);
      sm[(0 * TY + ty) * F + tx] = a.n;
      sm[(1 * TY + ty) * F + tx] = a.mean;
      sm[(2 * TY + ty) * F + tx] = a.m2;
    }
    __syncthreads();
  }
  if (ty == 0) {
    float* p = partial + (size_t)blockIdx.x * 3 * F;
    p[tx] = sm[tx];
    p[F + tx] = sm[(1 * TY) * F + tx];
    p[2 * F + tx] = sm[(2 * TY) * F + tx];
  }
}

// 16-byte variant (x 16-byte aligned).  The float4 stream has period P = F / gcd(F, 4): float4 number i holds features
// (4 i + j) % F, which depends only on i % P.  Thread (tx = tid % P, ty = tid / P) reads float4s {g * P + tx} for its
// period groups g, so each of its four lanes always sees ONE feature.  Lanes are merged through shared memory: a tree
// over ty, then the 4P/F replicas of every feature in index order.  Block 0 also takes the < 4P floats left over.
__global__ void moments_partial_vec_kernel(const float* __restrict__ x, long long n, int F, int P,
                                           float* __restrict__ partial) {
  extern __shared__ float sm[];  // [3][TY][4P]
  const int TY = blockDim.x / P, tx = threadIdx.x % P, ty = threadIdx.x / P, LP = 4 * P;
  const long long groups = n / LP;  // whole periods
  const long long per_block = (groups + gridDim.x - 1) / gridDim.x;
  const long long g_lo = (long long)blockIdx.x * per_block;
  const long long g_hi = g_lo + per_block < groups ? g_lo + per_block : groups;
  const float4* x4 = reinterpret_cast<const float4*>(x);
  if (ty < TY) {
    float K[4] = {0.f, 0.f, 0.f, 0.f}, s1[4] = {0.f, 0.f, 0.f, 0.f}, s2[4] = {0.f, 0.f, 0.f, 0.f}, cnt = 0.0f;
    if (g_lo + ty < g_hi) {
      const float4 v = x4[(g_lo + ty) * P + tx];
      K[0] = v.x, K[1] = v.y, K[2] = v.z, K[3] = v.w;
    }
#pragma unroll 4
    for (long long g = g_lo + ty; g < g_hi; g += TY) {
      const float4 v = __ldcs(x4 + g * P + tx);
      const float d0 = v.x - K[0], d1 = v.y - K[1], d2 = v.z - K[2], d3 = v.w - K[3];
      s1[0] += d0, s1[1] += d1, s1[2] += d2, s1[3] += d3;
      s2[0] = fmaf(d0, d0, s2[0]), s2[1] = fmaf(d1, d1, s2[1]), s2[2] = fmaf(d2, d2, s2[2]), s2[3] = fmaf(d3, d3, s2[3]);
      cnt += 1.0f;
    }
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int pos = ty * LP + 4 * tx + j;
      sm[pos] = cnt;
      sm[TY * LP + pos] = cnt > 0.0f ? K[j] + s1[j] / cnt : 0.0f;
      sm[2 * TY * LP + pos] = cnt > 0.0f ? fmaxf(s2[j] - s1[j] * s1[j] / cnt, 0.0f) : 0.0f;
    }
  }
  __syncthreads();
  int span = 1;
  while (span < TY) span <<= 1;
  for (int s = span >> 1; s >= 1; s >>= 1) {
    if (ty < s && ty + s < TY) {
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int a_ = ty * LP + 4 * tx + j, b_ = (ty + s) * LP + 4 * tx + j;
        Moments a = {sm[a_], sm[TY * LP + a_], sm[2 * TY * LP + a_]};
        a = merge_moments(a, Moments{sm[b_], sm[TY * LP + b_], sm[2 * TY * LP + b_]});
        sm[a_] = a.n, sm[TY * LP + a_] = a.mean, sm[2 * TY * LP + a_] = a.m2;
      }
    }
    __syncthreads();
  }
  for (int f = threadIdx.x; f < F; f += blockDim.x) {
    Moments a = {0.0f, 0.0f, 0.0f};
    for (int pos = f; pos < LP; pos += F) a = merge_moments(a, Moments{sm[pos], sm[TY * LP + pos], sm[2 * TY * LP + pos]});
    if (blockIdx.x == 0)  // leftover rows (fewer than 4P / F of them)
      for (long long i = groups * LP + f; i < n; i += F) a = merge_moments(a, Moments{1.0f, x[i], 0.0f});
    float* p = partial + (size_t)blockIdx.x * 3 * F;
    p[f] = a.n, p[F + f] = a.mean, p[2 * F + f] = a.m2;
  }
}

// stats (2F+1): mean[F], m2[F], count — the batch moments, blocks merged in index order
__global__ void __launch_bounds__(256) moments_finish_kernel(const float* __restrict__ partial, int n_blocks, int F,
                                                             float* __restrict__ stats) {
  // one block per feature: thread i folds partials i, i+256, ... (index order), then a fixed-order tree over threads —
  // a single thread walking all ~1200 partials is a 0.4 ms chain of dependent L2 round trips
  __shared__ float s_n[256], s_mean[256], s_m2[256];
  const int f = blockIdx.x, t = threadIdx.x;
  Moments a = {0.0f, 0.0f, 0.0f};
  for (int b = t; b < n_blocks; b += 256) {
    const float* p = partial + (size_t)b * 3 * F;
    a = merge_moments(a, Moments{p[f], p[F + f], p[2 * F + f]});
  }
  s_n[t] = a.n, s_mean[t] = a.mean, s_m2[t] = a.m2;
  __syncthreads();
  for (int s = 128; s >= 1; s >>= 1) {
    if (t < s) {
      a = merge_moments(Moments{s_n[t], s_mean[t], s_m2[t]}, Moments{s_n[t + s], s_mean[t + s], s_m2[t + s]});
      s_n[t] = a.n, s_mean[t] = a.mean, s_m2[t] = a.m2;
    }
    __syncthreads();
  }
  if (t == 0) {
    stats[f] = s_mean[0];
    stats[F + f] = s_m2[0];
    if (f == 0) stats[2 * F] = s_n[0];
  }
}

// running (mean, m2, count) <- merge(running, parts[0], parts[1], ...) in index order (standardize.py:153-170); parts
// are (n_parts, 2F+1) batch-moment records, e.g. one per data-parallel rank after an all-gather
__global__ void moments_merge_kernel(float* __restrict__ mean, float* __restrict__ m2, float* __restrict__ count,
                                     const float* __restrict__ parts, int n_parts, int F) {
  const float cnt0 = count[0];
  __syncthreads();
  float cnt = cnt0;
  for (int f = threadIdx.x; f < F; f += blockDim.x) {
    Moments a = {cnt0, mean[f], m2[f]};
    for (int p = 0; p < n_parts; ++p) {
      const float* s = parts + (size_t)p * (2 * F + 1);
      a = merge_moments(a, Moments{s[2 * F], s[f], s[F + f]});
    }
    mean[f] = a.mean;
    m2[f] = a.m2;
    cnt = a.n;
  }
  __syncthreads();
  if (threadIdx.x == 0) count[0] = cnt;
}

static inline int moments_blocks(long long rows, int F, int n_sms) {
  const int tpb = F <= 256 ? 256 : 1024, TY = tpb / F;
  long long want = (rows + (long long)TY * 8 - 1) / ((long long)TY * 8);  // >= 8 rows per thread
  long long cap = (long long)n_sms * (tpb == 256 ? 8 : 2);
  if (want < 1) want = 1;
  return (int)(want < cap ? want : cap);
}

// out = a*x + b*y + c*z over a flat fp32 buffer (y, z optional)
__global__ void __launch_bounds__(256) lincomb3_kernel(float* __restrict__ out, float a, const float* __restrict__ x, float b,
                                                       const float* __restrict__ y, float c, const float* __restrict__ z,
                                                       long long n) {
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    float v = a * x[i];
    if (y) v = fmaf(b, y[i], v);
    if (z) v = fmaf(c, z[i], v);
    out[i] = v;
  }
}


// FlowMatching.compute_metrics' sampling step (flow_matching.py:143-163) in ONE kernel: x0 ~ N(0, 1) (the Philox stream of
// bfk_normal_fill for the same seed: counter = GLOBAL element index / 4), u ~ U(0, 1) (a second stream keyed by
// seed ^ 0x5851F42D4C957F2D, one word per GLOBAL row), t = u^(1 + alpha), x = t x1 + (1 - t) x0, target = x1 - x0.
// A data-parallel rank passes the global index of its first row, so the batch is the single-GPU batch.
__global__ void __launch_bounds__(256) fm_train_batch_kernel(const float* __restrict__ x1, long long B, int D, long long row0,
                                                             unsigned long long seed, const uint32_t* __restrict__ seed_dev,
                                                             float alpha, float* __restrict__ x0_out, float* __restrict__ t_out,
                                                             float* __restrict__ x, float* __restrict__ target) {
  if (seed_dev) seed = ((unsigned long long)seed_dev[1] << 32) | seed_dev[0];
  const long long r = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= B) return;
  const unsigned long long gr = (unsigned long long)(row0 + r);
  uint32_t c[4] = {(uint32_t)(gr >> 2), (uint32_t)(gr >> 34), 0u, 0u};
  const unsigned long long s2 = seed ^ 0x5851F42D4C957F2Dull;
  philox4x32_10(c, (uint32_t)s2, (uint32_t)(s2 >> 32));
  const float u = u01(c[gr & 3]);
  const float t = alpha == 0.f ? u : powf(u, 1.f + alpha);
  t_out[r] = t;
  unsigned long long qprev = ~0ull;
  float z[4];
  for (int d = 0; d < D; ++d) {
    const unsigned long long e = gr * (unsigned long long)D + d;
    if ((e >> 2) != qprev) { qprev = e >> 2; normal_quad(qprev, seed, z); }
    const float a0 = z[e & 3], a1 = x1[r * D + d];
    if (x0_out) x0_out[r * D + d] = a0;
    x[r * D + d] = t * a1 + (1.f - t) * a0;
    target[r * D + d] = a1 - a0;
  }
}



// conditions (B, C) as the reference's Sampler materialises them (samplers.py:51-63: each dataset's row repeated
// num_samples times): find the repeat count WITHOUT a sort / unique pass.  out[0] = index of the first row that differs
// from row 0 (B if none); out[1] = 0 iff every row equals the first row of its length-out[0] group.
__global__ void first_diff_kernel(const float* __restrict__ c, long long B, int C, long long* __restrict__ out) {
  long long best = B;
  for (long long r = 1 + (long long)blockIdx.x * blockDim.x + threadIdx.x; r < B; r += (long long)gridDim.x * blockDim.x) {
    if (r >= best) break;
    bool diff = false;
    for (int j = 0; j < C; ++j) diff |= (c[r * C + j] != c[j]);
    if (diff) { best = r; break; }
  }
  if (best < B) atomicMin(reinterpret_cast<unsigned long long*>(out), (unsigned long long)best);
}
__global__ void verify_groups_kernel(const float* __restrict__ c, long long B, int C, long long* __restrict__ out) {
  const long long R = out[0];
  if (R <= 0 || R >= B || (B % R) != 0) { if (blockIdx.x == 0 && threadIdx.x == 0 && R < B) out[1] = 1; return; }
  bool bad = false;
  for (long long r = (long long)blockIdx.x * blockDim.x + threadIdx.x; r < B; r += (long long)gridDim.x * blockDim.x) {
    const long long g0 = (r / R) * R;
    if (g0 == r) continue;
    for (int j = 0; j < C; ++j) bad |= (c[r * C + j] != c[g0 * C + j]);
  }
  if (bad) out[1] = 1;
}

// ---- annealed-Langevin updates of the reverse-SDE samplers ------------------------------------------------------------
// utils/integrate.py:924-961 (_apply_corrector) and :1167-1251 (integrate_langevin), one launch per update.  The score is
// rebuilt from the probability-flow velocity the network kernels return: v_ode = fz*x - beta/2 * score (diffusion_model.py
// :539-552)  =>  score = 2 * (fz*x - v_ode) / beta.  One warp per row:
//   per_row != 0 (corrector):  e = 2 * a * (b * ||z|| / max(||score||, 1e-8))^2      (a = alpha_t, b = step_size_factor)
//   per_row == 0 (langevin):   e = a                                                 (a = effective_factor * sigma_t^2)
//   out = x + e * score + sqrt(2 e) * z
__global__ void __launch_bounds__(256) sde_langevin_kernel(const float* __restrict__ x, const float* __restrict__ v,
                                                           const float* __restrict__ z, long long B, int D, float fz,
                                                           float two_over_beta, float a, float b, int per_row,
                                                           float* __restrict__ out) {
  const int lane = threadIdx.x & 31;
  const long long warp = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const long long n_warps = ((long long)gridDim.x * blockDim.x) >> 5;
  for (long long r = warp; r < B; r += n_warps) {
    const float* xr = x + r * D;
    const float* vr = v + r * D;
    const float* zr = z + r * D;
    float e = a;
    if (per_row) {
      float zz = 0.f, ss = 0.f;
      for (int j = lane; j < D; j += 32) {
        const float sc = two_over_beta * (fz * xr[j] - vr[j]);
        zz = fmaf(zr[j], zr[j], zz);
        ss = fmaf(sc, sc, ss);
      }
#pragma unroll
      for (int o = 16; o; o >>= 1) {
        zz += __shfl_xor_sync(0xffffffffu, zz, o);
        ss += __shfl_xor_sync(0xffffffffu, ss, o);
      }
      const float q = b * sqrtf(zz) / fmaxf(sqrtf(ss), 1e-8f);
      e = 2.f * a * q * q;
    }
    const float s2e = sqrtf(2.f * e);
    for (int j = lane; j < D; j += 32) {
      const float sc = two_over_beta * (fz * xr[j] - vr[j]);
      out[r * D + j] = xr[j] + e * sc + s2e * zr[j];
    }
  }
}

}  // namespace bfk

// ---- C ABI ------------------------------------------------------------------------------------------------------------
extern "C" int bfk_normal_fill(float* out, int64_t rows, int32_t D, int64_t row_offset, const float* mean,
                               const float* stdv, uint64_t seed, uint32_t* raw_words, void* stream) {
  if (rows < 0 || D < 1 || row_offset < 0) return fail(-60, "bfk_normal_fill: bad rows/D/row_offset");
  if (rows == 0) return 0;
  if (!out) return fail(-61, "bfk_normal_fill: null output");
  const long long n = (long long)rows * D, e0 = (long long)row_offset * D;
  const long long quads = ((e0 + n - 1) >> 2) - (e0 >> 2) + 1;
  const int sms = bfk_num_sms();
  if (sms <= 0) return fail(-62, "bfk_normal_fill: no CUDA device");
  long long blocks = (quads + 255) / 256, cap = (long long)sms * 8;
  if (blocks > cap) blocks = cap;
  bfk::normal_fill_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(out, n, e0, D, mean, stdv,
                                                                             (unsigned long long)seed, raw_words);
  CK(cudaGetLastError());
  return 0;
}

extern "C" int bfk_standardize(const float* x, float* out, int64_t rows, int32_t F, const float* mean, const float* m2,
                               const float* count, int32_t forward, void* stream) {
  if (rows < 0 || F < 1 || F > 4096) return fail(-63, "bfk_standardize: bad rows/F (F <= 4096)");
  if (rows == 0) return 0;
  if (!x || !out || !mean || !m2 || !count) return fail(-64, "bfk_standardize: null pointer");
  const long long n = (long long)rows * F;
  const int sms = bfk_num_sms();
  if (sms <= 0) return fail(-62, "bfk_standardize: no CUDA device");
  long long blocks = ((n + 3) / 4 + 255) / 256, cap = (long long)sms * 8;
  if (blocks > cap) blocks = cap;
  const size_t smem = 2 * (size_t)F * sizeof(float);
  if (forward)
    bfk::standardize_kernel<true><<<(unsigned)blocks, 256, smem, (cudaStream_t)stream>>>(x, out, n, F, mean, m2, count);
  else
    bfk::standardize_kernel<false><<<(unsigned)blocks, 256, smem, (cudaStream_t)stream>>>(x, out, n, F, mean, m2, count);
  CK(cudaGetLastError());
  return 0;
}

extern "C" size_t bfk_moments_workspace_floats(int64_t rows, int32_t F) {
  if (rows < 0 || F < 1 || F > 1024) return 0;
  int sms = bfk_num_sms();
  if (sms <= 0) sms = 148;
  return (size_t)bfk::moments_blocks(rows, F, sms) * 3 * F;
}

extern "C" int bfk_batch_moments(const float* x, int64_t rows, int32_t F, float* workspace, float* stats, void* stream) {
  if (rows < 0 || F < 1 || F > 1024) return fail(-65, "bfk_batch_moments: bad rows/F (F <= 1024)");
  if (!stats || !workspace || (rows > 0 && !x)) return fail(-66, "bfk_batch_moments: null pointer");
  const int sms = bfk_num_sms();
  if (sms <= 0) return fail(-62, "bfk_batch_moments: no CUDA device");
  cudaStream_t st = (cudaStream_t)stream;
  const int blocks = bfk::moments_blocks(rows, F, sms);
  const int P = F / ((F % 4 == 0) ? 4 : (F % 2 == 0) ? 2 : 1);  // float4 period of the feature pattern
  if ((reinterpret_cast<uintptr_t>(x) & 15) == 0 && rows > 0) {
    const int tpb = P <= 256 ? 256 : 1024, TY = tpb / P;
    const size_t smem = 3 * (size_t)TY * 4 * P * sizeof(float);  // <= 48 KB: TY * P <= 1024
    bfk::moments_partial_vec_kernel<<<blocks, tpb, smem, st>>>(x, (long long)rows * F, F, P, workspace);
  } else {
    const int tpb = F <= 256 ? 256 : 1024, TY = tpb / F;
    const size_t smem = 3 * (size_t)TY * F * sizeof(float);
    bfk::moments_partial_kernel<<<blocks, tpb, smem, st>>>(x, rows, F, workspace);
  }
  CK(cudaGetLastError());
  bfk::moments_finish_kernel<<<F, 256, 0, st>>>(workspace, blocks, F, stats);
  CK(cudaGetLastError());
  return 0;
}

extern "C" int bfk_moments_merge(float* mean, float* m2, float* count, const float* parts, int32_t n_parts, int32_t F,
                                 void* stream) {
  if (n_parts < 0 || F < 1 || F > 1024) return fail(-67, "bfk_moments_merge: bad n_parts/F");
  if (n_parts == 0) return 0;
  if (!mean || !m2 || !count || !parts) return fail(-68, "bfk_moments_merge: null pointer");
  bfk::moments_merge_kernel<<<1, F < 256 ? 256 : 1024, 0, (cudaStream_t)stream>>>(mean, m2, count, parts, n_parts, F);
  CK(cudaGetLastError());
  return 0;
}

extern "C" int bfk_lincomb3(float* out, float a, const float* x, float b, const float* y, float c, const float* z,
                            int64_t numel, void* stream) {
  if (numel < 0) return fail(-69, "bfk_lincomb3: bad numel");
  if (numel == 0) return 0;
  if (!out || !x) return fail(-70, "bfk_lincomb3: null pointer");
  const int sms = bfk_num_sms();
  if (sms <= 0) return fail(-62, "bfk_lincomb3: no CUDA device");
  long long blocks = (numel + 255) / 256, cap = (long long)sms * 8;
  if (blocks > cap) blocks = cap;
  bfk::lincomb3_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(out, a, x, b, y, c, z, numel);
  CK(cudaGetLastError());
  return 0;
}

extern "C" int bfk_fm_train_batch(const float* x1, int64_t B, int32_t D, int64_t row_offset, uint64_t seed, const uint32_t* seed_dev,
                                  float alpha, float* x0_out, float* t_out, float* x, float* target, void* stream_) {
  if (!x1 || !t_out || !x || !target || B <= 0 || D <= 0 || row_offset < 0) return fail(-83, "bfk_fm_train_batch: bad arguments");
  bfk::fm_train_batch_kernel<<<cdiv((size_t)B, 256), 256, 0, (cudaStream_t)stream_>>>(x1, B, D, row_offset, (unsigned long long)seed, seed_dev,
                                                                               alpha, x0_out, t_out, x, target);
  CK(cudaGetLastError());
  return 0;
}

extern "C" int bfk_detect_repeats(const float* cond, int64_t B, int32_t C, int64_t* out2, void* stream_) {
  if (!cond || !out2 || B <= 0 || C <= 0) return fail(-84, "bfk_detect_repeats: bad arguments");
  cudaStream_t st = (cudaStream_t)stream_;
  const long long init[2] = {(long long)B, 0};
  CK(cudaMemcpyAsync(out2, init, sizeof(init), cudaMemcpyHostToDevice, st));
  int sms = bfk_num_sms();
  if (sms <= 0) sms = 148;
  const unsigned grid = (unsigned)((B + 255) / 256 < (long long)sms * 8 ? (B + 255) / 256 : (long long)sms * 8);
  bfk::first_diff_kernel<<<grid, 256, 0, st>>>(cond, B, C, (long long*)out2);
  bfk::verify_groups_kernel<<<grid, 256, 0, st>>>(cond, B, C, (long long*)out2);
  CK(cudaGetLastError());
  return 0;
}

extern "C" int bfk_sde_langevin_update(const float* x, const float* v_ode, const float* z, int64_t B, int32_t D, double fz,
                                       double beta, double a, double b, int32_t per_row, float* out, void* stream_) {
  if (B < 0 || D < 1) return fail(-66, "bfk_sde_langevin_update: bad B/D");
  if (B == 0) return 0;
  if (!x || !v_ode || !z || !out) return fail(-67, "bfk_sde_langevin_update: null pointer");
  if (!(beta > 0.0)) return fail(-68, "bfk_sde_langevin_update: beta = g(t)^2 must be positive");
  const int sms = bfk_num_sms();
  if (sms <= 0) return fail(-62, "bfk_sde_langevin_update: no CUDA device");
  long long blocks = (B + 7) / 8, cap = (long long)sms * 8;
  if (blocks > cap) blocks = cap;
  bfk::sde_langevin_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream_>>>(x, v_ode, z, (long long)B, D, (float)fz,
                                                                               (float)(2.0 / beta), (float)a, (float)b,
                                                                               per_row, out);
  CK(cudaGetLastError());
  return 0;
}

// ---- parameter staging for unequal hidden widths (bfk_net_config.block_width) -------------------------------------------
namespace bfk {
__global__ void unpad2d_kernel(const float* __restrict__ src, int scols, float* __restrict__ dst, int rows, int cols) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= rows * cols) return;
  dst[i] = src[(size_t)(i / cols) * scols + (i % cols)];
}
}  // namespace bfk
extern "C" int bfk_pad2d(const float* src, int32_t rows, int32_t cols, float* dst, int32_t rows_pad, int32_t cols_pad, void* stream) {
  if (!src || !dst || rows < 1 || cols < 1 || rows_pad < rows || cols_pad < cols) return fail(-93, "bfk_pad2d: bad arguments");
  const size_t n = (size_t)rows_pad * cols_pad;
  bfk::pad2d_kernel<<<(unsigned)((n + 255) / 256), 256, 0, (cudaStream_t)stream>>>(src, rows, cols, dst, rows_pad, cols_pad);
  CK(cudaGetLastError());
  return 0;
}
extern "C" int bfk_unpad2d(const float* src, int32_t cols_pad, float* dst, int32_t rows, int32_t cols, void* stream) {
  if (!src || !dst || rows < 1 || cols < 1 || cols_pad < cols) return fail(-94, "bfk_unpad2d: bad arguments");
  const size_t n = (size_t)rows * cols;
  bfk::unpad2d_kernel<<<(unsigned)((n + 255) / 256), 256, 0, (cudaStream_t)stream>>>(src, cols_pad, dst, rows, cols);
  CK(cudaGetLastError());
  return 0;
}
