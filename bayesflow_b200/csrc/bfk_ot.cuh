// bfk_ot.cuh — mini-batch optimal-transport pairing for FlowMatching training:
// utils/optimal_transport/log_sinkhorn.py:14-172 (log-stabilised Sinkhorn-Knopp plan + categorical assignment),
// ot_utils.py:6-34 (squared Euclidean / cosine cost) and :95-164 (conditional weight search).  The n x m matrices stay in
// HBM/L2 as fp32 (n = m = batch; 4096^2 floats = 64 MB, L2-resident on B200); every pass is a coalesced sweep with the
// reductions done in a fixed order (no float atomics except a MAX on non-negative values, which is order-independent).
// Included at the end of bfk_host.cu after bfk_edge.cuh (uses its Philox and the host fail()/CK helpers).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <math.h>

namespace bfk {

enum { OT_ERR = 0, OT_STEPS = 1, OT_DONE = 2, OT_ATOL = 3, OT_MAX_STEPS = 4, OT_ACC = 5, OT_PLAIN = 6, OT_SHIFT = 7,
       OT_CTRL_FLOATS = 8 };
// Plain Sinkhorn-Knopp (utils/optimal_transport/sinkhorn.py:117-160) on the same sweeps: the multiplicative scalings u, v
// are kept as logs (U = log u, V = log v; K = exp(S - max S)), so nothing under- or overflows, and the reference's `tiny`
// terms and its error metric are applied exactly:  sum_j K_ij v_j + tiny = exp(logaddexp(L_i - shift, log tiny)),
// err = max(max_i |log((u' + tiny)/(u + tiny))|, max_j |log((v' + tiny)/(v + tiny))|).
__device__ __forceinline__ float log_add_exp(float a, float b) {
  const float hi = fmaxf(a, b), lo = fminf(a, b);
  if (hi == -INFINITY) return -INFINITY;
  return hi + log1pf(__expf(lo - hi));
}
#define OT_LOG_TINY (-27.631021115928547f) /* log(1e-12) */

// M[i,j] = max(|a_i|^2 + |b_j|^2 - 2 a_i.b_j, 0)  (ot_utils.py:6-16);  cosine: 1 - clip(a_i.b_j / ((|a_i|+eps)(|b_j|+eps)))
template <bool COSINE>
__global__ void __launch_bounds__(256) ot_cost_kernel(const float* __restrict__ a, const float* __restrict__ b, long long n,
                                                      long long m, int D, float* __restrict__ M) {
  const long long j = (long long)blockIdx.x * 32 + threadIdx.x, i = (long long)blockIdx.y * 8 + threadIdx.y;
  if (i >= n || j >= m) return;
  const float* ai = a + i * D;
  const float* bj = b + j * D;
  float aa = 0.f, bb = 0.f, ab = 0.f;
  for (int d = 0; d < D; ++d) {
    const float x = ai[d], y = bj[d];
    aa = fmaf(x, x, aa);
    bb = fmaf(y, y, bb);
    ab = fmaf(x, y, ab);
  }
  float r;
  if (COSINE) {
    const float sim = ab / ((sqrtf(aa) + 1e-8f) * (sqrtf(bb) + 1e-8f));
    r = 1.0f - fminf(fmaxf(sim, -1.0f), 1.0f);
  } else {
    r = fmaxf(aa + bb - 2.0f * ab, 0.0f);
  }
  M[i * m + j] = r;
}

// count of (w * C[i,j] <= M[i,i] - M[i,j])  (ot_utils.py:117-121); integer atomics: exact and order-independent
__global__ void __launch_bounds__(256) ot_ratio_kernel(const float* __restrict__ M, const float* __restrict__ Cm, long long n,
                                                       float w, unsigned long long* __restrict__ count) {
  const long long total = n * n, stride = (long long)gridDim.x * blockDim.x;
  unsigned int local = 0;
  for (long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += stride) {
    const long long i = e / n;
    local += (w * Cm[e] <= M[i * n + i] - M[e]) ? 1u : 0u;
  }
  for (int o = 16; o > 0; o >>= 1) local += __shfl_down_sync(0xffffffffu, local, o);
  if ((threadIdx.x & 31) == 0 && local) atomicAdd(count, (unsigned long long)local);
}

// S = -(M + w * C) / regularization, in place over M  (log_sinkhorn.py:98-110)
__global__ void __launch_bounds__(256) ot_scale_kernel(float* __restrict__ M, const float* __restrict__ Cm, float w, float reg,
                                                       long long numel) {
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x; e < numel; e += stride) {
    const float c = Cm ? M[e] + w * Cm[e] : M[e];
    M[e] = -c / reg;
  }
}

__device__ __forceinline__ void lse_push(float& mx, float& sm, float v) {
  if (v > mx) {
    sm = sm * __expf(mx - v) + 1.0f;
    mx = v;
  } else if (v > -INFINITY) {
    sm += __expf(v - mx);
  }
}
// four values at once: one rescale of the running sum per group and four independent exponentials, so the loads of a group
// are in flight together and the loop-carried chain is one exp per four elements (a one-at-a-time loop is bound by
// L2 latency x row length: 47 us per 4096-wide sweep)
__device__ __forceinline__ void lse_push4(float& mx, float& sm, float a, float b, float c, float d) {
  const float g = fmaxf(fmaxf(a, b), fmaxf(c, d));
  if (g == -INFINITY) return;
  if (g > mx) {
    sm *= __expf(mx - g);  // mx = -inf: sm is 0 and exp(-inf) = 0
    mx = g;
  }
  sm += __expf(a - mx) + __expf(b - mx) + __expf(c - mx) + __expf(d - mx);
}
__device__ __forceinline__ void lse_merge(float& mx, float& sm, float mx2, float sm2) {
  if (mx2 == -INFINITY) return;
  if (mx == -INFINITY) { mx = mx2; sm = sm2; return; }
  if (mx2 > mx) { sm = sm * __expf(mx - mx2) + sm2; mx = mx2; }
  else sm += sm2 * __expf(mx2 - mx);
}

// u_next[i] = log_a - logsumexp_j(S[i,j] + v[j]);  err = max_i |u_next - u|   (log_sinkhorn.py:131-137).  One warp per row.
__global__ void __launch_bounds__(256) ot_row_lse_kernel(const float* __restrict__ S, long long n, long long m, float log_a,
                                                         const float* __restrict__ log_a_vec, float* __restrict__ u,
                                                         const float* __restrict__ v, float* __restrict__ ctrl) {
  if (ctrl[OT_DONE] != 0.f) return;
  const int lane = threadIdx.x & 31;
  const long long i = (long long)blockIdx.x * 8 + (threadIdx.x >> 5);
  if (i >= n) return;
  const float* row = S + i * m;
  float mx = -INFINITY, sm = 0.f;
  long long j = lane;
  for (; j + 96 < m; j += 128)
    lse_push4(mx, sm, row[j] + v[j], row[j + 32] + v[j + 32], row[j + 64] + v[j + 64], row[j + 96] + v[j + 96]);
  for (; j < m; j += 32) lse_push(mx, sm, row[j] + v[j]);
  for (int o = 16; o > 0; o >>= 1) {
    const float mx2 = __shfl_xor_sync(0xffffffffu, mx, o), sm2 = __shfl_xor_sync(0xffffffffu, sm, o);
    // symmetric merge so that every lane ends with the same value
    float a = mx, b = sm;
    lse_merge(a, b, mx2, sm2);
    mx = a, sm = b;
  }
  if (lane == 0) {
    const float la = log_a_vec ? log_a_vec[i] : log_a;
    float un, d;
    if (ctrl[OT_PLAIN] != 0.f) {
      un = la - log_add_exp(mx + logf(sm) - ctrl[OT_SHIFT], OT_LOG_TINY);
      d = fabsf(log_add_exp(un, OT_LOG_TINY) - log_add_exp(u[i], OT_LOG_TINY));
    } else {
      un = la - (mx + logf(sm));
      d = fabsf(un - u[i]);
    }
    u[i] = un;
    atomicMax(reinterpret_cast<unsigned int*>(ctrl + OT_ACC), __float_as_uint(d == d ? d : INFINITY));
  }
}

// column pass, stage 1: (max, sum) of S[i,j] + u[i] over the rows of chunk blockIdx.y -> part[chunk][2][m]
__global__ void __launch_bounds__(256) ot_col_partial_kernel(const float* __restrict__ S, long long n, long long m,
                                                             const float* __restrict__ u, float* __restrict__ part,
                                                             const float* __restrict__ ctrl) {
  if (ctrl[OT_DONE] != 0.f) return;
  const long long j = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= m) return;
  const long long per = (n + gridDim.y - 1) / gridDim.y, lo = (long long)blockIdx.y * per;
  const long long hi = lo + per < n ? lo + per : n;
  float mx = -INFINITY, sm = 0.f;
  long long i = lo;
  for (; i + 3 < hi; i += 4)
    lse_push4(mx, sm, S[i * m + j] + u[i], S[(i + 1) * m + j] + u[i + 1], S[(i + 2) * m + j] + u[i + 2],
              S[(i + 3) * m + j] + u[i + 3]);
  for (; i < hi; ++i) lse_push(mx, sm, S[i * m + j] + u[i]);
  part[((long long)blockIdx.y * 2) * m + j] = mx;
  part[((long long)blockIdx.y * 2 + 1) * m + j] = sm;
}
// stage 2: chunks merged in index order; v_next[j] = log_b - logsumexp  (log_sinkhorn.py:133)
__global__ void __launch_bounds__(256) ot_col_finish_kernel(const float* __restrict__ part, int chunks, long long m, float log_b,
                                                            const float* __restrict__ log_b_vec, float* __restrict__ v,
                                                            float* __restrict__ ctrl) {
  if (ctrl[OT_DONE] != 0.f) return;
  const long long j = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= m) return;
  float mx = -INFINITY, sm = 0.f;
  for (int c = 0; c < chunks; ++c) lse_merge(mx, sm, part[((long long)c * 2) * m + j], part[((long long)c * 2 + 1) * m + j]);
  const float lb = log_b_vec ? log_b_vec[j] : log_b;
  if (ctrl[OT_PLAIN] != 0.f) {
    const float vn = lb - log_add_exp(mx + logf(sm) - ctrl[OT_SHIFT], OT_LOG_TINY);
    const float d = fabsf(log_add_exp(vn, OT_LOG_TINY) - log_add_exp(v[j], OT_LOG_TINY));
    v[j] = vn;
    atomicMax(reinterpret_cast<unsigned int*>(ctrl + OT_ACC), __float_as_uint(d == d ? d : INFINITY));
  } else {
    v[j] = lb - (mx + logf(sm));
  }
}
// max of a matrix into ctrl[OT_SHIFT] and the plain-mode flag (the shift of K = exp(S - max S), sinkhorn.py:119)
__global__ void __launch_bounds__(256) ot_max_kernel(const float* __restrict__ S, long long numel, float* __restrict__ ctrl) {
  float mx = -INFINITY;
  for (long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x; e < numel; e += (long long)gridDim.x * blockDim.x)
    mx = fmaxf(mx, S[e]);
  for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
  if ((threadIdx.x & 31) == 0) {  // float max through the ordered-int trick (values may be negative)
    int* addr = reinterpret_cast<int*>(ctrl + OT_SHIFT);
    mx += 0.0f;  // -0.0f (a zero cost, S = -M / reg) would compare as INT_MIN below and lose against the -inf identity
    if (mx >= 0.f) atomicMax(addr, __float_as_int(mx));
    else atomicMin(reinterpret_cast<unsigned int*>(addr), __float_as_uint(mx));
  }
}
// loop bookkeeping of keras.ops.while_loop(cond, body, ..., maximum_iterations)  (log_sinkhorn.py:125-141)
__global__ void ot_ctrl_kernel(float* __restrict__ ctrl) {
  if (ctrl[OT_DONE] != 0.f) return;
  const float err = __uint_as_float(*reinterpret_cast<unsigned int*>(ctrl + OT_ACC));
  ctrl[OT_ERR] = err;
  ctrl[OT_STEPS] += 1.0f;
  *reinterpret_cast<unsigned int*>(ctrl + OT_ACC) = 0u;
  if (!(err > ctrl[OT_ATOL]) || ctrl[OT_STEPS] >= ctrl[OT_MAX_STEPS]) ctrl[OT_DONE] = 1.0f;
}

// assignments[i] ~ Categorical(softmax_j(S[i,j] + u[i] + v[j]))  (log_sinkhorn.py:33-35) by the Gumbel-max trick:
// argmax_j S[i,j] + v[j] - log(-log U_ij), U_ij = word (i*m+j) of the Philox stream keyed by seed; ties -> smallest j.
// u == nullptr: logits S_ij + v_j (+ a per-row constant).  u != nullptr (plain Sinkhorn, sinkhorn.py:60-62): logits
// log(plan_ij + 1e-10) with plan_ij = exp(u_i + S_ij - shift + v_j).
__global__ void __launch_bounds__(256) ot_sample_kernel(const float* __restrict__ S, long long ld, const float* __restrict__ v,
                                                        long long n, long long m, unsigned long long seed,
                                                        long long* __restrict__ out, const float* __restrict__ u = nullptr,
                                                        const float* __restrict__ ctrl = nullptr) {
  const int lane = threadIdx.x & 31;
  const long long i = (long long)blockIdx.x * 8 + (threadIdx.x >> 5);
  if (i >= n) return;
  float best = -INFINITY;
  long long arg = m;
  for (long long j = lane; j < m; j += 32) {
    const unsigned long long e = (unsigned long long)(i * m + j);
    uint32_t c[4] = {(uint32_t)(e >> 2), (uint32_t)(e >> 34), 0u, 0u};
    philox4x32_10(c, (uint32_t)seed, (uint32_t)(seed >> 32));
    const float g = -logf(-logf(u01(c[e & 3])));
    float val = S[i * ld + j] + v[j];
    if (u) val = logf(__expf(val + u[i] - ctrl[OT_SHIFT]) + 1e-10f);
    val += g;
    if (val > best) { best = val; arg = j; }
  }
  for (int o = 16; o > 0; o >>= 1) {
    const float b2 = __shfl_xor_sync(0xffffffffu, best, o);
    const long long a2 = __shfl_xor_sync(0xffffffffu, arg, o);
    if (b2 > best || (b2 == best && a2 < arg)) { best = b2; arg = a2; }
  }
  if (lane == 0) out[i] = arg < m ? arg : 0;
}

static inline int ot_chunks(long long n) {
  long long c = (n + 31) / 32;
  return (int)(c < 1 ? 1 : (c > 64 ? 64 : c));
}

}  // namespace bfk

// ---- C ABI ------------------------------------------------------------------------------------------------------------
extern "C" int bfk_ot_cost(const float* x1, const float* x2, int64_t n, int64_t m, int32_t D, int32_t cosine, float* M,
                           void* stream) {
  if (n < 1 || m < 1 || D < 1) return fail(-80, "bfk_ot_cost: bad n/m/D");
  if (!x1 || !x2 || !M) return fail(-81, "bfk_ot_cost: null pointer");
  dim3 grid((unsigned)((m + 31) / 32), (unsigned)((n + 7) / 8)), block(32, 8);
  if (grid.y > 65535) return fail(-82, "bfk_ot_cost: n too large (<= 524280)");
  if (cosine) bfk::ot_cost_kernel<true><<<grid, block, 0, (cudaStream_t)stream>>>(x1, x2, n, m, D, M);
  else bfk::ot_cost_kernel<false><<<grid, block, 0, (cudaStream_t)stream>>>(x1, x2, n, m, D, M);
  CK(cudaGetLastError());
  return 0;
}

extern "C" int bfk_ot_ratio(const float* M, const float* Cm, int64_t n, float w, unsigned long long* count, void* stream) {
  if (n < 1) return fail(-83, "bfk_ot_ratio: bad n");
  if (!M || !Cm || !count) return fail(-84, "bfk_ot_ratio: null pointer");
  cudaStream_t st = (cudaStream_t)stream;
  CK(cudaMemsetAsync(count, 0, sizeof(unsigned long long), st));
  const int sms = bfk_num_sms();
  long long blocks = (n * n + 255) / 256, cap = (long long)(sms > 0 ? sms : 148) * 8;
  if (blocks > cap) blocks = cap;
  bfk::ot_ratio_kernel<<<(unsigned)blocks, 256, 0, st>>>(M, Cm, n, w, count);
  CK(cudaGetLastError());
  return 0;
}

extern "C" int bfk_ot_scale(float* M, const float* Cm, float w, float regularization, int64_t numel, void* stream) {
  if (numel < 1 || !(regularization > 0.f)) return fail(-85, "bfk_ot_scale: bad numel / regularization must be positive");
  if (!M) return fail(-86, "bfk_ot_scale: null pointer");
  const int sms = bfk_num_sms();
  long long blocks = (numel + 255) / 256, cap = (long long)(sms > 0 ? sms : 148) * 8;
  if (blocks > cap) blocks = cap;
  bfk::ot_scale_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(M, Cm, w, regularization, numel);
  CK(cudaGetLastError());
  return 0;
}

extern "C" size_t bfk_ot_work_floats(int64_t n, int64_t m) {
  if (n < 1 || m < 1) return 0;
  return (size_t)bfk::ot_chunks(n) * 2 * (size_t)m;
}

extern "C" int bfk_ot_init(float* ctrl, float* u, float* v, int64_t n, int64_t m, float atol, int64_t max_steps,
                           void* stream) {
  if (!ctrl || !u || !v || n < 1 || m < 1) return fail(-87, "bfk_ot_init: bad arguments");
  cudaStream_t st = (cudaStream_t)stream;
  float h[bfk::OT_CTRL_FLOATS] = {1e30f, 0.f, max_steps > 0 ? 0.f : 1.f, atol, (float)max_steps, 0.f, 0.f, 0.f};
  CK(cudaMemcpyAsync(ctrl, h, sizeof(h), cudaMemcpyHostToDevice, st));
  CK(cudaStreamSynchronize(st));  // h is a stack buffer
  CK(cudaMemsetAsync(u, 0, (size_t)n * 4, st));
  CK(cudaMemsetAsync(v, 0, (size_t)m * 4, st));
  return 0;
}

extern "C" int bfk_ot_log_sinkhorn_step(const float* S, int64_t n, int64_t m, const float* log_a_vec, const float* log_b_vec,
                                        float* u, float* v, float* ctrl, float* work, int32_t n_steps, void* stream) {
  if (n < 1 || m < 1) return fail(-88, "bfk_ot_log_sinkhorn_step: bad n/m");
  if (!S || !u || !v || !ctrl || !work) return fail(-89, "bfk_ot_log_sinkhorn_step: null pointer");
  cudaStream_t st = (cudaStream_t)stream;
  const float log_a = -logf((float)n), log_b = -logf((float)m);
  const int chunks = bfk::ot_chunks(n);
  for (int it = 0; it < n_steps; ++it) {  // every kernel is a no-op once ctrl[done] is set
    bfk::ot_row_lse_kernel<<<(unsigned)((n + 7) / 8), 256, 0, st>>>(S, n, m, log_a, log_a_vec, u, v, ctrl);
    dim3 g2((unsigned)((m + 255) / 256), (unsigned)chunks);
    bfk::ot_col_partial_kernel<<<g2, 256, 0, st>>>(S, n, m, u, work, ctrl);
    bfk::ot_col_finish_kernel<<<(unsigned)((m + 255) / 256), 256, 0, st>>>(work, chunks, m, log_b, log_b_vec, v, ctrl);
    bfk::ot_ctrl_kernel<<<1, 1, 0, st>>>(ctrl);
  }
  CK(cudaGetLastError());
  return 0;
}

extern "C" int bfk_ot_plain_mode(const float* S, int64_t numel, float* ctrl, void* stream) {
  if (!S || !ctrl || numel < 1) return fail(-92, "bfk_ot_plain_mode: bad arguments");
  cudaStream_t st = (cudaStream_t)stream;
  const float init[2] = {1.0f, -INFINITY};  // ctrl[OT_PLAIN] = 1, ctrl[OT_SHIFT] = -inf (max identity)
  CK(cudaMemcpyAsync(ctrl + bfk::OT_PLAIN, init, sizeof(init), cudaMemcpyHostToDevice, st));
  CK(cudaStreamSynchronize(st));
  const int sms = bfk_num_sms();
  long long blocks = (numel + 255) / 256, cap = (long long)(sms > 0 ? sms : 148) * 8;
  if (blocks > cap) blocks = cap;
  bfk::ot_max_kernel<<<(unsigned)blocks, 256, 0, st>>>(S, numel, ctrl);
  CK(cudaGetLastError());
  return 0;
}

extern "C" int bfk_ot_sample_plan(const float* S, int64_t ld, const float* u, const float* v, const float* ctrl, int64_t n,
                                  int64_t m, uint64_t seed, int64_t* assignments, void* stream) {
  if (n < 1 || m < 1 || ld < m) return fail(-90, "bfk_ot_sample_plan: bad n/m/ld");
  if (!S || !u || !v || !ctrl || !assignments) return fail(-91, "bfk_ot_sample_plan: null pointer");
  bfk::ot_sample_kernel<<<(unsigned)((n + 7) / 8), 256, 0, (cudaStream_t)stream>>>(S, ld, v, n, m, (unsigned long long)seed,
                                                                                  (long long*)assignments, u, ctrl);
  CK(cudaGetLastError());
  return 0;
}

extern "C" int bfk_ot_sample(const float* S, int64_t ld, const float* v, int64_t n, int64_t m, uint64_t seed,
                             int64_t* assignments, void* stream) {
  if (n < 1 || m < 1 || ld < m) return fail(-90, "bfk_ot_sample: bad n/m/ld");
  if (!S || !v || !assignments) return fail(-91, "bfk_ot_sample: null pointer");
  bfk::ot_sample_kernel<<<(unsigned)((n + 7) / 8), 256, 0, (cudaStream_t)stream>>>(S, ld, v, n, m, (unsigned long long)seed,
                                                                                  (long long*)assignments);
  CK(cudaGetLastError());
  return 0;
}
