// Packed-f32x2 epilogue of the fused sampler for the reference's default block shape (LayerNorm, residual, FiLM shift
// only — networks/defaults.py:13-26): half the issued instructions per element of the generic path.
//   pass 1:  v = skip + act(acc + bd) + beta_l(t)          conditional_dense_block.py:117-127
//            row sum / sum of squares                        (keras LayerNormalization, eps 1e-3)
//   pass 2:  h = (v - mean) * rstd * gamma + beta  -> fp16/bf16 A tile of the next layer     :129-130
//   last block: no pass 2 — the output projector is folded through the LayerNorm affine:
//            pred_d = rstd * (sum_c v_c * g_c Wo[c,d] - mean * sum_c g_c Wo[c,d]) + sum_c b_c Wo[c,d] + bo_d
// FFMA2 / FADD2 / FMUL2 (sm_100 packed fp32) process two columns per instruction; TMEM loads are software-pipelined
// in 16-column chunks.  Used by bfk_sampler_fast.cuh (18 warps: 16 epilogue warps, each owning half the columns of 32 rows).
#pragma once
#include "bfk_prologue.cuh"

namespace bfk {

__device__ __forceinline__ float2 f2(float a) { return make_float2(a, a); }

template <int ACT>
__device__ __forceinline__ float2 act2(float2 x) {
  if (ACT == BFK_ACT_MISH) {
    // mish(x) = x * tanh(softplus(x)) = x * (1 - 2/((1+e^x)^2 + 1)); overflow-safe without a clamp: e=inf -> r=0 -> x
    const float2 t = __fmul2_rn(x, f2(1.4426950408889634f));
    float2 e;
    e.x = ex2_approx(t.x); e.y = ex2_approx(t.y);
    const float2 q = __fadd2_rn(e, f2(1.f));
    const float2 n = __ffma2_rn(q, q, f2(1.f));
    float2 r;
    r.x = rcp_approx(n.x); r.y = rcp_approx(n.y);
    const float2 w = __ffma2_rn(r, f2(-2.f), f2(1.f));
    return __fmul2_rn(x, w);
  } else if (ACT == BFK_ACT_SILU) {
    const float2 t = __fmul2_rn(x, f2(-1.4426950408889634f));
    float2 e;
    e.x = ex2_approx(t.x); e.y = ex2_approx(t.y);
    const float2 n = __fadd2_rn(e, f2(1.f));
    float2 r;
    r.x = rcp_approx(n.x); r.y = rcp_approx(n.y);
    return __fmul2_rn(x, r);
  } else {
    return make_float2(act_fn<ACT>(x.x), act_fn<ACT>(x.y));
  }
}

template <bool BF16>
__device__ __forceinline__ float2 unpack2f(uint32_t u) {
  if (BF16) {
    return make_float2(__uint_as_float(u << 16), __uint_as_float(u & 0xffff0000u));
  } else {
    return __half22float2(*reinterpret_cast<__half2*>(&u));
  }
}
template <bool BF16>
__device__ __forceinline__ uint32_t pack2f(float2 v) {
  if (BF16) {
    __nv_bfloat162 p = __floats2bfloat162_rn(v.x, v.y);
    return *reinterpret_cast<uint32_t*>(&p);
  } else {
    __half2 p = __floats2half2_rn(v.x, v.y);
    return *reinterpret_cast<uint32_t*>(&p);
  }
}

// Staged activation over 8 column pairs: every MUFU of the chunk is issued back to back (16 ex2, then 16 rcp) so that
// the quarter-rate XU pipe never waits on a dependent FMA and its latency hides behind its own throughput.
template <int ACT>
__device__ __forceinline__ void act2x8(float2* x) {
  if (ACT == BFK_ACT_MISH) {
    float2 e[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) e[i] = __fmul2_rn(x[i], f2(1.4426950408889634f));
#pragma unroll
    for (int i = 0; i < 8; ++i) { e[i].x = ex2_approx(e[i].x); e[i].y = ex2_approx(e[i].y); }
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const float2 q = __fadd2_rn(e[i], f2(1.f));
      e[i] = __ffma2_rn(q, q, f2(1.f));
    }
#pragma unroll
    for (int i = 0; i < 8; ++i) { e[i].x = rcp_approx(e[i].x); e[i].y = rcp_approx(e[i].y); }
#pragma unroll
    for (int i = 0; i < 8; ++i) x[i] = __fmul2_rn(x[i], __ffma2_rn(e[i], f2(-2.f), f2(1.f)));
  } else if (ACT == BFK_ACT_SILU) {
    float2 e[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) e[i] = __fmul2_rn(x[i], f2(-1.4426950408889634f));
#pragma unroll
    for (int i = 0; i < 8; ++i) { e[i].x = ex2_approx(e[i].x); e[i].y = ex2_approx(e[i].y); }
#pragma unroll
    for (int i = 0; i < 8; ++i) e[i] = __fadd2_rn(e[i], f2(1.f));
#pragma unroll
    for (int i = 0; i < 8; ++i) { e[i].x = rcp_approx(e[i].x); e[i].y = rcp_approx(e[i].y); }
#pragma unroll
    for (int i = 0; i < 8; ++i) x[i] = __fmul2_rn(x[i], e[i]);
  } else {
#pragma unroll
    for (int i = 0; i < 8; ++i) x[i] = make_float2(act_fn<ACT>(x[i].x), act_fn<ACT>(x[i].y));
  }
}

// 16 columns starting at column `col`: v (8 float2, TMEM accumulators) -> v = skip + act(v + bd) + fbeta; stats
template <int ACT, bool BF16>
__device__ __forceinline__ void fast_pass1_chunk(float2* v, int col, const float* __restrict__ bd,
                                                 const float* __restrict__ fbeta, const unsigned char* sA_row,
                                                 float2* sum2, float2* sq2) {
#pragma unroll
  for (int jj = 0; jj < 2; ++jj) {
    const float4 b0 = *reinterpret_cast<const float4*>(bd + col + jj * 8);
    const float4 b1 = *reinterpret_cast<const float4*>(bd + col + jj * 8 + 4);
    v[jj * 4 + 0] = __fadd2_rn(v[jj * 4 + 0], make_float2(b0.x, b0.y));
    v[jj * 4 + 1] = __fadd2_rn(v[jj * 4 + 1], make_float2(b0.z, b0.w));
    v[jj * 4 + 2] = __fadd2_rn(v[jj * 4 + 2], make_float2(b1.x, b1.y));
    v[jj * 4 + 3] = __fadd2_rn(v[jj * 4 + 3], make_float2(b1.z, b1.w));
  }
  act2x8<ACT>(v);
#pragma unroll
  for (int jj = 0; jj < 2; ++jj) {  // 8 columns = one 16-byte K chunk of the A tile
    const uint4 sk = *reinterpret_cast<const uint4*>(sA_row + (size_t)((col >> 3) + jj) * (TILE_M * 16));
    const uint32_t su[4] = {sk.x, sk.y, sk.z, sk.w};
    const float4 f0 = *reinterpret_cast<const float4*>(fbeta + col + jj * 8);
    const float4 f1 = *reinterpret_cast<const float4*>(fbeta + col + jj * 8 + 4);
    const float2 ff[4] = {make_float2(f0.x, f0.y), make_float2(f0.z, f0.w), make_float2(f1.x, f1.y), make_float2(f1.z, f1.w)};
#pragma unroll
    for (int p = 0; p < 4; ++p) {
      const int i = jj * 4 + p;
      const float2 u = __fadd2_rn(__fadd2_rn(v[i], ff[p]), unpack2f<BF16>(su[p]));
      sum2[p] = __fadd2_rn(sum2[p], u);
      sq2[p] = __ffma2_rn(u, u, sq2[p]);
      v[i] = u;
    }
  }
}

// 16 columns: h = v * rstd + nmr (= (v-mean)*rstd), * gamma + beta -> packed 16-bit A tile
template <bool BF16>
__device__ __forceinline__ void fast_pass2_chunk(const float2* v, int col, float rstd, float nmr,
                                                 const float* __restrict__ lng, const float* __restrict__ lnb,
                                                 unsigned char* sA_row) {
  const float2 rs = f2(rstd), nm = f2(nmr);
#pragma unroll
  for (int jj = 0; jj < 2; ++jj) {
    const float4 g0 = *reinterpret_cast<const float4*>(lng + col + jj * 8);
    const float4 g1 = *reinterpret_cast<const float4*>(lng + col + jj * 8 + 4);
    const float4 l0 = *reinterpret_cast<const float4*>(lnb + col + jj * 8);
    const float4 l1 = *reinterpret_cast<const float4*>(lnb + col + jj * 8 + 4);
    const float2 gg[4] = {make_float2(g0.x, g0.y), make_float2(g0.z, g0.w), make_float2(g1.x, g1.y), make_float2(g1.z, g1.w)};
    const float2 ll[4] = {make_float2(l0.x, l0.y), make_float2(l0.z, l0.w), make_float2(l1.x, l1.y), make_float2(l1.z, l1.w)};
    uint32_t pk[4];
#pragma unroll
    for (int p = 0; p < 4; ++p) {
      const float2 xh = __ffma2_rn(v[jj * 4 + p], rs, nm);
      pk[p] = pack2f<BF16>(__ffma2_rn(xh, gg[p], ll[p]));
    }
    *reinterpret_cast<uint4*>(sA_row + (size_t)((col >> 3) + jj) * (TILE_M * 16)) = make_uint4(pk[0], pk[1], pk[2], pk[3]);
  }
}

}  // namespace bfk
