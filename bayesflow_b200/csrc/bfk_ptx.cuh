// Thin inline-PTX layer for sm_100a: mbarrier, bulk async copy (TMA engine, UBLKCP), tcgen05 (TMEM alloc, UMMA,
// commit, ld/st), proxy fences.  No CUTLASS dependency.  Descriptor bit layouts follow the PTX ISA "tcgen05
// shared-memory matrix descriptor" / "instruction descriptor" tables.
#pragma once
#include <cuda_fp16.h>
#include <cuda_bf16.h>
#include <stdint.h>

namespace bfk {

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

// ------------------------------------------------------------------ mbarrier
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
  asm volatile("{\n\t.reg .b64 st;\n\tmbarrier.arrive.shared::cta.b64 st, [%0];\n\t}" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("{\n\t.reg .b64 st;\n\tmbarrier.arrive.expect_tx.shared::cta.b64 st, [%0], %1;\n\t}" ::"r"(bar), "r"(bytes)
               : "memory");
}
// The waiting warps (epilogue warps during their tile's MMA, issuer, producer) share their SM sub-partition's issue port
// with the working ones: without a suspend-time hint try_wait comes back every ~40 cycles and the spin loop was 46 % of
// all issued instructions (profiles/r02e); with it the kernel is 4.6 % faster (r02f; 500 ns ... 1 ms all the same).
#ifndef BFK_WAIT_HINT_NS
#define BFK_WAIT_HINT_NS 2000
#endif
__device__ __forceinline__ bool mbar_try_wait(uint32_t bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
#if BFK_WAIT_HINT_NS
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2, %3;\n\t"  // suspend-time hint: fewer spin iterations
#else
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
#endif
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(ok)
#if BFK_WAIT_HINT_NS
      : "r"(bar), "r"(parity), "r"((uint32_t)BFK_WAIT_HINT_NS)
#else
      : "r"(bar), "r"(parity)
#endif
      : "memory");
  return ok != 0;
}
// Bounded wait: a protocol bug must not hang the GPU.  On timeout: record the site and trap.
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity, int* err_flag, int site) {
  uint32_t spins = 0;
  while (!mbar_try_wait(bar, parity)) {
    if (++spins > (1u << 24)) {
      if (err_flag) atomicExch(err_flag, 1000 + site);
      __threadfence_system();
      __trap();
    }
  }
}

__device__ __forceinline__ void fence_barrier_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
// generic-proxy writes (st.shared) -> visible to the async proxy (UMMA / bulk copy reads)
__device__ __forceinline__ void fence_proxy_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

// ------------------------------------------------------------------ bulk async copy global -> shared (TMA engine)
__device__ __forceinline__ void bulk_g2s(uint32_t dst_smem, const void* src, uint32_t bytes, uint32_t bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst_smem),
               "l"(src), "r"(bytes), "r"(bar)
               : "memory");
}

// the same copy delivered to the same CTA-relative offset in every CTA of `cta_mask` (each one's mbarrier at `bar`'s
// offset receives the complete_tx)
__device__ __forceinline__ void bulk_g2s_multicast(uint32_t dst_smem, const void* src, uint32_t bytes, uint32_t bar, uint16_t cta_mask) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.multicast::cluster [%0], [%1], %2, [%3], %4;" ::"r"(dst_smem),
               "l"(src), "r"(bytes), "r"(bar), "h"(cta_mask)
               : "memory");
}
__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}

// ------------------------------------------------------------------ tcgen05
__device__ __forceinline__ void tmem_alloc(uint32_t dst_smem, uint32_t ncols) {  // one full warp
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(dst_smem), "r"(ncols) : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t ncols) {  // same warp that allocated
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

// D[tmem] (+)= A[smem] * B[smem]; kind::f16 covers fp16 and bf16 inputs, fp32 accumulate.  One thread issues.
__device__ __forceinline__ void umma_f16(uint32_t d_tmem, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(d_tmem),
      "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// all previously issued UMMAs of this thread -> arrive(1) on the mbarrier when they complete
__device__ __forceinline__ void umma_commit(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}

// the same arrive delivered to the mbarrier at `bar`'s offset in every CTA of `cta_mask`
__device__ __forceinline__ void umma_commit_multicast(uint32_t bar, uint16_t cta_mask) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(bar), "h"(cta_mask)
               : "memory");
}

// K-major, no-swizzle ("interleave") canonical layout: 8x(16B) core matrices, rows 16 B apart inside a core matrix,
//   SBO = byte distance between 8-row groups (M/N direction), LBO = byte distance between 16-byte K chunks.
__device__ __forceinline__ uint64_t umma_smem_desc(uint32_t smem_addr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
  uint64_t d = 0;
  d |= (uint64_t)((smem_addr >> 4) & 0x3FFF);
  d |= (uint64_t)((lbo_bytes >> 4) & 0x3FFF) << 16;
  d |= (uint64_t)((sbo_bytes >> 4) & 0x3FFF) << 32;
  d |= (uint64_t)1 << 46;  // descriptor version (Blackwell)
  return d;                // base_offset 0, lbo_mode 0, layout_type 0 (SWIZZLE_NONE)
}
// instruction descriptor, kind::f16: c=f32, a/b = f16 (0) or bf16 (1), both K-major, dense
__host__ __device__ constexpr uint32_t umma_idesc_f16(int M, int N, int ab_format) {
  return (1u << 4) | ((uint32_t)ab_format << 7) | ((uint32_t)ab_format << 10) | ((uint32_t)(N >> 3) << 17) |
         ((uint32_t)(M >> 4) << 24);
}

// TMEM <-> registers, 32 lanes x 32 bit, 32 consecutive columns: thread i of the warp owns lane (32*(warp%4)+i)
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, float* v) {
  uint32_t* r = reinterpret_cast<uint32_t*>(v);
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]),
        "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]),
        "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
      : "r"(taddr)
      : "memory");
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
// wait::ld that also carries the data dependency of the 32 loaded registers, so that no use of them can be scheduled
// above the wait when loads are software-pipelined
__device__ __forceinline__ void tmem_ld_wait32(float* v) {
  uint32_t* r = reinterpret_cast<uint32_t*>(v);
  asm volatile("tcgen05.wait::ld.sync.aligned;"
               : "+r"(r[0]), "+r"(r[1]), "+r"(r[2]), "+r"(r[3]), "+r"(r[4]), "+r"(r[5]), "+r"(r[6]), "+r"(r[7]), "+r"(r[8]),
                 "+r"(r[9]), "+r"(r[10]), "+r"(r[11]), "+r"(r[12]), "+r"(r[13]), "+r"(r[14]), "+r"(r[15]), "+r"(r[16]),
                 "+r"(r[17]), "+r"(r[18]), "+r"(r[19]), "+r"(r[20]), "+r"(r[21]), "+r"(r[22]), "+r"(r[23]), "+r"(r[24]),
                 "+r"(r[25]), "+r"(r[26]), "+r"(r[27]), "+r"(r[28]), "+r"(r[29]), "+r"(r[30]), "+r"(r[31])
               :
               : "memory");
}
__device__ __forceinline__ void tmem_st32(uint32_t taddr, const float* v) {
  const uint32_t* r = reinterpret_cast<const uint32_t*>(v);
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], "
      "{%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31,%32};" ::"r"(taddr),
      "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]), "r"(r[9]),
      "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15]), "r"(r[16]), "r"(r[17]), "r"(r[18]),
      "r"(r[19]), "r"(r[20]), "r"(r[21]), "r"(r[22]), "r"(r[23]), "r"(r[24]), "r"(r[25]), "r"(r[26]), "r"(r[27]),
      "r"(r[28]), "r"(r[29]), "r"(r[30]), "r"(r[31])
      : "memory");
}
__device__ __forceinline__ void tmem_st_wait() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }

// 16-column variants (the 18-warp fast kernel keeps two 16-column buffers in flight per thread)
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, float* v) {
  uint32_t* r = reinterpret_cast<uint32_t*>(v);
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
      : "r"(taddr)
      : "memory");
}
__device__ __forceinline__ void tmem_ld_wait16(float* v) {
  uint32_t* r = reinterpret_cast<uint32_t*>(v);
  asm volatile("tcgen05.wait::ld.sync.aligned;"
               : "+r"(r[0]), "+r"(r[1]), "+r"(r[2]), "+r"(r[3]), "+r"(r[4]), "+r"(r[5]), "+r"(r[6]), "+r"(r[7]), "+r"(r[8]),
                 "+r"(r[9]), "+r"(r[10]), "+r"(r[11]), "+r"(r[12]), "+r"(r[13]), "+r"(r[14]), "+r"(r[15])
               :
               : "memory");
}
__device__ __forceinline__ void tmem_st16(uint32_t taddr, const float* v) {
  const uint32_t* r = reinterpret_cast<const uint32_t*>(v);
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16};" ::"r"(taddr),
      "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]), "r"(r[9]),
      "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15])
      : "memory");
}
// named barrier among `nthreads` threads (a subset of warps of the CTA)
__device__ __forceinline__ void named_bar_sync(int id, int nthreads) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}

// ------------------------------------------------------------------ math
// BFK_EXP: compile-time timing experiments (tools/gpu_chase_exp.sh); results are garbage, only the time matters.
#ifndef BFK_EXP
#define BFK_EXP 0
#endif
__device__ __forceinline__ float ex2_approx(float x) {
#if (BFK_EXP & 1)
  return fmaf(x, 0.5f, 1.f);  // experiment: no MUFU
#else
  float y;
#ifdef BFK_MUFU_VOLATILE  // keeps the source order of the MUFU ops (the compiler otherwise serialises pair by pair)
  asm volatile("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
#else
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
#endif
  return y;
#endif
}
__device__ __forceinline__ float rcp_approx(float x) {
#if (BFK_EXP & 1)
  return fmaf(x, 0.25f, 0.5f);
#else
  float y;
#ifdef BFK_MUFU_VOLATILE
  asm volatile("rcp.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
#else
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
#endif
  return y;
#endif
}

// ---- MUFU relief: the epilogues are bound by the 16-lane XU pipe (ex2 + rcp per element), so a compile-time share of
// the element PAIRS takes the FMA pipe instead (the FlashAttention-4 remedy).  Masks are over the 8 float2 of a 16-column
// chunk; 0 = MUFU.
#ifndef BFK_RCP_NEWTON_MASK
#define BFK_RCP_NEWTON_MASK 0
#endif
#ifndef BFK_EX2_POLY_MASK
#define BFK_EX2_POLY_MASK 0
#endif
#ifndef BFK_EX2_POLY_DEG
#define BFK_EX2_POLY_DEG 4
#endif
// 1 / d for d in [1, 2^100]: exponent-flip seed (5 % off) and two Newton steps on packed FMAs -> 7e-6 relative
__device__ __forceinline__ float2 rcp_newton2(float2 d) {
  float2 y = make_float2(__int_as_float(0x7EF311C7 - __float_as_int(d.x)), __int_as_float(0x7EF311C7 - __float_as_int(d.y)));
  const float2 nd = make_float2(-d.x, -d.y);  // folded into the FFMA2 operand negation
  float2 r = __ffma2_rn(nd, y, make_float2(1.f, 1.f));
  y = __ffma2_rn(y, r, y);
  r = __ffma2_rn(nd, y, make_float2(1.f, 1.f));
  return __ffma2_rn(y, r, y);
}
// 2^t for t in [-120, 120]: round-to-nearest split by the 1.5 * 2^23 trick, minimax polynomial on [-0.5, 0.5] (degree 4:
// 4e-6 relative, degree 3: 1e-4), exponent added to the bit pattern ((bits(z) << 23) keeps only the integer part)
__device__ __forceinline__ float2 ex2_poly2(float2 t) {
  const float2 magic = make_float2(12582912.f, 12582912.f);
  const float2 z = __fadd2_rn(t, magic);
  const float2 fr = __fadd2_rn(t, __fadd2_rn(magic, make_float2(-z.x, -z.y)));  // t - (z - magic)
#if BFK_EX2_POLY_DEG == 3
  float2 p = __ffma2_rn(fr, make_float2(0.05592203512787819f, 0.05592203512787819f), make_float2(0.24264007806777954f, 0.24264007806777954f));
  p = __ffma2_rn(p, fr, make_float2(0.6931210160255432f, 0.6931210160255432f));
  p = __ffma2_rn(p, fr, make_float2(0.9999244809150696f, 0.9999244809150696f));
#else
  float2 p = __ffma2_rn(fr, make_float2(0.009676037356257439f, 0.009676037356257439f), make_float2(0.05592203512787819f, 0.05592203512787819f));
  p = __ffma2_rn(p, fr, make_float2(0.2402210682630539f, 0.2402210682630539f));
  p = __ffma2_rn(p, fr, make_float2(0.6931210160255432f, 0.6931210160255432f));
  p = __ffma2_rn(p, fr, make_float2(1.0000001192092896f, 1.0000001192092896f));
#endif
  return make_float2(__int_as_float(__float_as_int(p.x) + (__float_as_int(z.x) << 23)),
                     __int_as_float(__float_as_int(p.y) + (__float_as_int(z.y) << 23)));
}

}  // namespace bfk
