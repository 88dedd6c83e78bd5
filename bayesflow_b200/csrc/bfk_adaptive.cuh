// Device-side step-size controller of integrate_adaptive (bayesflow/utils/integrate.py:291-406).  One thread; every
// operation is the fp32 operation torch performs in the reference (0-dim fp32 tensors; python doubles are cast to fp32
// when they meet a tensor), written with explicit _rn intrinsics so that nothing is contracted into an FMA.
#pragma once
#include <cuda_runtime.h>

#include "bfk_layout.h"

namespace bfk {

struct StageC { float c[8]; };  // abscissae of the 6 evaluations of a trial: stages 2..6 and the FSAL evaluation at t+h

__device__ __forceinline__ float clipf(float x, float lo, float hi) { return fminf(fmaxf(x, lo), hi); }  // torch.clip
__device__ __forceinline__ float signf(float x) { return x > 0.f ? 1.f : (x < 0.f ? -1.f : (x == 0.f ? 0.f : x)); }

__device__ inline void plan_times(float* c, const StageC& sc) {
  const float t = c[BFK_CTRL_TIME], h = c[BFK_CTRL_H];
  for (int i = 0; i < 6; ++i) c[BFK_CTRL_TIMES + i] = (sc.c[i] == 1.f) ? __fadd_rn(t, h) : __fadd_rn(t, __fmul_rn(h, sc.c[i]));
  c[BFK_CTRL_ACC + 0] = 0.f; c[BFK_CTRL_ACC + 1] = 0.f; c[BFK_CTRL_ACC + 2] = 0.f; c[BFK_CTRL_ACC + 3] = 0.f;
}

__global__ void adaptive_init_kernel(float* c, float start, float stop, float min_steps, float max_steps, float atol,
                                     float rtol, float step_size) {
  if (threadIdx.x != 0 || blockIdx.x != 0) return;
  for (int i = 0; i < BFK_CTRL_FLOATS; ++i) c[i] = 0.f;
  c[BFK_CTRL_TIME] = start; c[BFK_CTRL_START] = start; c[BFK_CTRL_STOP] = stop;
  c[BFK_CTRL_STEP_SIZE] = step_size;
  c[BFK_CTRL_ATOL] = atol; c[BFK_CTRL_RTOL] = rtol;
  c[BFK_CTRL_MIN_STEPS] = min_steps; c[BFK_CTRL_MAX_STEPS] = max_steps;
  c[BFK_CTRL_SKIP] = 1.f;
}

__global__ void adaptive_plan_kernel(float* c, StageC sc) {
  if (threadIdx.x != 0 || blockIdx.x != 0) return;
  if (c[BFK_CTRL_DONE] != 0.f) return;
  const float stop = c[BFK_CTRL_STOP], minS = c[BFK_CTRL_MIN_STEPS], maxS = c[BFK_CTRL_MAX_STEPS];
  const float sign_dir = signf(__fsub_rn(stop, c[BFK_CTRL_START]));
  if (c[BFK_CTRL_HAVE_TRIAL] != 0.f) {
    // ---- digest the trial (integrate.py:353-378)
    const float h = c[BFK_CTRL_H], min_step = c[BFK_CTRL_MIN_STEP_SIZE];
    const float tr = fabsf(__fsub_rn(stop, c[BFK_CTRL_TIME]));
    const float max_step = __fdiv_rn(tr, fmaxf(__fsub_rn(minS, c[BFK_CTRL_STEP]), 1.f));
    float err = c[BFK_CTRL_ACC + 0], max_abs = c[BFK_CTRL_ACC + 1];
    if (c[BFK_CTRL_ACC + 2] != 0.f) { err = __int_as_float(0x7fc00000); }  // a NaN row makes torch.max NaN
    const float scale = __fadd_rn(c[BFK_CTRL_ATOL], __fmul_rn(c[BFK_CTRL_RTOL], max_abs));
    const float ratio = __fdiv_rn(err, scale);
    float fac = __fmul_rn(0.9f, powf(__fdiv_rn(1.0f, __fadd_rn(ratio, 1e-12f)), 0.2f));
    fac = (fac != fac) ? fac : clipf(fac, 0.2f, 5.0f);
    float ns = __fmul_rn(h, fac);
    ns = __fmul_rn(signf(ns), (ns != ns) ? ns : clipf(fabsf(ns), min_step, max_step));
    const bool too_big = ratio > 1.0f;
    const bool at_min = fabsf(h) <= fabsf(min_step);
    if (!too_big || at_min) {
      c[BFK_CTRL_SEL] = 1.f - c[BFK_CTRL_SEL];
      c[BFK_CTRL_TIME] = __fadd_rn(c[BFK_CTRL_TIME], h);
      c[BFK_CTRL_STEP] += 1.f;
      c[BFK_CTRL_ALL_NAN] = (c[BFK_CTRL_ACC + 3] == 0.f) ? 1.f : 0.f;
    } else {
      c[BFK_CTRL_REJECTED] += 1.f;
    }
    c[BFK_CTRL_STEP_SIZE] = ns;
    c[BFK_CTRL_HAVE_TRIAL] = 0.f;
  }
  // ---- loop condition (integrate.py:325-335)
  const float time = c[BFK_CTRL_TIME], step = c[BFK_CTRL_STEP], ss = c[BFK_CTRL_STEP_SIZE];
  const float rem_after = __fmul_rn(sign_dir, __fsub_rn(stop, __fadd_rn(time, ss)));
  const bool go = (c[BFK_CTRL_ALL_NAN] == 0.f) && ((step < minS) || ((rem_after > 0.f) && (step < maxS)));
  if (!go) { c[BFK_CTRL_DONE] = 1.f; c[BFK_CTRL_SKIP] = 1.f; return; }
  // ---- plan the next trial (integrate.py:339-342)
  const float tr = fabsf(__fsub_rn(stop, time));
  const float min_step = __fdiv_rn(tr, __fsub_rn(maxS, step));
  const float max_step = __fdiv_rn(tr, fmaxf(__fsub_rn(minS, step), 1.f));
  c[BFK_CTRL_MIN_STEP_SIZE] = min_step;
  c[BFK_CTRL_H] = __fmul_rn(signf(ss), (ss != ss) ? ss : clipf(fabsf(ss), min_step, max_step));
  c[BFK_CTRL_HAVE_TRIAL] = 1.f;
  c[BFK_CTRL_SKIP] = 0.f;
  c[BFK_CTRL_TRIALS] += 1.f;
  plan_times(c, sc);
}

// integrate.py:393-403: if time is still short of stop_time, one more step of exactly the remaining length
__global__ void adaptive_plan_final_kernel(float* c, StageC sc) {
  if (threadIdx.x != 0 || blockIdx.x != 0) return;
  c[BFK_CTRL_SKIP] = 1.f;
  c[BFK_CTRL_FINAL] = 0.f;
  if (c[BFK_CTRL_ALL_NAN] != 0.f) return;
  const float stop = c[BFK_CTRL_STOP];
  const float sign_dir = signf(__fsub_rn(stop, c[BFK_CTRL_START]));
  const float diff = __fsub_rn(stop, c[BFK_CTRL_TIME]);
  if (__fmul_rn(sign_dir, diff) > 0.f) {
    c[BFK_CTRL_H] = diff;
    c[BFK_CTRL_SKIP] = 0.f;
    c[BFK_CTRL_FINAL] = 1.f;
    c[BFK_CTRL_TRIALS] += 1.f;
    plan_times(c, sc);
  }
}
__global__ void adaptive_commit_kernel(float* c) {
  if (threadIdx.x != 0 || blockIdx.x != 0) return;
  if (c[BFK_CTRL_FINAL] != 0.f) {
    c[BFK_CTRL_SEL] = 1.f - c[BFK_CTRL_SEL];
    c[BFK_CTRL_TIME] = __fadd_rn(c[BFK_CTRL_TIME], c[BFK_CTRL_H]);
    c[BFK_CTRL_STEP] += 1.f;
    c[BFK_CTRL_FINAL] = 0.f;
  }
  c[BFK_CTRL_SKIP] = 1.f;
}

}  // namespace bfk
