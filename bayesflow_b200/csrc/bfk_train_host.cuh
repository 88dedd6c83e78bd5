// Host orchestration of one training step (forward + loss + backward) of the velocity network, and the optimizer.
// Included by bfk_host.cu (uses its fail()/CK helpers).  Reference: flow_matching.py:132-178 (compute_metrics),
// time_mlp.py:125-147, conditional_dense_block.py:114-131, torch_approximator.py:88-99 (autograd + optimizer step).
#pragma once
#include "bfk_train.cuh"

namespace {

struct Arena {
  unsigned char* base; size_t off, cap;
  void* take(size_t bytes) {
    off = (off + 255) / 256 * 256;
    void* p = base ? base + off : nullptr;
    off += bytes;
    return p;
  }
};

struct TrainBufs {
  int KU, KT, KTN, N1, NO, has_merge;
  long long Rp;
  float* fbb_all;  // [L * W]: the FiLM shift biases side by side (bias vector of the one batched FiLM GEMM)
  unsigned char *u_img, *te_img, *w1_pk, *a1_img, *wm_f, *wm_b, *h0_img, *wo_f, *wo_b, *dout_img, *dP_img, *dV_img, *dh0pre_img;
  float *w1, *b1, *h0pre, *m, *h0, *FB, *bo_pad, *out, *dH, *dte, *da1, *dW1, *db1;
  unsigned char *dP_img2, *dV_img2;  // second cotangent-image pair: lets the dW GEMMs of block l overlap block l-1's chain
  float* skip;                      // residual projector output of the current block (unequal widths), (B, W)
  unsigned char *wp_f, *wp_b;       // projector operand images (forward / dX)
  float* FG;                        // film_use_gamma: FiLM scale of the current block (recomputed in the backward)
  unsigned char *wg_f, *wg_b, *dG_img;  // film_use_gamma: to_gamma operand images (forward / dte), cotangent image of FG
  unsigned char *wd_f[BFK_MAX_BLOCKS], *wd_b[BFK_MAX_BLOCKS], *wb_f[BFK_MAX_BLOCKS], *wb_b[BFK_MAX_BLOCKS], *h_img[BFK_MAX_BLOCKS];
  float *P[BFK_MAX_BLOCKS], *xhat[BFK_MAX_BLOCKS], *rstd[BFK_MAX_BLOCKS];
};

inline int r32(int x) { return (x + 31) / 32 * 32; }
inline int r64(int x) { return (x + 63) / 64 * 64; }
inline int nt_of(int N) {
  static int cap = -1;  // BFK_NT_MAX=128 / 64: narrower column tiles (more, smaller output tiles per GEMM)
  if (cap < 0) { const char* e = getenv("BFK_NT_MAX"); cap = e ? atoi(e) : 256; }
  if (cap >= 256 && (N % 256) == 0) return 256;
  if (cap >= 128 && (N % 128) == 0) return 128;
  return 64;
}

void plan_train(const bfk_net_config* c, long long B, bool backward, Arena& ar, TrainBufs& b) {
  const int W = c->W;
  b.has_merge = (c->C > 0) ? 1 : 0;
  b.KU = r32(c->D + c->C);
  b.KT = r32(c->T);
  b.KTN = r64(c->T);
  b.N1 = (b.has_merge && c->merge == BFK_MERGE_CONCAT) ? 2 * W : W;
  b.NO = r64(c->D);
  b.Rp = (B + 127) / 128 * 128;
  const size_t Rp = (size_t)b.Rp, Bs = (size_t)B;
  b.u_img = (unsigned char*)ar.take(Rp * b.KU * 2);
  b.te_img = (unsigned char*)ar.take(Rp * b.KT * 2);
  b.w1 = (float*)ar.take((size_t)b.KU * b.N1 * 4);
  b.b1 = (float*)ar.take((size_t)b.N1 * 4);
  b.w1_pk = (unsigned char*)ar.take((size_t)b.N1 * b.KU * 2);
  b.h0pre = (float*)ar.take(Bs * b.N1 * 4);
  if (b.has_merge) {
    b.a1_img = (unsigned char*)ar.take(Rp * b.N1 * 2);
    b.wm_f = (unsigned char*)ar.take((size_t)W * b.N1 * 2);
    b.wm_b = (unsigned char*)ar.take((size_t)W * b.N1 * 2);
    b.m = (float*)ar.take(Bs * W * 4);
  }
  b.h0 = (float*)ar.take(Bs * W * 4);
  b.h0_img = (unsigned char*)ar.take(Rp * W * 2);
  // the FiLM shift images of all blocks are contiguous: together they are the B operand of ONE GEMM with N = L * W (an image
  // is W / NT column tiles of KT * NT * 2 bytes, so concatenated images are concatenated tiles)
  unsigned char* wb_all = (unsigned char*)ar.take((size_t)c->L * W * b.KT * 2);
  b.fbb_all = (float*)ar.take((size_t)c->L * W * 4);
  for (int l = 0; l < c->L; ++l) {
    b.wd_f[l] = (unsigned char*)ar.take((size_t)W * W * 2);
    b.wd_b[l] = (unsigned char*)ar.take((size_t)W * W * 2);
    b.wb_f[l] = wb_all ? wb_all + (size_t)l * W * b.KT * 2 : nullptr;
    b.wb_b[l] = (unsigned char*)ar.take((size_t)r64(c->T) * W * 2);
    b.P[l] = (float*)ar.take(Bs * W * 4);
    b.xhat[l] = (float*)ar.take(Bs * W * 4);
    b.rstd[l] = (float*)ar.take(Bs * 4);
    b.h_img[l] = (unsigned char*)ar.take(Rp * W * 2);
  }
  b.FB = (float*)ar.take(Bs * (size_t)c->L * W * 4);  // (B, L * W): block l reads columns l*W .. with row stride L*W
  bool ragged = false;
  for (int l = 0; l < c->L; ++l) ragged = ragged || (c->block_width[l] != 0);
  if (ragged) {  // a projector may sit in any block
    b.skip = (float*)ar.take(Bs * W * 4);
    b.wp_f = (unsigned char*)ar.take((size_t)W * W * 2);
    b.wp_b = (unsigned char*)ar.take((size_t)W * W * 2);
  }
  if (c->film_use_gamma) {
    b.FG = (float*)ar.take(Bs * W * 4);
    b.wg_f = (unsigned char*)ar.take((size_t)W * b.KT * 2);
    b.wg_b = (unsigned char*)ar.take((size_t)r64(c->T) * W * 2);
    if (backward) b.dG_img = (unsigned char*)ar.take(Rp * W * 2);
  }
  b.wo_f = (unsigned char*)ar.take((size_t)b.NO * W * 2);
  b.wo_b = (unsigned char*)ar.take((size_t)W * b.NO * 2);
  b.bo_pad = (float*)ar.take((size_t)b.NO * 4);
  b.out = (float*)ar.take(Bs * b.NO * 4);
  b.dout_img = (unsigned char*)ar.take(Rp * b.NO * 2);
  if (backward) {
    b.dH = (float*)ar.take(Bs * W * 4);
    b.dP_img = (unsigned char*)ar.take(Rp * W * 2);
    b.dV_img = (unsigned char*)ar.take(Rp * W * 2);
    b.dP_img2 = (unsigned char*)ar.take(Rp * W * 2);
    b.dV_img2 = (unsigned char*)ar.take(Rp * W * 2);
    b.dte = (float*)ar.take(Bs * b.KTN * 4);
    if (b.has_merge) b.da1 = (float*)ar.take(Bs * b.N1 * 4);
    b.dh0pre_img = (unsigned char*)ar.take(Rp * b.N1 * 2);
    b.dW1 = (float*)ar.take((size_t)b.KU * b.N1 * 4);
    b.db1 = (float*)ar.take((size_t)b.N1 * 4);
  }
}

int check_train_cfg(const bfk_net_config* c) {
  if (int r = check_cfg(c)) return r;
  if ((c->W % 64) != 0 || c->W > 1024) return fail(-70, "layer-wise path: W must be a multiple of 64, <= 1024");
  if (c->D > 64) return fail(-71, "layer-wise path: D <= 64");
  if (c->precision != BFK_PREC_FP16) return fail(-74, "layer-wise path: fp16 operands only (bf16 cannot meet the 2e-3 gradient gate)");
  for (int l = 0; l < c->L; ++l)
    if (c->block_width[l] < 0 || c->block_width[l] > c->W || (c->block_width[l] % 8) != 0)
      return fail(-69, "layer-wise path: block_width[l] must be 0 or a multiple of 8 that is <= W (the padded common width)");
  return 0;
}

int launch_gemm_dw(const unsigned char* X, int XK8, const unsigned char* Y, int YK8, float* Cm, long long ldc, int Mv, int Nv,
                   long long Rp, float alpha, int swap, cudaStream_t st) {
  GemmDwArgs a{};
  a.X = X; a.Y = Y; a.C = Cm; a.ldc = ldc; a.XK8 = XK8; a.YK8 = YK8; a.Mv = Mv; a.Nv = Nv; a.alpha = alpha; a.swap = swap;
  a.NT = (YK8 * 8 >= 128) ? 128 : 64;
  a.n_btiles = (int)(Rp / 128);
  const int mt = (XK8 + 15) / 16, nt = (YK8 * 8 + a.NT - 1) / a.NT;
  int sms = bfk_num_sms();
  if (sms <= 0) sms = 148;
  int splits = sms / (mt * nt);
  if (splits < 1) splits = 1;
  if (splits > a.n_btiles) splits = a.n_btiles;
  a.tiles_per_split = (a.n_btiles + splits - 1) / splits;
  splits = (a.n_btiles + a.tiles_per_split - 1) / a.tiles_per_split;
  a.idesc = umma_idesc_f16_mn(G_TM, a.NT, 0);
  static bool attr_set_dev[64] = {};
  bool& attr_set = *attr_flag(attr_set_dev);
  if (!attr_set) {
    CK(cudaFuncSetAttribute(gemm_dw_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)gemm_dw_smem_bytes(128)));
    attr_set = true;
  }
  gemm_dw_kernel<<<dim3(mt, nt, splits), G_THREADS, gemm_dw_smem_bytes(a.NT), st>>>(a);
  CK(cudaGetLastError());
  return 0;
}

// one resident wave: a grid that is not a multiple of what fits leaves a half-empty second wave (r01u: 592 blocks on 444
// slots cost 2 waves)
template <typename K>
unsigned resident_grid(K kernel, int threads, size_t smem) {
  int per_sm = 0;
  if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, threads, smem) != cudaSuccess || per_sm < 1) per_sm = 1;
  int sms = bfk_num_sms();
  return (unsigned)((sms > 0 ? sms : 148) * per_sm);
}
static unsigned row_grid(unsigned resident, long long rows) {  // 8 row-warps per CTA; no more CTAs than rows need
  const long long need = (rows + 7) / 8;
  return (unsigned)(need < (long long)resident ? (need < 1 ? 1 : need) : resident);
}
int launch_act_pack(const float* P, long long ld, long long B, long long Rp, int K8, int act, unsigned char* img, float* out_f32,
                    cudaStream_t st) {
  static unsigned res = 0;
  if (!res) res = resident_grid(act_pack_kernel, 256, 0);
  act_pack_kernel<<<row_grid(res, Rp), 256, 0, st>>>(P, ld, B, Rp, K8, act, img, out_f32);
  CK(cudaGetLastError());
  return 0;
}
int launch_act_bwd(const float* dY, const float* P, long long ld, long long B, long long Rp, int K8, int act, unsigned char* img,
                   float* dbias, float gscale, cudaStream_t st) {
  const int nj = (K8 + 31) / 32;
  const size_t sh = (size_t)K8 * 8 * 4;
  if (nj > 8) return fail(-79, "bfk_train_step: first-layer width above 2048 columns");
#define BFK_ACT_BWD(NJ)                                                                                         \
  {                                                                                                              \
    static unsigned res = 0;                                                                                     \
    if (!res) res = resident_grid(act_bwd_kernel<NJ>, 256, 8192);                                                \
    act_bwd_kernel<NJ><<<row_grid(res, Rp), 256, sh, st>>>(dY, P, ld, B, Rp, K8, act, img, dbias, gscale);       \
  }
  if (nj == 1) BFK_ACT_BWD(1) else if (nj == 2) BFK_ACT_BWD(2) else if (nj <= 4) BFK_ACT_BWD(4) else BFK_ACT_BWD(8)
#undef BFK_ACT_BWD
  CK(cudaGetLastError());
  return 0;
}
int launch_block_fwd(const BlockFwdArgs& a, cudaStream_t st) {
  const int nj = (a.W / 8 + 31) / 32;
  static unsigned grid[3] = {0, 0, 0};
  if (!grid[0]) {
    grid[0] = resident_grid(block_fwd_kernel<1>, 256, 0);
    grid[1] = resident_grid(block_fwd_wide_kernel<2>, 256, 0);
    grid[2] = resident_grid(block_fwd_wide_kernel<4>, 256, 0);
  }
  if (nj == 1) block_fwd_kernel<1><<<grid[0], 256, 0, st>>>(a);
  else if (nj == 2) block_fwd_wide_kernel<2><<<grid[1], 256, 0, st>>>(a);
  else block_fwd_wide_kernel<4><<<grid[2], 256, 0, st>>>(a);
  CK(cudaGetLastError());
  return 0;
}
int launch_block_bwd(const BlockBwdArgs& a, cudaStream_t st) {
  const int nj = (a.W / 8 + 31) / 32;
  const size_t sh = (size_t)(a.FG ? 5 : 4) * a.W * 4;
  unsigned grid;
  if (a.FG) {  // film_use_gamma
    if (nj == 1) { grid = resident_grid(block_bwd_kernel<1, true>, 256, sh); block_bwd_kernel<1, true><<<grid, 256, sh, st>>>(a); }
    else if (nj == 2) { grid = resident_grid(block_bwd_kernel<2, true>, 256, sh); block_bwd_kernel<2, true><<<grid, 256, sh, st>>>(a); }
    else { grid = resident_grid(block_bwd_kernel<4, true>, 256, sh); block_bwd_kernel<4, true><<<grid, 256, sh, st>>>(a); }
  } else if (nj == 1) { grid = resident_grid(block_bwd_kernel<1>, 256, sh); block_bwd_kernel<1><<<grid, 256, sh, st>>>(a); }
  else {
    // W > 256: a row shared by 2 / 4 warps (block_bwd_split_kernel); BFK_BWD_SPLIT=0 keeps the one-warp-per-row kernel
    static int split = -1;
    if (split < 0) { const char* e = getenv("BFK_BWD_SPLIT"); split = e ? atoi(e) : 1; }
    if (split && nj == 2) { grid = resident_grid(block_bwd_split_kernel<2>, 256, sh); block_bwd_split_kernel<2><<<grid, 256, sh, st>>>(a); }
    else if (split) { grid = resident_grid(block_bwd_split_kernel<4>, 256, sh); block_bwd_split_kernel<4><<<grid, 256, sh, st>>>(a); }
    else if (nj == 2) { grid = resident_grid(block_bwd_kernel<2>, 256, sh); block_bwd_kernel<2><<<grid, 256, sh, st>>>(a); }
    else { grid = resident_grid(block_bwd_kernel<4>, 256, sh); block_bwd_kernel<4><<<grid, 256, sh, st>>>(a); }
  }
  CK(cudaGetLastError());
  return 0;
}

// Side stream of the backward pass: the parameter-gradient GEMMs of a block (dW of Dense and FiLM, the time-embedding
// cotangent) do not feed the next block's chain (block_bwd -> dX GEMM), so they run beside it; fork / join with events,
// which is also how a stream capture of the step (training.py Trainer, CUDA graph) picks the side stream up.
struct SideCtx {
  cudaStream_t s = nullptr;
  cudaEvent_t main_ev[BFK_MAX_BLOCKS + 1] = {}, side_ev[BFK_MAX_BLOCKS + 1] = {};
  bool ok = false;
};
static SideCtx* side_ctx() {
  static SideCtx ctx[64];
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return nullptr;
  SideCtx& c = ctx[dev];
  if (!c.ok) {
    if (cudaStreamCreateWithFlags(&c.s, cudaStreamNonBlocking) != cudaSuccess) return nullptr;
    for (int i = 0; i <= BFK_MAX_BLOCKS; ++i)
      if (cudaEventCreateWithFlags(&c.main_ev[i], cudaEventDisableTiming) != cudaSuccess ||
          cudaEventCreateWithFlags(&c.side_ev[i], cudaEventDisableTiming) != cudaSuccess) return nullptr;
    c.ok = true;
  }
  return &c;
}

// fp16 weight images of one training step, collected and packed by ONE launch (pack_batch_kernel)
struct PackList {
  PackBatch pb{};
  long long max_n = 0;
  static long long rup(long long x, long long m) { return (x + m - 1) / m * m; }
  // bfk_pack_kmajor(src, ld, R, k_valid, k_pad, row_tile) -> dst
  void add(const float* src, long long ld, long long R, int k_valid, int k_pad, int RT, void* dst) {
    PackJob& q = pb.j[pb.n++];
    q.src = src; q.dst = (uint16_t*)dst; q.ld = ld; q.a = R; q.b = rup(R, RT); q.c = k_valid; q.d = k_pad; q.RT = RT; q.t = 0;
    q.n = q.b * (k_pad / 8);
    if (q.n > max_n) max_n = q.n;
  }
  // bfk_pack_kmajor_t(src, ld, k_valid, k_pad, F, row_tile) -> dst
  void add_t(const float* src, long long ld, long long k_valid, long long k_pad, int F, int RT, void* dst) {
    PackJob& q = pb.j[pb.n++];
    q.src = src; q.dst = (uint16_t*)dst; q.ld = ld; q.a = k_valid; q.b = k_pad; q.c = F; q.d = (int)rup(F, RT); q.RT = RT; q.t = 1;
    q.n = (long long)q.d * (k_pad / 8);
    if (q.n > max_n) max_n = q.n;
  }
  int launch(cudaStream_t st) {
    if (pb.n == 0) return 0;
    pack_batch_kernel<<<dim3((unsigned)((max_n + 255) / 256), (unsigned)pb.n), 256, 0, st>>>(pb);
    CK(cudaGetLastError());
    return 0;
  }
};

#define GP(p) const_cast<float*>(p) /* `grads` shares the layout of bfk_weights; its tensors are written */
#define RC(call)            \
  do {                      \
    int _r = (call);        \
    if (_r) return _r;      \
  } while (0)

}  // namespace

extern "C" size_t bfk_train_workspace_bytes(const bfk_net_config* c, int64_t B, int32_t backward) {
  if (check_train_cfg(c) || B <= 0) return 0;
  Arena ar{nullptr, 0, 0};
  TrainBufs b{};
  plan_train(c, B, backward != 0, ar, b);
  return ar.off + 256;
}

extern "C" int bfk_train_step(const bfk_net_config* c, const bfk_weights* w, const bfk_weights* g, const bfk_train_args* ta,
                              float* loss, float* pred_out, void* workspace, size_t workspace_bytes, void* stream_) {
  RC(check_train_cfg(c));
  if (!w || !ta || !workspace || !ta->x || !ta->t || ta->B <= 0) return fail(-75, "bfk_train_step: bad arguments");
  if (c->C > 0 && !ta->cond) return fail(-76, "bfk_train_step: conditional network needs conditions");
  if ((loss || g) && !ta->target) return fail(-77, "bfk_train_step: loss / gradients need a target");
  if (ta->dropout < 0.f || ta->dropout >= 1.f) return fail(-78, "bfk_train_step: dropout rate must be in [0, 1)");
  cudaStream_t st = (cudaStream_t)stream_;
  const long long B = ta->B;
  const int W = c->W, D = c->D, C = c->C, L = c->L, T = c->T, act = c->activation, P16 = BFK_PREC_FP16;
  const bool bwd = g != nullptr;
  Arena ar{(unsigned char*)workspace, 0, workspace_bytes};
  TrainBufs b{};
  plan_train(c, B, bwd, ar, b);
  if (ar.off > workspace_bytes) return fail(-79, "bfk_train_step: workspace too small (bfk_train_workspace_bytes)");
  const long long Rp = b.Rp;
  const int KU = b.KU, KT = b.KT, KTN = b.KTN, N1 = b.N1, NO = b.NO, NTW = nt_of(W), NT1 = nt_of(N1);
  const float gs = ta->grad_scale, inv_keep = 1.f / (1.f - ta->dropout);
  {
    // every weight image with a buffer of its own, forward and (when gradients are wanted) backward, in one launch; the
    // images that share a scratch buffer across blocks (FiLM scale, residual projector) and w1 (assembled below) stay inline
    PackList pl;
    if ((c->C > 0 ? 2 : 0) + 4 * L + 2 > PACK_MAX_JOBS) return fail(-79, "bfk_train_step: too many blocks for the pack batch");
    if (b.has_merge) pl.add_t(w->m_k, W, N1, N1, W, NTW, b.wm_f);
    for (int l = 0; l < L; ++l) {
      pl.add_t(w->dense_k[l], W, W, W, W, NTW, b.wd_f[l]);
      pl.add_t(w->film_b_k[l], W, T, KT, W, NTW, b.wb_f[l]);
    }
    pl.add_t(w->out_k, D, W, W, D, 64, b.wo_f);
    if (bwd) {
      pl.add(w->out_k, D, W, D, NO, NTW, b.wo_b);
      for (int l = 0; l < L; ++l) {
        if (T > 1) pl.add(w->film_b_k[l], W, T, W, W, 64, b.wb_b[l]);
        pl.add(w->dense_k[l], W, W, W, W, NTW, b.wd_b[l]);
      }
      if (b.has_merge) pl.add(w->m_k, W, N1, W, W, NT1, b.wm_b);
    }
    RC(pl.launch(st));
  }

  // ---- forward ---------------------------------------------------------------------------------------------------
  train_input_kernel<<<dim3((unsigned)((Rp + 127) / 128), (unsigned)(KU / 8 + KT / 8)), 128, 0, st>>>(ta->x, ta->cond, ta->t, w->freq, c->fourier_scale, B, Rp, D, C, T,
                                                                 KU / 8, KT / 8, b.u_img, b.te_img);
  // FiLM shifts of ALL blocks in one GEMM (N = L * W) beside the input chain: they depend on the time embedding only
  SideCtx* fsc = getenv("BFK_TRAIN_NO_OVERLAP") ? nullptr : side_ctx();
  {
    cudaStream_t fs = fsc ? fsc->s : st;
    if (fsc) {
      CK(cudaEventRecord(fsc->main_ev[L], st));
      CK(cudaStreamWaitEvent(fs, fsc->main_ev[L], 0));
    }
    for (int l = 0; l < L; ++l)
      CK(cudaMemcpyAsync(b.fbb_all + (size_t)l * W, w->film_b_b[l], (size_t)W * 4, cudaMemcpyDeviceToDevice, fs));
    CK(cudaMemsetAsync(b.bo_pad, 0, (size_t)NO * 4, fs));  // zero-padded output bias, off the dependent chain as well
    CK(cudaMemcpyAsync(b.bo_pad, w->out_b, (size_t)D * 4, cudaMemcpyDeviceToDevice, fs));
    RC(bfk_gemm(b.te_img, b.wb_f[0], b.FB, (long long)L * W, B, L * W, KT, NTW, 1, 0, b.fbb_all, 1.f, P16, fs));
    if (fsc) CK(cudaEventRecord(fsc->side_ev[L], fs));
  }
  build_w1_kernel<<<cdiv((size_t)KU * N1, 256), 256, 0, st>>>(w->x_k, w->x_b, w->c_k, w->c_b, D, C, W, b.has_merge ? c->merge : 0, KU, N1,
                                                             b.w1, b.b1);
  CK(cudaGetLastError());
  RC(bfk_pack_kmajor_t(b.w1, N1, D + C, KU, N1, NT1, P16, b.w1_pk, st));
  RC(bfk_gemm(b.u_img, b.w1_pk, b.h0pre, N1, B, N1, KU, NT1, 1, 0, b.b1, 1.f, P16, st));
  if (b.has_merge) {
    RC(launch_act_pack(b.h0pre, N1, B, Rp, N1 / 8, act, b.a1_img, nullptr, st));
    RC(bfk_gemm(b.a1_img, b.wm_f, b.m, W, B, W, N1, NTW, 1, 0, w->m_b, 1.f, P16, st));
    RC(launch_act_pack(b.m, W, B, Rp, W / 8, act, b.h0_img, b.h0, st));
  } else {
    RC(launch_act_pack(b.h0pre, W, B, Rp, W / 8, act, b.h0_img, b.h0, st));
  }
  CK(cudaGetLastError());
  for (int l = 0; l < L; ++l) {
    RC(bfk_gemm(l == 0 ? b.h0_img : b.h_img[l - 1], b.wd_f[l], b.P[l], W, B, W, W, NTW, 1, 0, w->dense_b[l], 1.f, P16, st));
    if (c->film_use_gamma) {  // FiLM scale to_gamma(t_emb): one more K = T GEMM per block
      RC(bfk_pack_kmajor_t(w->film_g_k[l], W, T, KT, W, NTW, P16, b.wg_f, st));
      RC(bfk_gemm(b.te_img, b.wg_f, b.FG, W, B, W, KT, NTW, 1, 0, w->film_g_b[l], 1.f, P16, st));
    }
    const bool proj = c->residual && w->proj_k[l] != nullptr;
    if (proj) {  // unequal in / out widths: skip = x Wp + bp (conditional_dense_block.py:125-127)
      RC(bfk_pack_kmajor_t(w->proj_k[l], W, W, W, W, NTW, P16, b.wp_f, st));
      RC(bfk_gemm(l == 0 ? b.h0_img : b.h_img[l - 1], b.wp_f, b.skip, W, B, W, W, NTW, 1, 0, w->proj_b[l], 1.f, P16, st));
    }
    BlockFwdArgs a{};
    a.P = b.P[l]; a.FB = b.FB + (size_t)l * W; a.FG = c->film_use_gamma ? b.FG : nullptr; a.fb_stride = (long long)L * W;
    a.fg_stride = W;
    if (l == 0 && fsc) CK(cudaStreamWaitEvent(st, fsc->side_ev[L], 0));  // join: the batched FiLM GEMM
    a.Weff = c->block_width[l];
    a.prev = c->residual ? (proj ? b.skip : (l == 0 ? b.h0 : b.xhat[l - 1])) : nullptr;
    a.pg = (!proj && l > 0 && c->norm != BFK_NORM_NONE) ? w->norm_g[l - 1] : nullptr;  // rms: scale only (pb stays null)
    a.pb = (!proj && l > 0 && c->norm == BFK_NORM_LAYER) ? w->norm_b[l - 1] : nullptr;
    a.gamma = c->norm != BFK_NORM_NONE ? w->norm_g[l] : nullptr;
    a.beta = c->norm == BFK_NORM_LAYER ? w->norm_b[l] : nullptr;
    a.xhat = b.xhat[l]; a.rstd = b.rstd[l]; a.img = b.h_img[l]; a.out_f32 = nullptr;
    a.B = B; a.R_pad = Rp; a.W = W; a.act = act; a.norm = c->norm; a.layer = l;
    a.drop_rate = ta->dropout; a.inv_keep = inv_keep; a.seed = ta->seed; a.seed_dev = ta->seed_dev; a.row0 = ta->row_offset;
    RC(launch_block_fwd(a, st));
  }
  RC(bfk_gemm(b.h_img[L - 1], b.wo_f, b.out, NO, B, NO, W, 64, 1, 0, b.bo_pad, 1.f, P16, st));
  if (pred_out) CK(cudaMemcpy2DAsync(pred_out, (size_t)D * 4, b.out, (size_t)NO * 4, (size_t)D * 4, (size_t)B, cudaMemcpyDeviceToDevice, st));
  if (!loss && !bwd) return 0;
  float* loss_dst = loss ? loss : b.FB;  // FB is free by now: scratch float when the caller does not want the loss
  loss_kernel<<<(unsigned)((Rp + 255) / 256), 256, 0, st>>>(b.out, NO, ta->target, ta->pred_scale, ta->row_weight, B, Rp, D, ta->loss,
                                                           b.dout_img, NO / 8, loss_dst, ta->loss_scale, bwd ? GP(g->out_b) : (float*)nullptr, gs, ta->huber_c,
                                                           ta->elem_mask);
  CK(cudaGetLastError());
  if (!bwd) return 0;

  // ---- backward --------------------------------------------------------------------------------------------------
  const int sw = ta->dw_swap;
  // overlap (default block shape only): ss = side stream, else everything stays on st
  bool ragged_cfg = false;
  for (int l = 0; l < L; ++l) ragged_cfg = ragged_cfg || (c->block_width[l] != 0);
  SideCtx* sc = (!c->film_use_gamma && !ragged_cfg && !getenv("BFK_TRAIN_NO_OVERLAP")) ? side_ctx() : nullptr;
  cudaStream_t ss = sc ? sc->s : st;
  if (sc) {  // fork: the side stream starts after everything enqueued so far (loss kernel, zeroed gradients)
    CK(cudaEventRecord(sc->main_ev[L], st));
    CK(cudaStreamWaitEvent(ss, sc->main_ev[L], 0));
  }
  // zeroed early, off the dependent chain: the accumulators of the input-side backward (first touched after the join)
  CK(cudaMemsetAsync(b.db1, 0, (size_t)N1 * 4, ss));
  CK(cudaMemsetAsync(b.dW1, 0, (size_t)KU * N1 * 4, ss));
  RC(launch_gemm_dw(b.h_img[L - 1], W / 8, b.dout_img, NO / 8, GP(g->out_k), D, W, D, Rp, gs, sw, ss));
  RC(bfk_gemm(b.dout_img, b.wo_b, b.dH, W, B, W, NO, NTW, 1, 0, nullptr, 1.f, P16, st));
  const bool want_dte = T > 1;
  for (int l = L - 1; l >= 0; --l) {
    // the cotangent images alternate between two buffer pairs; the pair of block l was last read by the side work of l + 2
    unsigned char* dP_img = (sc && (l & 1)) ? b.dP_img2 : b.dP_img;
    unsigned char* dV_img = (sc && (l & 1)) ? b.dV_img2 : b.dV_img;
    if (sc && l + 2 <= L - 1) CK(cudaStreamWaitEvent(st, sc->side_ev[l + 2], 0));
    BlockBwdArgs a{};
    a.dH = b.dH; a.xhat = b.xhat[l]; a.rstd = b.rstd[l]; a.P = b.P[l];
    a.gamma = c->norm != BFK_NORM_NONE ? w->norm_g[l] : nullptr;
    a.dP_img = dP_img; a.dV_img = dV_img;
    a.d_gamma = GP(g->norm_g[l]); a.d_beta = c->norm == BFK_NORM_LAYER ? GP(g->norm_b[l]) : nullptr;
    a.d_bd = GP(g->dense_b[l]); a.d_bfilm = GP(g->film_b_b[l]);
    a.B = B; a.R_pad = Rp; a.W = W; a.act = act; a.norm = c->norm; a.layer = l;
    a.drop_rate = ta->dropout; a.inv_keep = inv_keep; a.seed = ta->seed; a.seed_dev = ta->seed_dev; a.row0 = ta->row_offset; a.gscale = gs;
    const bool proj = c->residual && w->proj_k[l] != nullptr;
    a.Weff = c->block_width[l];
    a.d_bskip = proj ? GP(g->proj_b[l]) : nullptr;
    if (c->film_use_gamma) {  // the forward kept only the last block's FiLM scale: rebuild this block's
      RC(bfk_pack_kmajor_t(w->film_g_k[l], W, T, KT, W, NTW, P16, b.wg_f, st));
      RC(bfk_gemm(b.te_img, b.wg_f, b.FG, W, B, W, KT, NTW, 1, 0, w->film_g_b[l], 1.f, P16, st));
      a.FG = b.FG; a.dG_img = b.dG_img; a.d_bgfilm = GP(g->film_g_b[l]);
    }
    RC(launch_block_bwd(a, st));
    if (sc) {
      CK(cudaEventRecord(sc->main_ev[l], st));
      CK(cudaStreamWaitEvent(ss, sc->main_ev[l], 0));
    }
    RC(launch_gemm_dw(l == 0 ? b.h0_img : b.h_img[l - 1], W / 8, dP_img, W / 8, GP(g->dense_k[l]), W, W, W, Rp, gs, sw, ss));
    RC(launch_gemm_dw(b.te_img, KT / 8, dV_img, W / 8, GP(g->film_b_k[l]), W, T, W, Rp, gs, sw, ss));
    if (c->film_use_gamma) RC(launch_gemm_dw(b.te_img, KT / 8, b.dG_img, W / 8, GP(g->film_g_k[l]), W, T, W, Rp, gs, sw, st));
    if (want_dte) {
      RC(bfk_gemm(dV_img, b.wb_b[l], b.dte, KTN, B, KTN, W, 64, 1, l == L - 1 ? 0 : GEMM_ADD_EXISTING, nullptr, 1.f, P16, ss));
      if (c->film_use_gamma) {
        RC(bfk_pack_kmajor(w->film_g_k[l], W, T, W, W, 64, P16, b.wg_b, st));
        RC(bfk_gemm(b.dG_img, b.wg_b, b.dte, KTN, B, KTN, W, 64, 1, GEMM_ADD_EXISTING, nullptr, 1.f, P16, st));
      }
    }
    if (sc) CK(cudaEventRecord(sc->side_ev[l], ss));
    if (proj) RC(launch_gemm_dw(l == 0 ? b.h0_img : b.h_img[l - 1], W / 8, dV_img, W / 8, GP(g->proj_k[l]), W, W, W, Rp, gs, sw, st));
    if (!c->residual || proj) CK(cudaMemsetAsync(b.dH, 0, (size_t)B * W * 4, st));  // no identity skip into the input
    RC(bfk_gemm(dP_img, b.wd_b[l], b.dH, W, B, W, W, NTW, 1, GEMM_ADD_EXISTING, nullptr, 1.f, P16, st));
    if (proj) {  // the skip's cotangent dV goes back through the projector
      RC(bfk_pack_kmajor(w->proj_k[l], W, W, W, W, NTW, P16, b.wp_b, st));
      RC(bfk_gemm(dV_img, b.wp_b, b.dH, W, B, W, W, NTW, 1, GEMM_ADD_EXISTING, nullptr, 1.f, P16, st));
    }
  }
  // join: the input-side backward reuses the first image pair (last read by block 0's or block 1's side work) and the
  // frequency gradient needs the finished time-embedding cotangent
  if (sc) CK(cudaStreamWaitEvent(st, sc->side_ev[0], 0));
  // the frequency gradient only needs the time-embedding cotangent, complete on the side stream in stream order; the
  // merge layer's dW joins it there, so the dependent chain on `st` is act_bwd -> dX -> act_bwd -> dW1 -> scatter
  if (want_dte && GP(g->freq) && T > 1)
    freq_bwd_kernel<<<dim3((unsigned)((B + 255) / 256), (unsigned)((T - 1) / 2)), 256, 0, ss>>>(b.dte, KTN, ta->t, w->freq, c->fourier_scale, B, (T - 1) / 2, GP(g->freq), gs);
  CK(cudaGetLastError());
  if (b.has_merge) {
    RC(launch_act_bwd(b.dH, b.m, W, B, Rp, W / 8, act, b.dP_img, GP(g->m_b), gs, st));
    if (sc) {
      CK(cudaEventRecord(sc->main_ev[0], st));
      CK(cudaStreamWaitEvent(ss, sc->main_ev[0], 0));
    }
    RC(launch_gemm_dw(b.a1_img, N1 / 8, b.dP_img, W / 8, GP(g->m_k), W, N1, W, Rp, gs, sw, ss));
    RC(bfk_gemm(b.dP_img, b.wm_b, b.da1, N1, B, N1, W, NT1, 1, 0, nullptr, 1.f, P16, st));
    RC(launch_act_bwd(b.da1, b.h0pre, N1, B, Rp, N1 / 8, act, b.dh0pre_img, b.db1, 1.f, st));
  } else {
    RC(launch_act_bwd(b.dH, b.h0pre, N1, B, Rp, N1 / 8, act, b.dh0pre_img, b.db1, 1.f, st));
  }
  CK(cudaGetLastError());
  RC(launch_gemm_dw(b.u_img, KU / 8, b.dh0pre_img, N1 / 8, b.dW1, N1, D + C, N1, Rp, 1.f, sw, st));
  scatter_w1_grad_kernel<<<cdiv((size_t)(D + C) * W, 256), 256, 0, st>>>(b.dW1, b.db1, D, C, W, b.has_merge ? c->merge : 0, N1, gs,
                                                                        GP(g->x_k), GP(g->x_b), GP(g->c_k), GP(g->c_b));
  CK(cudaGetLastError());
  if (sc) {  // final join
    CK(cudaEventRecord(sc->side_ev[1], ss));
    CK(cudaStreamWaitEvent(st, sc->side_ev[1], 0));
  }
  return 0;
}

// One reverse pass of the velocity network w.r.t. its INPUT over the activations a preceding
// bfk_train_step(c, w, NULL, ta, NULL, pred, workspace /* sized with backward=1 */) left in `workspace`.
// Cotangent = dout (B,D) or, when dout == NULL, the one-hot e_col.  Writes dx (B,D) if non-NULL and adds dx[:,col] to
// trace (B,) if non-NULL.  No parameter gradients are touched.  Reference: utils/jacobian/vjp.py:44-68 (torch.autograd.grad
// through the subnet), utils/jacobian/jacobian_trace.py:44-55.
extern "C" int bfk_train_vjp(const bfk_net_config* c, const bfk_weights* w, const bfk_train_args* ta, const float* dout,
                             int32_t col, float* dx, float* trace, void* workspace, size_t workspace_bytes, void* stream_) {
  RC(check_train_cfg(c));
  if (!w || !ta || !workspace || ta->B <= 0) return fail(-82, "bfk_train_vjp: bad arguments");
  if (!dout && (col < 0 || col >= c->D)) return fail(-83, "bfk_train_vjp: one-hot column out of range");
  if (!dx && !trace) return fail(-84, "bfk_train_vjp: nothing to compute (dx and trace are NULL)");
  cudaStream_t st = (cudaStream_t)stream_;
  const long long B = ta->B;
  const int W = c->W, D = c->D, L = c->L, act = c->activation, P16 = BFK_PREC_FP16;
  Arena ar{(unsigned char*)workspace, 0, workspace_bytes};
  TrainBufs b{};
  plan_train(c, B, true, ar, b);
  if (ar.off > workspace_bytes) return fail(-79, "bfk_train_vjp: workspace too small (bfk_train_workspace_bytes, backward=1)");
  const long long Rp = b.Rp;
  const int N1 = b.N1, NO = b.NO, NTW = nt_of(W), NT1 = nt_of(N1);
  cotangent_pack_kernel<<<(unsigned)((Rp + 127) / 128), 128, 0, st>>>(dout, col, B, Rp, D, NO / 8, b.dout_img);
  CK(cudaGetLastError());
  // wo_b / wd_b / wm_b: the MN-major weight images; packed here so the pass does not depend on an earlier backward
  RC(bfk_pack_kmajor(w->out_k, D, W, D, NO, NTW, P16, b.wo_b, st));
  RC(bfk_gemm(b.dout_img, b.wo_b, b.dH, W, B, W, NO, NTW, 1, 0, nullptr, 1.f, P16, st));
  for (int l = L - 1; l >= 0; --l) {
    BlockBwdArgs a{};
    a.dH = b.dH; a.xhat = b.xhat[l]; a.rstd = b.rstd[l]; a.P = b.P[l];
    a.gamma = c->norm != BFK_NORM_NONE ? w->norm_g[l] : nullptr;
    a.dP_img = b.dP_img; a.dV_img = b.dV_img;
    a.d_gamma = a.d_beta = a.d_bd = a.d_bfilm = nullptr;
    a.B = B; a.R_pad = Rp; a.W = W; a.act = act; a.norm = c->norm; a.layer = l;
    a.drop_rate = 0.f; a.inv_keep = 1.f; a.seed = 0; a.seed_dev = nullptr; a.row0 = 0; a.gscale = 1.f;
    const bool proj = c->residual && w->proj_k[l] != nullptr;
    a.Weff = c->block_width[l];
    if (c->film_use_gamma) {  // dU = dV (1 + FG): rebuild this block's FiLM scale (time does not depend on x: no dG needed)
      RC(bfk_pack_kmajor_t(w->film_g_k[l], W, c->T, b.KT, W, NTW, P16, b.wg_f, st));
      RC(bfk_gemm(b.te_img, b.wg_f, b.FG, W, B, W, b.KT, NTW, 1, 0, w->film_g_b[l], 1.f, P16, st));
      a.FG = b.FG; a.dG_img = nullptr; a.d_bgfilm = nullptr;
    }
    RC(launch_block_bwd(a, st));
    if (!c->residual || proj) CK(cudaMemsetAsync(b.dH, 0, (size_t)B * W * 4, st));
    RC(bfk_pack_kmajor(w->dense_k[l], W, W, W, W, NTW, P16, b.wd_b[l], st));
    RC(bfk_gemm(b.dP_img, b.wd_b[l], b.dH, W, B, W, W, NTW, 1, GEMM_ADD_EXISTING, nullptr, 1.f, P16, st));
    if (proj) {
      RC(bfk_pack_kmajor(w->proj_k[l], W, W, W, W, NTW, P16, b.wp_b, st));
      RC(bfk_gemm(b.dV_img, b.wp_b, b.dH, W, B, W, W, NTW, 1, GEMM_ADD_EXISTING, nullptr, 1.f, P16, st));
    }
  }
  if (b.has_merge) {
    RC(launch_act_bwd(b.dH, b.m, W, B, Rp, W / 8, act, b.dP_img, nullptr, 1.f, st));
    RC(bfk_pack_kmajor(w->m_k, W, N1, W, W, NT1, P16, b.wm_b, st));
    RC(bfk_gemm(b.dP_img, b.wm_b, b.da1, N1, B, N1, W, NT1, 1, 0, nullptr, 1.f, P16, st));
    RC(launch_act_bwd(b.da1, b.h0pre, N1, B, Rp, N1 / 8, act, b.dh0pre_img, nullptr, 1.f, st));
  } else {
    RC(launch_act_bwd(b.dH, b.h0pre, N1, B, Rp, N1 / 8, act, b.dh0pre_img, nullptr, 1.f, st));
  }
  CK(cudaGetLastError());
  const unsigned gb = (unsigned)((B + 127) / 128);
  if (D <= 2) vjp_input_kernel<2><<<gb, 128, 0, st>>>(b.dh0pre_img, N1 / 8, w->x_k, B, D, W, dx, trace, col);
  else vjp_input_kernel<8><<<gb, 128, 0, st>>>(b.dh0pre_img, N1 / 8, w->x_k, B, D, W, dx, trace, col);
  CK(cudaGetLastError());
  return 0;
}

// The keep scales (0 or 1/(1-rate)) block `layer` of bfk_train_step multiplies its activation with, for rows
// row_offset .. row_offset+B-1 and columns 0 .. W-1: makes the counter-based mask observable (tests/test_gpu_train.py
// restates the hash in numpy and checks the statistics).  Same device function as the block kernels.
namespace bfk {
__global__ void drop_mask_kernel(unsigned long long seed, const uint32_t* seed_dev, int layer, long long row0, long long B, int W,
                                 float rate, float inv_keep, float* __restrict__ out) {
  const unsigned long long key = drop_key(seed, seed_dev, layer);
  const int K8 = W / 8;
  const long long n = B * K8;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    const long long r = i / K8;
    const int c0 = (int)(i % K8) * 8;
    float ds[8];
    drop_scale8(key, row0 + r, c0, rate, inv_keep, ds);
    st8(out + r * W + c0, ds);
  }
}
}  // namespace bfk
extern "C" int bfk_train_dropout_mask(uint64_t seed, const uint32_t* seed_dev, int32_t layer, int64_t row_offset, int64_t B,
                                      int32_t W, float rate, float* out, void* stream) {
  if (!out || B < 0 || W <= 0 || (W % 8) != 0 || W > 1024 || layer < 0 || !(rate >= 0.f) || !(rate < 1.f))
    return fail(-87, "bfk_train_dropout_mask: bad arguments (W a multiple of 8 <= 1024, 0 <= rate < 1)");
  if (B == 0) return 0;
  cudaStream_t st = (cudaStream_t)stream;
  const long long n = (long long)B * (W / 8);
  const unsigned grid = (unsigned)std::min<long long>((n + 255) / 256, 148 * 8);
  bfk::drop_mask_kernel<<<grid, 256, 0, st>>>(seed, seed_dev, layer, row_offset, B, W, rate, 1.f / (1.f - rate), out);
  CK(cudaGetLastError());
  return 0;
}

// Host-side evaluation of the same double-precision schedule algebra the evaluation tables use (density path of
// DiffusionModel: velocity = coefZ*z + coefP*pred, so trace(dv/dz) = D*coefZ + coefP*trace(dpred/dz)).
extern "C" int bfk_schedule_coefs(const bfk_schedule* sched, double t, double* coefZ, double* coefP, double* net_time) {
  if (!sched || !coefZ || !coefP || !net_time) return fail(-86, "bfk_schedule_coefs: bad arguments");
  bfk::schedule_coefs(*sched, t, coefZ, coefP, net_time);
  return 0;
}

// SDE sampling terms (diffusion_model.py:539-552 with stochastic_solver=True, :561-595): beta = g^2 and the coefficient
// fz of the linear drift f = fz * z (-beta/2 variance preserving, 0 variance exploding).  With the probability-flow
// velocity v_ode = f - beta/2 * score the reverse-SDE drift is f - beta * score = 2 * v_ode - fz * z.
extern "C" int bfk_schedule_sde_terms(const bfk_schedule* sched, double t, double* fz, double* beta) {
  if (!sched || !fz || !beta || sched->model != BFK_MODEL_DM) return fail(-87, "bfk_schedule_sde_terms: bad arguments");
  double lam;
  bfk::schedule_lam_beta(*sched, t, &lam, beta);
  *fz = sched->exploding ? 0.0 : -0.5 * *beta;
  return 0;
}

// alpha_t, sigma_t of the inference-time schedule (noise_schedule.py get_alpha_sigma; edm_noise_schedule.py:61-75,
// cosine_noise_schedule.py:60-63 for lambda(t)) - what the annealed-Langevin updates of utils/integrate.py:924-961 and
// :1167-1251 read per step.
extern "C" int bfk_schedule_alpha_sigma(const bfk_schedule* sched, double t, double* alpha, double* sigma) {
  if (!sched || !alpha || !sigma || sched->model != BFK_MODEL_DM) return fail(-87, "bfk_schedule_alpha_sigma: bad arguments");
  double lam, beta;
  bfk::schedule_lam_beta(*sched, t, &lam, &beta);
  if (sched->exploding) { *alpha = 1.0; *sigma = sqrt(exp(-lam)); }
  else { *alpha = sqrt(1.0 / (1.0 + exp(-lam))); *sigma = sqrt(1.0 / (1.0 + exp(lam))); }
  return 0;
}

extern "C" int bfk_rk_combine(const float* y, const float* const* k, const float* coef, int32_t n, float h, int64_t numel,
                              float* out, void* stream_) {
  if (!y || !k || !coef || !out || n < 0 || n > 8 || numel <= 0) return fail(-85, "bfk_rk_combine: bad arguments");
  RkCombineArgs a{};
  a.n = n;
  for (int j = 0; j < n; ++j) { a.k[j] = k[j]; a.c[j] = coef[j]; }
  rk_combine_kernel<<<cdiv((size_t)numel, 256), 256, 0, (cudaStream_t)stream_>>>(y, a, h, numel, out);
  CK(cudaGetLastError());
  return 0;
}

extern "C" int bfk_fm_interpolate(const float* x0, const float* x1, const float* t, int64_t B, int32_t D, float* x,
                                  float* target, void* stream_) {
  if (!x0 || !x1 || !t || !x || !target || B <= 0 || D <= 0) return fail(-80, "bfk_fm_interpolate: bad arguments");
  fm_interp_kernel<<<cdiv((size_t)B * D, 256), 256, 0, (cudaStream_t)stream_>>>(x0, x1, t, B, D, x, target);
  CK(cudaGetLastError());
  return 0;
}

static int adam_step_impl(float* params, const float* grads, float* m, float* v, const int64_t* seg_offsets, int32_t n_seg,
                          float* seg_sqnorm, float clipnorm, float lr_t, float beta1, float beta2, float eps, float wd_lr,
                          const float* hyper_dev, void* stream_) {
  if (!params || !grads || !m || !v || !seg_offsets || n_seg <= 0 || !seg_sqnorm) return fail(-81, "bfk_adam_step: bad arguments");
  cudaStream_t st = (cudaStream_t)stream_;
  static_assert(sizeof(long long) == sizeof(int64_t), "");
  if (clipnorm > 0.f) {
    CK(cudaMemsetAsync(seg_sqnorm, 0, (size_t)n_seg * 4, st));
    seg_sqnorm_kernel<<<dim3(n_seg, 64), 256, 0, st>>>(grads, (const long long*)seg_offsets, n_seg, seg_sqnorm);
  }
  adam_kernel<<<dim3(n_seg, 256), 256, 0, st>>>(params, grads, m, v, (const long long*)seg_offsets, seg_sqnorm, clipnorm, lr_t, beta1,
                                              beta2, eps, wd_lr, hyper_dev);
  CK(cudaGetLastError());
  return 0;
}
extern "C" int bfk_adam_step(float* params, const float* grads, float* m, float* v, const int64_t* seg_offsets, int32_t n_seg,
                             float* seg_sqnorm, float clipnorm, float lr_t, float beta1, float beta2, float eps, float wd_lr,
                             void* stream_) {
  return adam_step_impl(params, grads, m, v, seg_offsets, n_seg, seg_sqnorm, clipnorm, lr_t, beta1, beta2, eps, wd_lr, nullptr, stream_);
}
extern "C" int bfk_adam_step_dev(float* params, const float* grads, float* m, float* v, const int64_t* seg_offsets, int32_t n_seg,
                                 float* seg_sqnorm, float clipnorm, const float* hyper_dev, float beta1, float beta2, float eps,
                                 void* stream_) {
  if (!hyper_dev) return fail(-82, "bfk_adam_step_dev: null hyper-parameter buffer");
  return adam_step_impl(params, grads, m, v, seg_offsets, n_seg, seg_sqnorm, clipnorm, 0.f, beta1, beta2, eps, 0.f, hyper_dev, stream_);
}

