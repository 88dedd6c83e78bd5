// Host side of the layer-wise sampler (bfk_sample_layerwise / bfk_layerwise_workspace_bytes).  Included by bfk_host.cu
// after bfk_train_host.cuh (uses its Arena / RC / nt_of helpers and fill_tableau from bfk_host.cu).
#pragma once
#include "bfk_layerwise.cuh"
#include "bfk_lw_fused.cuh"

namespace {

struct LwBufs {
  int KU, NO, has_merge;
  long long chunk, Rp;
  unsigned char *wx_pk, *wm_pk, *wo_pk, *wd_pk[BFK_MAX_BLOCKS];
  unsigned char *x_img, *a_img, *h_img, *h_img2;
  float *bo_pad, *pre_add, *post_add, *table, *P, *y, *ys, *k[BFK_MAX_STAGES];
  float *stats[2], *gcol[BFK_MAX_BLOCKS], *ccol[BFK_MAX_BLOCKS], *gcol_o, *ccol_o;  // fused block kernels (bfk_lw_fused.cuh)
};

void plan_lw(const bfk_net_config* c, long long chunk, long long n_evals, long long G, Arena& ar, LwBufs& b) {
  const size_t W = c->W;
  b.has_merge = c->C > 0 ? 1 : 0;
  b.KU = r32(c->D);
  b.NO = r64(c->D);
  b.chunk = chunk;
  b.Rp = (chunk + 127) / 128 * 128;
  const size_t Rp = (size_t)b.Rp;
  b.wx_pk = (unsigned char*)ar.take(W * b.KU * 2);
  b.wm_pk = b.has_merge ? (unsigned char*)ar.take(W * W * 2) : nullptr;
  for (int l = 0; l < c->L; ++l) b.wd_pk[l] = (unsigned char*)ar.take(W * W * 2);
  b.wo_pk = (unsigned char*)ar.take((size_t)b.NO * W * 2);
  b.bo_pad = (float*)ar.take((size_t)b.NO * 4);
  b.pre_add = (b.has_merge && c->merge == BFK_MERGE_ADD) ? (float*)ar.take((size_t)G * W * 4) : nullptr;
  b.post_add = b.has_merge ? (float*)ar.take((size_t)G * W * 4) : nullptr;
  b.table = (float*)ar.take((size_t)n_evals * lw_table_stride(*c) * 4);
  b.x_img = (unsigned char*)ar.take(Rp * b.KU * 2);
  b.a_img = b.has_merge ? (unsigned char*)ar.take(Rp * W * 2) : nullptr;
  b.h_img = (unsigned char*)ar.take(Rp * W * 2);
  b.h_img2 = (unsigned char*)ar.take(Rp * W * 2);
  for (int i = 0; i < 2; ++i) b.stats[i] = (float*)ar.take(Rp * 2 * 4 * ((W + 63) / 64));  // one plane per 64-column panel
  for (int l = 0; l < c->L; ++l) { b.gcol[l] = (float*)ar.take(W * 4); b.ccol[l] = (float*)ar.take(W * 4); }
  b.gcol_o = (float*)ar.take((size_t)b.NO * 4);
  b.ccol_o = (float*)ar.take((size_t)b.NO * 4);
  b.P = (float*)ar.take(Rp * W * 4);
  b.y = (float*)ar.take(Rp * c->D * 4);
  b.ys = (float*)ar.take(Rp * c->D * 4);
  for (int s = 0; s < BFK_MAX_STAGES; ++s) b.k[s] = (float*)ar.take(Rp * b.NO * 4);
}

// launch as clusters of two CTAs (bfk_lw_fused.cuh, CL = 2)
template <typename Kern>
int launch_pairs(Kern kernel, unsigned grid, size_t smem, cudaStream_t st, const LwFusedArgs& fa) {
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(grid & ~1u); cfg.blockDim = dim3(LWF_THREADS); cfg.dynamicSmemBytes = smem; cfg.stream = st;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeClusterDimension;
  at[0].val.clusterDim.x = 2; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
  cfg.attrs = at; cfg.numAttrs = 1;
  CK(cudaLaunchKernelEx(&cfg, kernel, fa));
  return 0;
}

}  // namespace

// the layer-wise SAMPLER takes its FiLM shift from the per-evaluation table (one vector per block); the FiLM scale of
// film_use_gamma is only built on the fused sampler (W <= 256) and in training
static int check_lw_cfg(const bfk_net_config* c) {
  if (int r = check_train_cfg(c)) return r;
  if (c->film_use_gamma) return fail(-73, "layer-wise sampler: film_use_gamma is not supported (fused sampler and training are)");
  for (int l = 0; l < c->L; ++l)
    if (c->block_width[l] != 0)
      return fail(-69, "layer-wise sampler: unequal hidden widths are served by the training / density kernels (host-driven sampling)");
  if (c->norm == BFK_NORM_RMS) return fail(-72, "layer-wise sampler: norm must be 'layer' or None (fused sampler and training have 'rms')");
  return 0;
}

extern "C" size_t bfk_layerwise_workspace_bytes(const bfk_net_config* c, int64_t chunk_rows, int64_t n_evals, int64_t G) {
  if (check_lw_cfg(c) || chunk_rows <= 0 || n_evals <= 0) return 0;
  Arena ar{nullptr, 0, 0};
  LwBufs b{};
  plan_lw(c, chunk_rows, n_evals, G > 0 ? G : 1, ar, b);
  return ar.off + 256;
}

extern "C" int bfk_sample_layerwise(const bfk_net_config* c, const bfk_weights* w, const bfk_schedule* sched,
                                    const float* times, const float* step_h, int32_t method, int64_t n_steps,
                                    const float* conditions, int64_t G, int64_t rpg, const float* tmask, const float* x_in,
                                    float* x_out, int64_t B, int32_t velocity_only, int64_t chunk_rows, void* workspace,
                                    size_t workspace_bytes, void* stream_) {
  RC(check_lw_cfg(c));
  if (!w || !sched || !times || !x_in || !x_out || !workspace || B <= 0 || n_steps <= 0 || chunk_rows <= 0)
    return fail(-90, "bfk_sample_layerwise: bad arguments");
  if (!velocity_only && !step_h) return fail(-90, "bfk_sample_layerwise: step sizes missing");
  if (method < BFK_METHOD_EULER || method > BFK_METHOD_TSIT5) return fail(-42, "unknown method");
  if (c->C > 0 && (!conditions || G <= 0)) return fail(-91, "bfk_sample_layerwise: conditional network needs conditions (G, C)");
  if (rpg <= 0) rpg = 1;
  if (c->C > 0 && G * rpg < B) return fail(-92, "bfk_sample_layerwise: conditions rows x rows_per_group < batch");
  cudaStream_t st = (cudaStream_t)stream_;
  SamplerArgs tab{};
  fill_tableau(velocity_only ? BFK_METHOD_EULER : method, tab);
  const int S = velocity_only ? 1 : tab.S;
  const long long n_evals = (long long)n_steps * S;
  if (chunk_rows > B) chunk_rows = (B + 127) / 128 * 128;
  Arena ar{(unsigned char*)workspace, 0, workspace_bytes};
  LwBufs b{};
  plan_lw(c, chunk_rows, n_evals, G > 0 ? G : 1, ar, b);
  if (ar.off > workspace_bytes) return fail(-93, "bfk_sample_layerwise: workspace too small (bfk_layerwise_workspace_bytes)");
  const int W = c->W, D = c->D, L = c->L, act = c->activation, P16 = BFK_PREC_FP16, NTW = nt_of(W), KU = b.KU, NO = b.NO;
  const int stride = lw_table_stride(*c);
  const bool dm = sched->model == BFK_MODEL_DM;

  // ---- once per call: weights -> operand images, per-dataset and per-evaluation constants ---------------------------
  RC(bfk_pack_kmajor_t(w->x_k, W, D, KU, W, NTW, P16, b.wx_pk, st));
  if (b.has_merge) RC(bfk_pack_kmajor_t(w->m_k, W, W, W, W, NTW, P16, b.wm_pk, st));  // x half (concat) / whole (add)
  static int fused = -1;
  if (fused < 0) { const char* e = getenv("BFK_LW_FUSED"); fused = e ? atoi(e) : 1; }
  const bool ln = c->norm == BFK_NORM_LAYER;
  const char* cl_env = getenv("BFK_LW_CLUSTER");
  const bool pairs = (cl_env ? atoi(cl_env) : 1) == 2 && (act == BFK_ACT_MISH || act == BFK_ACT_SILU);
  if (fused) {
    // block l consumes the un-normalised output of block l-1: its LayerNorm is folded into W_l (bfk_lw_fused.cuh)
    for (int l = 0; l < L; ++l)
      lw_fold_pack_kernel<<<cdiv((size_t)W, 32), dim3(32, 32), 0, st>>>(
          w->dense_k[l], W, W, W, (l > 0 && ln) ? w->norm_g[l - 1] : nullptr, (l > 0 && ln) ? w->norm_b[l - 1] : nullptr,
          w->dense_b[l], NTW, b.wd_pk[l], b.gcol[l], b.ccol[l]);
    // output projector: consumes the last block's un-normalised output with that block's LayerNorm folded in
    lw_fold_pack_kernel<<<cdiv((size_t)NO, 32), dim3(32, 32), 0, st>>>(w->out_k, W, NO, D, ln ? w->norm_g[L - 1] : nullptr,
                                                                       ln ? w->norm_b[L - 1] : nullptr, w->out_b, 64, b.wo_pk,
                                                                       b.gcol_o, b.ccol_o);
    CK(cudaGetLastError());
    static bool attr_dev[64] = {};
    bool& attr = *attr_flag(attr_dev);
    if (!attr) {
      const int sm = (int)lw_fused_smem_bytes(256);
      CK(cudaFuncSetAttribute(lw_block_gemm_kernel<LWF_BLOCK, -1, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, sm));
      CK(cudaFuncSetAttribute(lw_block_gemm_kernel<LWF_BLOCK, BFK_ACT_MISH, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, sm));
      CK(cudaFuncSetAttribute(lw_block_gemm_kernel<LWF_BLOCK, BFK_ACT_SILU, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, sm));
      CK(cudaFuncSetAttribute(lw_block_gemm_kernel<LWF_ACT, -1, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, sm));
      CK(cudaFuncSetAttribute(lw_block_gemm_kernel<LWF_ACT, BFK_ACT_MISH, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, sm));
      CK(cudaFuncSetAttribute(lw_block_gemm_kernel<LWF_ACT, BFK_ACT_SILU, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, sm));
      CK(cudaFuncSetAttribute(lw_block_gemm_kernel<LWF_OUT, -1, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, sm));
      CK(cudaFuncSetAttribute(lw_block_gemm_kernel<LWF_BLOCK, BFK_ACT_MISH, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, sm));
      CK(cudaFuncSetAttribute(lw_block_gemm_kernel<LWF_BLOCK, BFK_ACT_SILU, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, sm));
      CK(cudaFuncSetAttribute(lw_block_gemm_kernel<LWF_ACT, BFK_ACT_MISH, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, sm));
      CK(cudaFuncSetAttribute(lw_block_gemm_kernel<LWF_ACT, BFK_ACT_SILU, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, sm));
      attr = true;
    }
  } else {
    for (int l = 0; l < L; ++l) RC(bfk_pack_kmajor_t(w->dense_k[l], W, W, W, W, NTW, P16, b.wd_pk[l], st));
  }
  if (!fused) RC(bfk_pack_kmajor_t(w->out_k, D, W, W, D, 64, P16, b.wo_pk, st));  // fused: folded pack above
  CK(cudaMemsetAsync(b.bo_pad, 0, (size_t)NO * 4, st));
  CK(cudaMemcpyAsync(b.bo_pad, w->out_b, (size_t)D * 4, cudaMemcpyDeviceToDevice, st));
  if (b.has_merge) {
    const unsigned grid = (unsigned)(G < 148 * 4 ? G : 148 * 4);
    lw_cond_prologue_kernel<<<grid, 256, (size_t)(c->C + W) * 4, st>>>(*c, *w, conditions, G, b.pre_add, b.post_add);
  }
  lw_eval_table_kernel<<<(unsigned)(n_evals < 148 * 8 ? n_evals : 148 * 8), 256, 0, st>>>(*c, *w, *sched, times, n_evals, b.table);
  CK(cudaGetLastError());

  // ---- chunks of rows: each runs its whole trajectory while its activations are L2-resident -----------------------------
  for (long long row0 = 0; row0 < B; row0 += chunk_rows) {
    const long long rows = (B - row0 < chunk_rows) ? B - row0 : chunk_rows;
    const long long n_tiles = (rows + 127) / 128;
    const unsigned tch = (unsigned)(n_tiles < 32 ? n_tiles : 32);
    const unsigned tch_c = (unsigned)(n_tiles < 148 ? n_tiles : 148);  // combine: few octet columns, so more row chunks
    CK(cudaMemcpyAsync(b.y, x_in + row0 * D, (size_t)rows * D * 4, cudaMemcpyDeviceToDevice, st));
    for (long long i = 0; i < n_steps; ++i) {
      for (int s = 0; s < S; ++s) {
        const float* trow = b.table + (size_t)(i * S + s) * stride;
        LwCombArgs ca{};
        ca.y = b.y; ca.ldy = D; ca.nk = s; ca.ldk = NO; ca.h_ptr = step_h ? step_h + i : nullptr;
        for (int j = 0; j < s; ++j) { ca.k[j] = b.k[j]; ca.coef[j] = tab.rk_a[s][j]; }
        if (!step_h) ca.nk = 0;
        ca.out = dm ? b.ys : nullptr; ca.ldo = D; ca.x_out = nullptr;
        ca.img = b.x_img; ca.KU8 = KU / 8; ca.rows = rows; ca.n_tiles = n_tiles; ca.D = D;
        lw_combine_kernel<<<dim3(KU / 8, tch_c), 128, 0, st>>>(ca);
        int sms = bfk_num_sms();
        if (sms <= 0) sms = 148;
        const long long wtiles = n_tiles * (W / NTW);
        const unsigned wgrid = (unsigned)(wtiles < sms ? wtiles : sms);
        if (fused && (KU % P_KC) == 0) {
          // x_proj (+ merge_proj) with their activations in the GEMM epilogue (time_mlp.py:131-140)
          LwFusedArgs fa{};
          fa.A = b.x_img; fa.B = b.wx_pk; fa.ccol = w->x_b; fa.out_img = b.has_merge ? b.a_img : b.h_img;
          fa.rowbias = b.has_merge ? b.pre_add : nullptr; fa.row0 = row0; fa.rpg = rpg;
          fa.M = rows; fa.N = W; fa.K = KU; fa.NT = NTW; fa.act = act; fa.idesc = umma_idesc_f16(G_TM, NTW, 0);
          auto launch_act = [&](const LwFusedArgs& f) -> int {
            const unsigned pgrid = (unsigned)(2 * ((n_tiles + 1) / 2) * (W / NTW) < sms ? 2 * ((n_tiles + 1) / 2) * (W / NTW) : sms);
            if (pairs && pgrid >= 2) {
              if (act == BFK_ACT_MISH) return launch_pairs(lw_block_gemm_kernel<LWF_ACT, BFK_ACT_MISH, 2>, pgrid, lw_fused_smem_bytes(NTW), st, f);
              return launch_pairs(lw_block_gemm_kernel<LWF_ACT, BFK_ACT_SILU, 2>, pgrid, lw_fused_smem_bytes(NTW), st, f);
            }
            if (act == BFK_ACT_MISH) lw_block_gemm_kernel<LWF_ACT, BFK_ACT_MISH, 1><<<wgrid, LWF_THREADS, lw_fused_smem_bytes(NTW), st>>>(f);
            else if (act == BFK_ACT_SILU) lw_block_gemm_kernel<LWF_ACT, BFK_ACT_SILU, 1><<<wgrid, LWF_THREADS, lw_fused_smem_bytes(NTW), st>>>(f);
            else lw_block_gemm_kernel<LWF_ACT, -1, 1><<<wgrid, LWF_THREADS, lw_fused_smem_bytes(NTW), st>>>(f);
            return 0;
          };
          RC(launch_act(fa));
          if (b.has_merge) {
            fa.A = b.a_img; fa.B = b.wm_pk; fa.ccol = nullptr; fa.out_img = b.h_img; fa.rowbias = b.post_add; fa.K = W;
            RC(launch_act(fa));
          }
          CK(cudaGetLastError());
        } else {
          RC(bfk_gemm(b.x_img, b.wx_pk, b.P, W, rows, W, KU, NTW, 1, 0, w->x_b, 1.f, P16, st));
          if (b.has_merge) {
            lw_act_kernel<<<dim3(W / 8, tch), 128, 0, st>>>(b.P, W, rows, n_tiles, W / 8, act, b.pre_add, row0, rpg, b.a_img);
            RC(bfk_gemm(b.a_img, b.wm_pk, b.P, W, rows, W, W, NTW, 1, 0, nullptr, 1.f, P16, st));
            lw_act_kernel<<<dim3(W / 8, tch), 128, 0, st>>>(b.P, W, rows, n_tiles, W / 8, act, b.post_add, row0, rpg, b.h_img);
          } else {
            lw_act_kernel<<<dim3(W / 8, tch), 128, 0, st>>>(b.P, W, rows, n_tiles, W / 8, act, nullptr, row0, rpg, b.h_img);
          }
        }
        const unsigned char* h_last = b.h_img;
        if (fused) {
          unsigned char* img[2] = {b.h_img, b.h_img2};
          for (int l = 0; l < L; ++l) {
            LwFusedArgs fa{};
            fa.A = img[l & 1]; fa.B = b.wd_pk[l]; fa.out_img = img[(l + 1) & 1];
            const bool fold = l > 0 && ln;
            fa.stats_prev = fold ? b.stats[l & 1] : nullptr;
            fa.gprev = fold ? w->norm_g[l - 1] : nullptr; fa.bprev = fold ? w->norm_b[l - 1] : nullptr;
            fa.gcol = b.gcol[l]; fa.ccol = b.ccol[l]; fa.film = trow + (size_t)l * W;
            fa.stats_out = ln ? b.stats[(l + 1) & 1] : nullptr;
            fa.M = rows; fa.N = W; fa.K = W; fa.NT = NTW; fa.act = act; fa.residual = c->residual ? 1 : 0;
            fa.idesc = umma_idesc_f16(G_TM, NTW, 0); fa.rows_pad = n_tiles * 128;
            const unsigned pgrid = (unsigned)(2 * ((n_tiles + 1) / 2) * (W / NTW) < sms ? 2 * ((n_tiles + 1) / 2) * (W / NTW) : sms);
            if (pairs && pgrid >= 2) {
              if (act == BFK_ACT_MISH) RC(launch_pairs(lw_block_gemm_kernel<LWF_BLOCK, BFK_ACT_MISH, 2>, pgrid, lw_fused_smem_bytes(NTW), st, fa));
              else RC(launch_pairs(lw_block_gemm_kernel<LWF_BLOCK, BFK_ACT_SILU, 2>, pgrid, lw_fused_smem_bytes(NTW), st, fa));
            } else
            if (act == BFK_ACT_MISH) lw_block_gemm_kernel<LWF_BLOCK, BFK_ACT_MISH, 1><<<wgrid, LWF_THREADS, lw_fused_smem_bytes(NTW), st>>>(fa);
            else if (act == BFK_ACT_SILU) lw_block_gemm_kernel<LWF_BLOCK, BFK_ACT_SILU, 1><<<wgrid, LWF_THREADS, lw_fused_smem_bytes(NTW), st>>>(fa);
            else lw_block_gemm_kernel<LWF_BLOCK, -1, 1><<<wgrid, LWF_THREADS, lw_fused_smem_bytes(NTW), st>>>(fa);
          }
          CK(cudaGetLastError());
          LwFusedArgs fo{};
          fo.A = img[L & 1]; fo.B = b.wo_pk; fo.stats_prev = ln ? b.stats[L & 1] : nullptr;
          fo.gcol = b.gcol_o; fo.ccol = b.ccol_o; fo.out_f32 = b.k[s]; fo.ldo = NO;
          fo.rows_pad = n_tiles * 128; fo.M = rows; fo.N = NO; fo.K = W; fo.NT = 64; fo.act = act; fo.idesc = umma_idesc_f16(G_TM, 64, 0);
          const long long otiles = n_tiles * (NO / 64);
          lw_block_gemm_kernel<LWF_OUT, -1, 1><<<(unsigned)(otiles < sms ? otiles : sms), LWF_THREADS, lw_fused_smem_bytes(64), st>>>(fo);
          CK(cudaGetLastError());
        } else {
          for (int l = 0; l < L; ++l) {
            RC(bfk_gemm(b.h_img, b.wd_pk[l], b.P, W, rows, W, W, NTW, 1, 0, w->dense_b[l], 1.f, P16, st));
            BlockFwdArgs a{};
            a.P = b.P; a.FB = trow + (size_t)l * W; a.FG = nullptr; a.fb_stride = 0;
            a.prev = nullptr; a.pg = a.pb = nullptr; a.prev_img = c->residual ? b.h_img : nullptr;
            a.gamma = c->norm == BFK_NORM_LAYER ? w->norm_g[l] : nullptr;
            a.beta = c->norm == BFK_NORM_LAYER ? w->norm_b[l] : nullptr;
            a.xhat = nullptr; a.rstd = nullptr; a.img = b.h_img; a.out_f32 = nullptr;
            a.B = rows; a.R_pad = n_tiles * 128; a.W = W; a.act = act; a.norm = c->norm; a.layer = l;
            a.drop_rate = 0.f; a.inv_keep = 1.f; a.seed = 0;
            RC(launch_block_fwd(a, st));
          }
        }
        if (!fused) RC(bfk_gemm(h_last, b.wo_pk, b.k[s], NO, rows, NO, W, 64, 1, 0, b.bo_pad, 1.f, P16, st));
        if (dm || tmask)
          lw_finish_kernel<<<cdiv((size_t)rows * D, 256), 256, 0, st>>>(b.k[s], NO, dm ? b.ys : nullptr, D, trow + (size_t)L * W, tmask, rows, D);
      }
      if (velocity_only) {
        CK(cudaMemcpy2DAsync(x_out + row0 * D, (size_t)D * 4, b.k[0], (size_t)NO * 4, (size_t)D * 4, (size_t)rows, cudaMemcpyDeviceToDevice, st));
        break;
      }
      LwCombArgs ua{};
      ua.y = b.y; ua.ldy = D; ua.nk = S; ua.ldk = NO; ua.h_ptr = step_h + i;
      for (int j = 0; j < S; ++j) { ua.k[j] = b.k[j]; ua.coef[j] = tab.rk_b[j]; }
      ua.out = b.y; ua.ldo = D;
      ua.x_out = (i == n_steps - 1) ? x_out + row0 * D : nullptr; ua.ldx = D;
      ua.img = nullptr; ua.KU8 = KU / 8; ua.rows = rows; ua.n_tiles = n_tiles; ua.D = D;
      lw_combine_kernel<<<dim3((D + 7) / 8, tch_c), 128, 0, st>>>(ua);
    }
    CK(cudaGetLastError());
  }
  return 0;
}
