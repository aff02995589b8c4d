// Host orchestration of the TransMIL bag-level forward (TransMIL.py:167-211):
// workspace layout, the GEMM-backed entry points (fc1, to_qkv, to_out) and the
// whole-path calls.  Everything is launched on the caller's stream; the only
// synchronising call is tm_transmil_fwd_host.
#include "common.cuh"

extern "C" int64_t tm_pinv_scratch_floats(int64_t B);

namespace tmil {

int nystrom_out_tc_tiles(const void* q16, const void* klm16, const void* mt16, const void* v16, const float* conv_w, int64_t B, int64_t Sp,
                         int64_t row_tiles, void* merged, uint32_t flags, cudaStream_t stream);
int attn3v_tc_fwd_budget(const void* qlm16, const void* k16, const void* v16, int64_t B, int64_t Sp, float* W3, float* scratch,
                         int reserved_sms, cudaStream_t st);
int prep_operands(const float* const* srcs, void* const* dsts, const long long* counts, const int* kinds, int njobs,
                  const tm_weights* w, float* packed, cudaStream_t st);
int ppeg_fwd_prepacked(const tm_weights* w, const float* h_in, int64_t B, int64_t G, float* h_out, float* packed, cudaStream_t st);
int ln_inplace(float* x, const float* gamma, const float* beta, long long rows, int width, int rpb, long long bag_rows,
               int off, bool round, cudaStream_t st);

static bool fused_lm_enabled() {
  static const bool on = [] { const char* e = getenv("TM_FUSED_LM"); return e && e[0] == '1'; }();
  return on;
}
static bool use_tc(uint32_t flags) { return !(flags & TM_FLAG_SIMT_ONLY) && tc_available(); }

static int gemm_auto(const GemmParams& p, uint32_t flags, cudaStream_t st) {
  if (use_tc(flags)) return gemm_tc(p, st);
  return gemm_simt(p, st);
}

// Which launches of a forward stay ordinary launches although their kernels support programmatic dependent launch
// (TM_PDL_X overrides; measured on B200 at 50k tokens, profiles/r1_pdl_ab.txt): bit 0 attn3.v behind the fork event,
// 1 the pinv chain on the side stream, 2 attn3.v + its combine, 3 landmarks, 4 nystrom_out, 5 ln_pad, 6 qkv, 7 to_out, 8 ppeg.
static int pdl_mask() {
  static const int m = [] { const char* e = getenv("TM_PDL_X"); return e ? atoi(e) : 61; }();
  return m;
}
static void pdl_gate(int bit) { if (pdl_mask() & bit) pdl_plain_next(); }

__global__ void cls_rows_kernel(const float* __restrict__ cls, long long bag_stride, float* __restrict__ h) {
  // h[b, 0, :] = cls_token  (TransMIL.py:184-186)
  pdl_sync();
  reinterpret_cast<float4*>(h + (long long)blockIdx.x * bag_stride)[threadIdx.x] = __ldg(reinterpret_cast<const float4*>(cls) + threadIdx.x);
}

static int fc1_width(const tm_weights* w) {
  switch (w->fc1_kind) {
    case TM_FC1_SINGLE: return 0;
    case TM_FC1_2048: return 1024;
    case TM_FC1_768: return 768;
    case TM_FC1_TOKENS: return 0;
    default: return -1;
  }
}

struct Layout {
  size_t total;  // floats
  size_t h[4], fc1s, xn, merged, q, k, v, small[2][6], pinv_s, a3_s, packed;
  size_t wr_fc1_0, wr_fc1_1, wr_qkv[2], wr_out[2];   // tf32-rounded copies of the GEMM weights
  size_t qlm16, klm16, mt16;                         // fp16 operands of the tensor-core attention kernels
  size_t lm_part;                                    // per-tile landmark partial sums of the q,k projection (l >= 128 only)
};
enum { SM_QLM = 0, SM_KLM, SM_ATTN2, SM_Z, SM_W3, SM_M };

static size_t bump(size_t& cur, size_t n) {
  size_t at = cur;
  cur += (n + 63) & ~(size_t)63;  // 256-byte granules
  return at;
}

static int make_layout(int64_t B, const tm_plan& p, const tm_weights* w, uint32_t flags, Layout* L) {
  const int f0 = fc1_width(w);
  TM_CHECK_ARG(f0 >= 0, TM_ERR_BAD_ARGUMENT, "unknown fc1_kind %d", w->fc1_kind);
  size_t cur = 0;
  const size_t hs = (size_t)B * p.S * TM_DIM;
  L->h[0] = bump(cur, hs);
  L->h[1] = bump(cur, hs);
  L->h[2] = bump(cur, hs);
  L->h[3] = (flags & TM_FLAG_KEEP_STAGES) ? bump(cur, hs) : L->h[0];
  L->fc1s = bump(cur, (size_t)B * p.n * f0);
  const size_t ps = (size_t)B * p.Sp * TM_DIM;
  L->xn = bump(cur, ps);
  L->merged = bump(cur, ps);
  L->q = bump(cur, ps);
  L->k = bump(cur, ps);
  L->v = bump(cur, ps);
  const size_t lm = (size_t)B * TM_HEADS * TM_LANDMARKS * TM_DHEAD, mm = (size_t)B * TM_HEADS * 65536;
  for (int l = 0; l < 2; ++l) {
    L->small[l][SM_QLM] = bump(cur, lm);
    L->small[l][SM_KLM] = bump(cur, lm);
    L->small[l][SM_ATTN2] = bump(cur, mm);
    L->small[l][SM_Z] = bump(cur, mm);
    L->small[l][SM_W3] = bump(cur, lm);
    L->small[l][SM_M] = bump(cur, lm);
  }
  L->pinv_s = bump(cur, (size_t)tm_pinv_scratch_floats(B));
  {
    const size_t a = (size_t)tm_attn3v_scratch_floats(B, p.Sp), t = (size_t)tm_attn3v_tc_scratch_floats(B, p.Sp);
    L->a3_s = bump(cur, a > t ? a : t);
  }
  L->qlm16 = bump(cur, lm / 2);
  L->klm16 = bump(cur, lm / 2);
  L->mt16 = bump(cur, lm / 2);
  L->packed = bump(cur, 50 * TM_DIM);
  L->lm_part = bump(cur, (fused_lm_enabled() && p.l >= 128) ? (size_t)B * (p.Sp / 128) * 8 * (2 * TM_DIM) : 0);
  L->wr_fc1_0 = bump(cur, w->fc1_kind == TM_FC1_TOKENS ? 0 : (size_t)(f0 ? f0 : TM_DIM) * w->in_features);
  L->wr_fc1_1 = bump(cur, (size_t)TM_DIM * f0);
  for (int l = 0; l < 2; ++l) {
    L->wr_qkv[l] = bump(cur, (size_t)3 * TM_DIM * TM_DIM);
    L->wr_out[l] = bump(cur, (size_t)TM_DIM * TM_DIM);
  }
  L->total = cur;
  return TM_OK;
}

static int check_weights(const tm_weights* w) {
  TM_CHECK_ARG(w != nullptr, TM_ERR_BAD_ARGUMENT, "weights struct is NULL");
  TM_CHECK_ARG(w->in_features > 0 && w->in_features % 4 == 0, TM_ERR_BAD_SHAPE, "in_features must be a positive multiple of 4 (got %d)", w->in_features);
  TM_CHECK_ARG(w->n_classes > 0, TM_ERR_BAD_SHAPE, "n_classes must be positive");
  TM_CHECK_ARG(w->cls_token && w->norm_w && w->norm_b && w->fc_w && w->fc_b, TM_ERR_BAD_ARGUMENT, "missing parameter pointer");
  if (w->fc1_kind == TM_FC1_TOKENS)
    TM_CHECK_ARG(w->in_features == TM_DIM, TM_ERR_BAD_SHAPE, "TM_FC1_TOKENS: the bag holds 512-d tokens (in_features=%d)", w->in_features);
  else
    TM_CHECK_ARG(w->fc1_w0 && w->fc1_b0, TM_ERR_BAD_ARGUMENT, "missing parameter pointer (_fc1.0)");
  if (w->fc1_kind == TM_FC1_2048)
    TM_CHECK_ARG(w->in_features == 2048 && w->fc1_ln_w && w->fc1_ln_b && w->fc1_w1 && w->fc1_b1, TM_ERR_BAD_ARGUMENT, "TM_FC1_2048 needs in_features=2048 and the LN/second Linear parameters");
  if (w->fc1_kind == TM_FC1_768)
    TM_CHECK_ARG(w->in_features == 768 && w->fc1_ln_w && w->fc1_ln_b && w->fc1_w1 && w->fc1_b1 && w->fc1_ln2_w && w->fc1_ln2_b, TM_ERR_BAD_ARGUMENT, "TM_FC1_768 needs in_features=768 and both LayerNorms");
  return TM_OK;
}

static int layer_fwd(const tm_layer_weights& lw, const tm_plan& p, int64_t B, const float* h_in, float* h_out, float* ws,
                     const Layout& L, int li, uint32_t flags, cudaStream_t st) {
  float* xn = ws + L.xn; float* q = ws + L.q; float* k = ws + L.k; float* v = ws + L.v;
  float* q_lm = ws + L.small[li][SM_QLM]; float* k_lm = ws + L.small[li][SM_KLM];
  float* attn2 = ws + L.small[li][SM_ATTN2]; float* Z = ws + L.small[li][SM_Z];
  float* W3 = ws + L.small[li][SM_W3]; float* M = ws + L.small[li][SM_M];
  float* merged = ws + L.merged;
  int64_t cls_tiles = 0;
  prof_mark(ST_LN_PAD, st);
  pdl_gate(32);
  TM_PROPAGATE(tm_ln_pad_fwd(h_in, lw.norm_w, lw.norm_b, B, p.S, p.P, xn, flags, st));
  if (use_tc(flags)) {
    // tensor-core engine: q, k, v live as fp16 (head-major) in the q/k/v regions; all contractions accumulate in fp32
    void* q16 = q; void* k16 = k; void* v16 = v;
    void* qlm16 = ws + L.qlm16; void* klm16 = ws + L.klm16; void* mt16 = ws + L.mt16;
    // sim2 + pinv (32 CTAs per bag, a latency-bound chain of 24 dependent products) and attn3.v (every other SM) are
    // independent: fork the pinv onto the library's high-priority side stream and join before Z.W3.  Serial when the
    // stage profiler is on (its events live on `st`), when asked (TM_FLAG_NO_OVERLAP), or when the pinv clusters alone
    // would take more than a third of the SMs (B > 1).
    const int pinv_sms = 4 * (int)(B * TM_HEADS);
    const bool overlap = !(flags & (TM_FLAG_NO_OVERLAP | TM_FLAG_PINV_GEMM_CHAIN)) && !prof_enabled() && pinv_sms * 3 <= 148;
    // With the fork, only q and k are projected before the landmarks; the v third of to_qkv runs after the fork, beside the
    // pinv chain, on the SMs attn3.v will use (the pinv chain is the longer branch, so v comes for free).
    static const bool split_v = [] { const char* e = getenv("TM_SPLIT_V"); return !(e && e[0] == '0'); }();
    const bool v_late = overlap && split_v;
    // Opt-in (TM_FUSED_LM=1), bags of >= 32 768 padded tokens (l >= 128): the projection's epilogue leaves per-tile column
    // sums of the q / k boxes it stores and the landmark means come from those 13 MB instead of a second pass over q and k
    // (103 MB at 50 k).  Measured on B200 at 50 k: landmarks 24.7 -> 10.2 us but qkv 69 -> 87 us (the epilogue is what
    // bounds that GEMM), 0.895 -> 0.908 ms per bag -- so the separate landmarks16 pass stays the default.
    const bool fused_lm = fused_lm_enabled() && p.l >= 128 && (flags & TM_FLAG_F16_ACT);
    float* lm_part = fused_lm ? ws + L.lm_part : nullptr;
    prof_mark(ST_QKV, st);
    pdl_gate(64);
    TM_PROPAGATE(tm_qkv16_cols_fwd(xn, lw.qkv_w, B, p.Sp, q16, k16, v16, 0, v_late ? 2 * TM_DIM : 3 * TM_DIM, 0, lm_part, flags, st));
    prof_mark(ST_LANDMARKS, st);
    pdl_gate(8);
    if (fused_lm) TM_PROPAGATE(tm_landmarks_parts_fwd(lm_part, B, p.Sp, p.l, q_lm, k_lm, qlm16, klm16, st));
    else TM_PROPAGATE(tm_landmarks16_fwd(q16, k16, B, p.Sp, p.l, q_lm, k_lm, qlm16, klm16, st));
    if (overlap) {
      SideStream ss;
      TM_PROPAGATE(side_stream(&ss));
      TM_CUDA(cudaEventRecord(ss.fork, st));
      TM_CUDA(cudaStreamWaitEvent(ss.stream, ss.fork, 0));
      const int pdl_x = pdl_mask();
      pdl_plain_next();   // first kernel behind a cross-stream wait: ordinary launch
      if (pdl_x & 2) pdl_suspend(true);
      const int pinv_rc = tm_pinv_fwd(q_lm, k_lm, B, attn2, Z, ws + L.pinv_s, flags, ss.stream);
      pdl_suspend(false);
      TM_PROPAGATE(pinv_rc);
      TM_CUDA(cudaEventRecord(ss.join, ss.stream));
      if (v_late) {
        int sms = 148, dev = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
        pdl_plain_next();   // an event record sits between landmarks and this launch
        TM_PROPAGATE(tm_qkv16_cols_fwd(xn, lw.qkv_w, B, p.Sp, q16, k16, v16, 2 * TM_DIM, TM_DIM, sms - pinv_sms, nullptr, flags, st));
      }
      if (pdl_x & 1) pdl_plain_next();   // an event record sits between landmarks and this launch
      if (pdl_x & 4) pdl_suspend(true);
      const int a3_rc = attn3v_tc_fwd_budget(qlm16, k16, v16, B, p.Sp, W3, ws + L.a3_s, pinv_sms, st);
      pdl_suspend(false);
      TM_PROPAGATE(a3_rc);
      TM_CUDA(cudaStreamWaitEvent(st, ss.join, 0));
      pdl_plain_next();
    } else {
      prof_mark(ST_PINV, st);
      TM_PROPAGATE(tm_pinv_fwd(q_lm, k_lm, B, attn2, Z, ws + L.pinv_s, flags, st));
      prof_mark(ST_ATTN3V, st);
      TM_PROPAGATE(tm_attn3v_tc_fwd(qlm16, k16, v16, B, p.Sp, W3, ws + L.a3_s, flags, st));
    }
    prof_mark(ST_ZW, st);
    TM_PROPAGATE(tm_zw16_fwd(Z, W3, B, M, mt16, st));
    prof_mark(ST_NYSTROM_OUT, st);
    pdl_gate(16);
    // second layer, only the cls row is read afterwards (TM_FLAG_CLS_ONLY_L2): the tiles up to and including the one that holds
    // padded row P (the cls token); one bag at a time would be B launches of to_out, so larger batches keep the full GEMM
    cls_tiles = (li == 1 && (flags & TM_FLAG_CLS_ONLY_L2) && !(flags & TM_FLAG_KEEP_STAGES)) ? p.P / 128 + 1 : 0;
    TM_PROPAGATE(nystrom_out_tc_tiles(q16, klm16, mt16, v16, lw.conv_w, B, p.Sp, cls_tiles, merged, flags, st));
  } else {
  prof_mark(ST_QKV, st);
  TM_PROPAGATE(tm_qkv_fwd(xn, lw.qkv_w, B, p.Sp, q, k, v, flags, st));
  prof_mark(ST_LANDMARKS, st);
  TM_PROPAGATE(tm_landmarks_fwd(q, k, B, p.Sp, p.l, q_lm, k_lm, st));
  prof_mark(ST_PINV, st);
  TM_PROPAGATE(tm_pinv_fwd(q_lm, k_lm, B, attn2, Z, ws + L.pinv_s, flags, st));
  prof_mark(ST_ATTN3V, st);
  TM_PROPAGATE(tm_attn3v_fwd(q_lm, k, v, B, p.Sp, W3, ws + L.a3_s, flags, st));
  prof_mark(ST_ZW, st);
  TM_PROPAGATE(tm_zw_fwd(Z, W3, B, M, st));
  prof_mark(ST_NYSTROM_OUT, st);
  TM_PROPAGATE(tm_nystrom_out_fwd(q, k_lm, M, v, lw.conv_w, B, p.Sp, merged, flags, st));
  }
  prof_mark(ST_TO_OUT, st);
  pdl_gate(128);
  if (cls_tiles > 0 && B == 1 && cls_tiles * 128 < p.Sp) {
    // rows [0, cls_tiles * 128) of the padded sequence: the same 128-row GEMM tiles as the full call, just fewer of them
    TM_PROPAGATE(tm_to_out_residual_fwd(merged, lw.out_w, lw.out_b, h_in, 1, cls_tiles * 128 - p.P, p.P, h_out, flags, st));
    return TM_OK;
  }
  TM_PROPAGATE(tm_to_out_residual_fwd(merged, lw.out_w, lw.out_b, h_in, B, p.S, p.P, h_out, flags, st));
  return TM_OK;
}

}  // namespace tmil

using namespace tmil;

// pad_cls: write the square duplicate pad and the cls row (whole bags).  Otherwise x is a plain run of `n` rows and h0 the
// matching run of output rows (token-sharded ranks).
static int fc1_run(const tm_weights* w, const float* x, int64_t B, int64_t n, float* h0, float* scratch, uint32_t flags,
                   cudaStream_t st, bool pad_cls) {
  TM_PROPAGATE(check_weights(w));
  tm_plan p;
  TM_PROPAGATE(tm_make_plan(n, &p));
  TM_CHECK_ARG(B > 0 && x && h0, TM_ERR_BAD_ARGUMENT, "tm_fc1_fwd: bad arguments");
  TM_CHECK_ARG(B * n < (1ll << 31) && B * p.Sp < (1ll << 31), TM_ERR_BAD_SHAPE, "tm_fc1_fwd: B*n exceeds 2^31 rows");
  const int D = w->in_features;
  if (w->fc1_kind == TM_FC1_TOKENS) {
    // the stem ran outside the library (conv stem of CTMIL.py:137-142): x already holds [B,n,512] tokens
    TM_CHECK_ARG(!(flags & TM_FLAG_X_F16), TM_ERR_BAD_ARGUMENT, "TM_FC1_TOKENS: fp32 tokens only");
    prof_mark(ST_FC1_TAIL, st);
    if (pad_cls) return tm_dup_cls_fwd(x, w->cls_token, B, n, p.dup, h0, st);
    TM_CUDA(cudaMemcpyAsync(h0, x, (size_t)B * n * TM_DIM * sizeof(float), cudaMemcpyDeviceToDevice, st));
    return TM_OK;
  }
  GemmParams g;
  memset(&g, 0, sizeof(g));
  g.batch = 1; g.epi = EPI_BIAS_GELU;
  const bool x16 = (flags & TM_FLAG_X_F16) != 0;   // x and fc1_w0 are fp16 (the caller passes the fp16 copy of the weight)
  TM_CHECK_ARG(!x16 || use_tc(flags), TM_ERR_BAD_ARGUMENT, "TM_FLAG_X_F16 needs the tensor-core engine");
  const float* a2 = x; int k2 = D; const float* w2 = w->fc1_w0; const float* b2 = w->fc1_b0;
  if (w->fc1_kind != TM_FC1_SINGLE) {
    const int F0 = fc1_width(w);
    TM_CHECK_ARG(scratch != nullptr, TM_ERR_BAD_ARGUMENT, "tm_fc1_fwd: scratch is required for the two-Linear stems");
    g.A = x; g.Bm = w->fc1_w0; g.C = scratch; g.bias = w->fc1_b0;
    g.M = (int)(B * n); g.N = F0; g.K = D; g.lda = D; g.ldb = D; g.ldc = F0;
    g.f16_in = x16 ? 1 : 0;
    prof_mark(ST_FC1_GEMM0, st);
    TM_PROPAGATE(gemm_auto(g, flags, st));
    g.f16_in = 0;
    prof_mark(ST_FC1_LN, st);
    TM_PROPAGATE(ln_inplace(scratch, w->fc1_ln_w, w->fc1_ln_b, B * n, F0, 0, 0, 0, (flags & TM_FLAG_TF32_OPERANDS) != 0, st));
    a2 = scratch; k2 = F0; w2 = w->fc1_w1; b2 = w->fc1_b1;
  }
  g.A = a2; g.Bm = w2; g.C = h0; g.bias = b2;
  g.f16_in = (x16 && w->fc1_kind == TM_FC1_SINGLE) ? 1 : 0;
  g.M = (int)(B * n); g.N = TM_DIM; g.K = k2; g.lda = k2; g.ldb = k2; g.ldc = TM_DIM;
  if (pad_cls) {
    g.rows_per_bag = (int)n; g.c_bag_rows = p.S; g.c_row_off = 1; g.c_skip = 0; g.dup = (int)p.dup;
  }
  prof_mark(ST_FC1_GEMM1, st);
  TM_PROPAGATE(gemm_auto(g, flags, st));
  prof_mark(ST_FC1_TAIL, st);
  if (!pad_cls) {
    if (w->fc1_kind == TM_FC1_768) TM_PROPAGATE(ln_inplace(h0, w->fc1_ln2_w, w->fc1_ln2_b, B * n, TM_DIM, 0, 0, 0, false, st));
    return TM_OK;
  }
  if (w->fc1_kind == TM_FC1_768)  // trailing LayerNorm(512) on every token row (row 0 is the cls token)
    TM_PROPAGATE(ln_inplace(h0, w->fc1_ln2_w, w->fc1_ln2_b, B * p.N, TM_DIM, (int)p.N, p.S, 1, false, st));
  launch_pdl(cls_rows_kernel, dim3((unsigned)B), dim3(TM_DIM / 4), 0, st, (const float*)w->cls_token, p.S * (long long)TM_DIM, h0);
  TM_LAUNCH_CHECK();
  return TM_OK;
}

extern "C" int tm_fc1_fwd(const tm_weights* w, const float* x, int64_t B, int64_t n, float* h0, float* scratch,
                          uint32_t flags, void* stream) {
  return fc1_run(w, x, B, n, h0, scratch, flags, (cudaStream_t)stream, true);
}

extern "C" int tm_fc1_rows_fwd(const tm_weights* w, const float* x_rows, int64_t rows, float* h_rows, float* scratch,
                               uint32_t flags, void* stream) {
  return fc1_run(w, x_rows, 1, rows, h_rows, scratch, flags, (cudaStream_t)stream, false);
}

extern "C" int tm_qkv_fwd(const float* xn, const float* qkv_w, int64_t B, int64_t Sp, float* q, float* k, float* v,
                          uint32_t flags, void* stream) {
  TM_CHECK_ARG(B > 0 && Sp > 0 && Sp % TM_LANDMARKS == 0 && B * Sp < (1ll << 31), TM_ERR_BAD_SHAPE, "tm_qkv_fwd: bad shape B=%lld Sp=%lld", (long long)B, (long long)Sp);
  TM_CHECK_ARG(xn && qkv_w && q && k && v, TM_ERR_BAD_ARGUMENT, "tm_qkv_fwd: NULL pointer");
  GemmParams g;
  memset(&g, 0, sizeof(g));
  g.A = xn; g.Bm = qkv_w; g.C = nullptr;
  g.M = (int)(B * Sp); g.N = 3 * TM_DIM; g.K = TM_DIM; g.lda = TM_DIM; g.ldb = TM_DIM; g.ldc = 3 * TM_DIM;
  g.batch = 1; g.epi = EPI_QKV; g.q = q; g.k = k; g.v = v; g.Sp = (int)Sp;
  g.round_out = (flags & TM_FLAG_TF32_OPERANDS) ? 1 : 0;
  return gemm_auto(g, flags, (cudaStream_t)stream);
}

extern "C" int tm_qkv16_cols_fwd(const void* xn, const void* qkv_w, int64_t B, int64_t Sp, void* q16, void* k16, void* v16,
                                 int64_t col0, int64_t ncols, int64_t max_ctas, float* lm_part, uint32_t flags, void* stream) {
  TM_CHECK_ARG(B > 0 && Sp > 0 && Sp % TM_LANDMARKS == 0 && B * Sp < (1ll << 31), TM_ERR_BAD_SHAPE, "tm_qkv16_fwd: bad shape B=%lld Sp=%lld", (long long)B, (long long)Sp);
  TM_CHECK_ARG(xn && qkv_w && q16 && k16 && v16, TM_ERR_BAD_ARGUMENT, "tm_qkv16_fwd: NULL pointer");
  TM_CHECK_ARG(col0 >= 0 && ncols > 0 && col0 % TM_DIM == 0 && ncols % TM_DIM == 0 && col0 + ncols <= 3 * TM_DIM, TM_ERR_BAD_SHAPE,
               "tm_qkv16_cols_fwd: the column slice must be whole q / k / v blocks (col0=%lld ncols=%lld)", (long long)col0, (long long)ncols);
  const bool f16 = (flags & TM_FLAG_F16_ACT) != 0;   // xn and qkv_w are fp16
  GemmParams g;
  memset(&g, 0, sizeof(g));
  g.A = (const float*)xn;
  g.Bm = (const float*)((const char*)qkv_w + (size_t)col0 * TM_DIM * (f16 ? 2 : 4));
  g.C = nullptr;
  g.M = (int)(B * Sp); g.N = (int)ncols; g.K = TM_DIM; g.lda = TM_DIM; g.ldb = TM_DIM; g.ldc = 3 * TM_DIM;
  g.batch = 1; g.epi = EPI_QKV16; g.q = (float*)q16; g.k = (float*)k16; g.v = (float*)v16; g.Sp = (int)Sp;
  g.col_off = (int)col0; g.max_ctas = (int)max_ctas;
  g.lm_part = lm_part; g.lm_l = (int)(Sp / TM_LANDMARKS);
  TM_CHECK_ARG(lm_part == nullptr || col0 == 0, TM_ERR_BAD_ARGUMENT, "tm_qkv16_cols_fwd: landmark partial sums need the slice to start at q");
  g.f16_in = f16 ? 1 : 0;
  return gemm_auto(g, flags, (cudaStream_t)stream);
}

extern "C" int tm_qkv16_fwd(const void* xn, const void* qkv_w, int64_t B, int64_t Sp, void* q16, void* k16, void* v16,
                            uint32_t flags, void* stream) {
  return tm_qkv16_cols_fwd(xn, qkv_w, B, Sp, q16, k16, v16, 0, 3 * TM_DIM, 0, nullptr, flags, stream);
}

extern "C" int tm_to_out_residual_fwd(const float* merged, const float* out_w, const float* out_b, const float* h_in,
                                      int64_t B, int64_t S, int64_t P, float* h_out, uint32_t flags, void* stream) {
  TM_CHECK_ARG(B > 0 && S > 0 && P >= 0 && B * (S + P) < (1ll << 31), TM_ERR_BAD_SHAPE, "tm_to_out_residual_fwd: bad shape");
  TM_CHECK_ARG(merged && out_w && out_b && h_in && h_out, TM_ERR_BAD_ARGUMENT, "tm_to_out_residual_fwd: NULL pointer");
  GemmParams g;
  memset(&g, 0, sizeof(g));
  g.A = merged; g.Bm = out_w; g.C = h_out; g.bias = out_b; g.L = h_in; g.ldl = TM_DIM;
  // all Sp padded rows go through the GEMM (identity A rows, TMA friendly); the epilogue drops the P front-pad rows
  g.M = (int)(B * (S + P)); g.N = TM_DIM; g.K = TM_DIM; g.lda = TM_DIM; g.ldb = TM_DIM; g.ldc = TM_DIM;
  g.batch = 1; g.epi = EPI_BIAS_RESID;
  g.f16_in = (flags & TM_FLAG_F16_ACT) ? 1 : 0;   // merged and out_w are fp16
  g.rows_per_bag = (int)(S + P); g.c_bag_rows = S; g.c_row_off = 0; g.c_skip = (int)P;
  return gemm_auto(g, flags, (cudaStream_t)stream);
}

extern "C" int tm_workspace_bytes(int64_t B, int64_t n, const tm_weights* w, uint32_t flags, size_t* bytes) {
  TM_CHECK_ARG(bytes != nullptr && B > 0, TM_ERR_BAD_ARGUMENT, "tm_workspace_bytes: bad arguments");
  TM_PROPAGATE(check_weights(w));
  tm_plan p;
  TM_PROPAGATE(tm_make_plan(n, &p));
  Layout L;
  TM_PROPAGATE(make_layout(B, p, w, flags, &L));
  *bytes = L.total * sizeof(float);
  return TM_OK;
}

extern "C" int tm_transmil_fwd(const tm_weights* w, const float* x, int64_t B, int64_t n, float* logits, void* workspace,
                               size_t workspace_bytes, uint32_t flags, void* stream) {
  TM_PROPAGATE(check_weights(w));
  TM_CHECK_ARG(B > 0 && x && logits && workspace, TM_ERR_BAD_ARGUMENT, "tm_transmil_fwd: bad arguments");
  TM_CHECK_ARG(aligned16(x) && ((uintptr_t)workspace & 255u) == 0, TM_ERR_BAD_ALIGNMENT, "tm_transmil_fwd: x must be 16-byte and workspace 256-byte aligned");
  tm_plan p;
  TM_PROPAGATE(tm_make_plan(n, &p));
  Layout L;
  TM_PROPAGATE(make_layout(B, p, w, flags, &L));
  TM_CHECK_ARG(workspace_bytes >= L.total * sizeof(float), TM_ERR_WORKSPACE_TOO_SMALL, "tm_transmil_fwd: workspace too small (%zu < %zu bytes)", workspace_bytes, L.total * sizeof(float));
  cudaStream_t st = (cudaStream_t)stream;
  float* ws = (float*)workspace;
  tm_weights wl = *w;
  bool ppeg_packed = false;
  if (use_tc(flags)) {
    // tensor-core engine: every tensor that feeds a tf32 contraction is rounded by its producer; the (static) GEMM
    // weights are rounded here into the workspace.  The user's bag x is read as is (operand truncation by the MMA).
    flags |= TM_FLAG_TF32_OPERANDS | TM_FLAG_F16_ACT;
    const int f0 = fc1_width(w);
    // every per-forward copy of the parameters in ONE launch: tf32-rounded fc1 weights, fp16 to_qkv / to_out weights (xn and
    // merged are produced as fp16 by their kernels) and the packed PPEG filter
    const float* srcs[6]; void* dsts[6]; long long cnt[6]; int kinds[6];
    int nj = 0;
    if (w->fc1_kind != TM_FC1_TOKENS) {
      srcs[nj] = w->fc1_w0; dsts[nj] = ws + L.wr_fc1_0; cnt[nj] = (long long)(f0 ? f0 : TM_DIM) * w->in_features;
      kinds[nj++] = (flags & TM_FLAG_X_F16) ? 1 : 0;   // fp16 copy when the bag itself is fp16
      wl.fc1_w0 = ws + L.wr_fc1_0;
    }
    if (f0) {
      srcs[nj] = w->fc1_w1; dsts[nj] = ws + L.wr_fc1_1; cnt[nj] = (long long)TM_DIM * f0; kinds[nj++] = 0;
      wl.fc1_w1 = ws + L.wr_fc1_1;
    }
    tm_layer_weights* lws[2] = {&wl.layer1, &wl.layer2};
    for (int l = 0; l < 2; ++l) {
      srcs[nj] = lws[l]->qkv_w; dsts[nj] = ws + L.wr_qkv[l]; cnt[nj] = (long long)3 * TM_DIM * TM_DIM; kinds[nj++] = 1;
      srcs[nj] = lws[l]->out_w; dsts[nj] = ws + L.wr_out[l]; cnt[nj] = (long long)TM_DIM * TM_DIM; kinds[nj++] = 1;
      lws[l]->qkv_w = ws + L.wr_qkv[l];
      lws[l]->out_w = ws + L.wr_out[l];
    }
    TM_PROPAGATE(prep_operands(srcs, dsts, cnt, kinds, nj, w, ws + L.packed, st));
    ppeg_packed = true;
  }
  TM_PROPAGATE(tm_fc1_fwd(&wl, x, B, n, ws + L.h[0], ws + L.fc1s, flags, st));
  TM_PROPAGATE(layer_fwd(wl.layer1, p, B, ws + L.h[0], ws + L.h[1], ws, L, 0, flags, st));
  prof_mark(ST_PPEG, st);
  pdl_gate(256);
  if (ppeg_packed) TM_PROPAGATE(ppeg_fwd_prepacked(&wl, ws + L.h[1], B, p.G, ws + L.h[2], ws + L.packed, st));
  else TM_PROPAGATE(tm_ppeg_fwd(&wl, ws + L.h[1], B, p.G, ws + L.h[2], ws + L.packed, st));
  TM_PROPAGATE(layer_fwd(wl.layer2, p, B, ws + L.h[2], ws + L.h[3], ws, L, 1, flags, st));
  prof_mark(ST_HEAD, st);
  TM_PROPAGATE(tm_head_fwd(&wl, ws + L.h[3], B, p.S, logits, st));
  prof_mark(ST_END, st);
  return TM_OK;
}

extern "C" int tm_transmil_fwd_host(const tm_weights* w, const float* x_host, int64_t B, int64_t n, float* logits_host,
                                    float* d_x, float* d_logits, void* workspace, size_t workspace_bytes, uint32_t flags,
                                    void* stream) {
  TM_PROPAGATE(check_weights(w));
  TM_CHECK_ARG(x_host && logits_host && d_x && d_logits, TM_ERR_BAD_ARGUMENT, "tm_transmil_fwd_host: NULL pointer");
  cudaStream_t st = (cudaStream_t)stream;
  const size_t esz = (flags & TM_FLAG_X_F16) ? 2 : sizeof(float);
  TM_CUDA(cudaMemcpyAsync(d_x, x_host, (size_t)B * n * w->in_features * esz, cudaMemcpyHostToDevice, st));
  TM_PROPAGATE(tm_transmil_fwd(w, d_x, B, n, d_logits, workspace, workspace_bytes, flags, st));
  TM_CUDA(cudaMemcpyAsync(logits_host, d_logits, (size_t)B * w->n_classes * sizeof(float), cudaMemcpyDeviceToHost, st));
  TM_CUDA(cudaStreamSynchronize(st));
  return TM_OK;
}

extern "C" int tm_stage_ptr(const char* name, int64_t B, int64_t n, const tm_weights* w, uint32_t flags, void* workspace,
                            void** ptr, int64_t* numel) {
  TM_CHECK_ARG(name && ptr && numel && workspace, TM_ERR_BAD_ARGUMENT, "tm_stage_ptr: bad arguments");
  TM_PROPAGATE(check_weights(w));
  tm_plan p;
  TM_PROPAGATE(tm_make_plan(n, &p));
  Layout L;
  TM_PROPAGATE(make_layout(B, p, w, flags, &L));
  float* ws = (float*)workspace;
  const int64_t hs = B * p.S * TM_DIM, ps = B * p.Sp * TM_DIM;
  const int64_t lm = B * TM_HEADS * TM_LANDMARKS * TM_DHEAD, mm = B * TM_HEADS * 65536;
  struct { const char* nm; size_t off; int64_t ne; } tab[] = {
      {"h0", L.h[0], hs}, {"h1", L.h[1], hs}, {"h2", L.h[2], hs}, {"h3", L.h[3], hs},
      {"xn", L.xn, ps}, {"merged", L.merged, ps}, {"q", L.q, ps}, {"k", L.k, ps}, {"v", L.v, ps},
      {"q_lm1", L.small[0][SM_QLM], lm}, {"k_lm1", L.small[0][SM_KLM], lm}, {"attn21", L.small[0][SM_ATTN2], mm},
      {"Z1", L.small[0][SM_Z], mm}, {"W31", L.small[0][SM_W3], lm}, {"M1", L.small[0][SM_M], lm},
      {"q_lm", L.small[1][SM_QLM], lm}, {"k_lm", L.small[1][SM_KLM], lm}, {"attn2", L.small[1][SM_ATTN2], mm},
      {"Z", L.small[1][SM_Z], mm}, {"W3", L.small[1][SM_W3], lm}, {"M", L.small[1][SM_M], lm},
  };
  for (auto& t : tab)
    if (strcmp(t.nm, name) == 0) {
      *ptr = ws + t.off;
      *numel = t.ne;
      return TM_OK;
    }
  set_error("tm_stage_ptr: unknown stage '%s'", name);
  return TM_ERR_BAD_ARGUMENT;
}

// 1 when the named stage buffer holds fp16 under `flags` on this device (q, k, v of the tensor-core engine).
extern "C" int tm_stage_is_f16(const char* name, uint32_t flags) {
  if (!name || !use_tc(flags)) return 0;
  return (strcmp(name, "q") == 0 || strcmp(name, "k") == 0 || strcmp(name, "v") == 0 || strcmp(name, "xn") == 0 ||
          strcmp(name, "merged") == 0) ? 1 : 0;
}
