/*
 * libtransmil_b200 -- C ABI of the B200-native TransMIL bag-level hot path.
 *
 * Every entry point is `extern "C"`, takes plain pointers and sizes (no torch
 * types), launches on the caller's stream and never synchronises the device
 * unless it says so.  All tensors are fp32, row-major, owned by the caller.
 * Return value: 0 on success, a negative tm_status otherwise; the text of the
 * last error on the calling thread is available from tm_last_error().
 *
 * Each function cites the reference interface it replaces; paths are relative
 * to the reference repository root (Ycblue/TransMIL-DeepGraft @ 78f530b).
 *
 * Notation: B bags per call, n patches per bag, D = in_features, C = 512,
 * h = 8 heads, d = 64, m = 256 landmarks, G = ceil(sqrt(n)), N = G*G,
 * dup = N - n, S = N + 1, l = ceil(S/m), Sp = l*m, P = Sp - S (front pad).
 */
#ifndef TRANSMIL_B200_H
#define TRANSMIL_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define TM_DIM 512
#define TM_HEADS 8
#define TM_DHEAD 64
#define TM_LANDMARKS 256
#define TM_PINV_ITERS 6
#define TM_CONV_TAPS 33

typedef enum tm_status {
  TM_OK = 0,
  TM_ERR_BAD_SHAPE = -1,
  TM_ERR_BAD_ALIGNMENT = -2,
  TM_ERR_UNSUPPORTED_ARCH = -3,
  TM_ERR_WORKSPACE_TOO_SMALL = -4,
  TM_ERR_CUDA = -5,
  TM_ERR_BAD_ARGUMENT = -6
} tm_status;

/* fc1 stem variants (code/models/TransMIL.py:100-133). */
typedef enum tm_fc1_kind {
  TM_FC1_SINGLE = 0,  /* Linear(D,512) GELU                      TransMIL.py:130-133, MDMIL.py:66 */
  TM_FC1_2048 = 1,    /* Linear(2048,1024) GELU LN(1024) Linear(1024,512) GELU   TransMIL.py:109-110 */
  TM_FC1_768 = 2,     /* Linear(768,768) GELU LN(768) Linear(768,512) GELU LN(512)  TransMIL.py:123-126 (eval) */
  TM_FC1_TOKENS = 3   /* no stem inside the library: x is [B,n,512] fp32 tokens produced by the caller (the conv stem of
                         CTMIL.py:137-142 / the commented forward of TransformerMIL.py:155-161); in_features = 512, the fc1_*
                         pointers are ignored; duplicate pad + cls (CTMIL.py:144-152) as for the other stems */
} tm_fc1_kind;

/* Flags for tm_transmil_fwd / tm_workspace_bytes. */
#define TM_FLAG_SIMT_ONLY 1u     /* force the fp32 CUDA-core kernels (on-device cross-check of the tensor-core path) */
#define TM_FLAG_KEEP_STAGES 2u   /* keep per-stage buffers addressable through tm_stage_ptr */
#define TM_FLAG_TF32_OPERANDS 4u /* producers round (cvt.rna) the tensors that feed tensor-core contractions to tf32, so the
                                    MMA's operand truncation is exact; set internally by tm_transmil_fwd when the tcgen05
                                    engine runs.  Callers of the per-kernel entry points pass it themselves. */

#define TM_FLAG_F16_ACT 16u       /* per-kernel entry points of the tensor-core engine: the activation that feeds the next fp16
                                    contraction (xn, merged) is read / written as fp16; set internally by tm_transmil_fwd */
#define TM_FLAG_PINV_GEMM_CHAIN 8u /* run the pinv chain as 24 batched tcgen05 GEMM launches (tf32 operands) instead of the
                                    single 4-CTA-cluster kernel (fp16 operands); kept for cross-checking */

#define TM_FLAG_NO_OVERLAP 32u      /* tm_transmil_fwd: keep every stage on the caller's stream (no fork of the pinv chain
                                    onto the library's side stream while attn3.v runs) */
#define TM_FLAG_X_F16 64u           /* tm_transmil_fwd / tm_transmil_fwd_host / tm_fc1_fwd: the bag x is fp16 [B,n,D] (feature files
                                       stored in half precision; the upcast x.float() of TransMIL.py:174 is exact, so is
                                       reading the halves directly).  The first fc1 contraction then runs fp16 x fp16 -> fp32 with an
                                       fp16 copy of its weight (same 10-bit mantissa as the tf32 operands of the fp32 path).
                                       Tensor-core engine only. */

#define TM_FLAG_CLS_ONLY_L2 128u    /* tm_transmil_fwd: of the second TransLayer's output only the cls row is read (TransMIL.py:197-204:
                                       h = self.norm(h)[:, 0]), so its out-projection (nystrom_out, to_out + residual) is run for the
                                       128-row tiles that hold the cls row only; q, k, v, landmarks, pinv and attn3.v are complete.
                                       The cls row, hence the logits, has the same bits as without the flag.  Off by default: the
                                       skipped rows are work the reference does.  Ignored with TM_FLAG_KEEP_STAGES / SIMT_ONLY. */

/* Index contract of one bag (TransMIL.py:177-179,184-186,190-193 and the
 * nystrom-attention 0.0.11 front pad / landmark grouping).  Pure host code. */
typedef struct tm_plan {
  int64_t n, G, N, dup, S, l, Sp, P, padding;
} tm_plan;

/* Parameters of one NystromAttention TransLayer (TransMIL.py:19-57).
 * Shapes are those of the reference state_dict. */
typedef struct tm_layer_weights {
  const float* norm_w;   /* layerX.norm.weight            [512]        */
  const float* norm_b;   /* layerX.norm.bias              [512]        */
  const float* qkv_w;    /* layerX.attn.to_qkv.weight     [1536,512]   */
  const float* out_w;    /* layerX.attn.to_out.0.weight   [512,512]    */
  const float* out_b;    /* layerX.attn.to_out.0.bias     [512]        */
  const float* conv_w;   /* layerX.attn.res_conv.weight   [8,1,33,1]   */
} tm_layer_weights;

/* All parameters of TransMIL (state_dict layout, TransMIL.py:79-155). */
typedef struct tm_weights {
  int32_t fc1_kind;      /* tm_fc1_kind */
  int32_t in_features;   /* D */
  int32_t n_classes;
  int32_t reserved;
  const float* fc1_w0;   /* _fc1.0.weight [F0,D]  (F0 = 512 | 1024 | 768) */
  const float* fc1_b0;   /* _fc1.0.bias   [F0]    */
  const float* fc1_ln_w; /* _fc1.2.* (2048) or _fc1.3.* (768): LayerNorm(F0) ; NULL for SINGLE */
  const float* fc1_ln_b;
  const float* fc1_w1;   /* _fc1.3.weight (2048) or _fc1.4.weight (768)  [512,F0] ; NULL for SINGLE */
  const float* fc1_b1;
  const float* fc1_ln2_w; /* _fc1.7.* (768 only): LayerNorm(512) ; else NULL */
  const float* fc1_ln2_b;
  const float* cls_token; /* cls_token [1,1,512] */
  tm_layer_weights layer1, layer2;
  const float* ppeg_w7;  /* pos_layer.proj.weight  [512,1,7,7] */
  const float* ppeg_b7;  /* pos_layer.proj.bias    [512]       */
  const float* ppeg_w5;  /* pos_layer.proj1.weight [512,1,5,5] */
  const float* ppeg_b5;
  const float* ppeg_w3;  /* pos_layer.proj2.weight [512,1,3,3] */
  const float* ppeg_b3;
  const float* norm_w;   /* norm.weight [512] */
  const float* norm_b;
  const float* fc_w;     /* _fc.weight [n_classes,512] */
  const float* fc_b;     /* _fc.bias   [n_classes]     */
} tm_weights;

/* ---- library ---------------------------------------------------------- */
const char* tm_version(void);
const char* tm_last_error(void);
/* 1 when the current device is sm_100 (tcgen05/TMA kernels usable). */
int tm_device_supported(void);
/* Number of kernel launches issued by this library in this process (any thread: torch.autograd
 * runs the backward on its own) since the last tm_reset_launch_count(). */
int64_t tm_launch_count(void);
void tm_reset_launch_count(void);

/* Per-stage CUDA-event profiler: while enabled, tm_transmil_fwd records an event on the
 * launch stream at every stage boundary; tm_profile_read synchronises them and returns,
 * per stage, the accumulated milliseconds and the number of intervals since the last read. */
int tm_profile_enable(int on);
int tm_profile_stage_count(void);
const char* tm_profile_stage_name(int i);
int tm_profile_read(double* ms_per_stage, int64_t* counts, int n_stages);

/* ---- index contract (host only) --------------------------------------- */
/* TransMIL.py:177-179 (G, N, dup), :184-186 (S), :190-193 (padding);
 * nystrom-attention forward (l, Sp, P). */
int tm_make_plan(int64_t n, tm_plan* out);
/* Source row of token t (0 <= t < N) after the duplicate pad: t < n ? t : t - n.  TransMIL.py:180 */
int64_t tm_dup_source(const tm_plan* p, int64_t t);
/* Landmark group of padded row r (0 <= r < Sp): r / l.  nystrom-attention reduce '(n l) d -> n d'. */
int64_t tm_landmark_group(const tm_plan* p, int64_t r);
/* Padded row that holds sequence position s (0 = cls, 1.. = tokens): P + s. */
int64_t tm_padded_row(const tm_plan* p, int64_t s);
/* PPEG pixel of token t (TransMIL.py:69): y = t / G, x = t % G. */
void tm_ppeg_pixel(const tm_plan* p, int64_t t, int64_t* y, int64_t* x);

/* ---- whole path -------------------------------------------------------- */
/* Bytes of device workspace tm_transmil_fwd needs for B bags of n patches. */
int tm_workspace_bytes(int64_t B, int64_t n, const tm_weights* w, uint32_t flags, size_t* bytes);

/* TransMIL.forward (TransMIL.py:167-211), eval mode, return_attn=False.
 * x [B,n,D] fp32 (device), logits [B,n_classes] fp32 (device).
 * The dense [B,8,Sp,Sp] attention is never built. */
int tm_transmil_fwd(const tm_weights* w, const float* x, int64_t B, int64_t n,
                    float* logits, void* workspace, size_t workspace_bytes,
                    uint32_t flags, void* stream);

/* Same call with HOST buffers: x_host [B,n,D] (pinned or pageable), logits_host
 * [B,n_classes].  Copies host->device, runs tm_transmil_fwd, copies the logits
 * back and synchronises `stream`.  d_x is a caller-owned device staging buffer
 * of B*n*D floats, d_logits one of B*n_classes floats.  This is the call
 * ModelInterface.step makes once per slide (code/models/model_interface.py:316-322). */
int tm_transmil_fwd_host(const tm_weights* w, const float* x_host, int64_t B, int64_t n,
                         float* logits_host, float* d_x, float* d_logits,
                         void* workspace, size_t workspace_bytes, uint32_t flags, void* stream);

/* Device pointer of a named stage buffer inside `workspace` after a
 * tm_transmil_fwd with TM_FLAG_KEEP_STAGES (for parity tests):
 * "h0" [B,S,512] fc1+pad+cls, "xn" [B,Sp,512] (layer 2's at the end),
 * "q","k","v" [B,8,Sp,64], "q_lm","k_lm" [B,8,256,64], "attn2","Z" [B,8,256,256],
 * "W3","M" [B,8,256,64], "merged" [B,Sp,512], "h1","h2","h3" [B,S,512]
 * (stages of the LAST layer executed unless suffixed "1", e.g. "Z1", "q_lm1"). */
int tm_stage_ptr(const char* name, int64_t B, int64_t n, const tm_weights* w, uint32_t flags,
                 void* workspace, void** ptr, int64_t* numel);
/* 1 when that stage buffer holds fp16 elements under `flags` on this device ("q","k","v" of the tensor-core engine). */
int tm_stage_is_f16(const char* name, uint32_t flags);

/* ---- per-kernel entry points (all device pointers) -------------------- */
/* K1+K2: _fc1 (TransMIL.py:175) fused with the duplicate pad (:177-180) and the
 * cls token (:184-186).  x [B,n,D] -> h0 [B,S,512].  scratch: B*n*F0 floats
 * (unused for TM_FC1_SINGLE).  With TM_FLAG_X_F16, x AND w->fc1_w0 point at fp16 data
 * (tm_transmil_fwd makes the weight copy itself). */
int tm_fc1_fwd(const tm_weights* w, const float* x, int64_t B, int64_t n, float* h0,
               float* scratch, uint32_t flags, void* stream);

/* K3: TransLayer.norm (TransMIL.py:23,47) + the FRONT zero pad of
 * NystromAttention.forward.  h [B,S,512] -> xn [B,Sp,512] (rows < P zero); xn is fp16 with TM_FLAG_F16_ACT. */
int tm_ln_pad_fwd(const float* h, const float* gamma, const float* beta, int64_t B, int64_t S,
                  int64_t P, float* xn, uint32_t flags, void* stream);

/* K4: to_qkv (no bias), head split 'b n (h d) -> b h n d', q *= 64^-0.5.
 * xn [B,Sp,512] -> q,k,v [B,8,Sp,64]. */
int tm_qkv_fwd(const float* xn, const float* qkv_w, int64_t B, int64_t Sp, float* q, float* k,
               float* v, uint32_t flags, void* stream);

/* K5: landmark means over l consecutive padded rows.  q,k [B,8,Sp,64] -> q_lm,k_lm [B,8,256,64]. */
int tm_landmarks_fwd(const float* q, const float* k, int64_t B, int64_t Sp, int64_t l,
                     float* q_lm, float* k_lm, void* stream);

/* K6+K7: attn2 = softmax(q_lm k_lm^T); Z = moore_penrose_iter_pinv(attn2, 6) with
 * the global (batch x heads) init scalar.  scratch: tm_pinv_scratch_floats(B) floats. */
int64_t tm_pinv_scratch_floats(int64_t B);
int tm_pinv_fwd(const float* q_lm, const float* k_lm, int64_t B, float* attn2, float* Z,
                float* scratch, uint32_t flags, void* stream);

/* K8: W3 = softmax_t(q_lm k^T) v over all Sp padded tokens, flash-style
 * (the [256,Sp] matrix never reaches HBM).  scratch: tm_attn3v_scratch_floats(). */
int64_t tm_attn3v_scratch_floats(int64_t B, int64_t Sp);
int tm_attn3v_fwd(const float* q_lm, const float* k, const float* v, int64_t B, int64_t Sp,
                  float* W3, float* scratch, uint32_t flags, void* stream);

/* M = Z @ W3  [B,8,256,64]  (re-association of (attn1 @ Z) @ W3). */
int tm_zw_fwd(const float* Z, const float* W3, int64_t B, float* M, void* stream);

/* K9+K10+K11: merged[b,t,h*64:(h+1)*64] = softmax_j(q_t . k_lm_j) @ M_h + res_conv(v)[t]
 * (33-tap depthwise conv along the padded token axis, zero outside [0,Sp)).
 * -> merged [B,Sp,512] ('b h n d -> b n (h d)'). */
int tm_nystrom_out_fwd(const float* q, const float* k_lm, const float* M, const float* v,
                       const float* conv_w, int64_t B, int64_t Sp, float* merged,
                       uint32_t flags, void* stream);

/* ---- fp16 tensor-core engine of the attention core (tcgen05 / TMEM / TMA) ----------
 * q16,k16,v16 [B,8,Sp,64], qlm16,klm16 [B,8,256,64], mt16 [B,8,64,256] are IEEE fp16 (device); all
 * contractions accumulate in fp32.  tm_transmil_fwd routes through these unless TM_FLAG_SIMT_ONLY. */
/* K4 with fp16 head-major outputs (saturating round-to-nearest): xn [B,Sp,512] -> q16 (x 64^-0.5), k16, v16.
 * xn and qkv_w are fp32 (tf32 contraction) or, with TM_FLAG_F16_ACT, fp16 (fp16 contraction, fp32 accumulate). */
int tm_qkv16_fwd(const void* xn, const void* qkv_w, int64_t B, int64_t Sp, void* q16, void* k16,
                 void* v16, uint32_t flags, void* stream);
/* The same projection restricted to columns [col0, col0+ncols) of to_qkv (whole q / k / v blocks: multiples of 512).
 * tm_transmil_fwd projects q and k first, forks the pinv chain, and projects v beside it on max_ctas SMs (0 = all).
 * qkv_w is the FULL weight; only the outputs the slice covers are written.
 * lm_part (may be NULL; needs col0 == 0, fp16 operands and l = Sp/256 >= 128): B*(Sp/128)*8*1024 floats that receive the
 * column sums of every stored 128-row q / k box, split at the landmark-group boundary inside the box and by row quarter
 * (layout in csrc/common.cuh GemmParams::lm_part); tm_landmarks_parts_fwd turns them into the landmark means, so q and k
 * are not read a second time ('... (n l) d -> ... n d' sum / l of nystrom-attention). */
int tm_qkv16_cols_fwd(const void* xn, const void* qkv_w, int64_t B, int64_t Sp, void* q16, void* k16,
                      void* v16, int64_t col0, int64_t ncols, int64_t max_ctas, float* lm_part, uint32_t flags,
                      void* stream);
int tm_landmarks_parts_fwd(const float* lm_part, int64_t B, int64_t Sp, int64_t l, float* q_lm, float* k_lm,
                           void* qlm16, void* klm16, void* stream);
/* K5 on fp16 q/k: fp32 landmark means (fp32 accumulation) plus their fp16 copies. */
int tm_landmarks16_fwd(const void* q16, const void* k16, int64_t B, int64_t Sp, int64_t l, float* q_lm,
                       float* k_lm, void* qlm16, void* klm16, void* stream);
/* M = Z @ W3 in fp32 with both layouts written at once: M [B,8,256,64] fp32 and mt16 [B,8,64,256] fp16. */
int tm_zw16_fwd(const float* Z, const float* W3, int64_t B, float* M, void* mt16, void* stream);
/* mt16[b,h,d,j] = half(M[b,h,j,d])  (K-major B operand of attn1 @ M). */
int tm_mt16_fwd(const float* M, void* mt16, int64_t B, void* stream);
/* K9+K10+K11, same contract as tm_nystrom_out_fwd (nystrom-attention 0.0.11 forward: attn1, (attn1 @ Z) @ W3, res_conv). */
int tm_nystrom_out_tc_fwd(const void* q16, const void* klm16, const void* mt16, const void* v16,
                          const float* conv_w, int64_t B, int64_t Sp, void* merged, uint32_t flags,
                          void* stream);   /* merged: fp32 [B,Sp,512], or fp16 with TM_FLAG_F16_ACT */
/* K8, same contract as tm_attn3v_fwd (attn3 @ v, flash style with lazy rescaling). */
int64_t tm_attn3v_tc_scratch_floats(int64_t B, int64_t Sp);
int tm_attn3v_tc_fwd(const void* qlm16, const void* k16, const void* v16, int64_t B, int64_t Sp,
                     float* W3, float* scratch, uint32_t flags, void* stream);
/* tm_attn3v_tc_fwd without the scratch argument, for callers that may not allocate device memory (CUDA-graph capture,
 * tight serving loops, harnesses that forbid cudaMalloc).  NO COUNTERPART IN THE REFERENCE.
 * The chunk partials live in a file-scope static device array of the library (10 002 432 bytes = 148
 * (bag, head, chunk) slots x 256 landmarks x 66 floats, held by every GPU that loads the library until the process exits);
 * the call issues no cudaMalloc / cudaFree / allocating library routine.  Same inputs, output and (measured) bits as
 * tm_attn3v_tc_fwd.  Limit: tm_attn3v_tc_scratch_floats(B, Sp) <= 2 500 608, i.e. B <= 18 on a 148-SM device, any Sp;
 * beyond it the call returns TM_ERR_BAD_SHAPE and tm_last_error() says so.  Every call on a device shares the one array
 * (it keeps the previous call's partials; nothing in it has to start at zero, each call overwrites what it reads), so two
 * calls that overlap on the GPU - different streams, or this entry point beside itself in two host threads - RACE:
 * order them on one stream or with events. */
int tm_attn3v_tc_fwd_static(const void* qlm16, const void* k16, const void* v16, int64_t B, int64_t Sp,
                            float* W3, uint32_t flags, void* stream);

/* K12: to_out Linear(512,512)+bias on rows [P,Sp) of merged (the `[:, -S:]` slice),
 * plus the TransLayer residual (TransMIL.py:53).  h_out may alias h_in.  With TM_FLAG_F16_ACT merged and out_w point
 * at fp16 data (tensor-core engine); h_in / h_out / out_b stay fp32. */
int tm_to_out_residual_fwd(const float* merged, const float* out_w, const float* out_b,
                           const float* h_in, int64_t B, int64_t S, int64_t P, float* h_out,
                           uint32_t flags, void* stream);

/* K13: PPEG.forward (TransMIL.py:60-75): x + dw7(x) + dw5(x) + dw3(x) on the GxG
 * token image, cls row passed through.  h_in,h_out [B,S,512], S = G*G+1.
 * packed: 512*50 floats of scratch (effective 7x7 filter + bias). */
int tm_ppeg_fwd(const tm_weights* w, const float* h_in, int64_t B, int64_t G, float* h_out,
                float* packed, void* stream);

/* ---- token-sharded single bag (the reference has no counterpart: it runs one bag on one GPU) ----------
 * Every rank holds a contiguous range [row0, row0+rows) of the PADDED token axis of ONE layer, rows % 256 == 0, and runs
 * the per-kernel entry points above on it with B = 1, Sp = rows (rank 0 passes P, the others P = 0).  These entry points
 * produce / consume what crosses rank boundaries.  Two hosts use them: tm_transmil_fwd_sharded below (the whole per-rank
 * schedule with exchanges over NVLink peer memory, one call per rank) and the Python-issued schedule of
 * transmil-deepgraft_b200/sharded.py with transport="nccl" (collectives through torch.distributed; the cross-check). */
/* K1 without the duplicate pad / cls row: x_rows [rows,D] -> h_rows [rows,512] (TransMIL.py:175 on a run of rows).
 * scratch: rows*F0 floats for the two-Linear stems. */
int tm_fc1_rows_fwd(const tm_weights* w, const float* x_rows, int64_t rows, float* h_rows, float* scratch,
                    uint32_t flags, void* stream);
/* K5, local half: sums [2][B*8][256][64] fp32 = per landmark group, the sum of this rank's rows of q16 / k16 that fall
 * into it (group of padded row r = r / l; zeros for groups with no local row).  allreduce(sum) across ranks, then: */
int tm_landmark_sums16_fwd(const void* q16, const void* k16, int64_t B, int64_t rows, int64_t row0, int64_t l,
                           float* sums, void* stream);
/* K5, finish: q_lm, k_lm = sums / l (fp32) and their fp16 copies (same outputs as tm_landmarks16_fwd). */
int tm_landmark_finish_fwd(const float* sums, int64_t B, int64_t l, float* q_lm, float* k_lm, void* qlm16, void* klm16,
                           void* stream);
/* K8, local half: (max [B*8,256], sum [B*8,256], numerator [B*8,256,64]) of softmax_t(q_lm k^T) v over this rank's rows;
 * partial holds tm_attn3v_partial_floats(B) floats; scratch as tm_attn3v_tc_fwd.  allgather across ranks, then: */
int64_t tm_attn3v_partial_floats(int64_t B);
int tm_attn3v_tc_partial_fwd(const void* qlm16, const void* k16, const void* v16, int64_t B, int64_t rows,
                             float* partial, float* scratch, void* stream);
/* K8, merge: gathered = R partials back to back in rank order -> W3 [B,8,256,64]; fixed visiting order, so every rank
 * obtains the same bits. */
int tm_attn3v_merge_fwd(const float* gathered, int64_t R, int64_t B, float* W3, void* stream);
/* K11 across a shard boundary: adds to the first / last 16 rows of merged16 [B,rows,512] (fp16) the res_conv taps that
 * read the previous rank's last / next rank's first 16 rows of v (v_prev16, v_next16: [B*8,16,64] fp16, NULL at the ends
 * of the whole sequence where the conv's zero padding applies). */
int tm_conv_halo_fix_fwd(const void* v_prev16, const void* v_next16, const float* conv_w, int64_t B, int64_t rows,
                         void* merged16, void* stream);
/* K13 on a token range of one bag: out rows = tokens [tok0,tok1); in_rows = whole image rows [y_in0,y_in1) of the GxG token
 * image (the rank's own rows plus the 3-row halo received from its neighbours; must cover
 * [max(0, tok0/G - 3), min(G, (tok1-1)/G + 4))).  packed as tm_ppeg_fwd.  The cls row is the caller's to copy. */
int tm_ppeg_rows_fwd(const tm_weights* w, const float* in_rows, int64_t y_in0, int64_t y_in1, int64_t G, int64_t tok0,
                     int64_t tok1, float* out, float* packed, void* stream);

/* ---- token-sharded single bag, whole schedule in one call -----------------------------------------
 * tm_transmil_fwd_sharded is TransMIL.forward (code/models/TransMIL.py:167-211) of ONE bag for a group of `world` GPUs, one
 * call per rank: the per-kernel entry points above plus every exchange between the ranks, all issued on the caller's stream
 * (asynchronous, capturable into a CUDA graph).  The exchanges run over NVLink peer memory: each rank owns a communication
 * region that the other ranks map (CUDA IPC for one process per GPU, device pointers for one thread per GPU); data moves
 * with flag-guarded pull kernels, no collective library on the data path.  Protocol for the host:
 *   tm_shard_comm_create(rank, world, max_n, &comm)         on the rank's device (cudaSetDevice first)
 *   tm_shard_comm_handle(comm, h)                            64 bytes; allgather them between the processes, any transport
 *   tm_shard_comm_attach(comm, handles[world][64])           once; barrier between the ranks before the first forward
 *   per bag:  tm_make_shard_plan(n, world, rank, &plan); tm_transmil_fwd_sharded(w, x_rows, &plan, comm, logits, ws, ...)
 * Every rank must issue the same sequence of forwards (it is a collective); a rank whose peer does not arrive within 4 s
 * records an error code readable with tm_shard_comm_status instead of hanging.  Barrier before tm_shard_comm_destroy. */
typedef struct tm_shard_plan {
  tm_plan bag;             /* the whole bag (tm_make_plan) */
  int64_t world, rank;
  int64_t row0, rows;      /* padded rows [row0, row0+rows), rows % 256 == 0 */
  int64_t P_local;         /* front-pad rows held here (rank 0: bag.P, else 0) */
  int64_t S_local;         /* sequence positions held here (cls = position 0, rank 0) */
  int64_t has_cls;
  int64_t tok0, ntok;      /* tokens [tok0, tok0+ntok) of the squared sequence (token t >= n is the duplicate pad x[t-n]) */
  int64_t n_x_ranges;      /* 1 or 2 */
  int64_t x_range[2][2];   /* rows [a,b) of the bag x this rank projects, concatenated in this order: x_rows [ntok, D] */
  int64_t y_in0, y_in1;    /* image rows of the G x G token image the PPEG of this rank reads (own rows + 3-row halo) */
  int64_t halo_lo, halo_hi;/* tokens received from rank-1 / rank+1 for the PPEG */
  int64_t n_up, n_dn;      /* tokens sent to rank+1 / rank-1 */
} tm_shard_plan;
typedef struct tm_shard_comm tm_shard_comm;
#define TM_SHARD_HANDLE_BYTES 64
#define TM_SHARD_MAX_WORLD 16

int tm_make_shard_plan(int64_t n, int64_t world, int64_t rank, tm_shard_plan* out);
int tm_shard_comm_create(int rank, int world, int64_t max_n, tm_shard_comm** out);
int64_t tm_shard_comm_bytes(const tm_shard_comm* comm);
int tm_shard_comm_handle(const tm_shard_comm* comm, void* handle64);
int tm_shard_comm_attach(tm_shard_comm* comm, const void* handles);
/* one process driving several GPUs: the peers' regions by pointer (tm_shard_comm_base) and their device ordinals */
void* tm_shard_comm_base(const tm_shard_comm* comm);
int tm_shard_comm_attach_ptrs(tm_shard_comm* comm, void* const* bases, const int* devices);
/* synchronises the device; *code = 0 or 1 + the index of the exchange that timed out since the last call */
int tm_shard_comm_status(tm_shard_comm* comm, int32_t* code);
int tm_shard_comm_destroy(tm_shard_comm* comm);
int tm_sharded_workspace_bytes(const tm_weights* w, const tm_shard_plan* plan, size_t* bytes);
/* x_rows: [plan->ntok, D] fp32 on the device (plan->x_range of the bag, concatenated); logits: [n_classes] on the device,
 * written on every rank.  flags: 0 or TM_FLAG_NO_OVERLAP. */
int tm_transmil_fwd_sharded(const tm_weights* w, const float* x_rows, const tm_shard_plan* plan, tm_shard_comm* comm,
                            float* logits, void* workspace, size_t workspace_bytes, uint32_t flags, void* stream);

/* K14: final LayerNorm on the cls row + _fc (TransMIL.py:202-204).  h [B,S,512] -> logits [B,n_classes]. */
int tm_head_fwd(const tm_weights* w, const float* h, int64_t B, int64_t S, float* logits,
                void* stream);

/* K15 (return_attn rows): rows of attn1 @ Z @ attn3 for one layer without the
 * dense matrix: out[b,h,i,:] = attn1[b,h,rows[i],:] @ Z @ attn3  -> [B,8,n_rows,Sp].
 * Consumers read attn[0,:,padding+1,padding+1:...] (code/visualize_mil.py:581). */
int tm_attn_rows_fwd(const float* q, const float* k, const float* q_lm, const float* k_lm,
                     const float* Z, int64_t B, int64_t Sp, const int64_t* rows_host,
                     int64_t n_rows, float* out, float* scratch, void* stream);
int64_t tm_attn_rows_scratch_floats(int64_t B, int64_t Sp, int64_t n_rows);

/* dst[i] = half(src[i]) (round to nearest, saturating): fp16 operand copies of the GEMM weights. */
int tm_cvt_f16(const float* src, void* dst, int64_t n, void* stream);

/* dst[i] = float(src[i]) for IEEE fp16 src (exact): the backward reads the tensor-core forward's fp16 stages (q, k, v, xn, merged)
 * through its fp32 contraction entry points. */
int tm_cvt_f32(const void* src, float* dst, int64_t n, void* stream);

/* dst[i] = tf32-rounded (round to nearest, ties away: cvt.rna.tf32.f32) src[i]; dst may alias src. */
int tm_round_tf32(const float* src, float* dst, int64_t n, void* stream);

/* Generic fp32 GEMM of the library (C[M,N] = A[M,K] W[N,K]^T), exposed for tests. */
int tm_gemm_nt(const float* A, const float* W, float* C, int64_t M, int64_t N, int64_t K,
               uint32_t flags, void* stream);

/* ---- backward (csrc/backward.cu) ------------------------------------------------------------------
 * The reference has no backward code of its own: `training_step` (code/models/model_interface.py:333-383) differentiates
 * TransMIL.forward (code/models/TransMIL.py:167-211) and nystrom-attention 0.0.11 with torch.autograd.  These entry points
 * are the contractions of that backward, in the order oracle/transmil_backward.py writes them out; the host side
 * (transmil-deepgraft_b200/training.py, a torch.autograd.Function) strings them together.  First correct build: fp32
 * CUDA-core kernels, deterministic two-stage reductions, no atomics.  All pointers are device pointers, fp32. */

/* Strided batched GEMM: C[z][m,n] = alpha * sum_k A[z][m*sam + k*sak] * B[z][k*sbk + n*sbn] + beta * L[z][m*ldl + n],
 * z < batch (pointer + z*stride).  Element strides make every transposition a view; L may be NULL, may alias C, and
 * ldl = 0 broadcasts a bias row.  Every dX = dY W, dW = dY^T X and attention contraction of the backward is one call. */
int tm_bgemm(const float* A, const float* B, float* C, const float* L, int64_t M, int64_t N, int64_t K, int64_t sam, int64_t sak,
             int64_t sbk, int64_t sbn, int64_t ldc, int64_t ldl, int64_t batch, int64_t strideA, int64_t strideB, int64_t strideC,
             int64_t strideL, float alpha, float beta, void* stream);
/* tm_bgemm with caller-owned scratch.  On sm_100 the product runs on the tensor cores (csrc/bgemm_tc.cu: tcgen05.mma kind::tf32,
 * fp32 accumulation in TMEM; operand tiles of any transposition are brought to the K-major shared-memory layout by cp.async
 * loader warps).  Weight gradients (small M x N, K = every token of the batch) are split along K into fp32 partials that a
 * second kernel sums in a fixed order; `scratch` holds tm_bgemm_scratch_floats(M, N, K, batch) floats for them (0 = this
 * shape is not split; a NULL / short scratch disables the split). */
int64_t tm_bgemm_scratch_floats(int64_t M, int64_t N, int64_t K, int64_t batch);
int tm_bgemm_ex(const float* A, const float* B, float* C, const float* L, int64_t M, int64_t N, int64_t K, int64_t sam, int64_t sak,
                int64_t sbk, int64_t sbn, int64_t ldc, int64_t ldl, int64_t batch, int64_t strideA, int64_t strideB, int64_t strideC,
                int64_t strideL, float alpha, float beta, float* scratch, int64_t scratch_floats, void* stream);
/* Engine of tm_bgemm / tm_bgemm_ex and of the products inside tm_pinv_bwd: 0 = tcgen05, tf32 operands (default on sm_100;
 * operand tiles by TMA when base and pitches are 16-byte aligned, else by cp.async loader warps), 1 = fp32 CUDA-core kernel
 * (the on-device cross-check of the former), 2 = tcgen05 fed by the loader warps only (cross-check of the TMA feed).
 * Returns the previous value; any other argument only queries.  TM_BGEMM=simt in the environment selects 1 at load time. */
int tm_bgemm_engine(int engine);
/* In-place row softmax (attn1 / attn3 are recomputed by the backward, not kept by the forward). */
int tm_softmax_rows(float* X, int64_t rows, int64_t len, void* stream);
/* dP <- P * (dP - sum(dP * P, -1)): backward of P = softmax(S, -1), in place on dP. */
int tm_softmax_rows_bwd(const float* P, float* dP, int64_t rows, int64_t len, void* stream);
/* LayerNorm(C) over rows (eps 1e-5, affine): the 2048-d stem's LN is recomputed by the backward. */
int tm_ln_rows_fwd(const float* x, const float* gamma, const float* beta, int64_t rows, int64_t C, float* y, void* stream);
/* LayerNorm backward.  x [rows,C]; the dy row of x row r is (r / rows_per_bag) * dy_bag_rows + dy_row_off + r % rows_per_bag
 * (TransLayer.norm: x is [B,S,512] while the gradient arrives on the front-padded axis [B,Sp,512], off = P).
 * dx = (add ? add : 0) + dLN (add = the residual branch's gradient, TransMIL.py:53); g_gamma, g_beta [C]. */
int64_t tm_ln_bwd_scratch_floats(int64_t rows, int64_t C);
int tm_ln_rows_bwd(const float* x, const float* gamma, const float* dy, int64_t rows, int64_t C, int64_t rows_per_bag,
                   int64_t dy_bag_rows, int64_t dy_row_off, const float* add, float* dx, float* g_gamma, float* g_beta,
                   float* scratch, void* stream);
/* out[c] = sum_r X[r*ld + c] (bias and cls_token gradients).  scratch: tm_colsum_scratch_floats(rows, cols). */
int64_t tm_colsum_scratch_floats(int64_t rows, int64_t cols);
int tm_colsum(const float* X, int64_t rows, int64_t cols, int64_t ld, float* out, float* scratch, void* stream);
/* nn.GELU() (erf form) and its derivative; dz may alias z or dy. */
int tm_gelu_fwd(const float* z, float* a, int64_t n, void* stream);
int tm_gelu_bwd(const float* z, const float* dy, float* dz, int64_t n, void* stream);
/* Training-mode Dropout(p) of to_out (nystrom-attention: to_out = Sequential(Linear, Dropout(dropout)); TransMIL.py:33 passes
 * 0.7), applied in place after tm_to_out_residual_fwd: y = h_in + to_out  ->  h_in + mask * to_out / (1-p).  The keep mask is
 * a counter-based function of (seed, element index), regenerated by tm_pad_front_bwd. */
int tm_dropout_resid_fwd(float* y, const float* h_in, int64_t n, uint64_t seed, float p, void* stream);
/* Dropout(0.6) inside the 768-d stem in training mode (TransMIL.py:123-126): y = mask(seed) * x / (1-p) with the same
 * counter-based mask; the backward applies the same call to the gradient.  x may alias y. */
int tm_dropout_fwd(const float* x, float* y, int64_t n, uint64_t seed, float p, void* stream);
/* dy [B,S,C] -> out [B,S+P,C]: P zero rows in front (backward of the `[:, -S:]` slice), times the dropout mask when p > 0. */
int tm_pad_front_bwd(const float* dy, int64_t B, int64_t S, int64_t P, int64_t C, uint64_t seed, float p, float* out, void* stream);
/* 'b n (h d) -> b h n d' on a [B,Sp,512] gradient. */
int tm_split_heads(const float* merged, int64_t B, int64_t Sp, float* out, void* stream);
/* dqkv [B,Sp,1536] = [ (dq + dq_lm[group]/l) * 64^-0.5 | dk + dk_lm[group]/l | dv ]: backward of the landmark means
 * ('... (n l) d -> ... n d' sum / l), of q *= 64^-0.5 and of the head split, in one pass. */
int tm_qkv_grad_merge(const float* dq, const float* dk, const float* dv, const float* dq_lm, const float* dk_lm, int64_t B, int64_t Sp,
                      int64_t l, float* dqkv, void* stream);
/* res_conv backward: dO, v [B,8,Sp,64] -> dv [B,8,Sp,64], g_w [8,1,33,1].  scratch: tm_conv33_bwd_scratch_floats(B, Sp) floats
 * (per-block tap sums, added in a fixed order by a second kernel). */
int64_t tm_conv33_bwd_scratch_floats(int64_t B, int64_t Sp);
int tm_conv33_bwd(const float* dO, const float* v, const float* conv_w, int64_t B, int64_t Sp, float* dv, float* g_w, float* scratch,
                  void* stream);
/* moore_penrose_iter_pinv backward (6 iterations, init scalar = global max row sum x max column sum over every bag and
 * head): A = attn2 [nb,256,256] with nb = B*8, dZ = gradient of Z (clobbered) -> dA.  The iterates are recomputed. */
int64_t tm_pinv_bwd_scratch_floats(int64_t nb);
int tm_pinv_bwd(const float* A, float* dZ, int64_t nb, float* dA, float* scratch, void* stream);
/* PPEG backward (TransMIL.py:60-75): h_in, dy -> dh [B,S,512] and the six parameter gradients. */
int64_t tm_ppeg_bwd_scratch_floats(int64_t B, int64_t G);
int tm_ppeg_bwd(const tm_weights* w, const float* h_in, const float* dy, int64_t B, int64_t G, float* dh, float* g_w7, float* g_b7,
                float* g_w5, float* g_b5, float* g_w3, float* g_b3, float* scratch, void* stream);
/* Head backward (TransMIL.py:202-204): h3 [B,S,512], dlogits [B,n_classes] -> dcls [B,512] (gradient of h3[:,0]; the other
 * rows of dh3 are zero), g_norm_w, g_norm_b [512], g_fc_w [n_classes,512], g_fc_b [n_classes]. */
int tm_head_bwd(const tm_weights* w, const float* h3, const float* dlogits, int64_t B, int64_t S, float* dcls, float* g_norm_w,
                float* g_norm_b, float* g_fc_w, float* g_fc_b, void* stream);
/* Backward of the duplicate pad and the cls concat (TransMIL.py:177-186): dh0 [B,1+n+dup,512] -> dhf [B,n,512]. */
/* Duplicate pad + cls concat (TransMIL.py:177-186; CTMIL.py:144-152) of tokens a stem produced outside tm_fc1_fwd:
 * hf [B,n,512], cls [512] -> h0 [B,1+n+dup,512]. */
int tm_dup_cls_fwd(const float* hf, const float* cls, int64_t B, int64_t n, int64_t dup, float* h0, void* stream);
int tm_dup_cls_bwd(const float* dh0, int64_t B, int64_t n, int64_t dup, float* dhf, void* stream);

/* ---- optimizer step ----------------------------------------------------------------------------------
 * One launch over the whole parameter set (29 tensors, 4.77 M floats) instead of one small kernel per tensor and operation. */
#define TM_OPT_MAX_TENSORS 48
typedef struct tm_opt_tensor {
  float* p;            /* parameter (fast weights) */
  const float* g;      /* gradient (unused by tm_lookahead_sync) */
  float* m;            /* exp_avg (tm_radam_step) / slow buffer (tm_lookahead_sync) */
  float* v;            /* exp_avg_sq (unused by tm_lookahead_sync) */
  int64_t n;           /* elements */
  float weight_decay;  /* L2 coefficient of this tensor's param group (0 for biases and 1-D tensors: add_weight_decay,
                          code/MyOptimizer/optim_factory.py:25-37) */
  float reserved;
} tm_opt_tensor;
/* torch.optim.RAdam step, the optimizer the reference's factory builds for opt 'radam' (code/MyOptimizer/optim_factory.py:78-80;
 * DeepGraft/TransMIL_feat_norm_rest.yaml:49-55: lr 1e-4, wd 0.01).  `tensors` is a HOST array; step counts from 1. */
int tm_radam_step(const tm_opt_tensor* tensors, int64_t count, int64_t step, float lr, float beta1, float beta2, float eps, void* stream);
/* Lookahead wrapper's slow-weight update ('lookahead_radam'; code/MyOptimizer/lookahead.py:28-41):
 * slow += alpha (fast - slow); fast = slow, with tensors[i].p = fast and tensors[i].m = slow. */
int tm_lookahead_sync(const tm_opt_tensor* tensors, int64_t count, float alpha, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* TRANSMIL_B200_H */
