// K9+K10+K11: merged[b,t,h*64:(h+1)*64] = softmax_j(q_t . k_lm_j) @ M_h + res_conv(v)[t]
// with M = Z @ W3 (re-association of (attn1 @ Z) @ W3), and
// res_conv the 33-tap depthwise conv along the padded token axis (cross-correlation,
// zero outside [0,Sp); one filter per head shared by its 64 channels).
// The [Sp,256] attn1 matrix never reaches HBM.
//
// v1 engine: fp32 CUDA cores, one token per thread with online softmax over the
// 256 landmarks; conv from a shared-memory halo tile; coalesced head-merged store.
#include "common.cuh"
#include "tc_common.cuh"

namespace tmil {

constexpr int NO_TOK = 256;                 // tokens per CTA
constexpr int NO_HALO = TM_CONV_TAPS / 2;   // 16

template <bool ROUND>
__global__ void __launch_bounds__(256, 1) nystrom_out_kernel(const float* __restrict__ q, const float* __restrict__ k_lm,
                                                             const float* __restrict__ M, const float* __restrict__ v,
                                                             const float* __restrict__ conv_w, int Sp,
                                                             float* __restrict__ merged) {
  extern __shared__ __align__(16) float sm[];
  float* ks = sm;                                   // [256][64]  (later: swizzled O tile)
  float* ms = ks + TM_LANDMARKS * TM_DHEAD;         // [256][64]
  float* vs = ms + TM_LANDMARKS * TM_DHEAD;         // [256+32][64]
  __shared__ float wsm[TM_CONV_TAPS];
  const int bh = blockIdx.y, tid = threadIdx.x;
  const int h = bh % TM_HEADS, b = bh / TM_HEADS;
  const int t0 = blockIdx.x * NO_TOK;
  {
    const float4* kg = reinterpret_cast<const float4*>(k_lm + (long long)bh * TM_LANDMARKS * TM_DHEAD);
    const float4* mg = reinterpret_cast<const float4*>(M + (long long)bh * TM_LANDMARKS * TM_DHEAD);
    float4* kd = reinterpret_cast<float4*>(ks);
    float4* md = reinterpret_cast<float4*>(ms);
    for (int i = tid; i < TM_LANDMARKS * 16; i += 256) { kd[i] = __ldg(kg + i); md[i] = __ldg(mg + i); }
    const float* vb = v + (long long)bh * Sp * TM_DHEAD;
    float4* vd = reinterpret_cast<float4*>(vs);
    for (int i = tid; i < (NO_TOK + 2 * NO_HALO) * 16; i += 256) {
      const int t = t0 - NO_HALO + (i >> 4);
      vd[i] = (t >= 0 && t < Sp) ? __ldg(reinterpret_cast<const float4*>(vb + (long long)t * TM_DHEAD) + (i & 15))
                                 : make_float4(0.f, 0.f, 0.f, 0.f);
    }
    if (tid < TM_CONV_TAPS) wsm[tid] = conv_w[h * TM_CONV_TAPS + tid];
  }
  float qr[TM_DHEAD], O[TM_DHEAD];
  {
    const float4* qg = reinterpret_cast<const float4*>(q + ((long long)bh * Sp + t0 + tid) * TM_DHEAD);
#pragma unroll
    for (int i = 0; i < 16; ++i) {
      float4 t = __ldg(qg + i);
      qr[4 * i] = t.x; qr[4 * i + 1] = t.y; qr[4 * i + 2] = t.z; qr[4 * i + 3] = t.w;
    }
#pragma unroll
    for (int i = 0; i < TM_DHEAD; ++i) O[i] = 0.f;
  }
  __syncthreads();
  float mx = -INFINITY, ssum = 0.f;
  for (int j0 = 0; j0 < TM_LANDMARKS; j0 += 32) {
    float s[32];
    float tmax = -INFINITY;
#pragma unroll
    for (int jj = 0; jj < 32; ++jj) {
      float a = 0.f;
      const float* kr = ks + (j0 + jj) * TM_DHEAD;
#pragma unroll
      for (int d = 0; d < TM_DHEAD; d += 4) {
        float4 kv = *reinterpret_cast<const float4*>(kr + d);
        a = fmaf(qr[d], kv.x, a); a = fmaf(qr[d + 1], kv.y, a); a = fmaf(qr[d + 2], kv.z, a); a = fmaf(qr[d + 3], kv.w, a);
      }
      s[jj] = a;
      tmax = fmaxf(tmax, a);
    }
    if (tmax > mx) {
      const float corr = expf(mx - tmax);
      ssum *= corr;
#pragma unroll
      for (int i = 0; i < TM_DHEAD; ++i) O[i] *= corr;
      mx = tmax;
    }
#pragma unroll
    for (int jj = 0; jj < 32; ++jj) {
      const float p = expf(s[jj] - mx);
      ssum += p;
      const float* mr = ms + (j0 + jj) * TM_DHEAD;
#pragma unroll
      for (int d = 0; d < TM_DHEAD; d += 4) {
        float4 mv = *reinterpret_cast<const float4*>(mr + d);
        O[d] = fmaf(p, mv.x, O[d]); O[d + 1] = fmaf(p, mv.y, O[d + 1]);
        O[d + 2] = fmaf(p, mv.z, O[d + 2]); O[d + 3] = fmaf(p, mv.w, O[d + 3]);
      }
    }
  }
  __syncthreads();  // everyone is done reading ks -> reuse it as the O tile
  {
    const float inv = 1.0f / ssum;
    float4* od = reinterpret_cast<float4*>(ks) + tid * 16;
#pragma unroll
    for (int i = 0; i < 16; ++i)
      od[i ^ (tid & 15)] = make_float4(O[4 * i] * inv, O[4 * i + 1] * inv, O[4 * i + 2] * inv, O[4 * i + 3] * inv);
  }
  __syncthreads();
  const int c4 = tid & 15;
  float* outb = merged + ((long long)b * Sp + t0) * TM_DIM + h * TM_DHEAD + c4 * 4;
  for (int tt = tid >> 4; tt < NO_TOK; tt += 16) {
    float4 acc = (reinterpret_cast<const float4*>(ks) + tt * 16)[c4 ^ (tt & 15)];
    const float4* vr = reinterpret_cast<const float4*>(vs) + tt * 16 + c4;  // row tt of the halo tile = token t0+tt-16
#pragma unroll
    for (int tau = 0; tau < TM_CONV_TAPS; ++tau) {
      const float w = wsm[tau];
      const float4 x = vr[tau * 16];
      acc.x = fmaf(w, x.x, acc.x); acc.y = fmaf(w, x.y, acc.y); acc.z = fmaf(w, x.z, acc.z); acc.w = fmaf(w, x.w, acc.w);
    }
    if (ROUND) { acc.x = round_tf32(acc.x); acc.y = round_tf32(acc.y); acc.z = round_tf32(acc.z); acc.w = round_tf32(acc.w); }
    *reinterpret_cast<float4*>(outb + (long long)tt * TM_DIM) = acc;
  }
}

}  // namespace tmil

extern "C" int tm_nystrom_out_fwd(const float* q, const float* k_lm, const float* M, const float* v, const float* conv_w,
                                  int64_t B, int64_t Sp, float* merged, uint32_t flags, void* stream) {
  using namespace tmil;
  TM_CHECK_ARG(B > 0 && Sp > 0 && Sp % TM_LANDMARKS == 0, TM_ERR_BAD_SHAPE, "tm_nystrom_out_fwd: Sp must be a multiple of 256 (Sp=%lld)", (long long)Sp);
  TM_CHECK_ARG(q && k_lm && M && v && conv_w && merged, TM_ERR_BAD_ARGUMENT, "tm_nystrom_out_fwd: NULL pointer");
  const size_t smem = (size_t)(2 * TM_LANDMARKS * TM_DHEAD + (NO_TOK + 2 * NO_HALO) * TM_DHEAD) * sizeof(float);
  static unsigned long long attr_done = 0;
  if (first_use_on_device(&attr_done)) {
    TM_CUDA(cudaFuncSetAttribute(nystrom_out_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    TM_CUDA(cudaFuncSetAttribute(nystrom_out_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  }
  dim3 grid((unsigned)(Sp / NO_TOK), (unsigned)(B * TM_HEADS));
  if (flags & TM_FLAG_TF32_OPERANDS)
    nystrom_out_kernel<true><<<grid, 256, smem, (cudaStream_t)stream>>>(q, k_lm, M, v, conv_w, (int)Sp, merged);
  else
    nystrom_out_kernel<false><<<grid, 256, smem, (cudaStream_t)stream>>>(q, k_lm, M, v, conv_w, (int)Sp, merged);
  TM_LAUNCH_CHECK();
  return TM_OK;
}

// M = Z @ W3  ([256,256] x [256,64] per (b,h)), exact fp32.
extern "C" int tm_zw_fwd(const float* Z, const float* W3, int64_t B, float* M, void* stream) {
  using namespace tmil;
  GemmParams p;
  memset(&p, 0, sizeof(p));
  p.A = Z; p.Bm = W3; p.C = M;
  p.M = TM_LANDMARKS; p.N = TM_DHEAD; p.K = TM_LANDMARKS;
  p.lda = TM_LANDMARKS; p.ldb = TM_DHEAD; p.ldc = TM_DHEAD;
  p.b_is_kn = 1; p.batch = (int)(B * TM_HEADS);
  p.strideA = 65536; p.strideB = TM_LANDMARKS * TM_DHEAD; p.strideC = TM_LANDMARKS * TM_DHEAD;
  p.epi = EPI_AXPBY; p.alpha = 1.f; p.beta = 0.f;
  return gemm_simt(p, (cudaStream_t)stream);
}
