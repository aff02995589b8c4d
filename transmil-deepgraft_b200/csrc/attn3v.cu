// K8: W3 = softmax_t(q_lm k^T) v over ALL Sp padded tokens (zero-score front-pad
// rows included, as in the reference: no mask is passed), flash-style: the
// [256, Sp] score/probability matrix never reaches HBM.  The token axis is split
// into chunks across CTAs; every chunk emits (max, sum, numerator) partials that a
// second kernel merges in a fixed order (deterministic log-sum-exp combine).  The
// same partial triple is what token-sharded ranks exchange.
//
// v1 engine: fp32 CUDA cores, one landmark row per thread.
#include "common.cuh"

namespace tmil {

constexpr int A3_TT = 32;  // tokens per smem tile

__global__ void __launch_bounds__(256, 1) attn3v_partial_kernel(const float* __restrict__ q_lm, const float* __restrict__ k,
                                                                const float* __restrict__ v, int Sp, int tiles_per_chunk,
                                                                int nchunks, float* __restrict__ pm, float* __restrict__ ps,
                                                                float* __restrict__ pO) {
  __shared__ __align__(16) float ks[A3_TT][TM_DHEAD];
  __shared__ __align__(16) float vs[A3_TT][TM_DHEAD];
  const int bh = blockIdx.y, chunk = blockIdx.x, j = threadIdx.x;
  const int ntiles = Sp / A3_TT;
  const int t_begin = chunk * tiles_per_chunk;
  const int t_end = min(ntiles, t_begin + tiles_per_chunk);
  float q[TM_DHEAD], O[TM_DHEAD];
  {
    const float4* qg = reinterpret_cast<const float4*>(q_lm + ((long long)bh * TM_LANDMARKS + j) * TM_DHEAD);
#pragma unroll
    for (int i = 0; i < 16; ++i) {
      float4 t = __ldg(qg + i);
      q[4 * i] = t.x; q[4 * i + 1] = t.y; q[4 * i + 2] = t.z; q[4 * i + 3] = t.w;
    }
#pragma unroll
    for (int i = 0; i < TM_DHEAD; ++i) O[i] = 0.f;
  }
  float m = -INFINITY, ssum = 0.f;
  const float* kb = k + (long long)bh * Sp * TM_DHEAD;
  const float* vb = v + (long long)bh * Sp * TM_DHEAD;
  for (int tile = t_begin; tile < t_end; ++tile) {
    __syncthreads();
    {
      const float4* kg = reinterpret_cast<const float4*>(kb + (long long)tile * A3_TT * TM_DHEAD);
      const float4* vg = reinterpret_cast<const float4*>(vb + (long long)tile * A3_TT * TM_DHEAD);
      float4* kd = reinterpret_cast<float4*>(&ks[0][0]);
      float4* vd = reinterpret_cast<float4*>(&vs[0][0]);
      kd[j] = __ldg(kg + j); kd[j + 256] = __ldg(kg + j + 256);
      vd[j] = __ldg(vg + j); vd[j + 256] = __ldg(vg + j + 256);
    }
    __syncthreads();
    float s[A3_TT];
    float tmax = -INFINITY;
#pragma unroll
    for (int t = 0; t < A3_TT; ++t) {
      float a = 0.f;
#pragma unroll
      for (int d = 0; d < TM_DHEAD; d += 4) {
        float4 kv = *reinterpret_cast<const float4*>(&ks[t][d]);
        a = fmaf(q[d], kv.x, a); a = fmaf(q[d + 1], kv.y, a); a = fmaf(q[d + 2], kv.z, a); a = fmaf(q[d + 3], kv.w, a);
      }
      s[t] = a;
      tmax = fmaxf(tmax, a);
    }
    if (tmax > m) {
      const float corr = expf(m - tmax);  // exp(-inf) = 0 on the first tile
      ssum *= corr;
#pragma unroll
      for (int i = 0; i < TM_DHEAD; ++i) O[i] *= corr;
      m = tmax;
    }
#pragma unroll
    for (int t = 0; t < A3_TT; ++t) {
      const float p = expf(s[t] - m);
      ssum += p;
#pragma unroll
      for (int d = 0; d < TM_DHEAD; d += 4) {
        float4 vv = *reinterpret_cast<const float4*>(&vs[t][d]);
        O[d] = fmaf(p, vv.x, O[d]); O[d + 1] = fmaf(p, vv.y, O[d + 1]);
        O[d + 2] = fmaf(p, vv.z, O[d + 2]); O[d + 3] = fmaf(p, vv.w, O[d + 3]);
      }
    }
  }
  const long long pi = ((long long)bh * nchunks + chunk) * TM_LANDMARKS + j;
  pm[pi] = m;
  ps[pi] = ssum;
  float4* od = reinterpret_cast<float4*>(pO + pi * TM_DHEAD);
#pragma unroll
  for (int i = 0; i < 16; ++i) od[i] = make_float4(O[4 * i], O[4 * i + 1], O[4 * i + 2], O[4 * i + 3]);
}

// W3[bh, j, :] = sum_c e^{m_c - M} O_c / sum_c e^{m_c - M} s_c, chunks visited in index order.
// Element (bh, c, j) of the max / sum arrays sits at bh*bs_ms + c*cs_ms + j, its numerator row at bh*bs_o + c*cs_o + j*64
// (floats): the same kernel merges the chunk partials of one launch and the per-rank partials gathered from token-sharded
// ranks.  With out_m != NULL nothing is normalised: W3 receives the merged numerator and out_m / out_s the merged max / sum
// (the (max, sum, numerator) triple a rank contributes to the cross-rank merge).
__global__ void __launch_bounds__(256) attn3v_combine_kernel(const float* __restrict__ pm, const float* __restrict__ ps,
                                                             const float* __restrict__ pO, int nchunks, long long cs_ms,
                                                             long long bs_ms, long long cs_o, long long bs_o,
                                                             float* __restrict__ W3, float* __restrict__ out_m,
                                                             float* __restrict__ out_s) {
  pdl_sync();
  const int bh = blockIdx.y;
  const int idx = blockIdx.x * 256 + threadIdx.x;  // (j, d4)
  const int j = idx >> 4, d4 = idx & 15;
  const long long base = (long long)bh * bs_ms + j, obase = (long long)bh * bs_o + (long long)j * TM_DHEAD;
  // chunks are visited in index order (fixed summation order), eight at a time so that their loads are all in flight
  float M = -INFINITY;
  for (int c0 = 0; c0 < nchunks; c0 += 8) {
    float m8[8];
#pragma unroll
    for (int u = 0; u < 8; ++u) m8[u] = c0 + u < nchunks ? pm[base + (c0 + u) * cs_ms] : -INFINITY;
#pragma unroll
    for (int u = 0; u < 8; ++u) M = fmaxf(M, m8[u]);
  }
  float den = 0.f;
  float4 num = make_float4(0.f, 0.f, 0.f, 0.f);
  for (int c0 = 0; c0 < nchunks; c0 += 8) {
    float m8[8], s8[8];
    float4 o8[8];
#pragma unroll
    for (int u = 0; u < 8; ++u) {
      const bool ok = c0 + u < nchunks;
      m8[u] = ok ? pm[base + (c0 + u) * cs_ms] : -INFINITY;
      s8[u] = ok ? ps[base + (c0 + u) * cs_ms] : 0.f;
      o8[u] = ok ? reinterpret_cast<const float4*>(pO + obase + (c0 + u) * cs_o)[d4] : make_float4(0.f, 0.f, 0.f, 0.f);
    }
#pragma unroll
    for (int u = 0; u < 8; ++u) {
      if (c0 + u < nchunks) {
        const float w = expf(m8[u] - M);
        den = fmaf(w, s8[u], den);
        num.x = fmaf(w, o8[u].x, num.x); num.y = fmaf(w, o8[u].y, num.y); num.z = fmaf(w, o8[u].z, num.z); num.w = fmaf(w, o8[u].w, num.w);
      }
    }
  }
  float4* dst = reinterpret_cast<float4*>(W3 + ((long long)bh * TM_LANDMARKS + j) * TM_DHEAD) + d4;
  if (out_m != nullptr) {
    *dst = num;
    if (d4 == 0) {
      out_m[(long long)bh * TM_LANDMARKS + j] = M;
      out_s[(long long)bh * TM_LANDMARKS + j] = den;
    }
  } else {
    *dst = make_float4(num.x / den, num.y / den, num.z / den, num.w / den);
  }
}

int attn3v_chunks(int64_t B, int64_t Sp) {
  const int64_t ntiles = Sp / A3_TT;
  int64_t want = (2 * 148 + B * TM_HEADS - 1) / (B * TM_HEADS);  // about two CTAs per SM in total
  if (want < 1) want = 1;
  const int64_t max_chunks = ceil_div(ntiles, 8);               // at least 256 tokens per chunk
  if (want > max_chunks) want = max_chunks;
  return (int)want;
}

}  // namespace tmil

extern "C" int64_t tm_attn3v_scratch_floats(int64_t B, int64_t Sp) {
  const int64_t nchunks = tmil::attn3v_chunks(B, Sp);
  return B * TM_HEADS * nchunks * TM_LANDMARKS * (TM_DHEAD + 2);
}

extern "C" int tm_attn3v_fwd(const float* q_lm, const float* k, const float* v, int64_t B, int64_t Sp, float* W3,
                             float* scratch, uint32_t flags, void* stream) {
  using namespace tmil;
  (void)flags;
  TM_CHECK_ARG(B > 0 && Sp > 0 && Sp % TM_LANDMARKS == 0, TM_ERR_BAD_SHAPE, "tm_attn3v_fwd: Sp must be a multiple of 256 (Sp=%lld)", (long long)Sp);
  TM_CHECK_ARG(q_lm && k && v && W3 && scratch, TM_ERR_BAD_ARGUMENT, "tm_attn3v_fwd: NULL pointer");
  cudaStream_t st = (cudaStream_t)stream;
  const int nbh = (int)(B * TM_HEADS);
  const int nchunks = attn3v_chunks(B, Sp);
  const int ntiles = (int)(Sp / A3_TT);
  const int tpc = (int)ceil_div(ntiles, nchunks);
  const int used = (int)ceil_div(ntiles, tpc);   // chunks that own at least one tile
  float* pm = scratch;
  float* ps = pm + (long long)nbh * nchunks * TM_LANDMARKS;
  float* pO = ps + (long long)nbh * nchunks * TM_LANDMARKS;
  attn3v_partial_kernel<<<dim3(used, nbh), 256, 0, st>>>(q_lm, k, v, (int)Sp, tpc, used, pm, ps, pO);
  TM_LAUNCH_CHECK();
  attn3v_combine_kernel<<<dim3(16, nbh), 256, 0, st>>>(pm, ps, pO, used, TM_LANDMARKS, (long long)used * TM_LANDMARKS, TM_LANDMARKS * TM_DHEAD,
                                                        (long long)used * TM_LANDMARKS * TM_DHEAD, W3, nullptr, nullptr);
  TM_LAUNCH_CHECK();
  return TM_OK;
}
