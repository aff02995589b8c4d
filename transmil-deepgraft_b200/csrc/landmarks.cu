// K5: landmark means.  q_lm[b,h,g,:] = (sum_{j<l} q[b,h,g*l+j,:]) / l over the PADDED
// token axis (nystrom-attention 0.0.11: reduce('... (n l) d -> ... n d', 'sum') / l).
// HBM-bound: q and k are each read exactly once, 16 lanes x float4 per 64-wide row.
#include "common.cuh"

namespace tmil {

__global__ void __launch_bounds__(256) landmarks_kernel(const float* __restrict__ q, const float* __restrict__ k,
                                                        long long groups_per_tensor, int l, float* __restrict__ q_lm,
                                                        float* __restrict__ k_lm) {
  // group index over [tensor(2)][B*8*256]; group g of a (b,h) slab starts at row g*l because Sp = 256*l
  const long long gid = (long long)blockIdx.x * 16 + (threadIdx.x >> 4);
  if (gid >= 2 * groups_per_tensor) return;
  const int lane16 = threadIdx.x & 15;
  const bool is_k = gid >= groups_per_tensor;
  const long long g = is_k ? gid - groups_per_tensor : gid;
  const float4* src = reinterpret_cast<const float4*>((is_k ? k : q) + g * (long long)l * TM_DHEAD) + lane16;
  float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
  int j = 0;
  for (; j + 4 <= l; j += 4) {
    float4 a = __ldg(src + (long long)(j + 0) * 16);
    float4 b = __ldg(src + (long long)(j + 1) * 16);
    float4 c = __ldg(src + (long long)(j + 2) * 16);
    float4 d = __ldg(src + (long long)(j + 3) * 16);
    acc.x += a.x; acc.y += a.y; acc.z += a.z; acc.w += a.w;
    acc.x += b.x; acc.y += b.y; acc.z += b.z; acc.w += b.w;
    acc.x += c.x; acc.y += c.y; acc.z += c.z; acc.w += c.w;
    acc.x += d.x; acc.y += d.y; acc.z += d.z; acc.w += d.w;
  }
  for (; j < l; ++j) {
    float4 a = __ldg(src + (long long)j * 16);
    acc.x += a.x; acc.y += a.y; acc.z += a.z; acc.w += a.w;
  }
  const float fl = (float)l;
  acc.x /= fl; acc.y /= fl; acc.z /= fl; acc.w /= fl;
  reinterpret_cast<float4*>((is_k ? k_lm : q_lm) + g * TM_DHEAD)[lane16] = acc;
}

}  // namespace tmil

extern "C" int tm_landmarks_fwd(const float* q, const float* k, int64_t B, int64_t Sp, int64_t l, float* q_lm,
                                float* k_lm, void* stream) {
  TM_CHECK_ARG(B > 0 && l > 0 && Sp == l * TM_LANDMARKS, TM_ERR_BAD_SHAPE,
               "tm_landmarks_fwd: Sp must equal 256*l (Sp=%lld l=%lld)", (long long)Sp, (long long)l);
  TM_CHECK_ARG(tmil::aligned16(q) && tmil::aligned16(k) && tmil::aligned16(q_lm) && tmil::aligned16(k_lm), TM_ERR_BAD_ALIGNMENT,
               "tm_landmarks_fwd: pointers must be 16-byte aligned");
  const long long groups = (long long)B * TM_HEADS * TM_LANDMARKS;
  tmil::landmarks_kernel<<<(unsigned)tmil::ceil_div(2 * groups, 16), 256, 0, (cudaStream_t)stream>>>(q, k, groups, (int)l, q_lm, k_lm);
  TM_LAUNCH_CHECK();
  return TM_OK;
}
