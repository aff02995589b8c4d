// K5: landmark means.  q_lm[b,h,g,:] = (sum_{j<l} q[b,h,g*l+j,:]) / l over the PADDED
// token axis (nystrom-attention 0.0.11: reduce('... (n l) d -> ... n d', 'sum') / l).
// HBM-bound: q and k are each read exactly once, 16 lanes x float4 per 64-wide row.
#include <cuda_fp16.h>

#include "common.cuh"
#include "tc_common.cuh"

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

// fp16 q/k (tensor-core engine): same means accumulated in fp32; emits the fp32 landmarks (sim2 / pinv run in fp32)
// and fp16 copies (operands of the tensor-core contractions).  One WARP per landmark group so that a 50k-token bag
// (4096 groups of ~200 rows) keeps enough loads in flight to reach HBM speed: lane = (row phase 0..3, 16-byte column
// chunk 0..7); the four phases are reduced with shuffles in a fixed order.
__global__ void __launch_bounds__(256) landmarks16_kernel(const __half* __restrict__ q, const __half* __restrict__ k,
                                                          long long groups_per_tensor, int l, float* __restrict__ q_lm,
                                                          float* __restrict__ k_lm, __half* __restrict__ qlm16,
                                                          __half* __restrict__ klm16) {
  pdl_sync();
  const long long gid = (long long)blockIdx.x * 8 + (threadIdx.x >> 5);
  if (gid >= 2 * groups_per_tensor) return;
  const int lane = threadIdx.x & 31, lane8 = lane & 7, phase = lane >> 3;
  const bool is_k = gid >= groups_per_tensor;
  const long long g = is_k ? gid - groups_per_tensor : gid;
  const uint4* src = reinterpret_cast<const uint4*>((is_k ? k : q) + g * (long long)l * TM_DHEAD) + lane8;
  float acc[8];
#pragma unroll
  for (int e = 0; e < 8; ++e) acc[e] = 0.f;
  auto add = [&](uint4 u) {
    const __half2* h2 = reinterpret_cast<const __half2*>(&u);
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      const float2 f = __half22float2(h2[e]);
      acc[2 * e] += f.x; acc[2 * e + 1] += f.y;
    }
  };
  int j = phase;
  for (; j + 12 < l; j += 16) {
    const uint4 a = __ldg(src + (long long)(j + 0) * 8), b = __ldg(src + (long long)(j + 4) * 8);
    const uint4 c = __ldg(src + (long long)(j + 8) * 8), d = __ldg(src + (long long)(j + 12) * 8);
    add(a); add(b); add(c); add(d);
  }
  for (; j < l; j += 4) add(__ldg(src + (long long)j * 8));
#pragma unroll
  for (int e = 0; e < 8; ++e) {
    acc[e] += __shfl_xor_sync(0xffffffffu, acc[e], 8);
    acc[e] += __shfl_xor_sync(0xffffffffu, acc[e], 16);
  }
  if (phase != 0) return;
  const float fl = (float)l;
#pragma unroll
  for (int e = 0; e < 8; ++e) acc[e] /= fl;
  float4* dst = reinterpret_cast<float4*>((is_k ? k_lm : q_lm) + g * TM_DHEAD) + lane8 * 2;
  dst[0] = make_float4(acc[0], acc[1], acc[2], acc[3]);
  dst[1] = make_float4(acc[4], acc[5], acc[6], acc[7]);
  uint4 o;
  o.x = pack_f16x2(acc[0], acc[1]); o.y = pack_f16x2(acc[2], acc[3]);
  o.z = pack_f16x2(acc[4], acc[5]); o.w = pack_f16x2(acc[6], acc[7]);
  reinterpret_cast<uint4*>((is_k ? klm16 : qlm16) + g * TM_DHEAD)[lane8] = o;
}

// Landmark means from the per-tile partial column sums the q,k projection left behind (GemmParams::lm_part): group j of
// l >= 128 rows meets the tiles floor(j l / 128) .. floor(((j+1) l - 1) / 128); of each tile it takes the "first group" or
// the "next group" half, four row quarters each, always in ascending row order (deterministic).
__global__ void __launch_bounds__(256) landmarks_from_parts_kernel(const float* __restrict__ part, int tiles_per_bag, int l,
                                                                   long long total, float* __restrict__ q_lm, float* __restrict__ k_lm,
                                                                   __half* __restrict__ qlm16, __half* __restrict__ klm16) {
  pdl_sync();
  const long long i = ((long long)blockIdx.x * 256 + threadIdx.x) * 2;   // (b, which, h, j, d pair)
  if (i >= total) return;
  const int d = (int)(i % TM_DHEAD);
  long long r = i / TM_DHEAD;
  const int j = (int)(r % TM_LANDMARKS); r /= TM_LANDMARKS;
  const int h = (int)(r % TM_HEADS); r /= TM_HEADS;
  const int which = (int)(r % 2);
  const long long b = r / 2;
  const int col = which * TM_DIM + h * TM_DHEAD + d;
  const int t0 = (int)(((long long)j * l) / 128), t1 = (int)((((long long)j + 1) * l - 1) / 128);
  float sx = 0.f, sy = 0.f;
  for (int t = t0; t <= t1; ++t) {
    const int g0 = (t * 128) / l;
    const int ab = j == g0 ? 0 : 1;
    const float* p = part + (((b * tiles_per_bag + t) * 4) * 2 + ab) * (long long)(2 * TM_DIM) + col;
#pragma unroll
    for (int rq = 0; rq < 4; ++rq) {
      const float2 v = __ldg(reinterpret_cast<const float2*>(p + (long long)rq * 2 * (2 * TM_DIM)));
      sx += v.x; sy += v.y;
    }
  }
  const float fl = (float)l;
  sx /= fl; sy /= fl;
  const long long o = ((b * TM_HEADS + h) * TM_LANDMARKS + j) * TM_DHEAD + d;
  *reinterpret_cast<float2*>((which ? k_lm : q_lm) + o) = make_float2(sx, sy);
  *reinterpret_cast<uint32_t*>((which ? klm16 : qlm16) + o) = pack_f16x2(sx, sy);
}

}  // namespace tmil

extern "C" int tm_landmarks_parts_fwd(const float* lm_part, int64_t B, int64_t Sp, int64_t l, float* q_lm, float* k_lm,
                                      void* qlm16, void* klm16, void* stream) {
  TM_CHECK_ARG(B > 0 && l >= 128 && Sp == l * TM_LANDMARKS, TM_ERR_BAD_SHAPE,
               "tm_landmarks_parts_fwd: needs Sp = 256*l with l >= 128 (Sp=%lld l=%lld)", (long long)Sp, (long long)l);
  TM_CHECK_ARG(lm_part && q_lm && k_lm && qlm16 && klm16, TM_ERR_BAD_ARGUMENT, "tm_landmarks_parts_fwd: NULL pointer");
  const long long total = (long long)B * 2 * TM_HEADS * TM_LANDMARKS * TM_DHEAD;
  tmil::launch_pdl(tmil::landmarks_from_parts_kernel, dim3((unsigned)tmil::ceil_div(total / 2, 256)), dim3(256), 0, (cudaStream_t)stream,
                   lm_part, (int)(Sp / 128), (int)l, total, q_lm, k_lm, (__half*)qlm16, (__half*)klm16);
  TM_LAUNCH_CHECK();
  return TM_OK;
}

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

extern "C" int tm_landmarks16_fwd(const void* q16, const void* k16, int64_t B, int64_t Sp, int64_t l, float* q_lm, float* k_lm,
                                  void* qlm16, void* klm16, void* stream) {
  TM_CHECK_ARG(B > 0 && l > 0 && Sp == l * TM_LANDMARKS, TM_ERR_BAD_SHAPE,
               "tm_landmarks16_fwd: Sp must equal 256*l (Sp=%lld l=%lld)", (long long)Sp, (long long)l);
  TM_CHECK_ARG(tmil::aligned16(q16) && tmil::aligned16(k16) && tmil::aligned16(q_lm) && tmil::aligned16(k_lm) &&
                   tmil::aligned16(qlm16) && tmil::aligned16(klm16), TM_ERR_BAD_ALIGNMENT,
               "tm_landmarks16_fwd: pointers must be 16-byte aligned");
  const long long groups = (long long)B * TM_HEADS * TM_LANDMARKS;
  tmil::launch_pdl(tmil::landmarks16_kernel, dim3((unsigned)tmil::ceil_div(2 * groups, 8)), dim3(256), 0, (cudaStream_t)stream,
                   (const __half*)q16, (const __half*)k16, groups, (int)l, q_lm, k_lm, (__half*)qlm16, (__half*)klm16);
  TM_LAUNCH_CHECK();
  return TM_OK;
}
