// Kernels that exist only for the token-sharded single-bag path: every rank holds a contiguous range of the
// PADDED token axis (a multiple of 256 rows) and runs the ordinary per-kernel entry points on it; these kernels produce
// and consume what crosses rank boundaries.
//   landmark partial sums   -> allreduce(sum) -> landmark_finish  (a landmark group of l rows may straddle two ranks)
//   conv halo fix           <- the neighbours' 16 boundary rows of v (res_conv, 33 taps, zero outside the WHOLE sequence only)
// The attn3.v partial / merge pair lives in attn3v_tc.cu, the row-range PPEG in ppeg.cu.
#include <cuda_fp16.h>

#include "common.cuh"
#include "tc_common.cuh"

namespace tmil {

// One warp per (tensor, bh, group): sum of the local rows that fall into global group g = (row0 + r) / l.
// sums [2][nbh][256][64] fp32; groups with no local row get zeros (so the cross-rank sum is the plain allreduce).
__global__ void __launch_bounds__(256) landmark_sums16_kernel(const __half* __restrict__ q, const __half* __restrict__ k, int nbh,
                                                              long long rows, long long row0, int l, float* __restrict__ sums) {
  const long long gid = (long long)blockIdx.x * 8 + (threadIdx.x >> 5);
  const long long per_tensor = (long long)nbh * TM_LANDMARKS;
  if (gid >= 2 * per_tensor) return;
  const int lane = threadIdx.x & 31, lane8 = lane & 7, phase = lane >> 3;
  const bool is_k = gid >= per_tensor;
  const long long rem = is_k ? gid - per_tensor : gid;
  const long long bh = rem / TM_LANDMARKS, g = rem % TM_LANDMARKS;
  long long lo = g * l - row0, hi = (g + 1) * l - row0;
  if (lo < 0) lo = 0;
  if (hi > rows) hi = rows;
  float acc[8];
#pragma unroll
  for (int e = 0; e < 8; ++e) acc[e] = 0.f;
  if (lo < hi) {
    const uint4* src = reinterpret_cast<const uint4*>((is_k ? k : q) + (bh * rows + lo) * TM_DHEAD) + lane8;
    const long long cnt = hi - lo;
    for (long long j = phase; j < cnt; j += 4) {
      const uint4 u = __ldg(src + j * 8);
      const __half2* h2 = reinterpret_cast<const __half2*>(&u);
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        const float2 f = __half22float2(h2[e]);
        acc[2 * e] += f.x; acc[2 * e + 1] += f.y;
      }
    }
  }
#pragma unroll
  for (int e = 0; e < 8; ++e) {
    acc[e] += __shfl_xor_sync(0xffffffffu, acc[e], 8);
    acc[e] += __shfl_xor_sync(0xffffffffu, acc[e], 16);
  }
  if (phase != 0) return;
  float4* dst = reinterpret_cast<float4*>(sums + gid * TM_DHEAD) + lane8 * 2;
  dst[0] = make_float4(acc[0], acc[1], acc[2], acc[3]);
  dst[1] = make_float4(acc[4], acc[5], acc[6], acc[7]);
}

// means = sums / l, plus the fp16 operand copies (same outputs as landmarks16_kernel).
__global__ void __launch_bounds__(256) landmark_finish_kernel(const float* __restrict__ sums, long long per_tensor_floats, float fl,
                                                              float* __restrict__ q_lm, float* __restrict__ k_lm,
                                                              __half* __restrict__ qlm16, __half* __restrict__ klm16) {
  const long long i = ((long long)blockIdx.x * 256 + threadIdx.x) * 2;
  if (i >= 2 * per_tensor_floats) return;
  const bool is_k = i >= per_tensor_floats;
  const long long o = is_k ? i - per_tensor_floats : i;
  const float2 s = *reinterpret_cast<const float2*>(sums + i);
  const float a = s.x / fl, b = s.y / fl;
  *reinterpret_cast<float2*>((is_k ? k_lm : q_lm) + o) = make_float2(a, b);
  *reinterpret_cast<uint32_t*>((is_k ? klm16 : qlm16) + o) = pack_f16x2(a, b);
}

// merged[b, t, h*64 + d] += sum over the taps whose source row lies in a neighbour's shard.
// grid (2 edges, 16 rows, B), 512 threads (one per channel).  v_prev / v_next: [B*8, 16, 64] fp16 = the previous rank's LAST
// and the next rank's FIRST 16 rows of v (NULL at the ends of the whole sequence: zero padding, nothing to add).
// Weights are rounded to fp16 like the Toeplitz operand of nystrom_out_tc_kernel.
__global__ void __launch_bounds__(512) conv_halo_fix_kernel(const __half* __restrict__ v_prev, const __half* __restrict__ v_next,
                                                            const float* __restrict__ conv_w, long long rows, __half* __restrict__ merged) {
  const int edge = blockIdx.x, t = blockIdx.y, b = blockIdx.z;
  const int c = threadIdx.x, h = c >> 6, d = c & 63;
  const __half* src = edge == 0 ? v_prev : v_next;
  if (src == nullptr) return;
  src += ((long long)(b * TM_HEADS + h) * 16) * TM_DHEAD + d;
  const float* w = conv_w + h * TM_CONV_TAPS;
  float add = 0.f;
  if (edge == 0) {
    // local row t, taps tau < 16 - t read local row t + tau - 16 < 0  ->  v_prev row t + tau
    for (int tau = 0; tau < 16 - t; ++tau) add = fmaf(__half2float(__float2half_rn(__ldg(w + tau))), __half2float(src[(t + tau) * TM_DHEAD]), add);
  } else {
    // local row rows - 16 + t, taps tau >= 32 - t read local row >= rows  ->  v_next row t + tau - 32
    for (int tau = 32 - t; tau <= 32; ++tau) add = fmaf(__half2float(__float2half_rn(__ldg(w + tau))), __half2float(src[(t + tau - 32) * TM_DHEAD]), add);
  }
  const long long row = edge == 0 ? t : rows - 16 + t;
  __half* m = merged + ((long long)b * rows + row) * TM_DIM + c;
  *m = __float2half_rn(__half2float(*m) + add);
}

}  // namespace tmil

using namespace tmil;

extern "C" int tm_landmark_sums16_fwd(const void* q16, const void* k16, int64_t B, int64_t rows, int64_t row0, int64_t l,
                                      float* sums, void* stream) {
  TM_CHECK_ARG(B > 0 && rows > 0 && row0 >= 0 && l > 0, TM_ERR_BAD_SHAPE, "tm_landmark_sums16_fwd: bad shape");
  TM_CHECK_ARG(aligned16(q16) && aligned16(k16) && aligned16(sums), TM_ERR_BAD_ALIGNMENT, "tm_landmark_sums16_fwd: pointers must be 16-byte aligned");
  const long long groups = 2ll * B * TM_HEADS * TM_LANDMARKS;
  landmark_sums16_kernel<<<(unsigned)ceil_div(groups, 8), 256, 0, (cudaStream_t)stream>>>((const __half*)q16, (const __half*)k16,
                                                                                         (int)(B * TM_HEADS), rows, row0, (int)l, sums);
  TM_LAUNCH_CHECK();
  return TM_OK;
}

extern "C" int tm_landmark_finish_fwd(const float* sums, int64_t B, int64_t l, float* q_lm, float* k_lm, void* qlm16, void* klm16,
                                      void* stream) {
  TM_CHECK_ARG(sums && q_lm && k_lm && qlm16 && klm16 && B > 0 && l > 0, TM_ERR_BAD_ARGUMENT, "tm_landmark_finish_fwd: bad arguments");
  const long long per = (long long)B * TM_HEADS * TM_LANDMARKS * TM_DHEAD;
  landmark_finish_kernel<<<(unsigned)ceil_div(per, 256), 256, 0, (cudaStream_t)stream>>>(sums, per, (float)l, q_lm, k_lm, (__half*)qlm16,
                                                                                        (__half*)klm16);
  TM_LAUNCH_CHECK();
  return TM_OK;
}

extern "C" int tm_conv_halo_fix_fwd(const void* v_prev16, const void* v_next16, const float* conv_w, int64_t B, int64_t rows,
                                    void* merged16, void* stream) {
  TM_CHECK_ARG(conv_w && merged16 && B > 0 && rows >= 32, TM_ERR_BAD_ARGUMENT, "tm_conv_halo_fix_fwd: bad arguments (rows >= 32)");
  if (!v_prev16 && !v_next16) return TM_OK;
  conv_halo_fix_kernel<<<dim3(2, 16, (unsigned)B), 512, 0, (cudaStream_t)stream>>>((const __half*)v_prev16, (const __half*)v_next16, conv_w,
                                                                                 rows, (__half*)merged16);
  TM_LAUNCH_CHECK();
  return TM_OK;
}
