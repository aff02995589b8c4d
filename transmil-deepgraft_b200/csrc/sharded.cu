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

// One block of 8 warps per (local landmark group, tensor, bh): sum of the local rows that fall into global group
// g = (row0 + r) / l.  A rank holds rows / l + 1 groups of l rows each (l = 782 for a 200 k bag), so the rows of a group are
// spread over the 8 warps (4 rows per warp and step: 8 lanes x 16 bytes per row) and reduced through shared memory in a fixed
// order.  The trailing `zero_blocks` blocks clear the groups without a local row (sums [2][nbh][256][64] fp32 is complete, so
// the cross-rank sum may also be a plain allreduce).
__global__ void __launch_bounds__(256) landmark_sums16_kernel(const __half* __restrict__ q, const __half* __restrict__ k, int nbh,
                                                              long long rows, long long row0, int l, int g_lo, int n_groups,
                                                              float* __restrict__ sums) {
  __shared__ float part[8][TM_DHEAD];
  const int active = n_groups * 2 * nbh;
  if ((int)blockIdx.x >= active) {
    // zero fill: groups outside [g_lo, g_lo + n_groups), 16 floats per thread step
    const long long total = 2ll * nbh * TM_LANDMARKS * (TM_DHEAD / 4);
    const int zb = gridDim.x - active;
    for (long long i = (long long)(blockIdx.x - active) * 256 + threadIdx.x; i < total; i += (long long)zb * 256) {
      const int g = (int)((i / (TM_DHEAD / 4)) % TM_LANDMARKS);
      if (g < g_lo || g >= g_lo + n_groups) reinterpret_cast<float4*>(sums)[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    }
    return;
  }
  const int gl = blockIdx.x / (2 * nbh), tb = blockIdx.x % (2 * nbh);
  const bool is_k = tb >= nbh;
  const int bh = is_k ? tb - nbh : tb;
  const long long g = g_lo + gl;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, lane8 = lane & 7, phase = lane >> 3;
  long long lo = g * l - row0, hi = (g + 1) * l - row0;
  if (lo < 0) lo = 0;
  if (hi > rows) hi = rows;
  float acc[8];
#pragma unroll
  for (int e = 0; e < 8; ++e) acc[e] = 0.f;
  if (lo < hi) {
    const uint4* src = reinterpret_cast<const uint4*>((is_k ? k : q) + ((long long)bh * rows + lo) * TM_DHEAD) + lane8;
    const long long cnt = hi - lo;
    for (long long j = warp * 4 + phase; j < cnt; j += 32) {
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
  if (phase == 0) {
#pragma unroll
    for (int e = 0; e < 8; ++e) part[warp][lane8 * 8 + e] = acc[e];
  }
  __syncthreads();
  if (threadIdx.x < TM_DHEAD) {
    float s = 0.f;
#pragma unroll
    for (int w = 0; w < 8; ++w) s += part[w][threadIdx.x];
    sums[(((long long)(is_k ? nbh : 0) + bh) * TM_LANDMARKS + g) * TM_DHEAD + threadIdx.x] = s;
  }
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
  TM_CHECK_ARG(B * TM_HEADS <= 16384, TM_ERR_BAD_SHAPE, "tm_landmark_sums16_fwd: too many bags");
  const int64_t g_lo = row0 / l;
  int64_t g_hi = (row0 + rows - 1) / l;
  if (g_hi > TM_LANDMARKS - 1) g_hi = TM_LANDMARKS - 1;
  const int n_groups = g_lo <= g_hi ? (int)(g_hi - g_lo + 1) : 0;
  const int nbh = (int)(B * TM_HEADS);
  landmark_sums16_kernel<<<(unsigned)(n_groups * 2 * nbh + 64), 256, 0, (cudaStream_t)stream>>>((const __half*)q16, (const __half*)k16, nbh,
                                                                                                 rows, row0, (int)l, (int)g_lo, n_groups, sums);
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
