// Row-wise kernels: LayerNorm (+ front zero pad), in-place LayerNorm for the fc1
// stems, and the cls-row head (final LayerNorm + _fc).  All HBM-bound: one warp
// per row, 128-bit loads, row kept in registers between the two passes.
#include <cuda_fp16.h>

#include "common.cuh"
#include "tc_common.cuh"

namespace tmil {

// One warp normalises one row of width W (multiple of 128, <= 1024).
template <int W, int OUT = 0>   // OUT 0: fp32, 1: fp32 rounded to tf32, 2: fp16 (dst then points at __half)
__device__ __forceinline__ void warp_layernorm_row(const float* src, float* dst,
                                                   const float* __restrict__ gamma, const float* __restrict__ beta, int lane) {
  constexpr int V = W / 128;  // float4 per lane
  float4 x[V];
  float s = 0.f;
#pragma unroll
  for (int i = 0; i < V; ++i) {
    x[i] = *reinterpret_cast<const float4*>(src + (i * 32 + lane) * 4);
    s += (x[i].x + x[i].y) + (x[i].z + x[i].w);
  }
  const float mean = warp_sum(s) * (1.0f / W);
  float q = 0.f;
#pragma unroll
  for (int i = 0; i < V; ++i) {
    float a = x[i].x - mean, b = x[i].y - mean, c = x[i].z - mean, d = x[i].w - mean;
    q += (a * a + b * b) + (c * c + d * d);
  }
  const float rstd = 1.0f / sqrtf(warp_sum(q) * (1.0f / W) + 1e-5f);  // eps of nn.LayerNorm
#pragma unroll
  for (int i = 0; i < V; ++i) {
    const int c = (i * 32 + lane) * 4;
    float4 g = __ldg(reinterpret_cast<const float4*>(gamma + c));
    float4 b = __ldg(reinterpret_cast<const float4*>(beta + c));
    float4 o;
    o.x = (x[i].x - mean) * rstd * g.x + b.x;
    o.y = (x[i].y - mean) * rstd * g.y + b.y;
    o.z = (x[i].z - mean) * rstd * g.z + b.z;
    o.w = (x[i].w - mean) * rstd * g.w + b.w;
    if (OUT == 1) { o.x = round_tf32(o.x); o.y = round_tf32(o.y); o.z = round_tf32(o.z); o.w = round_tf32(o.w); }
    if (OUT == 2) *reinterpret_cast<uint2*>(reinterpret_cast<__half*>(dst) + c) = make_uint2(pack_f16x2(o.x, o.y), pack_f16x2(o.z, o.w));
    else *reinterpret_cast<float4*>(dst + c) = o;
  }
}

// xn[b, t] = t < P ? 0 : LN(h[b, t - P])
template <int OUT>
__global__ void __launch_bounds__(256) ln_pad_kernel(const float* __restrict__ h, const float* __restrict__ gamma,
                                                     const float* __restrict__ beta, long long rows, int S, int Sp, int P,
                                                     float* __restrict__ xn) {
  pdl_sync();
  const int lane = threadIdx.x & 31;
  const long long row = (long long)blockIdx.x * 8 + (threadIdx.x >> 5);
  if (row >= rows) return;
  const long long b = row / Sp;
  const int t = (int)(row - b * Sp);
  float* dst = OUT == 2 ? reinterpret_cast<float*>(reinterpret_cast<__half*>(xn) + row * TM_DIM) : xn + row * TM_DIM;
  if (t < P) {
    if (OUT == 2) {
#pragma unroll
      for (int i = 0; i < 2; ++i) *reinterpret_cast<uint4*>(reinterpret_cast<__half*>(dst) + (i * 32 + lane) * 8) = make_uint4(0u, 0u, 0u, 0u);
    } else {
#pragma unroll
      for (int i = 0; i < 4; ++i) *reinterpret_cast<float4*>(dst + (i * 32 + lane) * 4) = make_float4(0.f, 0.f, 0.f, 0.f);
    }
    return;
  }
  warp_layernorm_row<TM_DIM, OUT>(h + (b * S + (t - P)) * TM_DIM, dst, gamma, beta, lane);
}

// In-place LayerNorm over mapped rows: row r -> (r / rpb) * bag_rows + off + r % rpb.
template <int W, int ROUND>
__global__ void __launch_bounds__(256) ln_inplace_kernel(float* __restrict__ x, const float* __restrict__ gamma,
                                                         const float* __restrict__ beta, long long rows, int rpb,
                                                         long long bag_rows, int off) {
  const int lane = threadIdx.x & 31;
  const long long row = (long long)blockIdx.x * 8 + (threadIdx.x >> 5);
  if (row >= rows) return;
  long long rr = row;
  if (rpb > 0) {
    long long b = row / rpb;
    rr = b * bag_rows + off + (row - b * rpb);
  }
  float* p = x + rr * W;
  warp_layernorm_row<W, ROUND>(p, p, gamma, beta, lane);
}

int ln_inplace(float* x, const float* gamma, const float* beta, long long rows, int width, int rpb, long long bag_rows,
               int off, bool round, cudaStream_t st) {
  if (rows <= 0) return TM_OK;
  unsigned grid = (unsigned)ceil_div(rows, 8);
  if (width == 512 && !round)
    ln_inplace_kernel<512, 0><<<grid, 256, 0, st>>>(x, gamma, beta, rows, rpb, bag_rows, off);
  else if (width == 768 && !round)
    ln_inplace_kernel<768, 0><<<grid, 256, 0, st>>>(x, gamma, beta, rows, rpb, bag_rows, off);
  else if (width == 1024 && !round)
    ln_inplace_kernel<1024, 0><<<grid, 256, 0, st>>>(x, gamma, beta, rows, rpb, bag_rows, off);
  else if (width == 768 && round)
    ln_inplace_kernel<768, 1><<<grid, 256, 0, st>>>(x, gamma, beta, rows, rpb, bag_rows, off);
  else if (width == 1024 && round)
    ln_inplace_kernel<1024, 1><<<grid, 256, 0, st>>>(x, gamma, beta, rows, rpb, bag_rows, off);
  else {
    set_error("ln_inplace: unsupported width %d", width);
    return TM_ERR_BAD_SHAPE;
  }
  TM_LAUNCH_CHECK();
  return TM_OK;
}

__global__ void __launch_bounds__(256) round_tf32_kernel(const float* __restrict__ src, float* __restrict__ dst, long long n4) {
  for (long long i = (long long)blockIdx.x * 256 + threadIdx.x; i < n4; i += (long long)gridDim.x * 256) {
    float4 v = reinterpret_cast<const float4*>(src)[i];
    reinterpret_cast<float4*>(dst)[i] = make_float4(round_tf32(v.x), round_tf32(v.y), round_tf32(v.z), round_tf32(v.w));
  }
}

__global__ void __launch_bounds__(256) cvt_f16_kernel(const float* __restrict__ src, __half* __restrict__ dst, long long n4) {
  for (long long i = (long long)blockIdx.x * 256 + threadIdx.x; i < n4; i += (long long)gridDim.x * 256) {
    const float4 v = reinterpret_cast<const float4*>(src)[i];
    reinterpret_cast<uint2*>(dst)[i] = make_uint2(pack_f16x2(v.x, v.y), pack_f16x2(v.z, v.w));
  }
}

__global__ void __launch_bounds__(256) cvt_f32_kernel(const __half* __restrict__ src, float* __restrict__ dst, long long n4) {
  for (long long i = (long long)blockIdx.x * 256 + threadIdx.x; i < n4; i += (long long)gridDim.x * 256) {
    const uint2 u = reinterpret_cast<const uint2*>(src)[i];
    const float2 a = __half22float2(*reinterpret_cast<const __half2*>(&u.x)), b = __half22float2(*reinterpret_cast<const __half2*>(&u.y));
    reinterpret_cast<float4*>(dst)[i] = make_float4(a.x, a.y, b.x, b.y);
  }
}

// One launch for every per-forward operand copy of the parameters: tf32-rounded / fp16 copies of the GEMM weights and the
// packed PPEG filter (effective 7x7 + bias, see ppeg.cu).  Block b serves job j with first_block[j] <= b < first_block[j+1];
// the blocks after the last job pack the PPEG filter.
struct PrepJobs {
  const float* src[6];
  void* dst[6];
  long long n4[6];
  int kind[6];            // 0: round to tf32 (fp32 out), 1: fp16 out
  int first_block[7];
  int njobs;
  const float *w7, *b7, *w5, *b5, *w3, *b3;
  float* packed;          // [50][512] or NULL
};

__global__ void __launch_bounds__(256) prep_operands_kernel(const PrepJobs J) {
  pdl_sync();
  const int b = blockIdx.x;
  if (b >= J.first_block[J.njobs]) {
    if (J.packed == nullptr) return;
    const int i = (b - J.first_block[J.njobs]) * 256 + threadIdx.x;   // (tap, channel), tap 49 = bias
    if (i >= 50 * TM_DIM) return;
    const int t = i / TM_DIM, c = i - t * TM_DIM;
    float v;
    if (t == 49) {
      v = (J.b7[c] + J.b5[c]) + J.b3[c];
    } else {
      const int dy = t / 7, dx = t - dy * 7;
      v = J.w7[c * 49 + t];
      if (dy >= 1 && dy <= 5 && dx >= 1 && dx <= 5) v += J.w5[c * 25 + (dy - 1) * 5 + (dx - 1)];
      if (dy >= 2 && dy <= 4 && dx >= 2 && dx <= 4) v += J.w3[c * 9 + (dy - 2) * 3 + (dx - 2)];
    }
    J.packed[i] = v;
    return;
  }
  int j = 0;
#pragma unroll
  for (int q = 1; q < 6; ++q)
    if (q < J.njobs && b >= J.first_block[q]) j = q;
  const long long nb = J.first_block[j + 1] - J.first_block[j];
  const float4* src = reinterpret_cast<const float4*>(J.src[j]);
  for (long long i = (long long)(b - J.first_block[j]) * 256 + threadIdx.x; i < J.n4[j]; i += nb * 256) {
    const float4 v = src[i];
    if (J.kind[j] == 0)
      reinterpret_cast<float4*>(J.dst[j])[i] = make_float4(round_tf32(v.x), round_tf32(v.y), round_tf32(v.z), round_tf32(v.w));
    else
      reinterpret_cast<uint2*>(J.dst[j])[i] = make_uint2(pack_f16x2(v.x, v.y), pack_f16x2(v.z, v.w));
  }
}

// srcs/dsts/counts/kinds: njobs <= 6 conversion jobs (counts in floats, multiples of 4); packed may be NULL
int prep_operands(const float* const* srcs, void* const* dsts, const long long* counts, const int* kinds, int njobs,
                  const tm_weights* w, float* packed, cudaStream_t st) {
  PrepJobs J;
  memset(&J, 0, sizeof(J));
  TM_CHECK_ARG(njobs >= 0 && njobs <= 6, TM_ERR_BAD_ARGUMENT, "prep_operands: at most 6 jobs");
  int blocks = 0;
  for (int j = 0; j < njobs; ++j) {
    TM_CHECK_ARG(counts[j] % 4 == 0 && aligned16(srcs[j]) && (reinterpret_cast<uintptr_t>(dsts[j]) & 7u) == 0, TM_ERR_BAD_ALIGNMENT,
                 "prep_operands: operand %d is not 4-float / 16-byte aligned", j);
    J.src[j] = srcs[j]; J.dst[j] = dsts[j]; J.n4[j] = counts[j] / 4; J.kind[j] = kinds[j];
    J.first_block[j] = blocks;
    long long nb = ceil_div(J.n4[j], 256 * 4);      // 4 float4 per thread
    if (nb < 1) nb = 1;
    if (nb > 256) nb = 256;
    blocks += (int)nb;
  }
  J.first_block[njobs] = blocks;
  J.njobs = njobs;
  if (packed != nullptr) {
    J.w7 = w->ppeg_w7; J.b7 = w->ppeg_b7; J.w5 = w->ppeg_w5; J.b5 = w->ppeg_b5; J.w3 = w->ppeg_w3; J.b3 = w->ppeg_b3;
    J.packed = packed;
    blocks += (50 * TM_DIM + 255) / 256;
  }
  if (blocks == 0) return TM_OK;
  launch_pdl(prep_operands_kernel, dim3(blocks), dim3(256), 0, st, J);
  TM_LAUNCH_CHECK();
  return TM_OK;
}

// cls row: logits[b] = _fc(LayerNorm(h[b, 0]))
__global__ void __launch_bounds__(128) head_kernel(const float* __restrict__ h, long long bag_stride,
                                                   const float* __restrict__ gamma, const float* __restrict__ beta,
                                                   const float* __restrict__ fc_w, const float* __restrict__ fc_b,
                                                   int n_classes, float* __restrict__ logits) {
  __shared__ __align__(16) float hn[TM_DIM];
  pdl_sync();
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int b = blockIdx.x;
  if (warp == 0) warp_layernorm_row<TM_DIM>(h + (long long)b * bag_stride, hn, gamma, beta, lane);
  __syncthreads();
  for (int c = warp; c < n_classes; c += 4) {
    const float* wr = fc_w + (long long)c * TM_DIM;
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int j = (i * 32 + lane) * 4;
      float4 w4 = __ldg(reinterpret_cast<const float4*>(wr + j));
      float4 x4 = *reinterpret_cast<const float4*>(hn + j);
      s += (w4.x * x4.x + w4.y * x4.y) + (w4.z * x4.z + w4.w * x4.w);
    }
    s = warp_sum(s);
    if (lane == 0) logits[(long long)b * n_classes + c] = s + fc_b[c];
  }
}

}  // namespace tmil

extern "C" int tm_round_tf32(const float* src, float* dst, int64_t n, void* stream) {
  TM_CHECK_ARG(src && dst && n >= 0 && n % 4 == 0 && tmil::aligned16(src) && tmil::aligned16(dst), TM_ERR_BAD_ALIGNMENT,
               "tm_round_tf32: n must be a multiple of 4 and the pointers 16-byte aligned");
  if (n == 0) return TM_OK;
  long long blocks = tmil::ceil_div(n / 4, 256);
  if (blocks > 148 * 8) blocks = 148 * 8;
  tmil::round_tf32_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(src, dst, n / 4);
  TM_LAUNCH_CHECK();
  return TM_OK;
}

extern "C" int tm_cvt_f16(const float* src, void* dst, int64_t n, void* stream) {
  TM_CHECK_ARG(src && dst && n >= 0 && n % 4 == 0 && tmil::aligned16(src) && (reinterpret_cast<uintptr_t>(dst) & 7u) == 0, TM_ERR_BAD_ALIGNMENT,
               "tm_cvt_f16: n must be a multiple of 4, src 16-byte and dst 8-byte aligned");
  if (n == 0) return TM_OK;
  long long blocks = tmil::ceil_div(n / 4, 256);
  if (blocks > 148 * 8) blocks = 148 * 8;
  tmil::cvt_f16_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(src, (__half*)dst, n / 4);
  TM_LAUNCH_CHECK();
  return TM_OK;
}

extern "C" int tm_cvt_f32(const void* src, float* dst, int64_t n, void* stream) {
  TM_CHECK_ARG(src && dst && n >= 0 && n % 4 == 0 && tmil::aligned16(dst) && (reinterpret_cast<uintptr_t>(src) & 7u) == 0, TM_ERR_BAD_ALIGNMENT,
               "tm_cvt_f32: n must be a multiple of 4, dst 16-byte and src 8-byte aligned");
  if (n == 0) return TM_OK;
  long long blocks = tmil::ceil_div(n / 4, 256);
  if (blocks > 148 * 8) blocks = 148 * 8;
  tmil::cvt_f32_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>((const __half*)src, dst, n / 4);
  TM_LAUNCH_CHECK();
  return TM_OK;
}

extern "C" int tm_ln_pad_fwd(const float* h, const float* gamma, const float* beta, int64_t B, int64_t S, int64_t P,
                             float* xn, uint32_t flags, void* stream) {
  TM_CHECK_ARG(B > 0 && S > 0 && P >= 0 && (S + P) % TM_LANDMARKS == 0, TM_ERR_BAD_SHAPE,
               "tm_ln_pad_fwd: S+P must be a multiple of 256 (S=%lld P=%lld)", (long long)S, (long long)P);
  TM_CHECK_ARG(tmil::aligned16(h) && tmil::aligned16(xn) && tmil::aligned16(gamma) && tmil::aligned16(beta), TM_ERR_BAD_ALIGNMENT,
               "tm_ln_pad_fwd: pointers must be 16-byte aligned");
  const long long rows = (long long)B * (S + P);
  const unsigned grid = (unsigned)tmil::ceil_div(rows, 8);
  cudaStream_t st = (cudaStream_t)stream;
  if (flags & TM_FLAG_F16_ACT)
    tmil::launch_pdl(tmil::ln_pad_kernel<2>, dim3(grid), dim3(256), 0, st, h, gamma, beta, rows, (int)S, (int)(S + P), (int)P, xn);
  else if (flags & TM_FLAG_TF32_OPERANDS)
    tmil::launch_pdl(tmil::ln_pad_kernel<1>, dim3(grid), dim3(256), 0, st, h, gamma, beta, rows, (int)S, (int)(S + P), (int)P, xn);
  else
    tmil::launch_pdl(tmil::ln_pad_kernel<0>, dim3(grid), dim3(256), 0, st, h, gamma, beta, rows, (int)S, (int)(S + P), (int)P, xn);
  TM_LAUNCH_CHECK();
  return TM_OK;
}

extern "C" int tm_head_fwd(const tm_weights* w, const float* h, int64_t B, int64_t S, float* logits, void* stream) {
  TM_CHECK_ARG(w != nullptr && B > 0 && S > 0 && w->n_classes > 0, TM_ERR_BAD_ARGUMENT, "tm_head_fwd: bad arguments");
  tmil::launch_pdl(tmil::head_kernel, dim3((unsigned)B), dim3(128), 0, (cudaStream_t)stream, h, (long long)S * TM_DIM, w->norm_w, w->norm_b,
                   w->fc_w, w->fc_b, (int)w->n_classes, logits);
  TM_LAUNCH_CHECK();
  return TM_OK;
}
