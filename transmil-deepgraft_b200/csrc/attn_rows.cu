// K15: rows of the Nystrom attention matrix attn1 @ Z @ attn3 on request
// (return_attn=True consumers read ONE row per head: code/visualize_mil.py:581),
// without ever forming the dense [Sp,Sp] matrix.
//   a1[r,:]  = softmax_j(q_r . k_lm_j)            [256]
//   u[r,:]   = a1[r,:] @ Z                        [256]
//   out[r,t] = sum_g u[r,g] * attn3[g,t],  attn3[g,t] = exp(q_lm_g . k_t - mx_g) / den_g
#include "common.cuh"

namespace tmil {

// pass 1: per (bh, landmark g): row max and denominator of softmax_t(q_lm_g . k_t) over all Sp tokens.
// grid (256/8, B*8), 256 threads: warp = landmark, lanes stride tokens.
__global__ void __launch_bounds__(256) attn3_stats_kernel(const float* __restrict__ q_lm, const float* __restrict__ k, int Sp,
                                                          float* __restrict__ mx_out, float* __restrict__ den_out) {
  const int bh = blockIdx.y, g = blockIdx.x * 8 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
  const float* qg = q_lm + ((long long)bh * TM_LANDMARKS + g) * TM_DHEAD;
  float qv[TM_DHEAD];
#pragma unroll
  for (int d = 0; d < TM_DHEAD; ++d) qv[d] = __ldg(qg + d);
  const float* kb = k + (long long)bh * Sp * TM_DHEAD;
  float m = -INFINITY, s = 0.f;
  for (int t = lane; t < Sp; t += 32) {
    const float4* kr = reinterpret_cast<const float4*>(kb + (long long)t * TM_DHEAD);
    float a = 0.f;
#pragma unroll
    for (int i = 0; i < 16; ++i) {
      float4 kv = __ldg(kr + i);
      a = fmaf(qv[4 * i], kv.x, a); a = fmaf(qv[4 * i + 1], kv.y, a); a = fmaf(qv[4 * i + 2], kv.z, a); a = fmaf(qv[4 * i + 3], kv.w, a);
    }
    if (a > m) { s *= expf(m - a); m = a; }
    s += expf(a - m);
  }
  const float M = warp_max(m);
  s *= expf(m - M);
  s = warp_sum(s);
  if (lane == 0) { mx_out[(long long)bh * TM_LANDMARKS + g] = M; den_out[(long long)bh * TM_LANDMARKS + g] = s; }
}

// pass 2: u[bh, i, :] = softmax(q_row . k_lm^T) @ Z, then scaled by 1/den.  grid (n_rows, B*8), 256 threads.
__global__ void __launch_bounds__(256) attn_rows_u_kernel(const float* __restrict__ q, const float* __restrict__ k_lm,
                                                          const float* __restrict__ Z, const float* __restrict__ den, int Sp,
                                                          const int* __restrict__ rows, int n_rows, float* __restrict__ u) {
  __shared__ float a1[TM_LANDMARKS];
  __shared__ float red[8];
  const int bh = blockIdx.y, i = blockIdx.x, j = threadIdx.x;
  const int r = rows[i];
  const float* qr = q + ((long long)bh * Sp + r) * TM_DHEAD;
  const float* kr = k_lm + ((long long)bh * TM_LANDMARKS + j) * TM_DHEAD;
  float s = 0.f;
#pragma unroll 8
  for (int d = 0; d < TM_DHEAD; ++d) s = fmaf(__ldg(qr + d), __ldg(kr + d), s);
  float m = warp_max(s);
  if ((j & 31) == 0) red[j >> 5] = m;
  __syncthreads();
  m = red[0];
#pragma unroll
  for (int w = 1; w < 8; ++w) m = fmaxf(m, red[w]);
  __syncthreads();
  const float e = expf(s - m);
  float t = warp_sum(e);
  if ((j & 31) == 0) red[j >> 5] = t;
  __syncthreads();
  t = 0.f;
#pragma unroll
  for (int w = 0; w < 8; ++w) t += red[w];
  a1[j] = e / t;
  __syncthreads();
  const float* Zb = Z + (long long)bh * 65536;
  float acc = 0.f;
  for (int g = 0; g < TM_LANDMARKS; ++g) acc = fmaf(a1[g], __ldg(Zb + g * TM_LANDMARKS + j), acc);
  u[((long long)bh * n_rows + i) * TM_LANDMARKS + j] = acc / den[(long long)bh * TM_LANDMARKS + j];
}

// pass 3: out[bh, i, t] = sum_g u[i,g] exp(q_lm_g . k_t - mx_g).  grid (Sp/256, B*8), thread = token.
__global__ void __launch_bounds__(256) attn_rows_out_kernel(const float* __restrict__ q_lm, const float* __restrict__ k,
                                                            const float* __restrict__ mx, const float* __restrict__ u, int Sp,
                                                            int n_rows, float* __restrict__ out) {
  extern __shared__ __align__(16) float sm[];
  float* qs = sm;                               // [256][64]
  float* ms = qs + TM_LANDMARKS * TM_DHEAD;     // [256]
  float* us = ms + TM_LANDMARKS;                // [n_rows][256]
  const int bh = blockIdx.y, t = blockIdx.x * 256 + threadIdx.x;
  for (int i = threadIdx.x; i < TM_LANDMARKS * 16; i += 256)
    reinterpret_cast<float4*>(qs)[i] = __ldg(reinterpret_cast<const float4*>(q_lm + (long long)bh * TM_LANDMARKS * TM_DHEAD) + i);
  ms[threadIdx.x] = mx[(long long)bh * TM_LANDMARKS + threadIdx.x];
  for (int i = threadIdx.x; i < n_rows * TM_LANDMARKS; i += 256) us[i] = u[(long long)bh * n_rows * TM_LANDMARKS + i];
  float kv[TM_DHEAD];
  const float4* kr = reinterpret_cast<const float4*>(k + ((long long)bh * Sp + t) * TM_DHEAD);
#pragma unroll
  for (int i = 0; i < 16; ++i) {
    float4 x = __ldg(kr + i);
    kv[4 * i] = x.x; kv[4 * i + 1] = x.y; kv[4 * i + 2] = x.z; kv[4 * i + 3] = x.w;
  }
  __syncthreads();
  for (int i0 = 0; i0 < n_rows; i0 += 4) {
    float acc[4] = {0.f, 0.f, 0.f, 0.f};
    for (int g = 0; g < TM_LANDMARKS; ++g) {
      float a = 0.f;
#pragma unroll
      for (int d = 0; d < TM_DHEAD; d += 4) {
        float4 qv = *reinterpret_cast<const float4*>(qs + g * TM_DHEAD + d);
        a = fmaf(kv[d], qv.x, a); a = fmaf(kv[d + 1], qv.y, a); a = fmaf(kv[d + 2], qv.z, a); a = fmaf(kv[d + 3], qv.w, a);
      }
      const float e = expf(a - ms[g]);
#pragma unroll
      for (int ii = 0; ii < 4; ++ii)
        if (i0 + ii < n_rows) acc[ii] = fmaf(us[(i0 + ii) * TM_LANDMARKS + g], e, acc[ii]);
    }
#pragma unroll
    for (int ii = 0; ii < 4; ++ii)
      if (i0 + ii < n_rows) out[((long long)bh * n_rows + i0 + ii) * Sp + t] = acc[ii];
  }
}

}  // namespace tmil

extern "C" int64_t tm_attn_rows_scratch_floats(int64_t B, int64_t Sp, int64_t n_rows) {
  (void)Sp;
  return B * TM_HEADS * TM_LANDMARKS * (2 + n_rows) + 64 + n_rows;
}

extern "C" int tm_attn_rows_fwd(const float* q, const float* k, const float* q_lm, const float* k_lm, const float* Z,
                                int64_t B, int64_t Sp, const int64_t* rows_host, int64_t n_rows, float* out, float* scratch,
                                void* stream) {
  using namespace tmil;
  TM_CHECK_ARG(B > 0 && Sp > 0 && Sp % TM_LANDMARKS == 0 && n_rows > 0 && n_rows <= 16, TM_ERR_BAD_SHAPE, "tm_attn_rows_fwd: bad shape (n_rows must be 1..16)");
  TM_CHECK_ARG(q && k && q_lm && k_lm && Z && rows_host && out && scratch, TM_ERR_BAD_ARGUMENT, "tm_attn_rows_fwd: NULL pointer");
  for (int64_t i = 0; i < n_rows; ++i)
    TM_CHECK_ARG(rows_host[i] >= 0 && rows_host[i] < Sp, TM_ERR_BAD_ARGUMENT, "tm_attn_rows_fwd: row %lld out of range [0,%lld)", (long long)rows_host[i], (long long)Sp);
  cudaStream_t st = (cudaStream_t)stream;
  const int nbh = (int)(B * TM_HEADS);
  float* mx = scratch;
  float* den = mx + (long long)nbh * TM_LANDMARKS;
  float* u = den + (long long)nbh * TM_LANDMARKS;
  int* rows_d = reinterpret_cast<int*>(u + (long long)nbh * TM_LANDMARKS * n_rows);
  int rows32[16];
  for (int64_t i = 0; i < n_rows; ++i) rows32[i] = (int)rows_host[i];
  TM_CUDA(cudaMemcpyAsync(rows_d, rows32, sizeof(int) * n_rows, cudaMemcpyHostToDevice, st));
  TM_CUDA(cudaStreamSynchronize(st));  // rows32 lives on this stack frame
  attn3_stats_kernel<<<dim3(TM_LANDMARKS / 8, nbh), 256, 0, st>>>(q_lm, k, (int)Sp, mx, den);
  TM_LAUNCH_CHECK();
  attn_rows_u_kernel<<<dim3((unsigned)n_rows, nbh), 256, 0, st>>>(q, k_lm, Z, den, (int)Sp, rows_d, (int)n_rows, u);
  TM_LAUNCH_CHECK();
  const size_t smem = (size_t)(TM_LANDMARKS * TM_DHEAD + TM_LANDMARKS + n_rows * TM_LANDMARKS) * sizeof(float);
  TM_CUDA(cudaFuncSetAttribute(attn_rows_out_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  attn_rows_out_kernel<<<dim3((unsigned)(Sp / 256), nbh), 256, smem, st>>>(q_lm, k, mx, u, (int)Sp, (int)n_rows, out);
  TM_LAUNCH_CHECK();
  return TM_OK;
}
