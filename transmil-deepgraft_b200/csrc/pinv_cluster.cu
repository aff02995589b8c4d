// K7 on the tensor cores: operand preparation for, and the launch of, the cluster kernel that runs the 6 Moore-Penrose
// iterations (24 dependent 256^3 products) of every (bag, head) inside a 4-CTA thread-block cluster (pinv_pipe.cu).
//
//   Z <- 1/4 Z (13 I - XZ (15 I - XZ (7 I - XZ))),  XZ = A Z        (nystrom-attention 0.0.11 moore_penrose_iter_pinv)
//   evaluated as  XZ = A Z;  T1 = 7 XZ - XZ XZ;  T2 = 15 XZ - XZ T1;  Z' = 3.25 Z - 0.25 Z T2   (same as pinv.cu)
//
// History of the cluster kernel (timelines under profiles/): generation 1 exchanged the operand blocks through global memory
// with TMA stores / loads behind one hardware cluster barrier per product; generation 2 pushed them through distributed
// shared memory (82 us per layer at B = 1, r1_pinv_bulk_push_timeline.txt); generation 3, the only one still compiled,
// pipelines the push by column half (r2/pinv_pipe_timeline_v1.txt).
#include <cuda_fp16.h>
#include <stdlib.h>

#include "common.cuh"
#include "tc_common.cuh"

namespace tmil {

// A16 = half(attn2);  Z0 = half(attn2^T / scale)      grid (8, 8, B*8), block (32, 8)
// scale = max_i rowsum * max_j colsum over ALL bags and heads of the call (moore_penrose_iter_pinv's global scalar): every
// CTA reduces the 2 x nbh x 256 sums itself (max is order independent, so all CTAs get the same bits) instead of waiting
// for a one-block kernel in between.
// With many (bag, head) pairs that per-CTA reduction would be quadratic work: the caller then passes `scale` computed by
// pinv_scale_kernel (scale != NULL) and the loop below is skipped.
__global__ void __launch_bounds__(256) pinv_prep16_kernel(const float* __restrict__ A, const float* __restrict__ rowsum,
                                                          const float* __restrict__ colpart, int nbh, int nparts,
                                                          const float* __restrict__ scale, __half* __restrict__ A16,
                                                          __half* __restrict__ Z16) {
  __shared__ float t[32][33];
  __shared__ float red[2][8];
  pdl_sync();
  const int tid = threadIdx.y * 32 + threadIdx.x;
  float mr = 0.f, mc = 0.f;
  if (scale == nullptr) {
    // global init scalar: every CTA needs it, so the scan is kept short -- four columns per thread, the 16 row-tile partials
    // of a head as 16 independent 16-byte loads (fixed summation order p = 0 .. 15: same bits in every CTA and every run)
    const int j4 = (tid & 63) * 4;
    for (int bh = tid >> 6; bh < nbh; bh += 4) {
      const float4 rs = __ldg(reinterpret_cast<const float4*>(rowsum + bh * TM_LANDMARKS + j4));
      mr = fmaxf(mr, fmaxf(fmaxf(rs.x, rs.y), fmaxf(rs.z, rs.w)));
      float4 c = make_float4(0.f, 0.f, 0.f, 0.f);
      if (nparts == 16) {
        float4 v[16];
#pragma unroll
        for (int p = 0; p < 16; ++p) v[p] = __ldg(reinterpret_cast<const float4*>(colpart + ((long long)bh * 16 + p) * TM_LANDMARKS + j4));
#pragma unroll
        for (int p = 0; p < 16; ++p) { c.x += v[p].x; c.y += v[p].y; c.z += v[p].z; c.w += v[p].w; }
      } else {
        for (int p = 0; p < nparts; ++p) {
          const float4 v = __ldg(reinterpret_cast<const float4*>(colpart + ((long long)bh * nparts + p) * TM_LANDMARKS + j4));
          c.x += v.x; c.y += v.y; c.z += v.z; c.w += v.w;
        }
      }
      mc = fmaxf(mc, fmaxf(fmaxf(c.x, c.y), fmaxf(c.z, c.w)));
    }
  }
  mr = warp_max(mr); mc = warp_max(mc);
  if (threadIdx.x == 0) { red[0][threadIdx.y] = mr; red[1][threadIdx.y] = mc; }
  const long long base = (long long)blockIdx.z * TM_LANDMARKS * TM_LANDMARKS;
  const int x0 = blockIdx.x * 32, y0 = blockIdx.y * 32;
  for (int i = threadIdx.y; i < 32; i += 8) {
    const long long o = base + (long long)(y0 + i) * TM_LANDMARKS + x0 + threadIdx.x;
    const float a = A[o];
    t[i][threadIdx.x] = a;
    A16[o] = __float2half_rn(a);
  }
  __syncthreads();
  mr = red[0][0]; mc = red[1][0];
#pragma unroll
  for (int w = 1; w < 8; ++w) { mr = fmaxf(mr, red[0][w]); mc = fmaxf(mc, red[1][w]); }
  const float sc = scale != nullptr ? scale[0] : mr * mc;
  for (int i = threadIdx.y; i < 32; i += 8)
    Z16[base + (long long)(x0 + i) * TM_LANDMARKS + y0 + threadIdx.x] = __float2half_rn(t[threadIdx.x][i] / sc);
}

int pinv_pipe_launch(__half* A16, __half* Za, __half* Zb, int nbh, float* Z, cudaStream_t st);   // pinv_pipe.cu

// scratch16: 6 fp16 matrices per (bag, head): A16, Za, Zb, XZ, T1, T2
int pinv_cluster_run(const float* attn2, const float* rowsum, const float* colpart, int nparts, const float* scale, int nbh,
                     void* scratch16, float* Z, cudaStream_t st) {
  TM_CHECK_ARG(tc_available(), TM_ERR_UNSUPPORTED_ARCH, "pinv_cluster_run needs an sm_100 device");
  const long long mat = (long long)TM_LANDMARKS * TM_LANDMARKS;
  __half* A16 = (__half*)scratch16;
  __half* Za = A16 + nbh * mat;
  __half* Zb = Za + nbh * mat;
  __half* XZ = Zb + nbh * mat;
  __half* T1 = XZ + nbh * mat;
  __half* T2 = T1 + nbh * mat;
  launch_pdl(pinv_prep16_kernel, dim3(8, 8, nbh), dim3(32, 8), 0, st, attn2, rowsum, colpart, nbh, nparts, scale, A16, Za);
  TM_LAUNCH_CHECK();
  return pinv_pipe_launch(A16, Za, Zb, nbh, Z, st);
}

}  // namespace tmil
