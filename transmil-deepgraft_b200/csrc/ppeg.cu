// K13: PPEG (code/models/TransMIL.py:60-75):  y = x + dw7(x) + dw5(x) + dw3(x) on the
// GxG token image, cls row passed through.  Tokens stay channels-last [S,512]
// (token t <-> pixel (t / G, t % G), channel fastest), so the reference's two
// transposes disappear and every global access is a full 128-byte line.
// The three depthwise filters are pre-summed into one effective 7x7 per channel
// (exact in real arithmetic; identity added as `+ x` like the reference).
#include "common.cuh"

namespace tmil {

constexpr int PP_T = 16;            // output tile edge
constexpr int PP_H = PP_T + 6;      // input tile edge with the 3-pixel halo
constexpr int PP_C = 32;            // channels per CTA (one 128-byte line per pixel)

// packed[tap*512 + c] (tap = dy*7+dx, 49 taps) and packed[49*512 + c] = b7+b5+b3
__global__ void ppeg_pack_kernel(const float* __restrict__ w7, const float* __restrict__ b7, const float* __restrict__ w5,
                                 const float* __restrict__ b5, const float* __restrict__ w3, const float* __restrict__ b3,
                                 float* __restrict__ packed) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= TM_DIM) return;
  for (int dy = 0; dy < 7; ++dy)
    for (int dx = 0; dx < 7; ++dx) {
      float w = w7[c * 49 + dy * 7 + dx];
      if (dy >= 1 && dy <= 5 && dx >= 1 && dx <= 5) w += w5[c * 25 + (dy - 1) * 5 + (dx - 1)];
      if (dy >= 2 && dy <= 4 && dx >= 2 && dx <= 4) w += w3[c * 9 + (dy - 2) * 3 + (dx - 2)];
      packed[(dy * 7 + dx) * TM_DIM + c] = w;
    }
  packed[49 * TM_DIM + c] = (b7[c] + b5[c]) + b3[c];
}

// grid (512/32, tiles_x*tiles_y, B), 256 threads: lane = channel, warp = pair of tile rows.  The channel slice is the
// FASTEST grid dimension: CTAs that run together then cover whole 2 KB token rows (a spatial tile at a time) instead of
// sweeping one 128-byte slice of every row, which thrashed HBM pages (1.2 TB/s measured with the other order).
__global__ void __launch_bounds__(256) ppeg_kernel(const float* __restrict__ hin, const float* __restrict__ packed, int G,
                                                   long long bag_stride, float* __restrict__ hout) {
  extern __shared__ __align__(16) float tile[];  // [PP_H][PP_H][PP_C]
  const int tiles_x = (G + PP_T - 1) / PP_T;
  const int ty0 = (blockIdx.y / tiles_x) * PP_T, tx0 = (blockIdx.y % tiles_x) * PP_T;
  const int c0 = blockIdx.x * PP_C;
  const float* xin = hin + (long long)blockIdx.z * bag_stride + TM_DIM;  // token 0 is row 1
  float* xout = hout + (long long)blockIdx.z * bag_stride + TM_DIM;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (blockIdx.y == 0 && tid < PP_C)  // cls token passes through (TransMIL.py:67,74)
    hout[(long long)blockIdx.z * bag_stride + c0 + tid] = hin[(long long)blockIdx.z * bag_stride + c0 + tid];
  // load the halo tile: 8 lanes x float4 per pixel, zero outside the image
  for (int i = tid; i < PP_H * PP_H * 8; i += 256) {
    const int pix = i >> 3, c4 = (i & 7) * 4;
    const int yy = ty0 - 3 + pix / PP_H, xx = tx0 - 3 + pix % PP_H;
    float4 val = make_float4(0.f, 0.f, 0.f, 0.f);
    if (yy >= 0 && yy < G && xx >= 0 && xx < G)
      val = __ldg(reinterpret_cast<const float4*>(xin + ((long long)yy * G + xx) * TM_DIM + c0 + c4));
    *reinterpret_cast<float4*>(tile + pix * PP_C + c4) = val;
  }
  float w[49];
#pragma unroll
  for (int t = 0; t < 49; ++t) w[t] = __ldg(packed + t * TM_DIM + c0 + lane);
  const float bias = __ldg(packed + 49 * TM_DIM + c0 + lane);
  __syncthreads();
#pragma unroll 1
  for (int ry = 0; ry < 2; ++ry) {
    const int oy = warp * 2 + ry;
    if (ty0 + oy >= G) break;
    // slide along x keeping the 7x7 window columns in registers
    float win[7][7];
#pragma unroll
    for (int dy = 0; dy < 7; ++dy)
#pragma unroll
      for (int dx = 0; dx < 6; ++dx) win[dy][dx + 1] = tile[((oy + dy) * PP_H + dx) * PP_C + lane];
#pragma unroll
    for (int ox = 0; ox < PP_T; ++ox) {
#pragma unroll
      for (int dy = 0; dy < 7; ++dy) {
#pragma unroll
        for (int dx = 0; dx < 6; ++dx) win[dy][dx] = win[dy][dx + 1];
        win[dy][6] = tile[((oy + dy) * PP_H + ox + 6) * PP_C + lane];
      }
      float a4[4] = {bias, 0.f, 0.f, 0.f};   // four independent FMA chains: the FMA pipe is the floor of this kernel
#pragma unroll
      for (int dy = 0; dy < 7; ++dy)
#pragma unroll
        for (int dx = 0; dx < 7; ++dx) a4[(dy * 7 + dx) & 3] = fmaf(w[dy * 7 + dx], win[dy][dx], a4[(dy * 7 + dx) & 3]);
      float acc = (a4[0] + a4[1]) + (a4[2] + a4[3]);
      acc += win[3][3];
      if (tx0 + ox < G) xout[((long long)(ty0 + oy) * G + tx0 + ox) * TM_DIM + c0 + lane] = acc;
    }
  }
}

}  // namespace tmil

extern "C" int tm_ppeg_fwd(const tm_weights* w, const float* h_in, int64_t B, int64_t G, float* h_out, float* packed,
                           void* stream) {
  using namespace tmil;
  TM_CHECK_ARG(w && h_in && h_out && packed && B > 0 && G > 0, TM_ERR_BAD_ARGUMENT, "tm_ppeg_fwd: bad arguments");
  TM_CHECK_ARG(h_in != h_out, TM_ERR_BAD_ARGUMENT, "tm_ppeg_fwd: in-place operation is not supported");
  cudaStream_t st = (cudaStream_t)stream;
  ppeg_pack_kernel<<<2, 256, 0, st>>>(w->ppeg_w7, w->ppeg_b7, w->ppeg_w5, w->ppeg_b5, w->ppeg_w3, w->ppeg_b3, packed);
  TM_LAUNCH_CHECK();
  const size_t smem = (size_t)PP_H * PP_H * PP_C * sizeof(float);
  static bool attr_done = false;
  if (!attr_done) {
    TM_CUDA(cudaFuncSetAttribute(ppeg_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    attr_done = true;
  }
  const int tiles = (int)(ceil_div(G, PP_T) * ceil_div(G, PP_T));
  ppeg_kernel<<<dim3(TM_DIM / PP_C, tiles, (unsigned)B), 256, smem, st>>>(h_in, packed, (int)G, (G * G + 1) * (long long)TM_DIM, h_out);
  TM_LAUNCH_CHECK();
  return TM_OK;
}
