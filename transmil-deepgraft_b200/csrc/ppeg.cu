// K13: PPEG (code/models/TransMIL.py:60-75):  y = x + dw7(x) + dw5(x) + dw3(x) on the
// GxG token image, cls row passed through.  Tokens stay channels-last [S,512]
// (token t <-> pixel (t / G, t % G), channel fastest), so the reference's two
// transposes disappear.  The three depthwise filters are pre-summed into one effective
// 7x7 per channel (exact in real arithmetic; identity added as `+ x` like the reference).
//
// Persistent kernel, one CTA per SM, 256 threads.  The token image is a 4-D TMA tensor (channel, x, y, bag): a
// [22 y][22 x][32 ch] fp32 halo box per 16x16 output tile lands in shared memory with the image border ZERO-FILLED by the
// TMA unit (= the convs' zero padding), double buffered so the next box streams in while the current one is consumed.
// A thread owns 2 channels x 8 consecutive x x 2 rows of outputs and runs the 49 taps as packed fp32 FMAs
// (fma.rn.f32x2 -> FFMA2: the only way to reach the 128 FMA/clk/SM fp32 rate on sm_100); the 49 weight pairs of its two
// channels stay in registers for the whole kernel because a CTA never changes its 32-channel slice.
// The same kernel serves the token-sharded path: it computes any token range [tok0, tok1) from an input buffer that
// holds whole image rows [y_in0, y_in1) (the rank's rows plus a 3-row halo).
#include "common.cuh"
#include "tc_common.cuh"

namespace tmil {

constexpr int PP_T = 16;            // output tile edge
constexpr int PP_H = PP_T + 6;      // input tile edge with the 3-pixel halo
constexpr int PP_C = 32;            // channels per CTA (one 128-byte line per pixel)
constexpr int PP_STAGE_FLOATS = PP_H * PP_H * PP_C;
constexpr uint32_t PP_STAGE_BYTES = PP_STAGE_FLOATS * sizeof(float);
constexpr size_t PP_SMEM = 2 * (size_t)PP_STAGE_BYTES + 64;

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

__device__ __forceinline__ unsigned long long ffma2(unsigned long long a, unsigned long long b, unsigned long long c) {
  unsigned long long d;
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c));
  return d;
}
__device__ __forceinline__ unsigned long long fadd2(unsigned long long a, unsigned long long b) {
  unsigned long long d;
  asm("add.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
  return d;
}

struct PpegArgs {
  int G;                  // image edge
  int y_in0;              // image row held at TMA y coordinate 0
  int ty0;                // first tile row origin (image row of the first output token)
  int tiles_x, tiles_y;   // spatial tiles per bag
  int B;
  long long tok0, tok1;   // output token range (per bag)
  long long out_bag_stride;  // floats
  float* out;             // row of token tok0 of bag 0
  const float* cls_in;    // row 0 of bag 0 (or NULL: no cls copy)
  float* cls_out;
  long long cls_in_stride, cls_out_stride;
};

// grid = 16 channel slices x W spatial workers; CTA (slice = blockIdx.x % 16, worker = blockIdx.x / 16) walks the spatial
// tiles worker, worker + W, ...  Consecutive CTAs are the 16 slices of ONE spatial tile, so the CTAs running together
// read whole 2 KB token rows.
__global__ void __launch_bounds__(256, 1) ppeg_kernel(const __grid_constant__ CUtensorMap tm_in, const float* __restrict__ packed,
                                                      const PpegArgs a) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  float* stage0 = reinterpret_cast<float*>(smem_raw);
  uint64_t* full = reinterpret_cast<uint64_t*>(smem_raw + 2 * (size_t)PP_STAGE_BYTES);
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int slice = blockIdx.x & 15, worker = blockIdx.x >> 4, nworkers = gridDim.x >> 4;
  const int c0 = slice * PP_C;
  const int cp = lane & 15, xg = lane >> 4;
  const int per_bag = a.tiles_x * a.tiles_y, ntiles = per_bag * a.B;

  if (tid == 0) {
    tma_prefetch_desc(&tm_in);
    mbar_init(&full[0], 1);
    mbar_init(&full[1], 1);
    fence_barrier_init();
  }
  __syncthreads();
  pdl_sync();
  if (a.cls_in != nullptr && worker == 0 && tid < PP_C)  // cls token passes through (TransMIL.py:67,74)
    for (int b = 0; b < a.B; ++b) a.cls_out[b * a.cls_out_stride + c0 + tid] = a.cls_in[b * a.cls_in_stride + c0 + tid];
  auto issue = [&](int tile, int s) {
    const int b = tile / per_bag, r = tile % per_bag;
    const int y0 = a.ty0 + (r / a.tiles_x) * PP_T, x0 = (r % a.tiles_x) * PP_T;
    mbar_arrive_expect_tx(&full[s], PP_STAGE_BYTES);
    tma_load_4d(stage0 + (size_t)s * PP_STAGE_FLOATS, &tm_in, &full[s], c0, x0 - 3, y0 - 3 - a.y_in0, b);
  };
  if (tid == 0 && worker < ntiles) issue(worker, 0);

  // the 49 weight pairs + bias pair of this thread's two channels
  unsigned long long w[49];
#pragma unroll
  for (int t = 0; t < 49; ++t) w[t] = __ldg(reinterpret_cast<const unsigned long long*>(packed + t * TM_DIM + c0 + 2 * cp));
  const unsigned long long bias = __ldg(reinterpret_cast<const unsigned long long*>(packed + 49 * TM_DIM + c0 + 2 * cp));

  int it = 0;
  for (int tile = worker; tile < ntiles; tile += nworkers, ++it) {
    const int s = it & 1;
    if (tid == 0 && tile + nworkers < ntiles) issue(tile + nworkers, s ^ 1);   // stage s^1 was released by the barrier below
    mbar_wait(&full[s], (it >> 1) & 1);
    const float* tl = stage0 + (size_t)s * PP_STAGE_FLOATS;
    const int b = tile / per_bag, r = tile % per_bag;
    const int y0 = a.ty0 + (r / a.tiles_x) * PP_T, x0 = (r % a.tiles_x) * PP_T;
    const int oy = warp * 2;                       // tile rows oy, oy+1
    unsigned long long acc0[8], acc1[8];
#pragma unroll
    for (int o = 0; o < 8; ++o) { acc0[o] = bias; acc1[o] = bias; }
#pragma unroll
    for (int iy = 0; iy < 8; ++iy) {
      unsigned long long in[14];
      const unsigned long long* src = reinterpret_cast<const unsigned long long*>(tl + ((oy + iy) * PP_H + xg * 8) * PP_C + 2 * cp);
#pragma unroll
      for (int j = 0; j < 14; ++j) in[j] = src[j * (PP_C / 2)];
      if (iy < 7) {
#pragma unroll
        for (int dx = 0; dx < 7; ++dx)
#pragma unroll
          for (int o = 0; o < 8; ++o) acc0[o] = ffma2(w[iy * 7 + dx], in[o + dx], acc0[o]);
      }
      if (iy >= 1) {
#pragma unroll
        for (int dx = 0; dx < 7; ++dx)
#pragma unroll
          for (int o = 0; o < 8; ++o) acc1[o] = ffma2(w[(iy - 1) * 7 + dx], in[o + dx], acc1[o]);
      }
      if (iy == 3) {
#pragma unroll
        for (int o = 0; o < 8; ++o) acc0[o] = fadd2(acc0[o], in[o + 3]);   // + x (TransMIL.py:71)
      }
      if (iy == 4) {
#pragma unroll
        for (int o = 0; o < 8; ++o) acc1[o] = fadd2(acc1[o], in[o + 3]);
      }
    }
    float* ob = a.out + (long long)b * a.out_bag_stride + c0 + 2 * cp;
#pragma unroll
    for (int ry = 0; ry < 2; ++ry) {
      const int y = y0 + oy + ry;
      if (y < a.G) {
#pragma unroll
        for (int o = 0; o < 8; ++o) {
          const int x = x0 + xg * 8 + o;
          const long long t = (long long)y * a.G + x;
          if (x < a.G && t >= a.tok0 && t < a.tok1)
            *reinterpret_cast<unsigned long long*>(ob + (t - a.tok0) * TM_DIM) = ry ? acc1[o] : acc0[o];
        }
      }
    }
    __syncthreads();   // every thread is done with stage s before thread 0 refills it two iterations later
  }
}

static int ppeg_launch(const tm_weights* w, const float* in_rows, int64_t y_in0, int64_t y_in1, long long in_bag_stride, int64_t B,
                       int64_t G, int64_t tok0, int64_t tok1, float* out, long long out_bag_stride, const float* cls_in,
                       long long cls_in_stride, float* cls_out, long long cls_out_stride, float* packed, cudaStream_t st,
                       bool packed_ready = false) {
  TM_CHECK_ARG(tc_available(), TM_ERR_UNSUPPORTED_ARCH, "tm_ppeg_fwd needs an sm_100 device (TMA)");
  if (!packed_ready) {
    ppeg_pack_kernel<<<2, 256, 0, st>>>(w->ppeg_w7, w->ppeg_b7, w->ppeg_w5, w->ppeg_b5, w->ppeg_w3, w->ppeg_b3, packed);
    TM_LAUNCH_CHECK();
  }
  CUtensorMap tm;
  const uint64_t dims[4] = {(uint64_t)TM_DIM, (uint64_t)G, (uint64_t)(y_in1 - y_in0), (uint64_t)B};
  if (in_bag_stride <= 0) in_bag_stride = (y_in1 - y_in0) * G * (long long)TM_DIM;   // single bag: any valid pitch
  const uint64_t strides[3] = {TM_DIM * sizeof(float), (uint64_t)G * TM_DIM * sizeof(float), (uint64_t)in_bag_stride * sizeof(float)};
  const uint32_t box[4] = {PP_C, PP_H, PP_H, 1};
  TM_PROPAGATE(make_tmap_f32_nd(&tm, in_rows, 4, dims, strides, box));
  static unsigned long long attr_done = 0;
  if (first_use_on_device(&attr_done)) {
    TM_CUDA(cudaFuncSetAttribute(ppeg_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)PP_SMEM));
  }
  PpegArgs a;
  a.G = (int)G; a.y_in0 = (int)y_in0;
  const int64_t y_first = tok0 / G, y_last = (tok1 - 1) / G;
  a.ty0 = (int)y_first;
  a.tiles_x = (int)ceil_div(G, PP_T);
  a.tiles_y = (int)ceil_div(y_last - y_first + 1, PP_T);
  a.B = (int)B; a.tok0 = tok0; a.tok1 = tok1; a.out_bag_stride = out_bag_stride; a.out = out;
  a.cls_in = cls_in; a.cls_out = cls_out; a.cls_in_stride = cls_in_stride; a.cls_out_stride = cls_out_stride;
  const long long ntiles = (long long)a.tiles_x * a.tiles_y * B;
  int sms = 148;
  {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  }
  long long workers = sms / 16;                  // 9 on a 148-SM B200: 144 CTAs, each with a fixed channel slice
  if (workers < 1) workers = 1;
  if (workers > ntiles) workers = ntiles;
  launch_pdl(ppeg_kernel, dim3((unsigned)(16 * workers)), dim3(256), PP_SMEM, st, tm, (const float*)packed, a);
  TM_LAUNCH_CHECK();
  return TM_OK;
}

// tm_ppeg_fwd with `packed` already filled by prep_operands (tm_transmil_fwd packs every parameter copy in one launch)
int ppeg_fwd_prepacked(const tm_weights* w, const float* h_in, int64_t B, int64_t G, float* h_out, float* packed, cudaStream_t st) {
  const long long bag = (G * G + 1) * (long long)TM_DIM;
  return ppeg_launch(w, h_in + TM_DIM, 0, G, bag, B, G, 0, G * G, h_out + TM_DIM, bag, h_in, bag, h_out, bag, packed, st, true);
}

// tm_ppeg_rows_fwd with `packed` already filled by prep_operands (sharded_fwd.cu)
int ppeg_rows_prepacked(const tm_weights* w, const float* in_rows, int64_t y_in0, int64_t y_in1, int64_t G, int64_t tok0, int64_t tok1,
                        float* out, float* packed, cudaStream_t st) {
  return ppeg_launch(w, in_rows, y_in0, y_in1, 0, 1, G, tok0, tok1, out, 0, nullptr, 0, nullptr, 0, packed, st, true);
}

}  // namespace tmil

extern "C" int tm_ppeg_fwd(const tm_weights* w, const float* h_in, int64_t B, int64_t G, float* h_out, float* packed,
                           void* stream) {
  using namespace tmil;
  TM_CHECK_ARG(w && h_in && h_out && packed && B > 0 && G > 0, TM_ERR_BAD_ARGUMENT, "tm_ppeg_fwd: bad arguments");
  TM_CHECK_ARG(h_in != h_out, TM_ERR_BAD_ARGUMENT, "tm_ppeg_fwd: in-place operation is not supported");
  const long long bag = (G * G + 1) * (long long)TM_DIM;
  return ppeg_launch(w, h_in + TM_DIM, 0, G, bag, B, G, 0, G * G, h_out + TM_DIM, bag, h_in, bag, h_out, bag, packed,
                     (cudaStream_t)stream);
}

// Token-sharded PPEG: output tokens [tok0, tok1) of ONE bag from `in_rows`, which holds the whole image rows
// [y_in0, y_in1) (y_in0 <= max(0, tok0/G - 3), y_in1 >= min(G, (tok1-1)/G + 4): the rank's own rows plus the 3-row halo
// received from its neighbours).  out points at the row of token tok0.
extern "C" int tm_ppeg_rows_fwd(const tm_weights* w, const float* in_rows, int64_t y_in0, int64_t y_in1, int64_t G, int64_t tok0,
                                int64_t tok1, float* out, float* packed, void* stream) {
  using namespace tmil;
  TM_CHECK_ARG(w && in_rows && out && packed && G > 0, TM_ERR_BAD_ARGUMENT, "tm_ppeg_rows_fwd: bad arguments");
  TM_CHECK_ARG(0 <= tok0 && tok0 < tok1 && tok1 <= G * G && 0 <= y_in0 && y_in0 < y_in1 && y_in1 <= G, TM_ERR_BAD_SHAPE,
               "tm_ppeg_rows_fwd: bad token / row range");
  const int64_t need0 = tok0 / G - 3 > 0 ? tok0 / G - 3 : 0, need1 = (tok1 - 1) / G + 4 < G ? (tok1 - 1) / G + 4 : G;
  TM_CHECK_ARG(y_in0 <= need0 && y_in1 >= need1, TM_ERR_BAD_SHAPE, "tm_ppeg_rows_fwd: input rows [%lld,%lld) do not cover the 3-row halo [%lld,%lld)",
               (long long)y_in0, (long long)y_in1, (long long)need0, (long long)need1);
  return ppeg_launch(w, in_rows, y_in0, y_in1, 0, 1, G, tok0, tok1, out, 0, nullptr, 0, nullptr, 0, packed, (cudaStream_t)stream);
}
