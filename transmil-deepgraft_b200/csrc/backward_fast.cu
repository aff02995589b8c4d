// Device-tuned versions of the HBM-bound kernels of the backward.  csrc/backward.cu holds the plain
// threadIdx/__shared__/__syncthreads versions (they also run on the host through tests/cuda_emu and stay selectable with
// tm_bgemm_engine(1) as the on-device cross-check); the kernels here do the same arithmetic with shared-memory tiles, warp
// shuffles and two-stage deterministic reductions (no atomics), and are what a training step launches on sm_100.
//   conv33_bwd_fast      res_conv backward: dv and the 33 tap gradients from ONE pass over dO and v (halo tiles in smem)
//   ppeg_wgrad_fast      PPEG filter gradients: 49 taps + bias per channel from smem halo tiles, fixed-order reduction
//   softmax_rows_*_fast  row softmax / its backward with one warp per row (rows of <= 2048 elements)
//   ln512_bwd_fast       LayerNorm(512) backward with one warp per row + per-block parameter-gradient partials
//   qkv_grad_merge_fast  landmark-mean / q-scale / head-merge pass, 16 bytes per thread
#include "common.cuh"

namespace tmil {

typedef long long i64;

// ---- res_conv backward ---------------------------------------------------------------------------------------------------
// forward: out[t] += sum_j w[h,j] v[t+j-16]   =>   dv[t] = sum_j w[h,j] dO[t+16-j],   g[h,j] = sum_{b,t,d} dO[t] v[t+j-16]
constexpr int CV_T = 64;                      // tokens per block
constexpr int CV_ROWS = CV_T + 32;            // + 16-row halo on both sides
constexpr int CV_SMEM = 2 * CV_ROWS * TM_DHEAD * 4;

// grid (Sp / 64 rounded up, B*8); block 256 = 4 token groups x 64 d.  partial[h][j][b * chunks + chunk]
__global__ void __launch_bounds__(256) conv33_bwd_fast_kernel(const float* __restrict__ dO, const float* __restrict__ v,
                                                              const float* __restrict__ w, i64 Sp, int chunks, float* __restrict__ dv,
                                                              float* __restrict__ partial, i64 per_tap) {
  extern __shared__ float cv_smem[];
  float* sdo = cv_smem;                         // [96][64]: rows t0-16 .. t0+80 of dO
  float* sv = cv_smem + CV_ROWS * TM_DHEAD;     // same rows of v
  __shared__ float sw[TM_CONV_TAPS];
  __shared__ float red[8][TM_CONV_TAPS];
  const int bh = blockIdx.y, h = bh % TM_HEADS, b = bh / TM_HEADS;
  const i64 t0 = (i64)blockIdx.x * CV_T;
  const float* dOb = dO + (i64)bh * Sp * TM_DHEAD;
  const float* vb = v + (i64)bh * Sp * TM_DHEAD;
  const int tid = threadIdx.x;
  if (tid < TM_CONV_TAPS) sw[tid] = w[h * TM_CONV_TAPS + tid];
  for (int i = tid; i < CV_ROWS * TM_DHEAD / 4; i += 256) {
    const int r = i / (TM_DHEAD / 4), c4 = i % (TM_DHEAD / 4);
    const i64 t = t0 - 16 + r;
    float4 a = make_float4(0.f, 0.f, 0.f, 0.f), c = a;
    if (t >= 0 && t < Sp) {
      a = __ldg(reinterpret_cast<const float4*>(dOb + t * TM_DHEAD) + c4);
      c = __ldg(reinterpret_cast<const float4*>(vb + t * TM_DHEAD) + c4);
    }
    reinterpret_cast<float4*>(sdo)[i] = a;
    reinterpret_cast<float4*>(sv)[i] = c;
  }
  __syncthreads();
  const int d = tid & 63, tg = tid >> 6;
  float acc[TM_CONV_TAPS];
#pragma unroll
  for (int j = 0; j < TM_CONV_TAPS; ++j) acc[j] = 0.f;
  for (int tt = tg; tt < CV_T; tt += 4) {
    const i64 t = t0 + tt;
    if (t >= Sp) break;
    const float o = sdo[(tt + 16) * TM_DHEAD + d];
    float dvv = 0.f;
#pragma unroll
    for (int j = 0; j < TM_CONV_TAPS; ++j) {
      acc[j] = fmaf(o, sv[(tt + j) * TM_DHEAD + d], acc[j]);              // v[t + j - 16]
      dvv = fmaf(sw[j], sdo[(tt + 32 - j) * TM_DHEAD + d], dvv);          // dO[t + 16 - j]
    }
    dv[((i64)bh * Sp + t) * TM_DHEAD + d] = dvv;
  }
  // taps: lanes (32 of the 64 d) -> warp, 8 warps -> block, fixed order
  const int warp = tid >> 5, lane = tid & 31;
#pragma unroll
  for (int j = 0; j < TM_CONV_TAPS; ++j) {
    const float s = warp_sum(acc[j]);
    if (lane == 0) red[warp][j] = s;
  }
  __syncthreads();
  if (tid < TM_CONV_TAPS) {
    float s = 0.f;
#pragma unroll
    for (int wi = 0; wi < 8; ++wi) s += red[wi][tid];
    partial[((i64)h * TM_CONV_TAPS + tid) * per_tap + (i64)b * chunks + blockIdx.x] = s;
  }
}
// g[i] = sum of partial[i][0 .. n)   (one block per (h, j))
__global__ void __launch_bounds__(256) rows_sum_kernel(const float* __restrict__ partial, i64 n, float* __restrict__ g) {
  __shared__ float red[8];
  const float* p = partial + (i64)blockIdx.x * n;
  float a = 0.f;
  for (i64 i = threadIdx.x; i < n; i += 256) a += p[i];
  a = warp_sum(a);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = a;
  __syncthreads();
  if (threadIdx.x == 0) {
    float s = 0.f;
    for (int i = 0; i < 8; ++i) s += red[i];
    g[blockIdx.x] = s;
  }
}

i64 conv33_bwd_fast_scratch_floats(i64 B, i64 Sp) { return (i64)TM_HEADS * TM_CONV_TAPS * B * ceil_div(Sp, CV_T); }

int conv33_bwd_fast(const float* dO, const float* v, const float* conv_w, i64 B, i64 Sp, float* dv, float* g_w, float* scratch,
                    cudaStream_t st) {
  static unsigned long long attr_done = 0;
  if (first_use_on_device(&attr_done))
    TM_CUDA(cudaFuncSetAttribute(conv33_bwd_fast_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, CV_SMEM));
  const int chunks = (int)ceil_div(Sp, CV_T);
  const i64 per_tap = B * chunks;
  TM_CHECK_ARG(B * TM_HEADS <= 65535, TM_ERR_BAD_SHAPE, "conv33_bwd: too many bags");
  conv33_bwd_fast_kernel<<<dim3((unsigned)chunks, (unsigned)(B * TM_HEADS)), 256, CV_SMEM, st>>>(dO, v, conv_w, Sp, chunks, dv, scratch, per_tap);
  TM_LAUNCH_CHECK();
  rows_sum_kernel<<<TM_HEADS * TM_CONV_TAPS, 256, 0, st>>>(scratch, per_tap, g_w);
  TM_LAUNCH_CHECK();
  return TM_OK;
}

// ---- row softmax with one warp per row (len <= 1024, len % 4 == 0) -------------------------------------------------------------
template <int VEC>   // float4 per lane: len <= 128 * VEC
__global__ void __launch_bounds__(256) softmax_rows_fast_kernel(float* __restrict__ X, i64 rows, int len) {
  const i64 row = (i64)blockIdx.x * 8 + (threadIdx.x >> 5);
  if (row >= rows) return;
  const int lane = threadIdx.x & 31;
  float4* x = reinterpret_cast<float4*>(X + row * len);
  const int n4 = len >> 2;
  float4 v[VEC];
  float m = -INFINITY;
#pragma unroll
  for (int i = 0; i < VEC; ++i) {
    const int c = lane + 32 * i;
    v[i] = c < n4 ? x[c] : make_float4(-INFINITY, -INFINITY, -INFINITY, -INFINITY);
    m = fmaxf(m, fmaxf(fmaxf(v[i].x, v[i].y), fmaxf(v[i].z, v[i].w)));
  }
  m = warp_max(m);
  float s = 0.f;
#pragma unroll
  for (int i = 0; i < VEC; ++i) {
    v[i].x = expf(v[i].x - m); v[i].y = expf(v[i].y - m); v[i].z = expf(v[i].z - m); v[i].w = expf(v[i].w - m);
    s += (v[i].x + v[i].y) + (v[i].z + v[i].w);
  }
  s = warp_sum(s);
  const float inv = 1.0f / s;
#pragma unroll
  for (int i = 0; i < VEC; ++i) {
    const int c = lane + 32 * i;
    if (c < n4) x[c] = make_float4(v[i].x * inv, v[i].y * inv, v[i].z * inv, v[i].w * inv);
  }
}
// dP <- P * (dP - sum(dP * P))
template <int VEC>
__global__ void __launch_bounds__(256) softmax_rows_bwd_fast_kernel(const float* __restrict__ P, float* __restrict__ dP, i64 rows, int len) {
  const i64 row = (i64)blockIdx.x * 8 + (threadIdx.x >> 5);
  if (row >= rows) return;
  const int lane = threadIdx.x & 31;
  const float4* p = reinterpret_cast<const float4*>(P + row * len);
  float4* d = reinterpret_cast<float4*>(dP + row * len);
  const int n4 = len >> 2;
  float4 pv[VEC], dv[VEC];
  float s = 0.f;
#pragma unroll
  for (int i = 0; i < VEC; ++i) {
    const int c = lane + 32 * i;
    pv[i] = c < n4 ? p[c] : make_float4(0.f, 0.f, 0.f, 0.f);
    dv[i] = c < n4 ? d[c] : make_float4(0.f, 0.f, 0.f, 0.f);
    s = fmaf(dv[i].x, pv[i].x, s); s = fmaf(dv[i].y, pv[i].y, s); s = fmaf(dv[i].z, pv[i].z, s); s = fmaf(dv[i].w, pv[i].w, s);
  }
  s = warp_sum(s);
#pragma unroll
  for (int i = 0; i < VEC; ++i) {
    const int c = lane + 32 * i;
    if (c < n4) d[c] = make_float4(pv[i].x * (dv[i].x - s), pv[i].y * (dv[i].y - s), pv[i].z * (dv[i].z - s), pv[i].w * (dv[i].w - s));
  }
}

bool softmax_rows_fast_ok(const void* a, const void* b, i64 len) {
  return len % 4 == 0 && len <= 2048 && aligned16(a) && (b == nullptr || aligned16(b));
}
int softmax_rows_fast(float* X, i64 rows, i64 len, cudaStream_t st) {
  const unsigned grid = (unsigned)ceil_div(rows, 8);
  if (len <= 256) softmax_rows_fast_kernel<2><<<grid, 256, 0, st>>>(X, rows, (int)len);
  else if (len <= 512) softmax_rows_fast_kernel<4><<<grid, 256, 0, st>>>(X, rows, (int)len);
  else if (len <= 1024) softmax_rows_fast_kernel<8><<<grid, 256, 0, st>>>(X, rows, (int)len);
  else softmax_rows_fast_kernel<16><<<grid, 256, 0, st>>>(X, rows, (int)len);
  TM_LAUNCH_CHECK();
  return TM_OK;
}
int softmax_rows_bwd_fast(const float* P, float* dP, i64 rows, i64 len, cudaStream_t st) {
  const unsigned grid = (unsigned)ceil_div(rows, 8);
  if (len <= 256) softmax_rows_bwd_fast_kernel<2><<<grid, 256, 0, st>>>(P, dP, rows, (int)len);
  else if (len <= 512) softmax_rows_bwd_fast_kernel<4><<<grid, 256, 0, st>>>(P, dP, rows, (int)len);
  else if (len <= 1024) softmax_rows_bwd_fast_kernel<8><<<grid, 256, 0, st>>>(P, dP, rows, (int)len);
  else softmax_rows_bwd_fast_kernel<16><<<grid, 256, 0, st>>>(P, dP, rows, (int)len);
  TM_LAUNCH_CHECK();
  return TM_OK;
}

// ---- LayerNorm(512) backward, one warp per row ------------------------------------------------------------------------------------
// Same contract as ln_rows_bwd_kernel + ln_param_grad_kernel of backward.cu for C = 512: dx = (add ? add : 0) + dLN(dy); the
// parameter gradients leave as per-block partials partial[block][2][512] (sum dy*xhat, sum dy), added by chunk_sum_kernel.
__global__ void __launch_bounds__(256) ln512_bwd_fast_kernel(const float* __restrict__ X, const float* __restrict__ gamma,
                                                            const float* __restrict__ dY, i64 rows, i64 rows_per_bag, i64 dy_bag_rows,
                                                            i64 dy_row_off, const float* __restrict__ add, float* __restrict__ dX,
                                                            float* __restrict__ partial, int rows_per_block) {
  __shared__ float red[8][2][TM_DIM];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  float4 gm[4];
#pragma unroll
  for (int i = 0; i < 4; ++i) gm[i] = __ldg(reinterpret_cast<const float4*>(gamma) + lane + 32 * i);
  float4 ag[4], ab[4];
#pragma unroll
  for (int i = 0; i < 4; ++i) ag[i] = ab[i] = make_float4(0.f, 0.f, 0.f, 0.f);
  const i64 r_begin = (i64)blockIdx.x * rows_per_block;
  const i64 r_end = r_begin + rows_per_block < rows ? r_begin + rows_per_block : rows;
  for (i64 r = r_begin + warp; r < r_end; r += 8) {
    const i64 bag = r / rows_per_bag, i0 = r - bag * rows_per_bag;
    const float4* x = reinterpret_cast<const float4*>(X + r * TM_DIM);
    const float4* dy = reinterpret_cast<const float4*>(dY + (bag * dy_bag_rows + dy_row_off + i0) * TM_DIM);
    float4 xv[4], dv[4];
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      xv[i] = x[lane + 32 * i];
      dv[i] = dy[lane + 32 * i];
      s += (xv[i].x + xv[i].y) + (xv[i].z + xv[i].w);
    }
    const float mean = warp_sum(s) * (1.0f / TM_DIM);
    float v = 0.f;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      xv[i].x -= mean; xv[i].y -= mean; xv[i].z -= mean; xv[i].w -= mean;
      v = fmaf(xv[i].x, xv[i].x, v); v = fmaf(xv[i].y, xv[i].y, v); v = fmaf(xv[i].z, xv[i].z, v); v = fmaf(xv[i].w, xv[i].w, v);
    }
    const float rstd = rsqrtf(warp_sum(v) * (1.0f / TM_DIM) + 1e-5f);
    float g1 = 0.f, g2 = 0.f;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      xv[i].x *= rstd; xv[i].y *= rstd; xv[i].z *= rstd; xv[i].w *= rstd;      // xhat
      ag[i].x = fmaf(dv[i].x, xv[i].x, ag[i].x); ag[i].y = fmaf(dv[i].y, xv[i].y, ag[i].y);
      ag[i].z = fmaf(dv[i].z, xv[i].z, ag[i].z); ag[i].w = fmaf(dv[i].w, xv[i].w, ag[i].w);
      ab[i].x += dv[i].x; ab[i].y += dv[i].y; ab[i].z += dv[i].z; ab[i].w += dv[i].w;
      dv[i].x *= gm[i].x; dv[i].y *= gm[i].y; dv[i].z *= gm[i].z; dv[i].w *= gm[i].w;   // g = dy * gamma
      g1 += (dv[i].x + dv[i].y) + (dv[i].z + dv[i].w);
      g2 = fmaf(dv[i].x, xv[i].x, g2); g2 = fmaf(dv[i].y, xv[i].y, g2); g2 = fmaf(dv[i].z, xv[i].z, g2); g2 = fmaf(dv[i].w, xv[i].w, g2);
    }
    g1 = warp_sum(g1) * (1.0f / TM_DIM);
    g2 = warp_sum(g2) * (1.0f / TM_DIM);
    float4* dx = reinterpret_cast<float4*>(dX + r * TM_DIM);
    const float4* ad = add != nullptr ? reinterpret_cast<const float4*>(add + r * TM_DIM) : nullptr;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      float4 o = make_float4(rstd * (dv[i].x - g1 - xv[i].x * g2), rstd * (dv[i].y - g1 - xv[i].y * g2),
                             rstd * (dv[i].z - g1 - xv[i].z * g2), rstd * (dv[i].w - g1 - xv[i].w * g2));
      if (ad != nullptr) { const float4 a4 = ad[lane + 32 * i]; o.x += a4.x; o.y += a4.y; o.z += a4.z; o.w += a4.w; }
      dx[lane + 32 * i] = o;
    }
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    reinterpret_cast<float4*>(red[warp][0])[lane + 32 * i] = ag[i];
    reinterpret_cast<float4*>(red[warp][1])[lane + 32 * i] = ab[i];
  }
  __syncthreads();
  for (int i = threadIdx.x; i < 2 * TM_DIM; i += 256) {
    float sacc = 0.f;
#pragma unroll
    for (int wi = 0; wi < 8; ++wi) sacc += (&red[wi][0][0])[i];
    partial[(i64)blockIdx.x * 2 * TM_DIM + i] = sacc;
  }
}
__global__ void __launch_bounds__(256) chunk_sum_kernel(const float* __restrict__ partial, int chunks, i64 width, float* __restrict__ out);

constexpr int LN_MAX_BLOCKS = 512;
i64 ln512_bwd_fast_scratch_floats() { return (i64)(LN_MAX_BLOCKS + 1) * 2 * TM_DIM; }
// scratch: ln512_bwd_fast_scratch_floats() floats
int ln512_bwd_fast(const float* x, const float* gamma, const float* dy, i64 rows, i64 rows_per_bag, i64 dy_bag_rows, i64 dy_row_off,
                   const float* add, float* dx, float* g_gamma, float* g_beta, float* scratch, cudaStream_t st) {
  i64 blocks = ceil_div(rows, 64);
  if (blocks > LN_MAX_BLOCKS) blocks = LN_MAX_BLOCKS;
  const int rpb = (int)ceil_div(rows, blocks);
  blocks = ceil_div(rows, rpb);
  ln512_bwd_fast_kernel<<<(unsigned)blocks, 256, 0, st>>>(x, gamma, dy, rows, rows_per_bag, dy_bag_rows, dy_row_off, add, dx, scratch, rpb);
  TM_LAUNCH_CHECK();
  float* both = scratch + (i64)LN_MAX_BLOCKS * 2 * TM_DIM;
  chunk_sum_kernel<<<(unsigned)ceil_div(2 * TM_DIM, 256), 256, 0, st>>>(scratch, (int)blocks, (i64)2 * TM_DIM, both);
  TM_LAUNCH_CHECK();
  TM_CUDA(cudaMemcpyAsync(g_gamma, both, TM_DIM * sizeof(float), cudaMemcpyDeviceToDevice, st));
  TM_CUDA(cudaMemcpyAsync(g_beta, both + TM_DIM, TM_DIM * sizeof(float), cudaMemcpyDeviceToDevice, st));
  return TM_OK;
}

// ---- dqkv[b,t,:] = [ (dq + dq_lm[group]/l) * 64^-0.5 | dk + dk_lm[group]/l | dv ], four columns per thread ---------------------------
__global__ void __launch_bounds__(256) qkv_grad_merge_fast_kernel(const float* __restrict__ dq, const float* __restrict__ dk,
                                                                 const float* __restrict__ dv, const float* __restrict__ dq_lm,
                                                                 const float* __restrict__ dk_lm, i64 B, i64 Sp, i64 l,
                                                                 float* __restrict__ dqkv) {
  const i64 i = (i64)blockIdx.x * 256 + threadIdx.x;        // float4 index into [B,Sp,1536]
  if (i >= B * Sp * (3 * TM_DIM / 4)) return;
  const int c4 = (int)(i % (3 * TM_DIM / 4));
  const i64 t = (i / (3 * TM_DIM / 4)) % Sp, b = i / ((3 * TM_DIM / 4) * Sp);
  const int c3 = c4 * 4, which = c3 / TM_DIM, cc = c3 % TM_DIM, h = cc / TM_DHEAD, d = cc % TM_DHEAD;
  const i64 hm = ((b * TM_HEADS + h) * Sp + t) * TM_DHEAD + d;
  const i64 lm = ((b * TM_HEADS + h) * TM_LANDMARKS + t / l) * TM_DHEAD + d;
  const float inv_l = 1.0f / (float)l;
  float4 o;
  if (which == 2) {
    o = __ldg(reinterpret_cast<const float4*>(dv + hm));
  } else {
    const float4 a = __ldg(reinterpret_cast<const float4*>((which == 0 ? dq : dk) + hm));
    const float4 m = __ldg(reinterpret_cast<const float4*>((which == 0 ? dq_lm : dk_lm) + lm));
    const float sc = which == 0 ? 0.125f : 1.0f;
    o = make_float4((a.x + m.x * inv_l) * sc, (a.y + m.y * inv_l) * sc, (a.z + m.z * inv_l) * sc, (a.w + m.w * inv_l) * sc);
  }
  reinterpret_cast<float4*>(dqkv)[i] = o;
}
int qkv_grad_merge_fast(const float* dq, const float* dk, const float* dv, const float* dq_lm, const float* dk_lm, i64 B, i64 Sp, i64 l,
                        float* dqkv, cudaStream_t st) {
  const i64 n4 = B * Sp * (3 * TM_DIM / 4);
  qkv_grad_merge_fast_kernel<<<(unsigned)ceil_div(n4, 256), 256, 0, st>>>(dq, dk, dv, dq_lm, dk_lm, B, Sp, l, dqkv);
  TM_LAUNCH_CHECK();
  return TM_OK;
}

// ---- PPEG filter gradients ---------------------------------------------------------------------------------------------------
// g[tap = ky*7+kx][c] = sum_{b,y,x} dy[b,y,x,c] h_in[b, y+ky-3, x+kx-3, c] (zero outside the GxG image); g[49][c] = sum dy.
// Tile = 8 image rows x 32 columns of one bag, 32 channels (lane = channel); the h_in tile with its 3-pixel halo sits in shared
// memory, every thread keeps the 50 sums of its channel over its warp's pixels in registers.  A block walks tiles
// blockIdx.x, blockIdx.x + gridDim.x, ... and leaves ONE partial [50][32]; a second kernel adds the partials in a fixed order.
constexpr int PW_TY = 8, PW_TX = 32, PW_HX = PW_TX + 6, PW_HY = PW_TY + 6;
constexpr int PW_SMEM = PW_HY * PW_HX * 32 * 4;   // 68 096 B

__global__ void __launch_bounds__(256) ppeg_wgrad_fast_kernel(const float* __restrict__ h_in, const float* __restrict__ dy, i64 B, i64 G,
                                                             i64 tiles_y, i64 tiles_x, float* __restrict__ partial) {
  extern __shared__ float pw_smem[];   // [PW_HY][PW_HX][32]
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int c0 = blockIdx.y * 32;
  const i64 S = G * G + 1;
  const i64 tiles = B * tiles_y * tiles_x;
  float acc[50];
#pragma unroll
  for (int i = 0; i < 50; ++i) acc[i] = 0.f;
  for (i64 tile = blockIdx.x; tile < tiles; tile += gridDim.x) {
    const i64 b = tile / (tiles_y * tiles_x), rem = tile - b * tiles_y * tiles_x;
    const i64 y0 = (rem / tiles_x) * PW_TY, x0 = (rem % tiles_x) * PW_TX;
    const float* hb = h_in + (b * S + 1) * TM_DIM + c0;
    const float* db = dy + (b * S + 1) * TM_DIM + c0;
    __syncthreads();   // previous tile fully consumed
    for (int i = warp; i < PW_HY * PW_HX; i += 8) {
      const i64 yy = y0 - 3 + i / PW_HX, xx = x0 - 3 + i % PW_HX;
      pw_smem[i * 32 + lane] = (yy >= 0 && yy < G && xx >= 0 && xx < G) ? __ldg(hb + (yy * G + xx) * TM_DIM + lane) : 0.f;
    }
    __syncthreads();
    // warp = image row of the tile, 32 pixels along x
    const i64 y = y0 + warp;
    if (y < G) {
      for (int px = 0; px < PW_TX; ++px) {
        const i64 x = x0 + px;
        if (x >= G) break;
        const float d = __ldg(db + (y * G + x) * TM_DIM + lane);
        const float* win = pw_smem + (warp * PW_HX + px) * 32 + lane;
#pragma unroll
        for (int ky = 0; ky < 7; ++ky)
#pragma unroll
          for (int kx = 0; kx < 7; ++kx) acc[ky * 7 + kx] = fmaf(d, win[(ky * PW_HX + kx) * 32], acc[ky * 7 + kx]);
        acc[49] += d;
      }
    }
  }
  // 8 warps -> one partial, fixed order; the tile buffer is reused as [8][50][32]
  __syncthreads();
#pragma unroll
  for (int i = 0; i < 50; ++i) pw_smem[(warp * 50 + i) * 32 + lane] = acc[i];
  __syncthreads();
  for (int i = threadIdx.x; i < 50 * 32; i += 256) {
    float s = 0.f;
#pragma unroll
    for (int wi = 0; wi < 8; ++wi) s += pw_smem[wi * 50 * 32 + i];
    partial[((i64)blockIdx.x * 50 + i / 32) * TM_DIM + c0 + (i & 31)] = s;
  }
}
// out[i] = sum_chunk partial[chunk][i], i < width
__global__ void __launch_bounds__(256) chunk_sum_kernel(const float* __restrict__ partial, int chunks, i64 width, float* __restrict__ out) {
  const i64 i = (i64)blockIdx.x * 256 + threadIdx.x;
  if (i >= width) return;
  float s = 0.f;
  for (int c = 0; c < chunks; ++c) s += partial[(i64)c * width + i];
  out[i] = s;
}

constexpr int PW_MAX_CHUNKS = 512;   // = COLSUM_MAX_CHUNKS of backward.cu: the scratch of tm_ppeg_bwd is sized for it
// partial: PW_MAX_CHUNKS * 50 * 512 floats; gsum: 50 * 512 floats ([tap][channel], tap 49 = bias)
int ppeg_wgrad_fast(const float* h_in, const float* dy, i64 B, i64 G, float* partial, float* gsum, cudaStream_t st) {
  static unsigned long long attr_done = 0;
  if (first_use_on_device(&attr_done))
    TM_CUDA(cudaFuncSetAttribute(ppeg_wgrad_fast_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, PW_SMEM));
  const i64 tiles_y = ceil_div(G, PW_TY), tiles_x = ceil_div(G, PW_TX);
  const i64 tiles = B * tiles_y * tiles_x;
  const int chunks = (int)(tiles < PW_MAX_CHUNKS ? tiles : PW_MAX_CHUNKS);
  ppeg_wgrad_fast_kernel<<<dim3((unsigned)chunks, TM_DIM / 32), 256, PW_SMEM, st>>>(h_in, dy, B, G, tiles_y, tiles_x, partial);
  TM_LAUNCH_CHECK();
  chunk_sum_kernel<<<(unsigned)ceil_div(50 * TM_DIM, 256), 256, 0, st>>>(partial, chunks, (i64)50 * TM_DIM, gsum);
  TM_LAUNCH_CHECK();
  return TM_OK;
}

// ---- PPEG data gradient = the forward PPEG kernel (ppeg.cu: FFMA2, TMA halo boxes) with the filter flipped -----------------------
// dh[y,x] = dy[y,x] + sum_{i,j} w[i,j] dy[y+3-i, x+3-j] = x + correlation of dy with w'[i,j] = w[6-i, 6-j]; no bias term.
int ppeg_fwd_prepacked(const tm_weights* w, const float* h_in, int64_t B, int64_t G, float* h_out, float* packed, cudaStream_t st);

__global__ void __launch_bounds__(256) ppeg_pack_flipped_kernel(const float* __restrict__ w7, const float* __restrict__ w5,
                                                               const float* __restrict__ w3, float* __restrict__ packed) {
  const int i = blockIdx.x * 256 + threadIdx.x;   // (tap', channel), tap' 49 = bias
  if (i >= 50 * TM_DIM) return;
  const int tp = i / TM_DIM, c = i - tp * TM_DIM;
  if (tp == 49) { packed[i] = 0.f; return; }
  const int t = 48 - tp, dy = t / 7, dx = t - dy * 7;
  float v = w7[c * 49 + t];
  if (dy >= 1 && dy <= 5 && dx >= 1 && dx <= 5) v += w5[c * 25 + (dy - 1) * 5 + (dx - 1)];
  if (dy >= 2 && dy <= 4 && dx >= 2 && dx <= 4) v += w3[c * 9 + (dy - 2) * 3 + (dx - 2)];
  packed[i] = v;
}
// packed: 50 * 512 floats of scratch
int ppeg_dgrad_fast(const tm_weights* w, const float* dy, i64 B, i64 G, float* dh, float* packed, cudaStream_t st) {
  ppeg_pack_flipped_kernel<<<(unsigned)ceil_div(50 * TM_DIM, 256), 256, 0, st>>>(w->ppeg_w7, w->ppeg_w5, w->ppeg_w3, packed);
  TM_LAUNCH_CHECK();
  pdl_plain_next();
  return ppeg_fwd_prepacked(w, dy, B, G, dh, packed, st);
}

}  // namespace tmil
