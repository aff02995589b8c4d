// Backward of the TransMIL bag-level path: the entry points, and the fp32 CUDA-core kernels behind them.
// When the tensor-core engine is on (the default on a B200) the contractions go to bgemm_tc.cu (tcgen05, tf32 operands) and
// the HBM-bound parts to backward_fast.cu; the plain kernels of this file are then the on-device cross-check
// (tm_bgemm_engine(1)) and, through tests/cuda_emu, the host-executable form of the same arithmetic.
//
// The reference has no backward code: gradients come from torch.autograd through code/models/TransMIL.py:167-211 and
// nystrom-attention 0.0.11.  The contractions and their order are those written out in oracle/transmil_backward.py; the
// host side (transmil-deepgraft_b200/training.py) strings the entry points below together.
//
// Every kernel here is plain SIMT (no warp intrinsics, no atomics: reductions are two-stage and deterministic), so the
// same source also compiles with g++ against tests/cuda_emu/cuda_emu.h (-DTM_HOST_EMU), which is how the CPU test suite
// checks indexing and arithmetic against the oracle in the GPU-less build container.  That harness is test
// infrastructure; the library built by nvcc contains device code only.
#ifdef TM_HOST_EMU
#include "cuda_emu.h"
#else
#include "common.cuh"
#endif

namespace tmil {

#ifdef TM_HOST_EMU
#define BW_LAUNCH(kern, grid, block, st, ...)                     \
  do {                                                            \
    (void)(st);                                                   \
    emu::launch(grid, block, [=]() { kern(__VA_ARGS__); });       \
  } while (0)
#else
#define BW_LAUNCH(kern, grid, block, st, ...)          \
  do {                                                 \
    kern<<<grid, block, 0, st>>>(__VA_ARGS__);         \
    TM_LAUNCH_CHECK();                                 \
  } while (0)
#endif

typedef long long i64;
constexpr int BT = 256;        // threads per block of every kernel here unless stated
constexpr int LMK = TM_LANDMARKS;
constexpr i64 MM = (i64)LMK * LMK;

static inline dim3 grid1d(i64 n) { return dim3((unsigned)ceil_div(n, BT)); }

// ---- block reductions (blockDim.x threads, shared scratch of blockDim.x floats) -------------------------------------
__device__ __forceinline__ float block_sum(float v, float* red) {
  const int t = threadIdx.x;
  red[t] = v;
  __syncthreads();
  for (int s = blockDim.x >> 1; s > 0; s >>= 1) {
    if (t < s) red[t] += red[t + s];
    __syncthreads();
  }
  const float r = red[0];
  __syncthreads();
  return r;
}
__device__ __forceinline__ float block_max(float v, float* red) {
  const int t = threadIdx.x;
  red[t] = v;
  __syncthreads();
  for (int s = blockDim.x >> 1; s > 0; s >>= 1) {
    if (t < s) red[t] = fmaxf(red[t], red[t + s]);
    __syncthreads();
  }
  const float r = red[0];
  __syncthreads();
  return r;
}

// ---- generic strided batched GEMM: C[z] = alpha * op(A[z]) op(B[z]) + beta * L[z] + diag * I ---------------------------
// Element strides make every transposition a view: A[m,k] = A[m*sam + k*sak], B[k,n] = B[k*sbk + n*sbn].
struct BGemm {
  const float* A; const float* B; const float* L; float* C;
  int M, N, K;
  i64 sam, sak, sbk, sbn, ldc, ldl;
  i64 strA, strB, strC, strL;
  float alpha, beta, diag;
  float* C2; float alpha2, diag2;   // optional second output from the same product: C2 = alpha2 * op(A) op(B) + diag2 * I (ldc / strC of C)
};
constexpr int GT = 64, GK = 16;

__global__ void __launch_bounds__(BT) bgemm_kernel(const BGemm p) {
  __shared__ float As[GK][GT + 1];
  __shared__ float Bs[GK][GT + 1];
  const int tid = threadIdx.x;
  const int m0 = blockIdx.x * GT, n0 = blockIdx.y * GT;
  const i64 z = blockIdx.z;
  const float* A = p.A + z * p.strA;
  const float* B = p.B + z * p.strB;
  const bool a_kfast = p.sak == 1, b_kfast = p.sbk == 1;
  const int tx = tid & 15, ty = tid >> 4;
  float acc[4][4];
  for (int i = 0; i < 4; ++i)
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;
  for (int k0 = 0; k0 < p.K; k0 += GK) {
    for (int idx = tid; idx < GT * GK; idx += BT) {
      int mm, kk;
      if (a_kfast) { kk = idx & (GK - 1); mm = idx >> 4; } else { mm = idx & (GT - 1); kk = idx >> 6; }
      int gm = m0 + mm, gk = k0 + kk;
      As[kk][mm] = (gm < p.M && gk < p.K) ? A[(i64)gm * p.sam + (i64)gk * p.sak] : 0.f;
      int nn;
      if (b_kfast) { kk = idx & (GK - 1); nn = idx >> 4; } else { nn = idx & (GT - 1); kk = idx >> 6; }
      int gn = n0 + nn;
      gk = k0 + kk;
      Bs[kk][nn] = (gn < p.N && gk < p.K) ? B[(i64)gk * p.sbk + (i64)gn * p.sbn] : 0.f;
    }
    __syncthreads();
    for (int kk = 0; kk < GK; ++kk) {
      float a[4], b[4];
      for (int i = 0; i < 4; ++i) a[i] = As[kk][ty * 4 + i];
      for (int j = 0; j < 4; ++j) b[j] = Bs[kk][tx * 4 + j];
      for (int i = 0; i < 4; ++i)
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
    }
    __syncthreads();
  }
  const float* L = p.L != nullptr ? p.L + z * p.strL : nullptr;
  float* C = p.C + z * p.strC;
  float* C2 = p.C2 != nullptr ? p.C2 + z * p.strC : nullptr;
  for (int i = 0; i < 4; ++i) {
    const int r = m0 + ty * 4 + i;
    if (r >= p.M) continue;
    for (int j = 0; j < 4; ++j) {
      const int c = n0 + tx * 4 + j;
      if (c >= p.N) continue;
      float o = p.alpha * acc[i][j];
      if (L != nullptr) o = fmaf(p.beta, L[(i64)r * p.ldl + c], o);
      if (r == c) o += p.diag;
      C[(i64)r * p.ldc + c] = o;
      if (C2 != nullptr) C2[(i64)r * p.ldc + c] = p.alpha2 * acc[i][j] + (r == c ? p.diag2 : 0.f);
    }
  }
}

#ifndef TM_HOST_EMU
// tcgen05 engine of the same contract (bgemm_tc.cu): tf32 operands, fp32 accumulate
i64 bgemm_tc_scratch_floats(i64 M, i64 N, i64 K, i64 batch);
int bgemm_tc(const float* A, const float* B, float* C, const float* L, i64 M, i64 N, i64 K, i64 sam, i64 sak, i64 sbk, i64 sbn, i64 ldc,
             i64 ldl, i64 batch, i64 strA, i64 strB, i64 strC, i64 strL, float alpha, float beta, float diag, float* C2, float alpha2,
             float diag2, float* scratch, i64 scratch_floats, cudaStream_t st);
// 0 = tensor cores when the device has them (default), 1 = fp32 CUDA-core kernel (the on-device cross-check); TM_BGEMM=simt
// in the environment sets the initial value.
static int g_bgemm_engine = [] { const char* e = getenv("TM_BGEMM"); return (e && strcmp(e, "simt") == 0) ? 1 : 0; }();
extern int g_bgemm_force_cpasync;
// device-tuned versions of the HBM-bound kernels (backward_fast.cu); the plain ones below stay the cross-check (engine 1)
i64 conv33_bwd_fast_scratch_floats(i64 B, i64 Sp);
int conv33_bwd_fast(const float* dO, const float* v, const float* conv_w, i64 B, i64 Sp, float* dv, float* g_w, float* scratch, cudaStream_t st);
bool softmax_rows_fast_ok(const void* a, const void* b, i64 len);
int softmax_rows_fast(float* X, i64 rows, i64 len, cudaStream_t st);
int softmax_rows_bwd_fast(const float* P, float* dP, i64 rows, i64 len, cudaStream_t st);
int ppeg_wgrad_fast(const float* h_in, const float* dy, i64 B, i64 G, float* partial, float* gsum, cudaStream_t st);
i64 ln512_bwd_fast_scratch_floats();
int ln512_bwd_fast(const float* x, const float* gamma, const float* dy, i64 rows, i64 rows_per_bag, i64 dy_bag_rows, i64 dy_row_off,
                   const float* add, float* dx, float* g_gamma, float* g_beta, float* scratch, cudaStream_t st);
int qkv_grad_merge_fast(const float* dq, const float* dk, const float* dv, const float* dq_lm, const float* dk_lm, i64 B, i64 Sp, i64 l,
                        float* dqkv, cudaStream_t st);
int ppeg_dgrad_fast(const tm_weights* w, const float* dy, i64 B, i64 G, float* dh, float* packed, cudaStream_t st);
static bool bgemm_use_tc() { return g_bgemm_engine != 1 && tc_available(); }
#endif

static int bgemm(const BGemm& p, int batch, cudaStream_t st, float* scratch = nullptr, i64 scratch_floats = 0) {
  TM_CHECK_ARG(p.M > 0 && p.N > 0 && p.K > 0 && batch > 0, TM_ERR_BAD_SHAPE, "tm_bgemm: empty problem %dx%dx%d x%d", p.M, p.N, p.K, batch);
  TM_CHECK_ARG(p.A && p.B && p.C, TM_ERR_BAD_ARGUMENT, "tm_bgemm: NULL operand");
#ifndef TM_HOST_EMU
  if (bgemm_use_tc())
    return bgemm_tc(p.A, p.B, p.C, p.L, p.M, p.N, p.K, p.sam, p.sak, p.sbk, p.sbn, p.ldc, p.ldl, batch, p.strA, p.strB, p.strC, p.strL,
                    p.alpha, p.beta, p.diag, p.C2, p.alpha2, p.diag2, scratch, scratch_floats, st);
#else
  (void)scratch; (void)scratch_floats;
#endif
  dim3 grid((unsigned)ceil_div(p.M, GT), (unsigned)ceil_div(p.N, GT), (unsigned)batch);
  TM_CHECK_ARG(grid.y <= 65535u && grid.z <= 65535u, TM_ERR_BAD_SHAPE, "tm_bgemm: grid too large (N=%d batch=%d)", p.N, batch);
  BW_LAUNCH(bgemm_kernel, grid, dim3(BT), st, p);
  return TM_OK;
}

// ---- row softmax (recomputing attn1 / attn3) and its backward ------------------------------------------------------------
__global__ void __launch_bounds__(BT) softmax_rows_kernel(float* X, int len) {
  __shared__ float red[BT];
  float* x = X + (i64)blockIdx.x * len;
  float m = -INFINITY;
  for (int i = threadIdx.x; i < len; i += BT) m = fmaxf(m, x[i]);
  m = block_max(m, red);
  float s = 0.f;
  for (int i = threadIdx.x; i < len; i += BT) s += expf(x[i] - m);
  s = block_sum(s, red);
  const float inv = 1.0f / s;
  for (int i = threadIdx.x; i < len; i += BT) x[i] = expf(x[i] - m) * inv;
}
// dP <- P * (dP - sum(dP * P))
__global__ void __launch_bounds__(BT) softmax_rows_bwd_kernel(const float* P, float* dP, int len) {
  __shared__ float red[BT];
  const float* p = P + (i64)blockIdx.x * len;
  float* d = dP + (i64)blockIdx.x * len;
  float s = 0.f;
  for (int i = threadIdx.x; i < len; i += BT) s = fmaf(d[i], p[i], s);
  s = block_sum(s, red);
  for (int i = threadIdx.x; i < len; i += BT) d[i] = p[i] * (d[i] - s);
}

// ---- LayerNorm over the last axis ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(BT) ln_rows_fwd_kernel(const float* X, const float* gamma, const float* beta, int C, float* Y) {
  __shared__ float red[BT];
  const float* x = X + (i64)blockIdx.x * C;
  float* y = Y + (i64)blockIdx.x * C;
  float s = 0.f;
  for (int i = threadIdx.x; i < C; i += BT) s += x[i];
  const float mean = block_sum(s, red) / C;
  float v = 0.f;
  for (int i = threadIdx.x; i < C; i += BT) { const float d = x[i] - mean; v = fmaf(d, d, v); }
  const float rstd = rsqrtf(block_sum(v, red) / C + 1e-5f);
  for (int i = threadIdx.x; i < C; i += BT) y[i] = (x[i] - mean) * rstd * gamma[i] + beta[i];
}
// Row r of x (bag = r / rows_per_bag, i = r % rows_per_bag) pairs with row bag*dy_bag_rows + dy_row_off + i of dy
// (the layer's dxn lives on the front-padded axis [B,Sp,512]; x is [B,S,512]).  dx = (add ? add : 0) + LN backward.
__global__ void __launch_bounds__(BT) ln_rows_bwd_kernel(const float* X, const float* gamma, const float* dY, int C, i64 rows_per_bag,
                                                         i64 dy_bag_rows, i64 dy_row_off, const float* add, float* dX, float* stats) {
  __shared__ float red[BT];
  const i64 r = blockIdx.x;
  const i64 bag = r / rows_per_bag, i0 = r - bag * rows_per_bag;
  const float* x = X + r * C;
  const float* dy = dY + (bag * dy_bag_rows + dy_row_off + i0) * C;
  float s = 0.f;
  for (int i = threadIdx.x; i < C; i += BT) s += x[i];
  const float mean = block_sum(s, red) / C;
  float v = 0.f;
  for (int i = threadIdx.x; i < C; i += BT) { const float d = x[i] - mean; v = fmaf(d, d, v); }
  const float rstd = rsqrtf(block_sum(v, red) / C + 1e-5f);
  float g1 = 0.f, g2 = 0.f;
  for (int i = threadIdx.x; i < C; i += BT) {
    const float g = dy[i] * gamma[i];
    g1 += g;
    g2 = fmaf(g, (x[i] - mean) * rstd, g2);
  }
  g1 = block_sum(g1, red) / C;
  g2 = block_sum(g2, red) / C;
  for (int i = threadIdx.x; i < C; i += BT) {
    const float xh = (x[i] - mean) * rstd;
    const float o = rstd * (dy[i] * gamma[i] - g1 - xh * g2);
    dX[r * C + i] = add != nullptr ? add[r * C + i] + o : o;
  }
  if (threadIdx.x == 0) { stats[2 * r] = mean; stats[2 * r + 1] = rstd; }
}
// partial[chunk][0][c] = sum_r dy*xhat, partial[chunk][1][c] = sum_r dy over the chunk's rows
__global__ void __launch_bounds__(BT) ln_param_grad_kernel(const float* X, const float* dY, const float* stats, i64 rows, int C,
                                                           i64 rows_per_bag, i64 dy_bag_rows, i64 dy_row_off, i64 rows_per_chunk,
                                                           float* partial) {
  const int c = blockIdx.x * BT + threadIdx.x;
  if (c >= C) return;
  const i64 r0 = (i64)blockIdx.y * rows_per_chunk;
  const i64 r1 = r0 + rows_per_chunk < rows ? r0 + rows_per_chunk : rows;
  float ag = 0.f, ab = 0.f;
  for (i64 r = r0; r < r1; ++r) {
    const i64 bag = r / rows_per_bag, i0 = r - bag * rows_per_bag;
    const float d = dY[(bag * dy_bag_rows + dy_row_off + i0) * C + c];
    ag = fmaf(d, (X[r * C + c] - stats[2 * r]) * stats[2 * r + 1], ag);
    ab += d;
  }
  partial[((i64)blockIdx.y * 2 + 0) * C + c] = ag;
  partial[((i64)blockIdx.y * 2 + 1) * C + c] = ab;
}

// ---- column sums: out[chunk][c] = sum over the chunk's rows of X[r*ld + c] ---------------------------------------------------
__global__ void __launch_bounds__(BT) colsum_kernel(const float* X, i64 rows, i64 cols, i64 ld, i64 rows_per_chunk, float* out) {
  const i64 c = (i64)blockIdx.x * BT + threadIdx.x;
  if (c >= cols) return;
  const i64 r0 = (i64)blockIdx.y * rows_per_chunk;
  const i64 r1 = r0 + rows_per_chunk < rows ? r0 + rows_per_chunk : rows;
  float a = 0.f;
  for (i64 r = r0; r < r1; ++r) a += X[r * ld + c];
  out[(i64)blockIdx.y * cols + c] = a;
}
constexpr i64 COLSUM_MAX_CHUNKS = 512;
static i64 colsum_chunks(i64 rows) {
  i64 n = ceil_div(rows, 128);
  return n > COLSUM_MAX_CHUNKS ? COLSUM_MAX_CHUNKS : (n < 1 ? 1 : n);
}
// scratch: colsum_chunks(rows) * cols floats (unused when a single chunk suffices)
static int colsum(const float* X, i64 rows, i64 cols, i64 ld, float* out, float* scratch, cudaStream_t st) {
  const i64 nch = colsum_chunks(rows);
  if (nch == 1) {
    BW_LAUNCH(colsum_kernel, dim3((unsigned)ceil_div(cols, BT), 1), dim3(BT), st, X, rows, cols, ld, rows, out);
    return TM_OK;
  }
  const i64 rpc = ceil_div(rows, nch);
  BW_LAUNCH(colsum_kernel, dim3((unsigned)ceil_div(cols, BT), (unsigned)nch), dim3(BT), st, X, rows, cols, ld, rpc, scratch);
  BW_LAUNCH(colsum_kernel, dim3((unsigned)ceil_div(cols, BT), 1), dim3(BT), st, (const float*)scratch, nch, cols, cols, nch, out);
  return TM_OK;
}

// ---- elementwise ---------------------------------------------------------------------------------------------------------
__device__ __forceinline__ float gelu_f(float x) { return 0.5f * x * (1.0f + erff(x * 0.70710678118654752440f)); }
__global__ void __launch_bounds__(BT) gelu_fwd_kernel(const float* z, float* a, i64 n) {
  const i64 i = (i64)blockIdx.x * BT + threadIdx.x;
  if (i < n) a[i] = gelu_f(z[i]);
}
// d/dz [z Phi(z)] = Phi(z) + z phi(z)   (nn.GELU default, erf form)
__global__ void __launch_bounds__(BT) gelu_bwd_kernel(const float* z, const float* dy, float* dz, i64 n) {
  const i64 i = (i64)blockIdx.x * BT + threadIdx.x;
  if (i >= n) return;
  const float x = z[i];
  const float cdf = 0.5f * (1.0f + erff(x * 0.70710678118654752440f));
  const float pdf = expf(-0.5f * x * x) * 0.39894228040143267794f;
  dz[i] = dy[i] * (cdf + x * pdf);
}

// Counter-based keep mask of the to_out Dropout (training mode): element `idx` of the [B,S,512] activation is kept when
// u(seed, idx) >= p.  splitmix64 finaliser; forward and backward regenerate the same mask from (seed, idx).
__device__ __forceinline__ float keep_scale(uint64_t seed, uint64_t idx, float p) {
  uint64_t x = seed + 0x9E3779B97F4A7C15ull * (idx + 1);
  x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
  x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
  x ^= x >> 31;
  const float u = (float)(x >> 40) * (1.0f / 16777216.0f);
  return u >= p ? 1.0f / (1.0f - p) : 0.f;
}
// y <- h_in + keep * (y - h_in)   (y = h_in + to_out(...) as written by tm_to_out_residual_fwd)
__global__ void __launch_bounds__(BT) dropout_resid_kernel(float* y, const float* h_in, i64 n, uint64_t seed, float p) {
  const i64 i = (i64)blockIdx.x * BT + threadIdx.x;
  if (i < n) y[i] = h_in[i] + keep_scale(seed, (uint64_t)i, p) * (y[i] - h_in[i]);
}
// y <- keep * x: the Dropout layers INSIDE the 768-d stem (TransMIL.py:123-126) in training mode, forward (activations) and
// backward (their gradients) alike; x may alias y
__global__ void __launch_bounds__(BT) dropout_kernel(const float* x, float* y, i64 n, uint64_t seed, float p) {
  const i64 i = (i64)blockIdx.x * BT + threadIdx.x;
  if (i < n) y[i] = keep_scale(seed, (uint64_t)i, p) * x[i];
}
// out[b, r, c] = r < P ? 0 : dy[b, r-P, c] (* keep mask when p > 0): gradient entering to_out on the front-padded axis
__global__ void __launch_bounds__(BT) pad_front_kernel(const float* dy, i64 B, i64 S, i64 P, int C, uint64_t seed, float p, float* out) {
  const i64 i = (i64)blockIdx.x * BT + threadIdx.x;
  const i64 Sp = S + P;
  if (i >= B * Sp * C) return;
  const int c = (int)(i % C);
  const i64 r = (i / C) % Sp, b = i / (C * Sp);
  if (r < P) { out[i] = 0.f; return; }
  const i64 src = (b * S + (r - P)) * C + c;
  float v = dy[src];
  if (p > 0.f) v *= keep_scale(seed, (uint64_t)src, p);
  out[i] = v;
}
// 'b n (h d) -> b h n d'
__global__ void __launch_bounds__(BT) split_heads_kernel(const float* m, i64 B, i64 Sp, float* o) {
  const i64 i = (i64)blockIdx.x * BT + threadIdx.x;
  if (i >= B * Sp * TM_DIM) return;
  const int d = (int)(i % TM_DHEAD);
  const i64 t = (i / TM_DHEAD) % Sp;
  const int h = (int)((i / (TM_DHEAD * Sp)) % TM_HEADS);
  const i64 b = i / (TM_DHEAD * Sp * TM_HEADS);
  o[i] = m[(b * Sp + t) * TM_DIM + h * TM_DHEAD + d];
}
// dqkv[b,t,:] = [ (dq + dq_lm[group]/l) * 64^-0.5 | dk + dk_lm[group]/l | dv ]  (landmark means + head merge + q scale)
__global__ void __launch_bounds__(BT) qkv_grad_merge_kernel(const float* dq, const float* dk, const float* dv, const float* dq_lm,
                                                            const float* dk_lm, i64 B, i64 Sp, i64 l, float* dqkv) {
  const i64 i = (i64)blockIdx.x * BT + threadIdx.x;
  if (i >= B * Sp * 3 * TM_DIM) return;
  const int c3 = (int)(i % (3 * TM_DIM));
  const i64 t = (i / (3 * TM_DIM)) % Sp, b = i / (3 * TM_DIM * Sp);
  const int which = c3 / TM_DIM, cc = c3 % TM_DIM, h = cc / TM_DHEAD, d = cc % TM_DHEAD;
  const i64 hm = ((b * TM_HEADS + h) * Sp + t) * TM_DHEAD + d;
  const i64 lm = ((b * TM_HEADS + h) * LMK + t / l) * TM_DHEAD + d;
  const float inv_l = 1.0f / (float)l;
  float o;
  if (which == 0) o = (dq[hm] + dq_lm[lm] * inv_l) * 0.125f;
  else if (which == 1) o = dk[hm] + dk_lm[lm] * inv_l;
  else o = dv[hm];
  dqkv[i] = o;
}

// ---- res_conv (33-tap depthwise conv along tokens) ----------------------------------------------------------------------------
// forward: out[t] += sum_j w[h,j] v[t+j-16]  =>  dv[t] = sum_j w[h,j] dO[t+16-j]
__global__ void __launch_bounds__(BT) conv33_bwd_data_kernel(const float* dO, const float* w, i64 B, i64 Sp, float* dv) {
  const i64 i = (i64)blockIdx.x * BT + threadIdx.x;
  if (i >= B * TM_HEADS * Sp * TM_DHEAD) return;
  const int d = (int)(i % TM_DHEAD);
  const i64 t = (i / TM_DHEAD) % Sp, bh = i / (TM_DHEAD * Sp);
  const int h = (int)(bh % TM_HEADS);
  float a = 0.f;
  for (int j = 0; j < TM_CONV_TAPS; ++j) {
    const i64 ts = t + 16 - j;
    if (ts >= 0 && ts < Sp) a = fmaf(w[h * TM_CONV_TAPS + j], dO[(bh * Sp + ts) * TM_DHEAD + d], a);
  }
  dv[i] = a;
}
// g[h,j] = sum_{b,t,d} dO[b,h,t,d] v[b,h,t+j-16,d]; one block per (h,j)
__global__ void __launch_bounds__(BT) conv33_bwd_weight_kernel(const float* dO, const float* v, i64 B, i64 Sp, float* g) {
  __shared__ float red[BT];
  const int h = blockIdx.x / TM_CONV_TAPS, j = blockIdx.x % TM_CONV_TAPS;
  const i64 per_bag = Sp * TM_DHEAD;
  float a = 0.f;
  for (i64 i = threadIdx.x; i < B * per_bag; i += BT) {
    const i64 b = i / per_bag, rem = i - b * per_bag;
    const i64 t = rem / TM_DHEAD, ts = t + j - 16;
    if (ts < 0 || ts >= Sp) continue;
    const int d = (int)(rem % TM_DHEAD);
    const i64 base = (b * TM_HEADS + h) * Sp;
    a = fmaf(dO[(base + t) * TM_DHEAD + d], v[(base + ts) * TM_DHEAD + d], a);
  }
  a = block_sum(a, red);
  if (threadIdx.x == 0) g[blockIdx.x] = a;
}

// ---- pinv backward helpers ------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(BT) rowcol_abs_sums_kernel(const float* A, float* rows, float* cols) {
  const float* a = A + (i64)blockIdx.x * MM;
  const int j = threadIdx.x;
  float cs = 0.f, rs = 0.f;
  for (int i = 0; i < LMK; ++i) { cs += fabsf(a[i * LMK + j]); rs += fabsf(a[j * LMK + i]); }
  rows[(i64)blockIdx.x * LMK + j] = rs;
  cols[(i64)blockIdx.x * LMK + j] = cs;
}
// sc = {r = max row sum, c = max col sum, r*c, #cols == c, #rows == r}  (maxima over every bag and head)
__global__ void __launch_bounds__(BT) pinv_scalars_kernel(const float* rows, const float* cols, i64 n, float* sc) {
  __shared__ float red[BT];
  float r = 0.f, c = 0.f;
  for (i64 i = threadIdx.x; i < n; i += BT) { r = fmaxf(r, rows[i]); c = fmaxf(c, cols[i]); }
  r = block_max(r, red);
  c = block_max(c, red);
  float nr = 0.f, nc = 0.f;
  for (i64 i = threadIdx.x; i < n; i += BT) { nr += rows[i] == r ? 1.f : 0.f; nc += cols[i] == c ? 1.f : 0.f; }
  nr = block_sum(nr, red);
  nc = block_sum(nc, red);
  if (threadIdx.x == 0) { sc[0] = r; sc[1] = c; sc[2] = r * c; sc[3] = nc; sc[4] = nr; }
}
__global__ void __launch_bounds__(BT) scale_transpose_kernel(const float* A, const float* sc, i64 n, float* Z0) {
  const i64 i = (i64)blockIdx.x * BT + threadIdx.x;
  if (i >= n) return;
  const i64 b = i / MM, rem = i - b * MM;
  const int r = (int)(rem / LMK), c = (int)(rem % LMK);
  Z0[i] = A[b * MM + (i64)c * LMK + r] / sc[2];
}
// out = a * X + diag * I on nb 256x256 matrices
__global__ void __launch_bounds__(BT) scale_add_diag_kernel(const float* X, float a, float diag, i64 n, float* out) {
  const i64 i = (i64)blockIdx.x * BT + threadIdx.x;
  if (i >= n) return;
  const i64 rem = i % MM;
  out[i] = a * X[i] + ((rem / LMK) == (rem % LMK) ? diag : 0.f);
}
__global__ void __launch_bounds__(BT) dot_partial_kernel(const float* X, const float* Y, i64 n, float* partial) {
  __shared__ float red[BT];
  float a = 0.f;
  for (i64 i = (i64)blockIdx.x * BT + threadIdx.x; i < n; i += (i64)gridDim.x * BT) a = fmaf(X[i], Y[i], a);
  a = block_sum(a, red);
  if (threadIdx.x == 0) partial[blockIdx.x] = a;
}
__global__ void __launch_bounds__(BT) sum_small_kernel(const float* partial, int n, float* out) {
  __shared__ float red[BT];
  float a = 0.f;
  for (int i = threadIdx.x; i < n; i += BT) a += partial[i];
  a = block_sum(a, red);
  if (threadIdx.x == 0) out[0] = a;
}
// Z0 = A^T / (r c):  dA += dZ0^T / (r c) + d(r c)-terms on the maximal rows / columns (torch spreads a max's gradient evenly
// over ties); A > 0 (softmax output), so sign(A) = 1.
__global__ void __launch_bounds__(BT) pinv_init_bwd_kernel(const float* dZ0, const float* rows, const float* cols, const float* sc,
                                                           const float* dot, i64 n, float* dA) {
  const i64 i = (i64)blockIdx.x * BT + threadIdx.x;
  if (i >= n) return;
  const i64 b = i / MM, rem = i - b * MM;
  const int r = (int)(rem / LMK), c = (int)(rem % LMK);
  const float scale = sc[2];
  const float dscale = -dot[0] / scale;
  float o = dA[i] + dZ0[b * MM + (i64)c * LMK + r] / scale;
  if (rows[b * LMK + r] == sc[0]) o += dscale * sc[1] / sc[4];
  if (cols[b * LMK + c] == sc[1]) o += dscale * sc[0] / sc[3];
  dA[i] = o;
}

// ---- PPEG ---------------------------------------------------------------------------------------------------------------------------
// effective 7x7 filter of x + dw7(x) + dw5(x) + dw3(x):  weff[c][i*7+j]
__global__ void __launch_bounds__(BT) ppeg_pack_kernel(const float* w7, const float* w5, const float* w3, float* weff) {
  const int i = blockIdx.x * BT + threadIdx.x;
  if (i >= TM_DIM * 49) return;
  const int c = i / 49, tap = i % 49, y = tap / 7, x = tap % 7;
  float o = w7[i];
  if (y >= 1 && y <= 5 && x >= 1 && x <= 5) o += w5[c * 25 + (y - 1) * 5 + (x - 1)];
  if (y >= 2 && y <= 4 && x >= 2 && x <= 4) o += w3[c * 9 + (y - 2) * 3 + (x - 2)];
  if (y == 3 && x == 3) o += 1.0f;
  weff[i] = o;
}
// dh[b,0] = dy[b,0] (cls row);  dh[b,1+t,c] = sum_{i,j} weff[c][i,j] dy[b, 1 + (y+3-i)*G + (x+3-j), c]
__global__ void __launch_bounds__(BT) ppeg_bwd_data_kernel(const float* dy, const float* weff, i64 B, i64 G, float* dh) {
  const i64 S = G * G + 1;
  const i64 i = (i64)blockIdx.x * BT + threadIdx.x;
  if (i >= B * S * TM_DIM) return;
  const int c = (int)(i % TM_DIM);
  const i64 s = (i / TM_DIM) % S, b = i / (TM_DIM * S);
  if (s == 0) { dh[i] = dy[i]; return; }
  const i64 t = s - 1, y = t / G, x = t % G;
  const float* w = weff + c * 49;
  float a = 0.f;
  for (int ky = 0; ky < 7; ++ky) {
    const i64 yy = y + 3 - ky;
    if (yy < 0 || yy >= G) continue;
    for (int kx = 0; kx < 7; ++kx) {
      const i64 xx = x + 3 - kx;
      if (xx < 0 || xx >= G) continue;
      a = fmaf(w[ky * 7 + kx], dy[(b * S + 1 + yy * G + xx) * TM_DIM + c], a);
    }
  }
  dh[i] = a;
}
// partial[chunk][tap][c] over the chunk's (bag, token) pairs; tap 49 = plain sum of dy (the three bias gradients)
__global__ void __launch_bounds__(TM_DIM) ppeg_bwd_weight_kernel(const float* h_in, const float* dy, i64 B, i64 G, i64 per_chunk,
                                                                 float* partial) {
  const int c = threadIdx.x, tap = blockIdx.x;
  const i64 N = G * G, S = N + 1;
  const i64 i0 = (i64)blockIdx.y * per_chunk;
  const i64 i1 = i0 + per_chunk < B * N ? i0 + per_chunk : B * N;
  const int ky = tap / 7, kx = tap % 7;
  float a = 0.f;
  for (i64 i = i0; i < i1; ++i) {
    const i64 b = i / N, t = i - b * N;
    const float d = dy[(b * S + 1 + t) * TM_DIM + c];
    if (tap == 49) { a += d; continue; }
    const i64 yy = t / G + ky - 3, xx = t % G + kx - 3;
    if (yy < 0 || yy >= G || xx < 0 || xx >= G) continue;
    a = fmaf(d, h_in[(b * S + 1 + yy * G + xx) * TM_DIM + c], a);
  }
  partial[((i64)blockIdx.y * 50 + tap) * TM_DIM + c] = a;
}
__global__ void __launch_bounds__(BT) ppeg_unpack_grad_kernel(const float* g, float* g_w7, float* g_b7, float* g_w5, float* g_b5,
                                                              float* g_w3, float* g_b3) {
  const int i = blockIdx.x * BT + threadIdx.x;
  if (i >= TM_DIM * 49) return;
  const int c = i / 49, tap = i % 49, y = tap / 7, x = tap % 7;
  const float v = g[tap * TM_DIM + c];
  g_w7[i] = v;
  if (y >= 1 && y <= 5 && x >= 1 && x <= 5) g_w5[c * 25 + (y - 1) * 5 + (x - 1)] = v;
  if (y >= 2 && y <= 4 && x >= 2 && x <= 4) g_w3[c * 9 + (y - 2) * 3 + (x - 2)] = v;
  if (tap == 0) { const float bsum = g[49 * TM_DIM + c]; g_b7[c] = bsum; g_b5[c] = bsum; g_b3[c] = bsum; }
}

// ---- head: logits = _fc(LayerNorm(h3[:,0])) -- one block of 512 threads walks the bags ------------------------------------------
__global__ void __launch_bounds__(TM_DIM) head_bwd_kernel(const float* h3, const float* dlogits, const float* gamma, const float* beta,
                                                          const float* fc_w, i64 B, i64 S, int n_classes, float* dcls, float* g_gamma,
                                                          float* g_beta, float* g_fc_w, float* g_fc_b) {
  __shared__ float red[TM_DIM];
  const int c = threadIdx.x;
  float ag = 0.f, ab = 0.f;
  for (int k = 0; k < n_classes; ++k) g_fc_w[k * TM_DIM + c] = 0.f;
  float afb = 0.f;
  for (i64 b = 0; b < B; ++b) {
    const float x = h3[b * S * TM_DIM + c];
    const float mean = block_sum(x, red) / TM_DIM;
    const float xc = x - mean;
    const float rstd = rsqrtf(block_sum(xc * xc, red) / TM_DIM + 1e-5f);
    const float xh = xc * rstd;
    const float cn = xh * gamma[c] + beta[c];
    float dcn = 0.f;
    for (int k = 0; k < n_classes; ++k) {
      const float dl = dlogits[b * n_classes + k];
      dcn = fmaf(dl, fc_w[k * TM_DIM + c], dcn);
      g_fc_w[k * TM_DIM + c] = fmaf(dl, cn, g_fc_w[k * TM_DIM + c]);
    }
    if (c < n_classes) afb += dlogits[b * n_classes + c];
    ag = fmaf(dcn, xh, ag);
    ab += dcn;
    const float g = dcn * gamma[c];
    const float m1 = block_sum(g, red) / TM_DIM;
    const float m2 = block_sum(g * xh, red) / TM_DIM;
    dcls[b * TM_DIM + c] = rstd * (g - m1 - xh * m2);
  }
  g_gamma[c] = ag;
  g_beta[c] = ab;
  if (c < n_classes) g_fc_b[c] = afb;
}

// dhf[b,i] = dh0[b,1+i] + (i < dup ? dh0[b,1+n+i] : 0)   (TransMIL.py:177-186 backward: cls row aside, duplicate pad folded)
__global__ void __launch_bounds__(BT) dup_cls_bwd_kernel(const float* dh0, i64 B, i64 n, i64 dup, float* dhf) {
  const i64 i = (i64)blockIdx.x * BT + threadIdx.x;
  if (i >= B * n * TM_DIM) return;
  const int c = (int)(i % TM_DIM);
  const i64 t = (i / TM_DIM) % n, b = i / (TM_DIM * n);
  const i64 S = n + dup + 1;
  float o = dh0[(b * S + 1 + t) * TM_DIM + c];
  if (t < dup) o += dh0[(b * S + 1 + n + t) * TM_DIM + c];
  dhf[i] = o;
}

// h0[b,0] = cls, h0[b,1+i] = hf[b,i] (i < n), h0[b,1+n+j] = hf[b,j] (j < dup)   (TransMIL.py:177-186: duplicate pad + cls concat
// for stems that produce their tokens outside tm_fc1_fwd: the training-mode 768-d stem, the conv stem of CTMIL.py:137-152)
__global__ void __launch_bounds__(BT) dup_cls_fwd_kernel(const float* hf, const float* cls, i64 B, i64 n, i64 dup, float* h0) {
  const i64 i = (i64)blockIdx.x * BT + threadIdx.x;
  const i64 S = n + dup + 1;
  if (i >= B * S * TM_DIM) return;
  const int c = (int)(i % TM_DIM);
  const i64 r = (i / TM_DIM) % S, b = i / (TM_DIM * S);
  h0[i] = r == 0 ? cls[c] : hf[(b * n + (r - 1 < n ? r - 1 : r - 1 - n)) * TM_DIM + c];
}

// ---- optimizer step over the whole parameter set: torch.optim.RAdam as selected by the reference's factory
// (code/MyOptimizer/optim_factory.py:78-80) + the Lookahead wrapper (code/MyOptimizer/lookahead.py:28-41), one launch each --------
struct OptBatch { tm_opt_tensor t[TM_OPT_MAX_TENSORS]; };
__global__ void __launch_bounds__(BT) radam_kernel(const OptBatch b, float lr, float beta1, float beta2, float eps, float bc1,
                                                   float bc2_sqrt, float rect, int rectified) {
  const tm_opt_tensor t = b.t[blockIdx.y];
  for (i64 i = (i64)blockIdx.x * BT + threadIdx.x; i < t.n; i += (i64)gridDim.x * BT) {
    float p = t.p[i];
    const float g = t.weight_decay != 0.f ? fmaf(t.weight_decay, p, t.g[i]) : t.g[i];
    float m = t.m[i], v = t.v[i];
    m = m + (1.0f - beta1) * (g - m);
    v = beta2 * v + (1.0f - beta2) * g * g;
    const float mh = m / bc1;
    if (rectified) p -= mh * lr * (bc2_sqrt / (sqrtf(v) + eps)) * rect;
    else p -= mh * lr;
    t.p[i] = p; t.m[i] = m; t.v[i] = v;
  }
}
// slow += alpha (fast - slow); fast = slow   (t.p = fast weights, t.m = slow buffer)
__global__ void __launch_bounds__(BT) lookahead_kernel(const OptBatch b, float alpha) {
  const tm_opt_tensor t = b.t[blockIdx.y];
  for (i64 i = (i64)blockIdx.x * BT + threadIdx.x; i < t.n; i += (i64)gridDim.x * BT) {
    const float s = t.m[i] + alpha * (t.p[i] - t.m[i]);
    t.m[i] = s;
    t.p[i] = s;
  }
}
static int opt_batch(const tm_opt_tensor* ts, int64_t count, OptBatch* b, unsigned* gx) {
  TM_CHECK_ARG(ts && count > 0 && count <= TM_OPT_MAX_TENSORS, TM_ERR_BAD_ARGUMENT, "optimizer step: 1..%d tensors per call (got %lld)",
               TM_OPT_MAX_TENSORS, (long long)count);
  i64 nmax = 0;
  for (int64_t i = 0; i < count; ++i) {
    TM_CHECK_ARG(ts[i].p && ts[i].m && ts[i].n > 0, TM_ERR_BAD_ARGUMENT, "optimizer step: tensor %lld has a NULL pointer or no elements", (long long)i);
    b->t[i] = ts[i];
    if (ts[i].n > nmax) nmax = ts[i].n;
  }
  i64 g = ceil_div(nmax, (i64)BT * 4);
  *gx = (unsigned)(g < 1 ? 1 : (g > 296 ? 296 : g));
  return TM_OK;
}

// one product of the pinv chain: C = alpha * op(A) op(B) + beta * L + diag * I on nb 256x256 matrices
static int mm256(const float* A, bool tA, const float* B, bool tB, float* C, float alpha, float beta, const float* L, float diag,
                 int nb, cudaStream_t st, float* C2 = nullptr, float alpha2 = 0.f, float diag2 = 0.f) {
  BGemm p;
  memset(&p, 0, sizeof(p));
  p.A = A; p.B = B; p.C = C; p.L = L;
  p.C2 = C2; p.alpha2 = alpha2; p.diag2 = diag2;
  p.M = p.N = p.K = LMK;
  p.sam = tA ? 1 : LMK; p.sak = tA ? LMK : 1;
  p.sbk = tB ? 1 : LMK; p.sbn = tB ? LMK : 1;
  p.ldc = p.ldl = LMK;
  p.strA = p.strB = p.strC = p.strL = MM;
  p.alpha = alpha; p.beta = beta; p.diag = diag;
  return bgemm(p, nb, st);
}

}  // namespace tmil

using namespace tmil;

extern "C" {

#ifdef TM_HOST_EMU
const char* tm_last_error(void) { return tmil::g_err; }
#endif

int tm_bgemm(const float* A, const float* B, float* C, const float* L, int64_t M, int64_t N, int64_t K, int64_t sam, int64_t sak,
             int64_t sbk, int64_t sbn, int64_t ldc, int64_t ldl, int64_t batch, int64_t strideA, int64_t strideB, int64_t strideC,
             int64_t strideL, float alpha, float beta, void* stream) {
  TM_CHECK_ARG(M < (1ll << 31) && N < (1ll << 31) && K < (1ll << 31), TM_ERR_BAD_SHAPE, "tm_bgemm: dimension too large");
  BGemm p;
  memset(&p, 0, sizeof(p));
  p.A = A; p.B = B; p.C = C; p.L = L;
  p.M = (int)M; p.N = (int)N; p.K = (int)K;
  p.sam = sam; p.sak = sak; p.sbk = sbk; p.sbn = sbn; p.ldc = ldc; p.ldl = ldl;
  p.strA = strideA; p.strB = strideB; p.strC = strideC; p.strL = strideL;
  p.alpha = alpha; p.beta = beta; p.diag = 0.f;
  return bgemm(p, (int)batch, (cudaStream_t)stream);
}

/* tm_bgemm with caller-owned scratch for split-K (weight gradients: small M x N, K = every token of the batch):
 * scratch holds tm_bgemm_scratch_floats(M, N, K, batch) floats (0 = this shape is not split). */
int64_t tm_bgemm_scratch_floats(int64_t M, int64_t N, int64_t K, int64_t batch) {
#ifndef TM_HOST_EMU
  if (bgemm_use_tc() && M > 0 && N > 0 && K > 0 && batch > 0) return bgemm_tc_scratch_floats(M, N, K, batch);
#endif
  (void)M; (void)N; (void)K; (void)batch;
  return 0;
}
int tm_bgemm_ex(const float* A, const float* B, float* C, const float* L, int64_t M, int64_t N, int64_t K, int64_t sam, int64_t sak,
                int64_t sbk, int64_t sbn, int64_t ldc, int64_t ldl, int64_t batch, int64_t strideA, int64_t strideB, int64_t strideC,
                int64_t strideL, float alpha, float beta, float* scratch, int64_t scratch_floats, void* stream) {
  TM_CHECK_ARG(M < (1ll << 31) && N < (1ll << 31) && K < (1ll << 31), TM_ERR_BAD_SHAPE, "tm_bgemm: dimension too large");
  BGemm p;
  memset(&p, 0, sizeof(p));
  p.A = A; p.B = B; p.C = C; p.L = L;
  p.M = (int)M; p.N = (int)N; p.K = (int)K;
  p.sam = sam; p.sak = sak; p.sbk = sbk; p.sbn = sbn; p.ldc = ldc; p.ldl = ldl;
  p.strA = strideA; p.strB = strideB; p.strC = strideC; p.strL = strideL;
  p.alpha = alpha; p.beta = beta; p.diag = 0.f;
  return bgemm(p, (int)batch, (cudaStream_t)stream, scratch, scratch_floats);
}
/* Engine of tm_bgemm / tm_bgemm_ex / tm_pinv_bwd's products: 0 = tcgen05 (tf32 operands; default on sm_100; operands fed by TMA
 * when they obey the tensor-map rules, by cp.async loader warps otherwise), 1 = fp32 CUDA cores, 2 = tcgen05 with the loader
 * warps only (cross-check of the TMA feed).  Returns the previous value. */
int tm_bgemm_engine(int engine) {
#ifndef TM_HOST_EMU
  const int prev = g_bgemm_engine;
  if (engine >= 0 && engine <= 2) {
    g_bgemm_engine = engine;
    g_bgemm_force_cpasync = engine == 2;
  }
  return prev;
#else
  (void)engine;
  return 1;
#endif
}

int tm_softmax_rows(float* X, int64_t rows, int64_t len, void* stream) {
  TM_CHECK_ARG(X && rows > 0 && len > 0 && rows < (1ll << 31) && len < (1ll << 31), TM_ERR_BAD_SHAPE, "tm_softmax_rows: bad shape");
#ifndef TM_HOST_EMU
  if (g_bgemm_engine != 1 && softmax_rows_fast_ok(X, nullptr, len)) return softmax_rows_fast(X, rows, len, (cudaStream_t)stream);
#endif
  BW_LAUNCH(softmax_rows_kernel, dim3((unsigned)rows), dim3(BT), (cudaStream_t)stream, X, (int)len);
  return TM_OK;
}
int tm_softmax_rows_bwd(const float* P, float* dP, int64_t rows, int64_t len, void* stream) {
  TM_CHECK_ARG(P && dP && rows > 0 && len > 0 && rows < (1ll << 31) && len < (1ll << 31), TM_ERR_BAD_SHAPE, "tm_softmax_rows_bwd: bad shape");
#ifndef TM_HOST_EMU
  if (g_bgemm_engine != 1 && softmax_rows_fast_ok(P, dP, len)) return softmax_rows_bwd_fast(P, dP, rows, len, (cudaStream_t)stream);
#endif
  BW_LAUNCH(softmax_rows_bwd_kernel, dim3((unsigned)rows), dim3(BT), (cudaStream_t)stream, P, dP, (int)len);
  return TM_OK;
}

int tm_ln_rows_fwd(const float* x, const float* gamma, const float* beta, int64_t rows, int64_t C, float* y, void* stream) {
  TM_CHECK_ARG(x && gamma && beta && y && rows > 0 && C > 0 && rows < (1ll << 31), TM_ERR_BAD_SHAPE, "tm_ln_rows_fwd: bad shape");
  BW_LAUNCH(ln_rows_fwd_kernel, dim3((unsigned)rows), dim3(BT), (cudaStream_t)stream, x, gamma, beta, (int)C, y);
  return TM_OK;
}

int64_t tm_ln_bwd_scratch_floats(int64_t rows, int64_t C) {
  int64_t n = 2 * rows + 2 * C * (colsum_chunks(rows) + 1);
#ifndef TM_HOST_EMU
  if (C == TM_DIM && n < ln512_bwd_fast_scratch_floats()) n = ln512_bwd_fast_scratch_floats();
#endif
  return n;
}
/* LayerNorm backward over rows of x [rows,C]; dy row of x row r = (r / rows_per_bag) * dy_bag_rows + dy_row_off + r % rows_per_bag.
 * dx = (add ? add : 0) + dLN; g_gamma, g_beta [C].  scratch: tm_ln_bwd_scratch_floats(rows, C). */
int tm_ln_rows_bwd(const float* x, const float* gamma, const float* dy, int64_t rows, int64_t C, int64_t rows_per_bag,
                   int64_t dy_bag_rows, int64_t dy_row_off, const float* add, float* dx, float* g_gamma, float* g_beta,
                   float* scratch, void* stream) {
  TM_CHECK_ARG(x && gamma && dy && dx && g_gamma && g_beta && scratch, TM_ERR_BAD_ARGUMENT, "tm_ln_rows_bwd: NULL pointer");
  TM_CHECK_ARG(rows > 0 && C > 0 && rows < (1ll << 31) && rows_per_bag > 0 && rows % rows_per_bag == 0, TM_ERR_BAD_SHAPE,
               "tm_ln_rows_bwd: bad shape rows=%lld C=%lld rows_per_bag=%lld", (long long)rows, (long long)C, (long long)rows_per_bag);
  cudaStream_t st = (cudaStream_t)stream;
#ifndef TM_HOST_EMU
  if (g_bgemm_engine != 1 && C == TM_DIM && aligned16(x) && aligned16(dy) && aligned16(dx) && aligned16(gamma) && (add == nullptr || aligned16(add)))
    return ln512_bwd_fast(x, gamma, dy, rows, rows_per_bag, dy_bag_rows, dy_row_off, add, dx, g_gamma, g_beta, scratch, st);
#endif
  float* stats = scratch;
  float* partial = scratch + 2 * rows;
  const i64 nch = colsum_chunks(rows);
  float* both = partial + 2 * C * nch;   // [2][C]
  BW_LAUNCH(ln_rows_bwd_kernel, dim3((unsigned)rows), dim3(BT), st, x, gamma, dy, (int)C, (i64)rows_per_bag, (i64)dy_bag_rows,
            (i64)dy_row_off, add, dx, stats);
  const i64 rpc = ceil_div(rows, nch);
  BW_LAUNCH(ln_param_grad_kernel, dim3((unsigned)ceil_div(C, BT), (unsigned)nch), dim3(BT), st, x, dy, (const float*)stats, (i64)rows,
            (int)C, (i64)rows_per_bag, (i64)dy_bag_rows, (i64)dy_row_off, rpc, partial);
  BW_LAUNCH(colsum_kernel, dim3((unsigned)ceil_div(2 * C, BT), 1), dim3(BT), st, (const float*)partial, nch, (i64)(2 * C), (i64)(2 * C), nch,
            both);
  BW_LAUNCH(colsum_kernel, dim3((unsigned)ceil_div(C, BT), 1), dim3(BT), st, (const float*)both, (i64)1, (i64)C, (i64)C, (i64)1, g_gamma);
  BW_LAUNCH(colsum_kernel, dim3((unsigned)ceil_div(C, BT), 1), dim3(BT), st, (const float*)(both + C), (i64)1, (i64)C, (i64)C, (i64)1, g_beta);
  return TM_OK;
}

int64_t tm_colsum_scratch_floats(int64_t rows, int64_t cols) { return colsum_chunks(rows) * cols; }
/* out[c] = sum_r X[r*ld + c]  (bias gradients, cls_token gradient) */
int tm_colsum(const float* X, int64_t rows, int64_t cols, int64_t ld, float* out, float* scratch, void* stream) {
  TM_CHECK_ARG(X && out && scratch && rows > 0 && cols > 0, TM_ERR_BAD_ARGUMENT, "tm_colsum: bad argument");
  return colsum(X, rows, cols, ld, out, scratch, (cudaStream_t)stream);
}

int tm_gelu_fwd(const float* z, float* a, int64_t n, void* stream) {
  TM_CHECK_ARG(z && a && n > 0, TM_ERR_BAD_ARGUMENT, "tm_gelu_fwd: bad argument");
  BW_LAUNCH(gelu_fwd_kernel, grid1d(n), dim3(BT), (cudaStream_t)stream, z, a, (i64)n);
  return TM_OK;
}
int tm_gelu_bwd(const float* z, const float* dy, float* dz, int64_t n, void* stream) {
  TM_CHECK_ARG(z && dy && dz && n > 0, TM_ERR_BAD_ARGUMENT, "tm_gelu_bwd: bad argument");
  BW_LAUNCH(gelu_bwd_kernel, grid1d(n), dim3(BT), (cudaStream_t)stream, z, dy, dz, (i64)n);
  return TM_OK;
}

/* Training-mode Dropout(p) of to_out (nystrom-attention to_out = Sequential(Linear, Dropout); TransMIL.py:33 passes 0.7)
 * applied after tm_to_out_residual_fwd: y = h_in + to_out  ->  y = h_in + mask * to_out / (1-p), in place. */
int tm_dropout_resid_fwd(float* y, const float* h_in, int64_t n, uint64_t seed, float p, void* stream) {
  TM_CHECK_ARG(y && h_in && n > 0 && p >= 0.f && p < 1.f, TM_ERR_BAD_ARGUMENT, "tm_dropout_resid_fwd: bad argument");
  BW_LAUNCH(dropout_resid_kernel, grid1d(n), dim3(BT), (cudaStream_t)stream, y, h_in, (i64)n, seed, p);
  return TM_OK;
}
/* Training-mode Dropout(p) inside the 768-d stem (TransMIL.py:123-126: Linear GELU Dropout(0.6) LN Linear GELU Dropout(0.6) LN):
 * y = mask(seed) * x / (1-p), the same counter-based mask as the to_out Dropout; the backward applies the same call to the
 * gradient.  x may alias y. */
int tm_dropout_fwd(const float* x, float* y, int64_t n, uint64_t seed, float p, void* stream) {
  TM_CHECK_ARG(x && y && n > 0 && p >= 0.f && p < 1.f, TM_ERR_BAD_ARGUMENT, "tm_dropout_fwd: bad argument");
  BW_LAUNCH(dropout_kernel, grid1d(n), dim3(BT), (cudaStream_t)stream, x, y, (i64)n, seed, p);
  return TM_OK;
}
/* dy [B,S,C] -> out [B,S+P,C] with P zero rows in front (the `[:, -S:]` slice backward), times the dropout mask if p > 0. */
int tm_pad_front_bwd(const float* dy, int64_t B, int64_t S, int64_t P, int64_t C, uint64_t seed, float p, float* out, void* stream) {
  TM_CHECK_ARG(dy && out && B > 0 && S > 0 && P >= 0 && C > 0 && p >= 0.f && p < 1.f, TM_ERR_BAD_ARGUMENT, "tm_pad_front_bwd: bad argument");
  BW_LAUNCH(pad_front_kernel, grid1d(B * (S + P) * C), dim3(BT), (cudaStream_t)stream, dy, (i64)B, (i64)S, (i64)P, (int)C, seed, p, out);
  return TM_OK;
}
int tm_split_heads(const float* merged, int64_t B, int64_t Sp, float* out, void* stream) {
  TM_CHECK_ARG(merged && out && B > 0 && Sp > 0, TM_ERR_BAD_ARGUMENT, "tm_split_heads: bad argument");
  BW_LAUNCH(split_heads_kernel, grid1d(B * Sp * TM_DIM), dim3(BT), (cudaStream_t)stream, merged, (i64)B, (i64)Sp, out);
  return TM_OK;
}
int tm_qkv_grad_merge(const float* dq, const float* dk, const float* dv, const float* dq_lm, const float* dk_lm, int64_t B, int64_t Sp,
                      int64_t l, float* dqkv, void* stream) {
  TM_CHECK_ARG(dq && dk && dv && dq_lm && dk_lm && dqkv, TM_ERR_BAD_ARGUMENT, "tm_qkv_grad_merge: NULL pointer");
  TM_CHECK_ARG(B > 0 && Sp > 0 && l > 0 && Sp == l * LMK, TM_ERR_BAD_SHAPE, "tm_qkv_grad_merge: Sp=%lld l=%lld", (long long)Sp, (long long)l);
#ifndef TM_HOST_EMU
  if (g_bgemm_engine != 1 && aligned16(dq) && aligned16(dk) && aligned16(dv) && aligned16(dq_lm) && aligned16(dk_lm) && aligned16(dqkv))
    return qkv_grad_merge_fast(dq, dk, dv, dq_lm, dk_lm, B, Sp, l, dqkv, (cudaStream_t)stream);
#endif
  BW_LAUNCH(qkv_grad_merge_kernel, grid1d(B * Sp * 3 * TM_DIM), dim3(BT), (cudaStream_t)stream, dq, dk, dv, dq_lm, dk_lm, (i64)B, (i64)Sp,
            (i64)l, dqkv);
  return TM_OK;
}

int64_t tm_conv33_bwd_scratch_floats(int64_t B, int64_t Sp) {
#ifndef TM_HOST_EMU
  if (B > 0 && Sp > 0) return conv33_bwd_fast_scratch_floats(B, Sp);
#endif
  (void)B; (void)Sp;
  return 1;
}
/* res_conv backward (Conv2d(8,8,(33,1),groups=8) along the padded token axis): dO, v [B,8,Sp,64]; dv [B,8,Sp,64]; g_w [8,1,33,1];
 * scratch: tm_conv33_bwd_scratch_floats(B, Sp) floats (per-block tap sums of the device-tuned kernel; may be NULL: plain kernels) */
int tm_conv33_bwd(const float* dO, const float* v, const float* conv_w, int64_t B, int64_t Sp, float* dv, float* g_w, float* scratch,
                  void* stream) {
  TM_CHECK_ARG(dO && v && conv_w && dv && g_w && B > 0 && Sp > 0, TM_ERR_BAD_ARGUMENT, "tm_conv33_bwd: bad argument");
  cudaStream_t st = (cudaStream_t)stream;
#ifndef TM_HOST_EMU
  if (g_bgemm_engine != 1 && scratch != nullptr && aligned16(dO) && aligned16(v))
    return conv33_bwd_fast(dO, v, conv_w, B, Sp, dv, g_w, scratch, st);
#else
  (void)scratch;
#endif
  BW_LAUNCH(conv33_bwd_data_kernel, grid1d(B * TM_HEADS * Sp * TM_DHEAD), dim3(BT), st, dO, conv_w, (i64)B, (i64)Sp, dv);
  BW_LAUNCH(conv33_bwd_weight_kernel, dim3(TM_HEADS * TM_CONV_TAPS), dim3(BT), st, dO, v, (i64)B, (i64)Sp, g_w);
  return TM_OK;
}

int64_t tm_pinv_bwd_scratch_floats(int64_t nb) { return 14 * nb * MM + 2 * nb * LMK + 8 + 1024 + 8; }
/* Backward of moore_penrose_iter_pinv (6 iterations, global init scalar): A = attn2 [nb,256,256] (nb = B*8), dZ = gradient of
 * the returned Z (clobbered), dA out.  The iterates are recomputed.  oracle/transmil_backward.py pinv_backward. */
int tm_pinv_bwd(const float* A, float* dZ, int64_t nb64, float* dA, float* scratch, void* stream) {
  TM_CHECK_ARG(A && dZ && dA && scratch && nb64 > 0 && nb64 <= 65535, TM_ERR_BAD_ARGUMENT, "tm_pinv_bwd: bad argument");
  cudaStream_t st = (cudaStream_t)stream;
  const int nb = (int)nb64;
  const i64 n = nb * MM;
  float* Zs = scratch;                    // 6 iterates Z0..Z5
  float* XZ = Zs + 6 * n;
  float* T1 = XZ + n;
  float* T2 = T1 + n;
  float* T3 = T2 + n;
  float* dT3 = T3 + n;
  float* dT2 = dT3 + n;
  float* dXZ = dT2 + n;
  float* dZn = dXZ + n;
  float* rows = dZn + n;
  float* cols = rows + (i64)nb * LMK;
  float* sc = cols + (i64)nb * LMK;
  float* partial = sc + 8;
  float* dot = partial + 1024;
  BW_LAUNCH(rowcol_abs_sums_kernel, dim3((unsigned)nb), dim3(LMK), st, A, rows, cols);
  BW_LAUNCH(pinv_scalars_kernel, dim3(1), dim3(BT), st, (const float*)rows, (const float*)cols, (i64)nb * LMK, sc);
  BW_LAUNCH(scale_transpose_kernel, grid1d(n), dim3(BT), st, A, (const float*)sc, n, Zs);
  auto chain = [&](const float* Z) -> int {   // XZ, T1, T2, T3 of one iteration
    TM_PROPAGATE(mm256(A, false, Z, false, XZ, 1.f, 0.f, nullptr, 0.f, nb, st, T1, -1.f, 7.f));      // XZ = A Z and T1 = 7I - XZ
    TM_PROPAGATE(mm256(XZ, false, T1, false, T2, -1.f, 0.f, nullptr, 15.f, nb, st));    // T2 = 15I - XZ T1
    TM_PROPAGATE(mm256(XZ, false, T2, false, T3, -1.f, 0.f, nullptr, 13.f, nb, st));    // T3 = 13I - XZ T2
    return TM_OK;
  };
  for (int it = 0; it + 1 < TM_PINV_ITERS; ++it) {
    TM_PROPAGATE(chain(Zs + it * n));
    TM_PROPAGATE(mm256(Zs + it * n, false, T3, false, Zs + (it + 1) * n, 0.25f, 0.f, nullptr, 0.f, nb, st));
  }
  float* dZc = dZ;      // current dZ
  float* dZo = dZn;     // the other buffer
  bool first = true;
  for (int it = TM_PINV_ITERS - 1; it >= 0; --it) {
    const float* Z = Zs + it * n;
    TM_PROPAGATE(chain(Z));
    TM_PROPAGATE(mm256(Z, true, dZc, false, dT3, 0.25f, 0.f, nullptr, 0.f, nb, st));       // dT3 = 0.25 Z^T dZ
    TM_PROPAGATE(mm256(dZc, false, T3, true, dZo, 0.25f, 0.f, nullptr, 0.f, nb, st));      // dZ' = 0.25 dZ T3^T
    TM_PROPAGATE(mm256(dT3, false, T2, true, dXZ, -1.f, 0.f, nullptr, 0.f, nb, st));       // dXZ = -dT3 T2^T
    TM_PROPAGATE(mm256(XZ, true, dT3, false, dT2, -1.f, 0.f, nullptr, 0.f, nb, st));       // dT2 = -XZ^T dT3
    TM_PROPAGATE(mm256(dT2, false, T1, true, dXZ, -1.f, 1.f, dXZ, 0.f, nb, st));           // dXZ -= dT2 T1^T
    TM_PROPAGATE(mm256(XZ, true, dT2, false, dXZ, 1.f, 1.f, dXZ, 0.f, nb, st));            // dXZ -= dT1, dT1 = -XZ^T dT2
    TM_PROPAGATE(mm256(dXZ, false, Z, true, dA, 1.f, first ? 0.f : 1.f, first ? nullptr : dA, 0.f, nb, st));   // dA += dXZ Z^T
    TM_PROPAGATE(mm256(A, true, dXZ, false, dZo, 1.f, 1.f, dZo, 0.f, nb, st));             // dZ = dZ' + A^T dXZ
    float* t = dZc; dZc = dZo; dZo = t;
    first = false;
  }
  // Z0 = A^T / (r c)
  const int nblk = (int)(ceil_div(n, BT) < 1024 ? ceil_div(n, BT) : 1024);
  BW_LAUNCH(dot_partial_kernel, dim3((unsigned)nblk), dim3(BT), st, (const float*)dZc, (const float*)Zs, n, partial);
  BW_LAUNCH(sum_small_kernel, dim3(1), dim3(BT), st, (const float*)partial, nblk, dot);
  BW_LAUNCH(pinv_init_bwd_kernel, grid1d(n), dim3(BT), st, (const float*)dZc, (const float*)rows, (const float*)cols, (const float*)sc,
            (const float*)dot, n, dA);
  return TM_OK;
}

int64_t tm_ppeg_bwd_scratch_floats(int64_t B, int64_t G) {
  (void)B; (void)G;
  return (int64_t)TM_DIM * 49 + (COLSUM_MAX_CHUNKS + 1) * 50 * TM_DIM + 8 * 50 * TM_DIM;
}
/* PPEG backward (TransMIL.py:60-75): h_in, dy, dh [B,S,512], S = G*G+1; g_w7 [512,1,7,7], g_w5 [512,1,5,5], g_w3 [512,1,3,3],
 * g_b* [512].  scratch: tm_ppeg_bwd_scratch_floats(). */
int tm_ppeg_bwd(const tm_weights* w, const float* h_in, const float* dy, int64_t B, int64_t G, float* dh, float* g_w7, float* g_b7,
                float* g_w5, float* g_b5, float* g_w3, float* g_b3, float* scratch, void* stream) {
  TM_CHECK_ARG(w && h_in && dy && dh && g_w7 && g_b7 && g_w5 && g_b5 && g_w3 && g_b3 && scratch, TM_ERR_BAD_ARGUMENT, "tm_ppeg_bwd: NULL pointer");
  TM_CHECK_ARG(B > 0 && G > 0, TM_ERR_BAD_SHAPE, "tm_ppeg_bwd: bad shape");
  cudaStream_t st = (cudaStream_t)stream;
  float* weff = scratch;
  float* partial = weff + TM_DIM * 49;
  const i64 total = B * G * G;
  i64 nch = ceil_div(total, 64);
  if (nch > COLSUM_MAX_CHUNKS) nch = COLSUM_MAX_CHUNKS;
  const i64 per_chunk = ceil_div(total, nch);
  float* gsum = partial + nch * 50 * TM_DIM;
  float* cs_scratch = gsum + 50 * TM_DIM;
#ifndef TM_HOST_EMU
  if (g_bgemm_engine != 1 && tc_available() && aligned16(dy) && aligned16(dh)) {
    // flipped packed filter: behind the COLSUM_MAX_CHUNKS partials and their sum (the tail region is free on this path)
    TM_PROPAGATE(ppeg_dgrad_fast(w, dy, B, G, dh, partial + (COLSUM_MAX_CHUNKS + 1) * 50 * TM_DIM, st));
  } else
#endif
  {
  BW_LAUNCH(ppeg_pack_kernel, grid1d(TM_DIM * 49), dim3(BT), st, w->ppeg_w7, w->ppeg_w5, w->ppeg_w3, weff);
  BW_LAUNCH(ppeg_bwd_data_kernel, grid1d(B * (G * G + 1) * TM_DIM), dim3(BT), st, dy, (const float*)weff, (i64)B, (i64)G, dh);
  }
#ifndef TM_HOST_EMU
  if (g_bgemm_engine != 1) {
    // the device-tuned kernel leaves at most COLSUM_MAX_CHUNKS partials [50][512] at `partial` and their sum at its end
    gsum = partial + COLSUM_MAX_CHUNKS * 50 * TM_DIM;
    TM_PROPAGATE(ppeg_wgrad_fast(h_in, dy, B, G, partial, gsum, st));
  } else
#endif
  {
  BW_LAUNCH(ppeg_bwd_weight_kernel, dim3(50, (unsigned)nch), dim3(TM_DIM), st, h_in, dy, (i64)B, (i64)G, per_chunk, partial);
  TM_PROPAGATE(colsum(partial, nch, 50 * TM_DIM, 50 * TM_DIM, gsum, cs_scratch, st));
  }
  BW_LAUNCH(ppeg_unpack_grad_kernel, grid1d(TM_DIM * 49), dim3(BT), st, (const float*)gsum, g_w7, g_b7, g_w5, g_b5, g_w3, g_b3);
  return TM_OK;
}

/* Head backward (TransMIL.py:202-204): h3 [B,S,512], dlogits [B,n_classes] -> dcls [B,512] (gradient of h3[:,0]; every other
 * row of dh3 is zero), g_norm_w, g_norm_b [512], g_fc_w [n_classes,512], g_fc_b [n_classes]. */
int tm_head_bwd(const tm_weights* w, const float* h3, const float* dlogits, int64_t B, int64_t S, float* dcls, float* g_norm_w,
                float* g_norm_b, float* g_fc_w, float* g_fc_b, void* stream) {
  TM_CHECK_ARG(w && h3 && dlogits && dcls && g_norm_w && g_norm_b && g_fc_w && g_fc_b, TM_ERR_BAD_ARGUMENT, "tm_head_bwd: NULL pointer");
  TM_CHECK_ARG(B > 0 && S > 0 && w->n_classes > 0 && w->n_classes <= TM_DIM, TM_ERR_BAD_SHAPE, "tm_head_bwd: bad shape");
  BW_LAUNCH(head_bwd_kernel, dim3(1), dim3(TM_DIM), (cudaStream_t)stream, h3, dlogits, w->norm_w, w->norm_b, w->fc_w, (i64)B, (i64)S,
            (int)w->n_classes, dcls, g_norm_w, g_norm_b, g_fc_w, g_fc_b);
  return TM_OK;
}

/* Duplicate pad + cls concat (TransMIL.py:177-186) of tokens a stem produced outside tm_fc1_fwd: hf [B,n,512], cls [512]
 * -> h0 [B,1+n+dup,512]. */
int tm_dup_cls_fwd(const float* hf, const float* cls, int64_t B, int64_t n, int64_t dup, float* h0, void* stream) {
  TM_CHECK_ARG(hf && cls && h0 && B > 0 && n > 0 && dup >= 0 && dup <= n, TM_ERR_BAD_ARGUMENT, "tm_dup_cls_fwd: bad argument");
  BW_LAUNCH(dup_cls_fwd_kernel, grid1d(B * (1 + n + dup) * TM_DIM), dim3(BT), (cudaStream_t)stream, hf, cls, (i64)B, (i64)n, (i64)dup, h0);
  return TM_OK;
}

/* Backward of the duplicate pad + cls concat (TransMIL.py:177-186): dh0 [B,1+n+dup,512] -> dhf [B,n,512]. */
int tm_dup_cls_bwd(const float* dh0, int64_t B, int64_t n, int64_t dup, float* dhf, void* stream) {
  TM_CHECK_ARG(dh0 && dhf && B > 0 && n > 0 && dup >= 0 && dup <= n, TM_ERR_BAD_ARGUMENT, "tm_dup_cls_bwd: bad argument");
  BW_LAUNCH(dup_cls_bwd_kernel, grid1d(B * n * TM_DIM), dim3(BT), (cudaStream_t)stream, dh0, (i64)B, (i64)n, (i64)dup, dhf);
  return TM_OK;
}

/* torch.optim.RAdam step (the optimizer code/MyOptimizer/optim_factory.py:78-80 builds for opt 'radam'; L2 weight decay folded
 * into the gradient per tensor) on up to TM_OPT_MAX_TENSORS tensors in one launch.  `step` is the 1-based step count. */
int tm_radam_step(const tm_opt_tensor* tensors, int64_t count, int64_t step, float lr, float beta1, float beta2, float eps, void* stream) {
  OptBatch b;
  unsigned gx = 1;
  TM_PROPAGATE(opt_batch(tensors, count, &b, &gx));
  for (int64_t i = 0; i < count; ++i) TM_CHECK_ARG(tensors[i].g && tensors[i].v, TM_ERR_BAD_ARGUMENT, "tm_radam_step: NULL g / v");
  TM_CHECK_ARG(step >= 1, TM_ERR_BAD_ARGUMENT, "tm_radam_step: step counts from 1");
  const double b1t = pow((double)beta1, (double)step), b2t = pow((double)beta2, (double)step);
  const double bc1 = 1.0 - b1t, bc2 = 1.0 - b2t;
  const double rho_inf = 2.0 / (1.0 - (double)beta2) - 1.0;
  const double rho_t = rho_inf - 2.0 * (double)step * b2t / bc2;
  const int rectified = rho_t > 5.0;
  const double rect = rectified ? sqrt((rho_t - 4.0) * (rho_t - 2.0) * rho_inf / ((rho_inf - 4.0) * (rho_inf - 2.0) * rho_t)) : 0.0;
  BW_LAUNCH(radam_kernel, dim3(gx, (unsigned)count), dim3(BT), (cudaStream_t)stream, b, lr, beta1, beta2, eps, (float)bc1, (float)sqrt(bc2),
            (float)rect, rectified);
  return TM_OK;
}
/* Lookahead slow-weight update (code/MyOptimizer/lookahead.py:28-41): tensors[i].p = fast weights, tensors[i].m = slow buffer. */
int tm_lookahead_sync(const tm_opt_tensor* tensors, int64_t count, float alpha, void* stream) {
  OptBatch b;
  unsigned gx = 1;
  TM_PROPAGATE(opt_batch(tensors, count, &b, &gx));
  BW_LAUNCH(lookahead_kernel, dim3(gx, (unsigned)count), dim3(BT), (cudaStream_t)stream, b, alpha);
  return TM_OK;
}

}  // extern "C"
