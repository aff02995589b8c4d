// K6+K7: attn2 = softmax(q_lm k_lm^T) and Z = moore_penrose_iter_pinv(attn2, 6)
// (nystrom-attention 0.0.11).  The init scalar max(colsum)*max(rowsum) is taken
// over ALL bags and heads of the call, exactly like the reference; it is computed
// with a fixed reduction order so the result is reproducible run to run.
//
// v1 iteration engine: 24 batched fp32 CUDA-core GEMMs with the polynomial terms
// fused into their epilogues:
//   XZ = A Z;  T1 = 7 XZ - XZ XZ;  T2 = 15 XZ - XZ T1;  Z' = 3.25 Z - 0.25 Z T2
// which equals 1/4 Z (13 I - XZ (15 I - XZ (7 I - XZ))) in exact arithmetic.
#include "common.cuh"
#include "tc_common.cuh"

namespace tmil {

constexpr int LM = TM_LANDMARKS;   // 256
constexpr int ROWS = 16;           // rows of attn2 per CTA (16 threads per row: 128 CTAs per bag)
constexpr int KLD = 68;            // padded row stride of k_lm in smem (floats)

// grid (LM/ROWS, B*8).  smem: k_lm [256][68] + q rows [16][68] + P tile [16][257]
template <bool ROUND>
__global__ void __launch_bounds__(256) attn2_kernel(const float* __restrict__ q_lm, const float* __restrict__ k_lm,
                                                    float* __restrict__ attn2, float* __restrict__ rowsum,
                                                    float* __restrict__ colpart) {
  extern __shared__ __align__(16) float sm[];
  pdl_sync();
  float* ks = sm;                       // 256*68
  float* qs = ks + LM * KLD;            // 32*68
  float* ps = qs + ROWS * KLD;          // 32*257
  const int bh = blockIdx.y, tile = blockIdx.x, tid = threadIdx.x;
  const float* kg = k_lm + (long long)bh * LM * TM_DHEAD;
  const float* qg = q_lm + ((long long)bh * LM + tile * ROWS) * TM_DHEAD;
  {
    float4 kr[16];   // all 16 loads of the 64 KB k_lm tile in flight before the first store (the kernel is on the pinv critical path)
#pragma unroll
    for (int u = 0; u < 16; ++u) kr[u] = __ldg(reinterpret_cast<const float4*>(kg) + tid + 256 * u);
#pragma unroll
    for (int u = 0; u < 16; ++u) {
      const int i = tid + 256 * u, r = i >> 4, c = (i & 15) * 4;
      *reinterpret_cast<float4*>(ks + r * KLD + c) = kr[u];
    }
  }
  for (int i = tid; i < ROWS * 16; i += 256) {
    int r = i >> 4, c = (i & 15) * 4;
    *reinterpret_cast<float4*>(qs + r * KLD + c) = __ldg(reinterpret_cast<const float4*>(qg + r * TM_DHEAD + c));
  }
  __syncthreads();
  const int r = tid >> 4, cg = tid & 15;   // 16 threads per row, columns cg, cg+16, ...
  float s[16];
#pragma unroll
  for (int jj = 0; jj < 16; ++jj) s[jj] = 0.f;
#pragma unroll 4
  for (int d = 0; d < TM_DHEAD; d += 4) {
    float4 qv = *reinterpret_cast<const float4*>(qs + r * KLD + d);
#pragma unroll
    for (int jj = 0; jj < 16; ++jj) {
      float4 kv = *reinterpret_cast<const float4*>(ks + (jj * 16 + cg) * KLD + d);
      s[jj] = fmaf(qv.x, kv.x, s[jj]); s[jj] = fmaf(qv.y, kv.y, s[jj]);
      s[jj] = fmaf(qv.z, kv.z, s[jj]); s[jj] = fmaf(qv.w, kv.w, s[jj]);
    }
  }
  float mx = s[0];
#pragma unroll
  for (int jj = 1; jj < 16; ++jj) mx = fmaxf(mx, s[jj]);
#pragma unroll
  for (int o = 1; o < 16; o <<= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
  float sum = 0.f;
#pragma unroll
  for (int jj = 0; jj < 16; ++jj) { s[jj] = expf(s[jj] - mx); sum += s[jj]; }
#pragma unroll
  for (int o = 1; o < 16; o <<= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
  float rs = 0.f;
#pragma unroll
  for (int jj = 0; jj < 16; ++jj) { float p = s[jj] / sum; ps[r * 257 + jj * 16 + cg] = p; rs += p; }
#pragma unroll
  for (int o = 1; o < 16; o <<= 1) rs += __shfl_xor_sync(0xffffffffu, rs, o);
  if (cg == 0) rowsum[(long long)bh * LM + tile * ROWS + r] = rs;   // |x|.sum(-1) (x >= 0)
  __syncthreads();
  float* ag = attn2 + ((long long)bh * LM + tile * ROWS) * LM;
  float cs = 0.f;
#pragma unroll 8
  for (int rr = 0; rr < ROWS; ++rr) {
    float p = ps[rr * 257 + tid];
    ag[rr * LM + tid] = ROUND ? round_tf32(p) : p;
    cs += p;
  }
  colpart[((long long)bh * (LM / ROWS) + tile) * LM + tid] = cs;     // partial |x|.sum(-2)
}

// One block: scale = max(rowsum) * max(colsum) over the whole call.
__global__ void __launch_bounds__(1024) pinv_scale_kernel(const float* __restrict__ rowsum, const float* __restrict__ colpart,
                                                          long long nbh, float* __restrict__ scale_out) {
  __shared__ float red[2][32];
  float mr = 0.f, mc = 0.f;
  for (long long i = threadIdx.x; i < nbh * LM; i += 1024) {
    mr = fmaxf(mr, rowsum[i]);
    const long long bh = i / LM; const int j = (int)(i - bh * LM);
    float c = 0.f;
#pragma unroll
    for (int t = 0; t < LM / ROWS; ++t) c += colpart[(bh * (LM / ROWS) + t) * LM + j];
    mc = fmaxf(mc, c);
  }
  mr = warp_max(mr); mc = warp_max(mc);
  if ((threadIdx.x & 31) == 0) { red[0][threadIdx.x >> 5] = mr; red[1][threadIdx.x >> 5] = mc; }
  __syncthreads();
  if (threadIdx.x < 32) {
    mr = warp_max(red[0][threadIdx.x]); mc = warp_max(red[1][threadIdx.x]);
    if (threadIdx.x == 0) { scale_out[0] = mr * mc; scale_out[1] = mr; scale_out[2] = mc; }
  }
}

// Z0 = A^T / scale, 32x32 smem transpose tiles.  grid (8, 8, B*8), block (32, 8)
template <bool TC>
__global__ void __launch_bounds__(256) pinv_z0_kernel(const float* __restrict__ A, const float* __restrict__ scale,
                                                      float* __restrict__ Z, float* __restrict__ ZT) {
  __shared__ float t[32][33];
  const long long base = (long long)blockIdx.z * LM * LM;
  const int x0 = blockIdx.x * 32, y0 = blockIdx.y * 32;
  for (int i = threadIdx.y; i < 32; i += 8) t[i][threadIdx.x] = A[base + (long long)(y0 + i) * LM + x0 + threadIdx.x];
  __syncthreads();
  const float sc = scale[0];
  for (int i = threadIdx.y; i < 32; i += 8) {
    const float z = t[threadIdx.x][i] / sc;
    Z[base + (long long)(x0 + i) * LM + y0 + threadIdx.x] = TC ? round_tf32(z) : z;
  }
  if (TC)   // Z0^T = A / scale: the B operand of the first A @ Z product
    for (int i = threadIdx.y; i < 32; i += 8) ZT[base + (long long)(y0 + i) * LM + x0 + threadIdx.x] = round_tf32(t[i][threadIdx.x] / sc);
}

static int pinv_gemm(const float* L, const float* R, float* C, float alpha, float beta, const float* addL, int nbh,
                     cudaStream_t st) {
  GemmParams p;
  memset(&p, 0, sizeof(p));
  p.A = L; p.Bm = R; p.C = C;
  p.M = p.N = p.K = LM; p.lda = p.ldb = p.ldc = p.ldl = LM;
  p.b_is_kn = 1; p.batch = nbh;
  p.strideA = p.strideB = p.strideC = p.strideL = (long long)LM * LM;
  p.epi = EPI_AXPBY; p.alpha = alpha; p.beta = beta; p.L = addL;
  return gemm_simt(p, st);
}

// One product of the chain on the tensor cores: every iterate is kept in both orientations, so X @ Y is the NT GEMM
// X . (Y^T)^T with tf32 operands (all stored iterates are tf32-rounded by their producer) and fp32 accumulation;
// out = alpha * X@Y + beta * L, with L supplied transposed (LT) and the result written as C and/or C^T.
int pinv_cluster_run(const float* attn2, const float* rowsum, const float* colpart, int nparts, const float* scale, int nbh,
                     void* scratch16, float* Z, cudaStream_t st);   // pinv_cluster.cu

static int pinv_gemm_tc(const float* X, const float* YT, float alpha, float beta, const float* LT, float* C, float* CT, bool round,
                        int nbh, cudaStream_t st) {
  GemmParams p;
  memset(&p, 0, sizeof(p));
  p.A = X; p.Bm = YT; p.C = C; p.CT = CT;
  p.M = p.N = p.K = LM; p.lda = p.ldb = p.ldc = p.ldl = LM;
  p.batch = nbh; p.bn128 = 1;
  p.strideA = p.strideB = p.strideC = p.strideL = (long long)LM * LM;
  p.epi = EPI_AXPBY_T; p.alpha = alpha; p.beta = beta; p.L = LT; p.round_out = round ? 1 : 0;
  return gemm_tc(p, st);
}

}  // namespace tmil

extern "C" int64_t tm_pinv_scratch_floats(int64_t B) {
  const int64_t nbh = B * TM_HEADS;
  return 7 * nbh * 65536 + nbh * 256 * (1 + 256 / 16) + 64;   // + rowsum, column partials (one per 16-row CTA), scale   // tensor-core chain: XZ, XZ^T, T1^T, T2^T, Z', Z^T, Z'^T
}

extern "C" int tm_pinv_fwd(const float* q_lm, const float* k_lm, int64_t B, float* attn2, float* Z, float* scratch,
                           uint32_t flags, void* stream) {
  using namespace tmil;
  TM_CHECK_ARG(B > 0 && q_lm && k_lm && attn2 && Z && scratch, TM_ERR_BAD_ARGUMENT, "tm_pinv_fwd: bad arguments");
  cudaStream_t st = (cudaStream_t)stream;
  const int nbh = (int)(B * TM_HEADS);
  const long long mat = 65536;
  const bool tc = !(flags & TM_FLAG_SIMT_ONLY) && tc_available();
  float* XZ = scratch;
  float* T1 = XZ + nbh * mat;
  float* T2 = T1 + nbh * mat;
  float* XZT = T2 + nbh * mat;
  float* Zb = XZT + nbh * mat;
  float* ZTa = Zb + nbh * mat;
  float* ZTb = ZTa + nbh * mat;
  float* rowsum = ZTb + nbh * mat;
  float* colpart = rowsum + (long long)nbh * LM;
  float* scale = colpart + (long long)nbh * LM * (LM / ROWS);

  static unsigned long long attr_done = 0;
  const size_t smem = (size_t)(LM * KLD + ROWS * KLD + ROWS * 257) * sizeof(float);
  if (first_use_on_device(&attr_done)) {
    TM_CUDA(cudaFuncSetAttribute(attn2_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    TM_CUDA(cudaFuncSetAttribute(attn2_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  }
  const bool cluster = tc && !(flags & TM_FLAG_PINV_GEMM_CHAIN);
  if (tc && !cluster) launch_pdl(attn2_kernel<true>, dim3(LM / ROWS, nbh), dim3(256), smem, st, q_lm, k_lm, attn2, rowsum, colpart);
  else launch_pdl(attn2_kernel<false>, dim3(LM / ROWS, nbh), dim3(256), smem, st, q_lm, k_lm, attn2, rowsum, colpart);
  TM_LAUNCH_CHECK();
  // one 4-CTA-cluster kernel for the whole chain (fp16 operands); attn2 stays exact fp32.  For a few heads the global init
  // scalar is reduced inside the operand-preparation kernel (one launch less on the latency-critical path); with many
  // heads that per-CTA reduction would be quadratic, so the one-block kernel computes it once.
  if (cluster && nbh <= 16) return pinv_cluster_run(attn2, rowsum, colpart, LM / ROWS, nullptr, nbh, XZ, Z, st);
  pinv_scale_kernel<<<1, 1024, 0, st>>>(rowsum, colpart, nbh, scale);
  TM_LAUNCH_CHECK();
  if (cluster) return pinv_cluster_run(attn2, rowsum, colpart, LM / ROWS, scale, nbh, XZ, Z, st);
  if (tc) {
    pinv_z0_kernel<true><<<dim3(8, 8, nbh), dim3(32, 8), 0, st>>>(attn2, scale, Z, ZTa);
    TM_LAUNCH_CHECK();
    float* Zc = Z; float* Zn = Zb; float* ZTc = ZTa; float* ZTn = ZTb;
    float* T1T = T1; float* T2T = T2;
    for (int it = 0; it < TM_PINV_ITERS; ++it) {
      const bool last = it == TM_PINV_ITERS - 1;
      TM_PROPAGATE(pinv_gemm_tc(attn2, ZTc, 1.f, 0.f, nullptr, XZ, XZT, true, nbh, st));          // XZ = A Z (+ XZ^T)
      TM_PROPAGATE(pinv_gemm_tc(XZ, XZT, -1.f, 7.f, XZT, nullptr, T1T, true, nbh, st));           // T1^T,  T1 = 7 XZ - XZ XZ
      TM_PROPAGATE(pinv_gemm_tc(XZ, T1T, -1.f, 15.f, XZT, nullptr, T2T, true, nbh, st));          // T2^T,  T2 = 15 XZ - XZ T1
      TM_PROPAGATE(pinv_gemm_tc(Zc, T2T, -0.25f, 3.25f, ZTc, Zn, last ? nullptr : ZTn, !last, nbh, st));  // Z' = 3.25 Z - 0.25 Z T2
      float* t = Zc; Zc = Zn; Zn = t;
      t = ZTc; ZTc = ZTn; ZTn = t;
    }
    return TM_OK;   // 6 swaps: the final iterate is back in the caller's Z
  }
  pinv_z0_kernel<false><<<dim3(8, 8, nbh), dim3(32, 8), 0, st>>>(attn2, scale, Z, nullptr);
  TM_LAUNCH_CHECK();
  float* Zc = Z;      // current iterate
  float* Zn = T1;     // T1's storage is free again when Z' is formed
  for (int it = 0; it < TM_PINV_ITERS; ++it) {
    float* t1 = Zn;   // T1 lives where the next iterate will be written
    TM_PROPAGATE(pinv_gemm(attn2, Zc, XZ, 1.f, 0.f, nullptr, nbh, st));       // XZ = A Z
    TM_PROPAGATE(pinv_gemm(XZ, XZ, t1, -1.f, 7.f, XZ, nbh, st));              // T1 = 7 XZ - XZ XZ
    TM_PROPAGATE(pinv_gemm(XZ, t1, T2, -1.f, 15.f, XZ, nbh, st));             // T2 = 15 XZ - XZ T1
    TM_PROPAGATE(pinv_gemm(Zc, T2, Zn, -0.25f, 3.25f, Zc, nbh, st));          // Z' = 3.25 Z - 0.25 Z T2
    float* tmp = Zc; Zc = Zn; Zn = tmp;
  }
  // 6 swaps: the final iterate is back in the caller's Z
  return TM_OK;
}
