// fp32 CUDA-core GEMM with the path's fused epilogues.
//
// This is the exact-fp32 kernel family: it backs the small latency-bound
// contractions (pinv steps, Z@W3) and serves as the on-device cross-check of the
// tcgen05/TMA GEMMs in gemm_tc.cu (TM_FLAG_SIMT_ONLY).  128x128x16 CTA tile,
// 256 threads, 8x8 register micro-tile, register-staged double buffering.
#include "common.cuh"
#include "gemm_epilogue.cuh"

namespace tmil {

constexpr int BM = 128, BN = 128, BK = 16, NT = 256;

template <bool B_IS_KN>
__global__ void __launch_bounds__(NT, 2) gemm_simt_kernel(const GemmParams p) {
  __shared__ __align__(16) float As[2][BK][BM + 4];
  __shared__ __align__(16) float Bs[2][BK][BN + 4];
  const int tid = threadIdx.x;
  const int m0 = blockIdx.y * BM, n0 = blockIdx.x * BN;
  const int z = blockIdx.z;
  const float* __restrict__ Ab = p.A + (long long)z * p.strideA;
  const float* __restrict__ Bb = p.Bm + (long long)z * p.strideB;
  float* __restrict__ Cb = p.C != nullptr ? p.C + (long long)z * p.strideC : nullptr;
  const float* __restrict__ Lb = (p.L != nullptr && p.epi == EPI_AXPBY) ? p.L + (long long)z * p.strideL : nullptr;

  // global->register staging: A tile 128 rows x 16 k = 512 float4, 2 per thread
  const float* a_ptr[2]; bool a_ok[2]; int a_r[2], a_k[2];
#pragma unroll
  for (int i = 0; i < 2; ++i) {
    int idx = tid + i * NT;
    a_r[i] = idx >> 2; a_k[i] = (idx & 3) * 4;
    int gr = m0 + a_r[i];
    a_ok[i] = gr < p.M;
    a_ptr[i] = Ab + (long long)(a_ok[i] ? gr : 0) * p.lda + a_k[i];
  }
  const float* b_ptr[2]; bool b_ok[2]; int b_r[2], b_c[2];
#pragma unroll
  for (int i = 0; i < 2; ++i) {
    int idx = tid + i * NT;
    if (B_IS_KN) {          // tile 16 k x 128 n, float4 along n
      b_r[i] = idx >> 5; b_c[i] = (idx & 31) * 4;
      b_ok[i] = (n0 + b_c[i]) < p.N;
      b_ptr[i] = Bb + (long long)b_r[i] * p.ldb + n0 + b_c[i];
    } else {                // tile 128 n x 16 k, float4 along k
      b_r[i] = idx >> 2; b_c[i] = (idx & 3) * 4;
      b_ok[i] = (n0 + b_r[i]) < p.N;
      b_ptr[i] = Bb + (long long)(n0 + b_r[i]) * p.ldb + b_c[i];
    }
  }
  float4 ra[2], rb[2];
  auto load_tiles = [&](int k0) {
#pragma unroll
    for (int i = 0; i < 2; ++i) {
      ra[i] = (a_ok[i] && (k0 + a_k[i]) < p.K) ? __ldg(reinterpret_cast<const float4*>(a_ptr[i] + k0)) : make_float4(0, 0, 0, 0);
      if (B_IS_KN)
        rb[i] = (b_ok[i] && (k0 + b_r[i]) < p.K) ? __ldg(reinterpret_cast<const float4*>(b_ptr[i] + (long long)k0 * p.ldb)) : make_float4(0, 0, 0, 0);
      else
        rb[i] = (b_ok[i] && (k0 + b_c[i]) < p.K) ? __ldg(reinterpret_cast<const float4*>(b_ptr[i] + k0)) : make_float4(0, 0, 0, 0);
    }
  };
  auto store_tiles = [&](int buf) {
#pragma unroll
    for (int i = 0; i < 2; ++i) {
      As[buf][a_k[i] + 0][a_r[i]] = ra[i].x; As[buf][a_k[i] + 1][a_r[i]] = ra[i].y;
      As[buf][a_k[i] + 2][a_r[i]] = ra[i].z; As[buf][a_k[i] + 3][a_r[i]] = ra[i].w;
      if (B_IS_KN) {
        *reinterpret_cast<float4*>(&Bs[buf][b_r[i]][b_c[i]]) = rb[i];
      } else {
        Bs[buf][b_c[i] + 0][b_r[i]] = rb[i].x; Bs[buf][b_c[i] + 1][b_r[i]] = rb[i].y;
        Bs[buf][b_c[i] + 2][b_r[i]] = rb[i].z; Bs[buf][b_c[i] + 3][b_r[i]] = rb[i].w;
      }
    }
  };

  const int tx = tid & 15, ty = tid >> 4;
  float acc[8][8];
#pragma unroll
  for (int i = 0; i < 8; ++i)
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[i][j] = 0.f;

  const int nk = (p.K + BK - 1) / BK;
  load_tiles(0);
  store_tiles(0);
  __syncthreads();
  for (int kt = 0; kt < nk; ++kt) {
    const int buf = kt & 1;
    if (kt + 1 < nk) load_tiles((kt + 1) * BK);
#pragma unroll
    for (int kk = 0; kk < BK; ++kk) {
      float4 a0 = *reinterpret_cast<const float4*>(&As[buf][kk][ty * 4]);
      float4 a1 = *reinterpret_cast<const float4*>(&As[buf][kk][64 + ty * 4]);
      float4 b0 = *reinterpret_cast<const float4*>(&Bs[buf][kk][tx * 4]);
      float4 b1 = *reinterpret_cast<const float4*>(&Bs[buf][kk][64 + tx * 4]);
      float av[8] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
      float bv[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
#pragma unroll
      for (int i = 0; i < 8; ++i)
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[i][j] = fmaf(av[i], bv[j], acc[i][j]);
    }
    if (kt + 1 < nk) {
      store_tiles(buf ^ 1);
      __syncthreads();
    }
  }

#pragma unroll
  for (int i = 0; i < 8; ++i) {
    int r = m0 + (i < 4 ? ty * 4 + i : 64 + ty * 4 + (i - 4));
    if (r >= p.M) continue;
#pragma unroll
    for (int jh = 0; jh < 2; ++jh) {
      int c = n0 + jh * 64 + tx * 4;
      if (c >= p.N) continue;
      epilogue_store4(p, r, c, make_float4(acc[i][jh * 4 + 0], acc[i][jh * 4 + 1], acc[i][jh * 4 + 2], acc[i][jh * 4 + 3]), Lb, Cb);
    }
  }
}

int gemm_simt(const GemmParams& p, cudaStream_t st) {
  TM_CHECK_ARG(p.M > 0 && p.N > 0 && p.K > 0, TM_ERR_BAD_SHAPE, "gemm: empty problem %dx%dx%d", p.M, p.N, p.K);
  TM_CHECK_ARG(p.K % 4 == 0 && p.N % 4 == 0 && p.lda % 4 == 0 && p.ldb % 4 == 0, TM_ERR_BAD_ALIGNMENT,
               "gemm: K, N, lda, ldb must be multiples of 4 (K=%d N=%d lda=%d ldb=%d)", p.K, p.N, p.lda, p.ldb);
  TM_CHECK_ARG(aligned16(p.A) && aligned16(p.Bm), TM_ERR_BAD_ALIGNMENT, "gemm: operands must be 16-byte aligned");
  dim3 grid((unsigned)ceil_div(p.N, BN), (unsigned)ceil_div(p.M, BM), (unsigned)(p.batch > 0 ? p.batch : 1));
  TM_CHECK_ARG(grid.y <= 65535u && grid.z <= 65535u, TM_ERR_BAD_SHAPE, "gemm: grid too large");
  if (p.b_is_kn)
    gemm_simt_kernel<true><<<grid, NT, 0, st>>>(p);
  else
    gemm_simt_kernel<false><<<grid, NT, 0, st>>>(p);
  TM_LAUNCH_CHECK();
  return TM_OK;
}

}  // namespace tmil

extern "C" int tm_gemm_nt(const float* A, const float* W, float* C, int64_t M, int64_t N, int64_t K,
                          uint32_t flags, void* stream) {
  tmil::GemmParams p;
  memset(&p, 0, sizeof(p));
  p.A = A; p.Bm = W; p.C = C;
  p.M = (int)M; p.N = (int)N; p.K = (int)K;
  p.lda = (int)K; p.ldb = (int)K; p.ldc = (int)N;
  p.batch = 1; p.epi = tmil::EPI_AXPBY; p.alpha = 1.f; p.beta = 0.f;
  if (!(flags & TM_FLAG_SIMT_ONLY) && tmil::tc_available()) return tmil::gemm_tc(p, (cudaStream_t)stream);
  return tmil::gemm_simt(p, (cudaStream_t)stream);
}
