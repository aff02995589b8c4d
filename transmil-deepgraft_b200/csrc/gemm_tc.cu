// tcgen05 / TMA GEMM:  C[M,N] = epilogue(A[M,K] . W[N,K]^T), tf32 operands, fp32 accumulate.
//
// Persistent, warp-specialised, one CTA per SM:
//   warp 0    TMA producer: A box [128 x 32 fp32] + W box [256 x 32 fp32] per stage (SWIZZLE_128B), 4-stage mbarrier ring
//   warp 1    MMA issuer: 4 x tcgen05.mma.kind::tf32 (M128 N256 K8) per stage, accumulators in TMEM, double buffered
//             (2 x 256 columns) so the epilogue of tile i overlaps the main loop of tile i+1
//   warps 2-5 epilogue: tcgen05.ld (32 lanes x 32 columns) -> per-warp smem transpose -> coalesced 128-bit global stores
//             through the path's fused epilogues (bias+GELU+duplicate-pad rows, head-major q/k/v split, bias+residual)
// Operands are expected tf32-representable (producers round with cvt.rna), so the MMA's truncation is exact.
#include <mutex>

#include "common.cuh"
#include "gemm_epilogue.cuh"
#include "tc_common.cuh"

namespace tmil {

// ---- host: driver entry point for tensor-map encoding (no link-time libcuda dependency) -----------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static EncodeTiledFn g_encode = nullptr;
static int g_tc_state = -1;  // -1 unknown, 0 unavailable, 1 available
static std::mutex g_tc_mutex;

bool tc_available() {
  std::lock_guard<std::mutex> lk(g_tc_mutex);
  if (g_tc_state >= 0) return g_tc_state == 1;
  g_tc_state = 0;
  int dev = 0, major = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return false;
  if (cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev) != cudaSuccess || major != 10) return false;
  void* fn = nullptr;
  cudaDriverEntryPointQueryResult qres;
  if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres) != cudaSuccess || fn == nullptr ||
      qres != cudaDriverEntryPointSuccess)
    return false;
  g_encode = (EncodeTiledFn)fn;
  g_tc_state = 1;
  return true;
}

int make_tmap_2d(CUtensorMap* out, const float* base, uint64_t rows, uint64_t cols, uint64_t ld, uint32_t box_rows, uint32_t box_cols) {
  TM_CHECK_ARG(tc_available(), TM_ERR_UNSUPPORTED_ARCH, "TMA tensor maps need an sm_100 device and driver");
  TM_CHECK_ARG(aligned16(base) && (ld * sizeof(float)) % 16 == 0, TM_ERR_BAD_ALIGNMENT, "TMA operand must be 16-byte aligned with a 16-byte row pitch");
  cuuint64_t dims[2] = {cols, rows};
  cuuint64_t strides[1] = {ld * sizeof(float)};
  cuuint32_t box[2] = {box_cols, box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = g_encode(out, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, (void*)base, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                        CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  TM_CHECK_ARG(r == CUDA_SUCCESS, TM_ERR_CUDA, "cuTensorMapEncodeTiled(2d) failed with %d (rows=%llu cols=%llu ld=%llu)", (int)r,
               (unsigned long long)rows, (unsigned long long)cols, (unsigned long long)ld);
  return TM_OK;
}

int make_tmap_3d(CUtensorMap* out, const float* base, uint64_t batch, uint64_t rows, uint64_t cols, uint64_t batch_stride, uint64_t ld,
                 uint32_t box_rows, uint32_t box_cols) {
  TM_CHECK_ARG(tc_available(), TM_ERR_UNSUPPORTED_ARCH, "TMA tensor maps need an sm_100 device and driver");
  TM_CHECK_ARG(aligned16(base) && (ld * sizeof(float)) % 16 == 0 && (batch_stride * sizeof(float)) % 16 == 0, TM_ERR_BAD_ALIGNMENT,
               "TMA operand must be 16-byte aligned with 16-byte pitches");
  cuuint64_t dims[3] = {cols, rows, batch};
  cuuint64_t strides[2] = {ld * sizeof(float), batch_stride * sizeof(float)};
  cuuint32_t box[3] = {box_cols, box_rows, 1};
  cuuint32_t estr[3] = {1, 1, 1};
  CUresult r = g_encode(out, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, (void*)base, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                        CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  TM_CHECK_ARG(r == CUDA_SUCCESS, TM_ERR_CUDA, "cuTensorMapEncodeTiled(3d) failed with %d", (int)r);
  return TM_OK;
}

int make_tmap_3d_f16(CUtensorMap* out, const void* base, uint64_t batch, uint64_t rows, uint64_t cols, uint64_t batch_stride, uint64_t ld,
                     uint32_t box_rows, uint32_t box_cols) {
  TM_CHECK_ARG(tc_available(), TM_ERR_UNSUPPORTED_ARCH, "TMA tensor maps need an sm_100 device and driver");
  TM_CHECK_ARG(aligned16(base) && (ld * 2) % 16 == 0 && (batch_stride * 2) % 16 == 0 && box_cols * 2 == 128 && box_rows <= 256,
               TM_ERR_BAD_ALIGNMENT, "fp16 TMA operand: 16-byte aligned, 16-byte pitches, 128-byte box rows, <= 256 box rows");
  cuuint64_t dims[3] = {cols, rows, batch};
  cuuint64_t strides[2] = {ld * 2, batch_stride * 2};
  cuuint32_t box[3] = {box_cols, box_rows, 1};
  cuuint32_t estr[3] = {1, 1, 1};
  CUresult r = g_encode(out, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 3, const_cast<void*>(base), dims, strides, box, estr,
                        CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                        CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  TM_CHECK_ARG(r == CUDA_SUCCESS, TM_ERR_CUDA, "cuTensorMapEncodeTiled(3d f16) failed with %d (batch=%llu rows=%llu cols=%llu)", (int)r,
               (unsigned long long)batch, (unsigned long long)rows, (unsigned long long)cols);
  return TM_OK;
}

// ---- kernel ------------------------------------------------------------------------------------------------
constexpr int TC_BM = 128, TC_BN = 256, TC_BK = 32, TC_STAGES = 4;
constexpr int TC_A_BYTES = TC_BM * TC_BK * 4;   // 16 KB
constexpr int TC_B_BYTES = TC_BN * TC_BK * 4;   // 32 KB
constexpr int TC_STG_LD = 36;                   // floats per staged row (32 + 4: conflict-free 128-bit access)
constexpr int TC_STG_BYTES = 32 * TC_STG_LD * 4;
constexpr int TC_THREADS = 192;
constexpr size_t TC_SMEM = 1024 + TC_STAGES * (TC_A_BYTES + TC_B_BYTES) + 4 * TC_STG_BYTES + 256;

__global__ void __launch_bounds__(TC_THREADS, 1)
gemm_tc_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, const GemmParams p, int m_tiles,
               int n_tiles, int k_blocks) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  uint8_t* sA = smem;
  uint8_t* sB = sA + TC_STAGES * TC_A_BYTES;
  float* stg = reinterpret_cast<float*>(sB + TC_STAGES * TC_B_BYTES);
  uint64_t* bars = reinterpret_cast<uint64_t*>(reinterpret_cast<uint8_t*>(stg) + 4 * TC_STG_BYTES);
  uint64_t* full = bars;                 // [TC_STAGES]  TMA -> MMA
  uint64_t* empty = bars + TC_STAGES;    // [TC_STAGES]  MMA -> TMA
  uint64_t* tfull = bars + 2 * TC_STAGES;   // [2]       MMA -> epilogue
  uint64_t* tempty = tfull + 2;             // [2]       epilogue -> MMA
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty + 2);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
  }
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < TC_STAGES; ++i) { mbar_init(&full[i], 1); mbar_init(&empty[i], 1); }
    for (int i = 0; i < 2; ++i) { mbar_init(&tfull[i], 1); mbar_init(&tempty[i], 4); }
    fence_barrier_init();
  }
  if (warp == 2) tmem_alloc(tmem_slot, 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const int total = m_tiles * n_tiles;

  if (warp == 0) {
    if (lane == 0) {
      int stage = 0; uint32_t phase = 0;
      for (int tile = blockIdx.x; tile < total; tile += gridDim.x) {
        const int m_blk = tile / n_tiles, n_blk = tile - m_blk * n_tiles;
        for (int kb = 0; kb < k_blocks; ++kb) {
          mbar_wait(&empty[stage], phase ^ 1);
          mbar_arrive_expect_tx(&full[stage], TC_A_BYTES + TC_B_BYTES);
          tma_load_2d(sA + stage * TC_A_BYTES, &tmA, &full[stage], kb * TC_BK, m_blk * TC_BM);
          tma_load_2d(sB + stage * TC_B_BYTES, &tmB, &full[stage], kb * TC_BK, n_blk * TC_BN);
          if (++stage == TC_STAGES) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      constexpr uint32_t idesc = umma_idesc_tf32(TC_BM, TC_BN);
      int stage = 0; uint32_t phase = 0;
      int it = 0;
      for (int tile = blockIdx.x; tile < total; tile += gridDim.x, ++it) {
        const int as = it & 1;
        const uint32_t aphase = (it >> 1) & 1;
        mbar_wait(&tempty[as], aphase ^ 1);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + as * TC_BN;
        for (int kb = 0; kb < k_blocks; ++kb) {
          mbar_wait(&full[stage], phase);
          tc_fence_after();
          const uint64_t a_desc = umma_desc_k_sw128(smem_u32(sA + stage * TC_A_BYTES));
          const uint64_t b_desc = umma_desc_k_sw128(smem_u32(sB + stage * TC_B_BYTES));
#pragma unroll
          for (int k = 0; k < TC_BK / 8; ++k)  // 8 tf32 = 32 bytes per MMA: advance the start address inside the swizzle atom
            umma_tf32_ss(d_tmem, a_desc + (uint64_t)(k * 2), b_desc + (uint64_t)(k * 2), idesc, (kb | k) != 0 ? 1u : 0u);
          umma_commit(&empty[stage]);
          if (++stage == TC_STAGES) { stage = 0; phase ^= 1; }
        }
        umma_commit(&tfull[as]);
      }
    }
  } else {
    const int q = warp & 3;   // TMEM lane quarter this warp may access
    float* my_stg = stg + (warp - 2) * (32 * TC_STG_LD);
    const float* Lb = (p.L != nullptr && p.epi == EPI_AXPBY) ? p.L : nullptr;
    int it = 0;
    for (int tile = blockIdx.x; tile < total; tile += gridDim.x, ++it) {
      const int m_blk = tile / n_tiles, n_blk = tile - m_blk * n_tiles;
      const int as = it & 1;
      const uint32_t aphase = (it >> 1) & 1;
      mbar_wait(&tfull[as], aphase);
      tc_fence_after();
      const uint32_t t_row = tmem_base + ((uint32_t)(q * 32) << 16) + as * TC_BN;
#pragma unroll 1
      for (int chunk = 0; chunk < TC_BN / 32; ++chunk) {
        float v[32];
        tmem_ld_32x32(t_row + chunk * 32, v);
        if (chunk == TC_BN / 32 - 1) {   // accumulator fully read: hand the TMEM buffer back to the MMA warp
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(&tempty[as]);
        }
#pragma unroll
        for (int i = 0; i < 8; ++i)
          *reinterpret_cast<float4*>(my_stg + lane * TC_STG_LD + i * 4) = make_float4(v[4 * i], v[4 * i + 1], v[4 * i + 2], v[4 * i + 3]);
        __syncwarp();
#pragma unroll
        for (int rr = 0; rr < 32; rr += 4) {
          const int rl = rr + (lane >> 3), cl = (lane & 7) * 4;
          const float4 acc = *reinterpret_cast<const float4*>(my_stg + rl * TC_STG_LD + cl);
          const int r = m_blk * TC_BM + q * 32 + rl;
          const int c = n_blk * TC_BN + chunk * 32 + cl;
          if (r < p.M && c < p.N) epilogue_store4(p, r, c, acc, Lb, p.C);
        }
        __syncwarp();
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 512);
  }
}

int gemm_tc(const GemmParams& p, cudaStream_t st) {
  TM_CHECK_ARG(p.M > 0 && p.N > 0 && p.K > 0, TM_ERR_BAD_SHAPE, "gemm_tc: empty problem");
  TM_CHECK_ARG(!p.b_is_kn && (p.batch <= 1), TM_ERR_BAD_ARGUMENT, "gemm_tc: only single NT GEMMs (weights [N,K]) are supported");
  TM_CHECK_ARG(p.N % 4 == 0 && p.K % 4 == 0, TM_ERR_BAD_ALIGNMENT, "gemm_tc: N and K must be multiples of 4");
  CUtensorMap tmA, tmB;
  TM_PROPAGATE(make_tmap_2d(&tmA, p.A, (uint64_t)p.M, (uint64_t)p.K, (uint64_t)p.lda, TC_BM, TC_BK));
  TM_PROPAGATE(make_tmap_2d(&tmB, p.Bm, (uint64_t)p.N, (uint64_t)p.K, (uint64_t)p.ldb, TC_BN, TC_BK));
  static int sms = 0;
  static bool attr_done = false;
  if (!attr_done) {
    int dev = 0;
    TM_CUDA(cudaGetDevice(&dev));
    TM_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    TM_CUDA(cudaFuncSetAttribute(gemm_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)TC_SMEM));
    attr_done = true;
  }
  const int m_tiles = (int)ceil_div(p.M, TC_BM), n_tiles = (int)ceil_div(p.N, TC_BN), k_blocks = (int)ceil_div(p.K, TC_BK);
  const int total = m_tiles * n_tiles;
  const int grid = total < sms ? total : sms;
  gemm_tc_kernel<<<grid, TC_THREADS, TC_SMEM, st>>>(tmA, tmB, p, m_tiles, n_tiles, k_blocks);
  TM_LAUNCH_CHECK();
  return TM_OK;
}

}  // namespace tmil
