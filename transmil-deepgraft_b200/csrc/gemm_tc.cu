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
#include <cuda_fp16.h>

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

// fp32 [batch, rows, cols] (cols contiguous) with the 32-byte-atom 128B swizzle: the shared-memory form of an MN-major tf32
// tcgen05 operand (rows = k, cols = m or n).  box = [1, box_rows, 32 floats].
int make_tmap_3d_mn32(CUtensorMap* out, const float* base, uint64_t batch, uint64_t rows, uint64_t cols, uint64_t batch_stride, uint64_t ld,
                      uint32_t box_rows) {
  TM_CHECK_ARG(tc_available(), TM_ERR_UNSUPPORTED_ARCH, "TMA tensor maps need an sm_100 device and driver");
  TM_CHECK_ARG(aligned16(base) && (ld * sizeof(float)) % 16 == 0 && (batch_stride * sizeof(float)) % 16 == 0, TM_ERR_BAD_ALIGNMENT,
               "TMA operand must be 16-byte aligned with 16-byte pitches");
  cuuint64_t dims[3] = {cols, rows, batch};
  cuuint64_t strides[2] = {ld * sizeof(float), batch_stride * sizeof(float)};
  cuuint32_t box[3] = {32, box_rows, 1};
  cuuint32_t estr[3] = {1, 1, 1};
  CUresult r = g_encode(out, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, (void*)base, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                        CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  TM_CHECK_ARG(r == CUDA_SUCCESS, TM_ERR_CUDA, "cuTensorMapEncodeTiled(3d, 128B swizzle / 32B atom) failed with %d", (int)r);
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

// fp32 tensor of rank <= 5 with arbitrary (16-byte multiple) byte strides, unswizzled box, zero fill outside the extents.
int make_tmap_f32_nd(CUtensorMap* out, const float* base, int rank, const uint64_t* dims, const uint64_t* strides_bytes,
                     const uint32_t* box) {
  TM_CHECK_ARG(tc_available(), TM_ERR_UNSUPPORTED_ARCH, "TMA tensor maps need an sm_100 device and driver");
  TM_CHECK_ARG(rank >= 2 && rank <= 5 && aligned16(base), TM_ERR_BAD_ALIGNMENT, "TMA operand must be 16-byte aligned (rank 2..5)");
  cuuint64_t d[5], s[4];
  cuuint32_t b[5], e[5];
  for (int i = 0; i < rank; ++i) { d[i] = dims[i]; b[i] = box[i]; e[i] = 1; }
  for (int i = 0; i + 1 < rank; ++i) {
    TM_CHECK_ARG(strides_bytes[i] % 16 == 0, TM_ERR_BAD_ALIGNMENT, "TMA strides must be multiples of 16 bytes");
    s[i] = strides_bytes[i];
  }
  CUresult r = g_encode(out, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, (cuuint32_t)rank, (void*)base, d, s, b, e, CU_TENSOR_MAP_INTERLEAVE_NONE,
                        CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  TM_CHECK_ARG(r == CUDA_SUCCESS, TM_ERR_CUDA, "cuTensorMapEncodeTiled(nd) failed with %d", (int)r);
  return TM_OK;
}

// ---- kernel ------------------------------------------------------------------------------------------------
// Template parameters: BN = N tile (256: big weight GEMMs, 128: the batched 256^3 products of the pinv chain),
// F16 = operand type (false: tf32 in fp32 storage, BK = 32; true: fp16, BK = 64).  Either way a k-block row is 128 bytes
// (one SWIZZLE_128B atom row) and holds 4 UMMA K-steps of 32 bytes.
constexpr int TC_BM = 128;
constexpr int TC_A_BYTES = TC_BM * 128;         // 16 KB per stage
constexpr int TC_STG_LD = 36;                   // floats per staged row (32 + 4: conflict-free 128-bit access)
constexpr int TC_STG_BYTES = 32 * TC_STG_LD * 4;
constexpr int TC_THREADS = 192;
constexpr int TC_OUT_BYTES = 2 * 16384;         // epilogue staging: two [128 rows][128 bytes] SWIZZLE_128B boxes (TMA store) or the
                                                // four per-warp transpose tiles of the generic epilogue
static_assert(4 * TC_STG_BYTES <= TC_OUT_BYTES, "generic staging must fit the output staging region");

// Output tensor maps of the TMA-store epilogues (see epilogue_fast)
struct TcOutMaps { CUtensorMap m[4]; };
enum TcFast : int { FAST_NONE = 0, FAST_QKV16 = 1, FAST_BIAS_RESID = 2, FAST_BIAS_GELU = 3 };

template <int BN, int FAST = 0> struct TcCfg {
  static constexpr int B_BYTES = BN * 128;
  // the bias+residual epilogue streams the residual through four in-place [128][128 B] boxes (TMA load -> add -> TMA store)
  // and gives up one operand stage for them
  // (four boxes + three stages for the other TMA-store epilogues too was measured slower: fc1 99 -> 109 us, qkv 74 -> 77 us,
  // profiles/r1_gemm_epilogue_study.txt -- the operand pipeline needs its fourth stage more than the stores need boxes)
  static constexpr int STAGES = BN == 256 ? (FAST == 2 ? 3 : 4) : 6;
  static constexpr int OUT_BYTES = FAST == 2 ? 4 * 16384 : TC_OUT_BYTES;
  static constexpr int THREADS = FAST == 2 ? 224 : TC_THREADS;   // + one warp that loads the residual boxes
  static constexpr size_t SMEM = 1024 + STAGES * (TC_A_BYTES + B_BYTES) + OUT_BYTES + 256;
};

// ---- TMA-store epilogues ---------------------------------------------------------------------------------------
// The accumulator layout (thread = output row, 32 consecutive columns per tcgen05.ld) is finished in registers, written to
// a [128 rows][128 bytes] SWIZZLE_128B box in smem (conflict-free: the 16-byte chunk index is XOR-ed with the row) and
// stored with ONE cp.async.bulk.tensor per box: full-line writes, no per-element address arithmetic, and the row maps of
// the path become tensor-map coordinates -- rows with a negative / out-of-extent coordinate are simply not written
// (front-pad rows of to_out, rows past the bag in fc1).  Two boxes rotate so the store of box k overlaps the math of k+1.
//   FAST_QKV16       q/k/v head-major fp16 boxes [128 tokens][64 d]         (two 32-column chunks per box)
//   FAST_BIAS_RESID  h_out = h_in + acc + bias, fp32 boxes [128][32], row coordinate i - P
//   FAST_BIAS_GELU   C = gelu(acc + bias), fp32 boxes [128][32], plus the duplicate-pad rows as a second store
template <int BN, int FAST>
__device__ __forceinline__ void epilogue_fast(const GemmParams& p, const TcOutMaps& om, uint32_t tmem_base, uint64_t* tfull,
                                              uint64_t* tempty, uint64_t* lfull, uint64_t* lempty, uint8_t* obuf, int n_tiles,
                                              int total) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  constexpr int NBUF = FAST == FAST_BIAS_RESID ? 4 : 2;

  if (FAST == FAST_BIAS_RESID && warp == 6) {
    // ---- residual loader: h_in boxes [128 rows][32 cols] -> smem ring, one per (tile, 32-column chunk), in the order the
    // epilogue consumes them.  A thread-per-row read of the residual (16 bytes per lane at a 2 KB stride) ran at 1.3 TB/s;
    // the TMA box is a coalesced full-line read and lands in the layout the TMA store takes back out.
    if (lane == 0) {
      uint32_t unit = 0;
      for (int tile = blockIdx.x; tile < total; tile += gridDim.x) {
        const int m_blk = tile / n_tiles, n_blk = tile - m_blk * n_tiles;
        const int r0 = m_blk * TC_BM;
        const int bag = r0 / p.rows_per_bag, i0 = r0 - bag * p.rows_per_bag;
        const int shift = i0 < p.c_skip ? p.c_skip - i0 : 0;
        if (shift >= TC_BM) continue;
        for (int chunk = 0; chunk < BN / 32; ++chunk) {
          const int c0 = n_blk * BN + chunk * 32;
          if (c0 >= p.N) continue;
          const int b = unit % NBUF;
          mbar_wait(&lempty[b], ((unit / NBUF) & 1) ^ 1);
          mbar_arrive_expect_tx(&lfull[b], (TC_BM - shift) * 128);
          if (shift > 0) tma_load_3d(obuf + b * 16384, &om.m[3], &lfull[b], c0, p.c_row_off, bag);
          else tma_load_3d(obuf + b * 16384, &om.m[2], &lfull[b], c0, p.c_row_off + i0 - p.c_skip, bag);
          ++unit;
        }
      }
    }
    return;
  }

  const int q = warp & 3;
  const int rt = q * 32 + lane;                 // row inside the 128-row tile = TMEM lane
  const bool leader = threadIdx.x == 64;        // warp 2, lane 0 issues the stores
  uint32_t unit = 0;                            // boxes written so far (buffer = unit % NBUF)
  int it = 0;
  for (int tile = blockIdx.x; tile < total; tile += gridDim.x, ++it) {
    const int m_blk = tile / n_tiles, n_blk = tile - m_blk * n_tiles;
    const int as = it & 1;
    const uint32_t aphase = (it >> 1) & 1;
    const int r0 = m_blk * TC_BM;
    int bag = 0, i0 = r0;
    if (p.rows_per_bag > 0) { bag = r0 / p.rows_per_bag; i0 = r0 - bag * p.rows_per_bag; }
    if (FAST == FAST_QKV16) { bag = r0 / p.Sp; i0 = r0 - bag * p.Sp; }
    // FAST_BIAS_RESID: the first c_skip rows of a bag (front pad) are not stored.  TMA stores take no negative coordinates
    // here, so a tile of pad rows only is skipped and the tile that straddles the boundary moves a shorter box (maps m[1] /
    // m[3], 128 - shift rows) whose rows sit `shift` rows higher in smem.
    const int shift = (FAST == FAST_BIAS_RESID && i0 < p.c_skip) ? p.c_skip - i0 : 0;
    const int srow = rt - shift;                // row of the smem box
    mbar_wait(&tfull[as], aphase);
    tc_fence_after();
    if (p.dbg == 1) {   // measurement: release the accumulator, do nothing else
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&tempty[as]);
      continue;
    }
    const uint32_t t_row = tmem_base + ((uint32_t)(q * 32) << 16) + as * BN;
#pragma unroll 1
    for (int chunk = 0; chunk < BN / 32; ++chunk) {
      float v[32];
      tmem_ld_32x32(t_row + chunk * 32, v);
      if (p.dbg == 2) {   // measurement: TMEM loads and a stand-in for the math, nothing stored
        float acc = 0.f;
#pragma unroll
        for (int e = 0; e < 32; ++e) acc += v[e] * 0.125f;
        if (acc == 12345.678f) obuf[0] = 1;
        if (chunk == BN / 32 - 1) {
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(&tempty[as]);
        }
        continue;
      }
      if (chunk == BN / 32 - 1) {   // accumulator fully read: hand the TMEM buffer back to the MMA warp
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&tempty[as]);
      }
      const int c0 = n_blk * BN + chunk * 32;
      if (c0 >= p.N || shift >= TC_BM) continue;      // uniform over the CTA
      const bool last = FAST == FAST_QKV16 ? (chunk & 1) == 1 : true;
      const int b = unit % NBUF;
      uint8_t* rowp = obuf + b * 16384 + srow * 128;
      if (FAST == FAST_BIAS_GELU) {
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const float4 bb = __ldg(reinterpret_cast<const float4*>(p.bias + c0) + j);
          v[4 * j] = gelu_erf(v[4 * j] + bb.x); v[4 * j + 1] = gelu_erf(v[4 * j + 1] + bb.y);
          v[4 * j + 2] = gelu_erf(v[4 * j + 2] + bb.z); v[4 * j + 3] = gelu_erf(v[4 * j + 3] + bb.w);
        }
        if (p.round_out) {
#pragma unroll
          for (int e = 0; e < 32; ++e) v[e] = round_tf32(v[e]);
        }
      }
      if (FAST == FAST_BIAS_RESID) {
        mbar_wait(&lfull[b], (unit / NBUF) & 1);   // the residual box of this chunk has landed in the buffer
        if (srow >= 0) {
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            float4* cell = reinterpret_cast<float4*>(rowp + ((j ^ (srow & 7)) << 4));
            const float4 l = *cell;
            const float4 bb = __ldg(reinterpret_cast<const float4*>(p.bias + c0) + j);
            *cell = make_float4(l.x + (v[4 * j] + bb.x), l.y + (v[4 * j + 1] + bb.y), l.z + (v[4 * j + 2] + bb.z), l.w + (v[4 * j + 3] + bb.w));
          }
        }
      } else {
        // (no barrier here: the buffer of this unit was freed before the previous unit's barrier, see below)
        if (FAST == FAST_QKV16) {
          const float sc = c0 + p.col_off < TM_DIM ? 0.125f : 1.0f;   // q * dim_head^-0.5 (exact in fp16)
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            uint4 u;
            u.x = pack_f16x2(v[8 * j] * sc, v[8 * j + 1] * sc); u.y = pack_f16x2(v[8 * j + 2] * sc, v[8 * j + 3] * sc);
            u.z = pack_f16x2(v[8 * j + 4] * sc, v[8 * j + 5] * sc); u.w = pack_f16x2(v[8 * j + 6] * sc, v[8 * j + 7] * sc);
            *reinterpret_cast<uint4*>(rowp + ((((chunk & 1) * 4 + j) ^ (rt & 7)) << 4)) = u;
          }
        } else {
#pragma unroll
          for (int j = 0; j < 8; ++j)
            *reinterpret_cast<float4*>(rowp + ((j ^ (rt & 7)) << 4)) = make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
        }
      }
      if (last) {
        if (p.dbg == 3) { ++unit; continue; }   // measurement: smem writes only
        fence_proxy_async();         // generic-proxy smem writes -> visible to the TMA engine
        // ONE barrier per box.  Two boxes rotate: before arriving, the leader waits until the store of the PREVIOUS box
        // (issued a whole unit ago) has drained its buffer -- that is the buffer the next unit writes after this barrier.
        if (FAST != FAST_BIAS_RESID && leader) tma_store_wait_read<0>();
        named_bar_sync(1, 128);
        if (p.dbg == 4) { ++unit; continue; }   // measurement: everything but the TMA store
        if (FAST == FAST_QKV16 && p.lm_part != nullptr && c0 + p.col_off - 32 < 2 * TM_DIM) {
          // landmark partial sums of this box (q or k head slice), read back from the fp16 values the box holds: lane =
          // column pair, warp = row quarter; rows before `rb` belong to the tile's first landmark group, the rest to the next
          const uint8_t* box = obuf + b * 16384;
          const int g0 = i0 / p.lm_l;
          const int rb = (g0 + 1) * p.lm_l - i0;                    // first row of the next group (may lie beyond the tile)
          const int split = min(32, max(0, rb - q * 32));           // rows of this quarter that fall before the boundary
          const uint32_t coff = (uint32_t)(lane & 3) * 4;
          const int c16 = lane >> 2;
          float ax = 0.f, ay = 0.f, bx = 0.f, by = 0.f;
          int rr = 0;
          for (; rr < split; ++rr) {
            const int r = q * 32 + rr;
            const uint32_t u = *reinterpret_cast<const uint32_t*>(box + r * 128 + ((c16 ^ (r & 7)) << 4) + coff);
            const float2 f = __half22float2(*reinterpret_cast<const __half2*>(&u));
            ax += f.x; ay += f.y;
          }
          for (; rr < 32; ++rr) {
            const int r = q * 32 + rr;
            const uint32_t u = *reinterpret_cast<const uint32_t*>(box + r * 128 + ((c16 ^ (r & 7)) << 4) + coff);
            const float2 f = __half22float2(*reinterpret_cast<const __half2*>(&u));
            bx += f.x; by += f.y;
          }
          const long long trow = ((long long)bag * (p.Sp / TC_BM) + i0 / TC_BM) * 4 + q;
          float* dst = p.lm_part + (trow * 2) * (2 * TM_DIM) + (c0 + p.col_off - 32) + 2 * lane;
          *reinterpret_cast<float2*>(dst) = make_float2(ax, ay);
          *reinterpret_cast<float2*>(dst + 2 * TM_DIM) = make_float2(bx, by);
        }
        if (leader) {
          const uint8_t* src = obuf + b * 16384;
          if (FAST == FAST_QKV16) {
            const int cb = c0 + p.col_off - 32;  // first column of the 64-wide head slice
            const int which = cb / TM_DIM, hh = (cb % TM_DIM) / TM_DHEAD;
            tma_store_3d(&om.m[which], src, 0, i0, bag * TM_HEADS + hh);
          } else if (FAST == FAST_BIAS_RESID) {
            if (shift > 0) tma_store_3d(&om.m[1], src, c0, p.c_row_off, bag);
            else tma_store_3d(&om.m[0], src, c0, p.c_row_off + i0 - p.c_skip, bag);
          } else {
            tma_store_3d(&om.m[0], src, c0, p.c_row_off + i0, bag);
            if (i0 < p.dup) tma_store_3d(&om.m[1], src, c0, p.c_row_off + p.rows_per_bag + i0, bag);   // TransMIL.py:180
          }
          tma_store_commit();
          if (FAST == FAST_BIAS_RESID) {   // the store issued two units ago has read its box: give it back to the loader
            tma_store_wait_read<2>();
            if (unit >= 2) mbar_arrive(&lempty[(unit - 2) % NBUF]);
          }
        }
        ++unit;
      }
    }
  }
  if (leader) tma_store_wait_all<0>();
}

template <int BN, bool F16, int FAST>
__global__ void __launch_bounds__(224, 1)
gemm_tc_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, const __grid_constant__ TcOutMaps om,
               const GemmParams p, int m_tiles, int n_tiles, int k_blocks, int total) {
  constexpr int STAGES = TcCfg<BN, FAST>::STAGES;
  constexpr int B_BYTES = TcCfg<BN, FAST>::B_BYTES;
  constexpr int BK = F16 ? 64 : 32;             // elements per k-block
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  uint8_t* sA = smem;
  uint8_t* sB = sA + STAGES * TC_A_BYTES;
  float* stg = reinterpret_cast<float*>(sB + STAGES * B_BYTES);
  uint64_t* bars = reinterpret_cast<uint64_t*>(reinterpret_cast<uint8_t*>(stg) + TcCfg<BN, FAST>::OUT_BYTES);
  uint64_t* full = bars;                 // [STAGES]  TMA -> MMA
  uint64_t* empty = bars + STAGES;       // [STAGES]  MMA -> TMA
  uint64_t* tfull = bars + 2 * STAGES;   // [2]       MMA -> epilogue
  uint64_t* tempty = tfull + 2;          // [2]       epilogue -> MMA
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty + 2);
  uint64_t* lfull = tempty + 3;          // [4] residual box landed      (FAST_BIAS_RESID)
  uint64_t* lempty = lfull + 4;          // [4] residual box stored back (FAST_BIAS_RESID)

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
  }
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < STAGES; ++i) { mbar_init(&full[i], 1); mbar_init(&empty[i], 1); }
    for (int i = 0; i < 2; ++i) { mbar_init(&tfull[i], 1); mbar_init(&tempty[i], 4); }
    if (FAST == FAST_BIAS_RESID)
      for (int i = 0; i < 4; ++i) { mbar_init(&lfull[i], 1); mbar_init(&lempty[i], 1); }
    fence_barrier_init();
  }
  if (warp == 2) tmem_alloc(tmem_slot, 2 * BN);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  pdl_sync();   // barriers, TMEM and descriptors are ready: from here on the predecessor's output is read
  const uint32_t tmem_base = *tmem_slot;
  const int tiles_per_z = m_tiles * n_tiles;

  if (warp == 0) {
    if (lane == 0) {
      int stage = 0; uint32_t phase = 0;
      for (int tile = blockIdx.x; tile < total; tile += gridDim.x) {
        const int z = tile / tiles_per_z, rem = tile - z * tiles_per_z;
        const int m_blk = rem / n_tiles, n_blk = rem - m_blk * n_tiles;
        for (int kb = 0; kb < k_blocks; ++kb) {
          mbar_wait(&empty[stage], phase ^ 1);
          mbar_arrive_expect_tx(&full[stage], TC_A_BYTES + B_BYTES);
          tma_load_3d(sA + stage * TC_A_BYTES, &tmA, &full[stage], kb * BK, m_blk * TC_BM, z);
          tma_load_3d(sB + stage * B_BYTES, &tmB, &full[stage], kb * BK, n_blk * BN, z);
          if (++stage == STAGES) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      constexpr uint32_t idesc = F16 ? umma_idesc_f16(TC_BM, BN) : umma_idesc_tf32(TC_BM, BN);
      int stage = 0; uint32_t phase = 0;
      int it = 0;
      for (int tile = blockIdx.x; tile < total; tile += gridDim.x, ++it) {
        const int as = it & 1;
        const uint32_t aphase = (it >> 1) & 1;
        mbar_wait(&tempty[as], aphase ^ 1);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + as * BN;
        for (int kb = 0; kb < k_blocks; ++kb) {
          mbar_wait(&full[stage], phase);
          tc_fence_after();
          const uint64_t a_desc = umma_desc_k_sw128(smem_u32(sA + stage * TC_A_BYTES));
          const uint64_t b_desc = umma_desc_k_sw128(smem_u32(sB + stage * B_BYTES));
#pragma unroll
          for (int k = 0; k < 4; ++k) {  // 32 bytes of K per MMA: advance the start address inside the swizzle atom
            if (F16) umma_f16_ss(d_tmem, a_desc + (uint64_t)(k * 2), b_desc + (uint64_t)(k * 2), idesc, (kb | k) != 0 ? 1u : 0u);
            else umma_tf32_ss(d_tmem, a_desc + (uint64_t)(k * 2), b_desc + (uint64_t)(k * 2), idesc, (kb | k) != 0 ? 1u : 0u);
          }
          umma_commit(&empty[stage]);
          if (++stage == STAGES) { stage = 0; phase ^= 1; }
        }
        umma_commit(&tfull[as]);
      }
    }
  } else if (FAST != FAST_NONE) {
    epilogue_fast<BN, FAST>(p, om, tmem_base, tfull, tempty, lfull, lempty, reinterpret_cast<uint8_t*>(stg), n_tiles, total);
  } else {
    // ---- epilogue warps: TMEM -> registers -> per-warp smem transpose -> coalesced 128-bit global accesses ----
    const int q = warp & 3;   // TMEM lane quarter this warp may access
    float* my_stg = stg + (warp - 2) * (32 * TC_STG_LD);
    const int sub = lane >> 3, cl = (lane & 7) * 4;   // store phase: lane covers rows sub, sub+4, ... and 4 columns at cl
    int it = 0;
    for (int tile = blockIdx.x; tile < total; tile += gridDim.x, ++it) {
      const int z = tile / tiles_per_z, rem = tile - z * tiles_per_z;
      const int m_blk = rem / n_tiles, n_blk = rem - m_blk * n_tiles;
      const int as = it & 1;
      const uint32_t aphase = (it >> 1) & 1;
      // row bookkeeping of the store phase, once per tile (the integer divisions stay out of the column loop)
      long long crow[8];
      int irow[8];
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const int r = m_blk * TC_BM + q * 32 + i * 4 + sub;
        irow[i] = 0;
        if (r >= p.M) { crow[i] = -1; continue; }
        switch (p.epi) {
          case EPI_BIAS_GELU:
          case EPI_BIAS_RESID: crow[i] = map_c_row(p, r, &irow[i]); break;
          case EPI_QKV:
          case EPI_QKV16: {
            const int bag = r / p.Sp;
            crow[i] = (long long)bag * TM_HEADS * p.Sp + (r - bag * p.Sp);   // row inside head 0 of the bag
            break;
          }
          default: crow[i] = r; break;
        }
      }
      const float* Lz = p.L != nullptr ? p.L + (long long)z * p.strideL : nullptr;
      float* Cz = p.C != nullptr ? p.C + (long long)z * p.strideC : nullptr;
      float* CTz = (p.epi == EPI_AXPBY_T && p.CT != nullptr) ? p.CT + (long long)z * p.strideC : nullptr;
      mbar_wait(&tfull[as], aphase);
      tc_fence_after();
      const uint32_t t_row = tmem_base + ((uint32_t)(q * 32) << 16) + as * BN;
#pragma unroll 1
      for (int chunk = 0; chunk < BN / 32; ++chunk) {
        float v[32];
        tmem_ld_32x32(t_row + chunk * 32, v);
        if (chunk == BN / 32 - 1) {   // accumulator fully read: hand the TMEM buffer back to the MMA warp
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(&tempty[as]);
        }
        const int c0 = n_blk * BN + chunk * 32;
        if (p.epi == EPI_AXPBY_T) {
          // accumulator layout: this lane owns row r_own, columns c0..c0+31 -> the transposed accesses are coalesced
          const int r_own = m_blk * TC_BM + q * 32 + lane;
          const bool ok = r_own < p.M;
          if (Lz != nullptr) {   // all 32 second-term loads in flight before the first store (CT may alias nothing, but the
                                 // compiler cannot know): one L2 round trip per chunk instead of 32
            float lv[32];
#pragma unroll
            for (int e = 0; e < 32; ++e) lv[e] = (ok && (c0 + e) < p.N) ? Lz[(long long)(c0 + e) * p.ldl + r_own] : 0.f;
#pragma unroll
            for (int e = 0; e < 32; ++e) v[e] = fmaf(p.beta, lv[e], p.alpha * v[e]);
          } else {
#pragma unroll
            for (int e = 0; e < 32; ++e) v[e] = p.alpha * v[e];
          }
          if (p.round_out) {
#pragma unroll
            for (int e = 0; e < 32; ++e) v[e] = round_tf32(v[e]);
          }
          if (CTz != nullptr) {
#pragma unroll
            for (int e = 0; e < 32; ++e)
              if (ok && (c0 + e) < p.N) CTz[(long long)(c0 + e) * p.ldc + r_own] = v[e];
          }
        }
#pragma unroll
        for (int i = 0; i < 8; ++i)
          *reinterpret_cast<float4*>(my_stg + lane * TC_STG_LD + i * 4) = make_float4(v[4 * i], v[4 * i + 1], v[4 * i + 2], v[4 * i + 3]);
        __syncwarp();
        const int c = c0 + cl;
        if (c < p.N) {
          float4 acc[8];
#pragma unroll
          for (int i = 0; i < 8; ++i) acc[i] = *reinterpret_cast<const float4*>(my_stg + (i * 4 + sub) * TC_STG_LD + cl);
          switch (p.epi) {
            case EPI_AXPBY: {
              float4 l[8];
              if (Lz != nullptr) {
#pragma unroll
                for (int i = 0; i < 8; ++i)
                  l[i] = crow[i] >= 0 ? *reinterpret_cast<const float4*>(Lz + crow[i] * p.ldl + c) : make_float4(0.f, 0.f, 0.f, 0.f);
              }
#pragma unroll
              for (int i = 0; i < 8; ++i) {
                if (crow[i] < 0) continue;
                float4 o = make_float4(p.alpha * acc[i].x, p.alpha * acc[i].y, p.alpha * acc[i].z, p.alpha * acc[i].w);
                if (Lz != nullptr) {
                  o.x = fmaf(p.beta, l[i].x, o.x); o.y = fmaf(p.beta, l[i].y, o.y);
                  o.z = fmaf(p.beta, l[i].z, o.z); o.w = fmaf(p.beta, l[i].w, o.w);
                }
                if (p.round_out) { o.x = round_tf32(o.x); o.y = round_tf32(o.y); o.z = round_tf32(o.z); o.w = round_tf32(o.w); }
                *reinterpret_cast<float4*>(Cz + crow[i] * p.ldc + c) = o;
              }
              break;
            }
            case EPI_AXPBY_T: {
              if (Cz != nullptr) {
#pragma unroll
                for (int i = 0; i < 8; ++i)
                  if (crow[i] >= 0) *reinterpret_cast<float4*>(Cz + crow[i] * p.ldc + c) = acc[i];
              }
              break;
            }
            case EPI_BIAS_GELU: {
              const float4 b = __ldg(reinterpret_cast<const float4*>(p.bias + c));
#pragma unroll
              for (int i = 0; i < 8; ++i) {
                if (crow[i] < 0) continue;
                float4 o = make_float4(gelu_erf(acc[i].x + b.x), gelu_erf(acc[i].y + b.y), gelu_erf(acc[i].z + b.z), gelu_erf(acc[i].w + b.w));
                if (p.round_out) { o.x = round_tf32(o.x); o.y = round_tf32(o.y); o.z = round_tf32(o.z); o.w = round_tf32(o.w); }
                *reinterpret_cast<float4*>(Cz + crow[i] * p.ldc + c) = o;
                if (irow[i] < p.dup) *reinterpret_cast<float4*>(Cz + (crow[i] + p.rows_per_bag) * p.ldc + c) = o;  // TransMIL.py:180
              }
              break;
            }
            case EPI_QKV:
            case EPI_QKV16: {
              const int cl = c + p.col_off;               // column of the full [q | k | v] projection
              const int which = cl / TM_DIM;              // uniform over the 32-column chunk
              const int cc = cl - which * TM_DIM;
              const int hh = cc / TM_DHEAD, d = cc % TM_DHEAD;
              const float s = which == 0 ? 0.125f : 1.0f;  // q * dim_head^-0.5
              float* base = which == 0 ? p.q : (which == 1 ? p.k : p.v);
#pragma unroll
              for (int i = 0; i < 8; ++i) {
                if (crow[i] < 0) continue;
                const long long off = (crow[i] + (long long)hh * p.Sp) * TM_DHEAD + d;
                if (p.epi == EPI_QKV16) {
                  uint2 o;
                  o.x = pack_f16x2(acc[i].x * s, acc[i].y * s);
                  o.y = pack_f16x2(acc[i].z * s, acc[i].w * s);
                  *reinterpret_cast<uint2*>(reinterpret_cast<__half*>(base) + off) = o;
                } else {
                  float4 o = make_float4(acc[i].x * s, acc[i].y * s, acc[i].z * s, acc[i].w * s);
                  if (p.round_out) { o.x = round_tf32(o.x); o.y = round_tf32(o.y); o.z = round_tf32(o.z); o.w = round_tf32(o.w); }
                  *reinterpret_cast<float4*>(base + off) = o;
                }
              }
              break;
            }
            case EPI_BIAS_RESID: {
              const float4 b = __ldg(reinterpret_cast<const float4*>(p.bias + c));
              float4 l[8];
#pragma unroll
              for (int i = 0; i < 8; ++i)   // all residual loads in flight before the first store (h_out may alias h_in)
                l[i] = crow[i] >= 0 ? *reinterpret_cast<const float4*>(p.L + crow[i] * p.ldl + c) : make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
              for (int i = 0; i < 8; ++i) {
                if (crow[i] < 0) continue;
                const float4 o = make_float4(l[i].x + (acc[i].x + b.x), l[i].y + (acc[i].y + b.y), l[i].z + (acc[i].z + b.z),
                                             l[i].w + (acc[i].w + b.w));
                *reinterpret_cast<float4*>(Cz + crow[i] * p.ldc + c) = o;
              }
              break;
            }
          }
        }
        __syncwarp();
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 2 * BN);
  }
}

template <int BN, bool F16, int FAST>
static int launch_tc(const CUtensorMap& tmA, const CUtensorMap& tmB, const TcOutMaps& om, const GemmParams& p, int batch, cudaStream_t st) {
  static int sms = 0;
  static unsigned long long attr_done = 0;
  if (first_use_on_device(&attr_done)) {
    int dev = 0;
    TM_CUDA(cudaGetDevice(&dev));
    TM_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    TM_CUDA(cudaFuncSetAttribute(gemm_tc_kernel<BN, F16, FAST>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)TcCfg<BN, FAST>::SMEM));
  }
  const int BK = F16 ? 64 : 32;
  const int m_tiles = (int)ceil_div(p.M, TC_BM), n_tiles = (int)ceil_div(p.N, BN), k_blocks = (int)ceil_div(p.K, BK);
  const long long total = (long long)m_tiles * n_tiles * batch;
  TM_CHECK_ARG(total < (1ll << 31), TM_ERR_BAD_SHAPE, "gemm_tc: too many tiles");
  int grid = total < sms ? (int)total : sms;
  if (p.max_ctas > 0 && grid > p.max_ctas) grid = p.max_ctas;
  static const int dbg = [] { const char* e = getenv("TM_GEMM_DBG"); return e ? atoi(e) : 0; }();
  GemmParams pd = p;
  pd.dbg = FAST == FAST_BIAS_RESID ? 0 : dbg;   // the residual ring has its own loader handshake: never switched off
  launch_pdl(gemm_tc_kernel<BN, F16, FAST>, dim3(grid), dim3(TcCfg<BN, FAST>::THREADS), TcCfg<BN, FAST>::SMEM, st, tmA, tmB, om, pd, m_tiles,
             n_tiles, k_blocks, (int)total);
  TM_LAUNCH_CHECK();
  return TM_OK;
}

// Chooses the TMA-store epilogue when the row map of the call is expressible as tensor-map coordinates.
static int pick_fast(const GemmParams& p, int batch, TcOutMaps* om) {
  if (batch != 1 || p.N % 32 != 0 || p.bn128) return FAST_NONE;
  if (p.epi == EPI_QKV16 && p.Sp % TC_BM == 0 && p.M % p.Sp == 0 && p.N + p.col_off <= 3 * TM_DIM && p.N % TM_DIM == 0 &&
      p.col_off % TM_DIM == 0) {
    const uint64_t nb = (uint64_t)(p.M / p.Sp) * TM_HEADS;
    const void* base[3] = {p.q, p.k, p.v};
    for (int i = 0; i < 3; ++i)
      if (make_tmap_3d_f16(&om->m[i], base[i], nb, (uint64_t)p.Sp, TM_DHEAD, (uint64_t)p.Sp * TM_DHEAD, TM_DHEAD, TC_BM, TM_DHEAD) != TM_OK) return FAST_NONE;
    return FAST_QKV16;
  }
  if (p.epi == EPI_BIAS_RESID && p.rows_per_bag > 0 && p.rows_per_bag % TC_BM == 0 && p.M % p.rows_per_bag == 0 && p.ldc % 4 == 0) {
    const uint64_t nb = (uint64_t)(p.M / p.rows_per_bag);
    if (make_tmap_3d(&om->m[0], p.C, nb, (uint64_t)p.c_bag_rows, (uint64_t)p.N, (uint64_t)p.c_bag_rows * p.ldc, (uint64_t)p.ldc, TC_BM, 32) != TM_OK)
      return FAST_NONE;
    if (p.ldl % 4 != 0 ||
        make_tmap_3d(&om->m[2], p.L, nb, (uint64_t)p.c_bag_rows, (uint64_t)p.N, (uint64_t)p.c_bag_rows * p.ldl, (uint64_t)p.ldl, TC_BM, 32) != TM_OK)
      return FAST_NONE;
    if (p.c_skip % TC_BM != 0) {
      const uint32_t rows = TC_BM - p.c_skip % TC_BM;
      if (make_tmap_3d(&om->m[1], p.C, nb, (uint64_t)p.c_bag_rows, (uint64_t)p.N, (uint64_t)p.c_bag_rows * p.ldc, (uint64_t)p.ldc, rows, 32) != TM_OK ||
          make_tmap_3d(&om->m[3], p.L, nb, (uint64_t)p.c_bag_rows, (uint64_t)p.N, (uint64_t)p.c_bag_rows * p.ldl, (uint64_t)p.ldl, rows, 32) != TM_OK)
        return FAST_NONE;
    }
    return FAST_BIAS_RESID;
  }
  if (p.epi == EPI_BIAS_GELU && p.ldc % 4 == 0) {
    // one bag, or bags whose row count is a multiple of the tile (a tile never straddles two bags)
    const int rpb = p.rows_per_bag > 0 ? p.rows_per_bag : p.M;
    if (p.M % rpb != 0 || (p.M != rpb && rpb % TC_BM != 0)) return FAST_NONE;
    const uint64_t nb = (uint64_t)(p.M / rpb);
    const uint64_t bag_rows = p.rows_per_bag > 0 ? (uint64_t)p.c_bag_rows : (uint64_t)p.M;
    if (make_tmap_3d(&om->m[0], p.C, nb, (uint64_t)(p.c_row_off + rpb), (uint64_t)p.N, bag_rows * p.ldc, (uint64_t)p.ldc, TC_BM, 32) != TM_OK)
      return FAST_NONE;
    if (p.dup > 0 && make_tmap_3d(&om->m[1], p.C, nb, bag_rows, (uint64_t)p.N, bag_rows * p.ldc, (uint64_t)p.ldc, TC_BM, 32) != TM_OK)
      return FAST_NONE;
    return FAST_BIAS_GELU;
  }
  return FAST_NONE;
}

int gemm_tc(const GemmParams& p, cudaStream_t st) {
  TM_CHECK_ARG(p.M > 0 && p.N > 0 && p.K > 0, TM_ERR_BAD_SHAPE, "gemm_tc: empty problem");
  TM_CHECK_ARG(!p.b_is_kn, TM_ERR_BAD_ARGUMENT, "gemm_tc: only NT GEMMs (B operand stored [N,K]) are supported");
  TM_CHECK_ARG(p.N % 4 == 0 && p.K % 4 == 0, TM_ERR_BAD_ALIGNMENT, "gemm_tc: N and K must be multiples of 4");
  const int batch = p.batch > 0 ? p.batch : 1;
  const uint64_t sA = batch > 1 ? (uint64_t)p.strideA : (uint64_t)p.M * p.lda;
  const uint64_t sB = batch > 1 ? (uint64_t)p.strideB : (uint64_t)p.N * p.ldb;
  CUtensorMap tmA, tmB;
  TcOutMaps om;
  memset(&om, 0, sizeof(om));
  const int fast = pick_fast(p, batch, &om);
  TM_CHECK_ARG(p.lm_part == nullptr || (fast == FAST_QKV16 && p.f16_in && !p.bn128 && p.lm_l >= TC_BM && p.lm_l * TM_LANDMARKS == p.Sp),
               TM_ERR_BAD_SHAPE, "gemm_tc: landmark partial sums need the fp16 TMA-store qkv epilogue and l >= 128");
  if (p.f16_in) {
    TM_PROPAGATE(make_tmap_3d_f16(&tmA, p.A, batch, (uint64_t)p.M, (uint64_t)p.K, sA, (uint64_t)p.lda, TC_BM, 64));
    if (p.bn128) {
      TM_PROPAGATE(make_tmap_3d_f16(&tmB, p.Bm, batch, (uint64_t)p.N, (uint64_t)p.K, sB, (uint64_t)p.ldb, 128, 64));
      return launch_tc<128, true, FAST_NONE>(tmA, tmB, om, p, batch, st);
    }
    TM_PROPAGATE(make_tmap_3d_f16(&tmB, p.Bm, batch, (uint64_t)p.N, (uint64_t)p.K, sB, (uint64_t)p.ldb, 256, 64));
    switch (fast) {
      case FAST_QKV16: return launch_tc<256, true, FAST_QKV16>(tmA, tmB, om, p, batch, st);
      case FAST_BIAS_RESID: return launch_tc<256, true, FAST_BIAS_RESID>(tmA, tmB, om, p, batch, st);
      case FAST_BIAS_GELU: return launch_tc<256, true, FAST_BIAS_GELU>(tmA, tmB, om, p, batch, st);
      default: return launch_tc<256, true, FAST_NONE>(tmA, tmB, om, p, batch, st);
    }
  }
  TM_PROPAGATE(make_tmap_3d(&tmA, p.A, batch, (uint64_t)p.M, (uint64_t)p.K, sA, (uint64_t)p.lda, TC_BM, 32));
  if (p.bn128) {
    TM_PROPAGATE(make_tmap_3d(&tmB, p.Bm, batch, (uint64_t)p.N, (uint64_t)p.K, sB, (uint64_t)p.ldb, 128, 32));
    return launch_tc<128, false, FAST_NONE>(tmA, tmB, om, p, batch, st);
  }
  TM_PROPAGATE(make_tmap_3d(&tmB, p.Bm, batch, (uint64_t)p.N, (uint64_t)p.K, sB, (uint64_t)p.ldb, 256, 32));
  switch (fast) {
    case FAST_QKV16: return launch_tc<256, false, FAST_QKV16>(tmA, tmB, om, p, batch, st);
    case FAST_BIAS_RESID: return launch_tc<256, false, FAST_BIAS_RESID>(tmA, tmB, om, p, batch, st);
    case FAST_BIAS_GELU: return launch_tc<256, false, FAST_BIAS_GELU>(tmA, tmB, om, p, batch, st);
    default: return launch_tc<256, false, FAST_NONE>(tmA, tmB, om, p, batch, st);
  }
}

}  // namespace tmil
