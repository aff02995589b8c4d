// tcgen05 strided batched GEMM of the BACKWARD:  C[z] = alpha * op(A[z]) op(B[z]) + beta * L[z] + diag * I,
// fp32 tensors in HBM, tf32 tensor-core contraction, fp32 accumulation in TMEM.
//
// The backward's contractions come in all four transposition forms (dX = dY W is NN, dW = dY^T X is TN, the attention
// score gradients are NT, the pinv chain has TT).  tcgen05 takes tf32 operands K-major (k contiguous) in the plain
// SWIZZLE_128B shared-memory layout; MN-major tf32 would need the 32-byte-atom swizzle variant.  So the operand tiles are
// always laid out K-major in shared memory and the transposition is done by the LOADER on the way in:
//   * source k-contiguous:  16-byte cp.async per (row, 4 k) chunk straight to its swizzled place;
//   * source m-contiguous:  4-byte cp.async per element, lanes arranged 8 rows x 4 k so that the global side reads whole
//     32-byte sectors and the shared side hits 32 different banks (chunk index XOR row, as the TMA unit would write it).
// Out-of-range rows / k are zero-filled by the copy (src-size 0), so ragged M, N, K need no separate path.
//
// One persistent CTA per SM, 13 warps:
//   warp 0      MMA issuer (one lane): 4 x tcgen05.mma.kind::tf32 (M128 x Nt x K8) per 32-wide k block, accumulators double
//               buffered in TMEM (2 x Nt columns) so the epilogue of tile i runs under the main loop of tile i+1
//   warps 1-4   epilogue: tcgen05.ld -> per-warp smem transpose -> coalesced 128-bit stores (alpha, beta*L, diag applied)
//   warps 5-12  loaders: 4-stage ring, 3 cp.async groups in flight per thread; a stage is published with
//               cp.async.wait_group -> fence.proxy.async -> mbarrier.arrive (256 arrivals per stage)
// Split-K (weight gradients: M, N small, K = all tokens of the batch): the K range is cut into `splits` slices that become
// extra tiles writing fp32 partials; a second kernel sums them in a fixed order (deterministic, no atomics).
#include <cuda_fp16.h>

#include "common.cuh"
#include "tc_common.cuh"

namespace tmil {

typedef long long i64;

struct BGemmTc {
  const float* A; const float* B; const float* L; float* C;
  int M, N, K;
  i64 sam, sak, sbk, sbn, ldc, ldl;
  i64 strA, strB, strC, strL;
  float alpha, beta, diag;
  int a_km, b_km;          // operand is k-contiguous in memory
  int bn;                  // N tile (64 / 128 / 256)
  int m_tiles, n_tiles;
  int splits, kb_per_split, k_blocks;   // k blocks of 32
  int total;               // tiles = batch * splits * m_tiles * n_tiles
  uint32_t idesc;
};

constexpr int BG_BM = 128;
constexpr int BG_BK = 32;
constexpr int BG_STAGES = 4;
constexpr int BG_LOOKAHEAD = 3;                 // cp.async groups in flight per loader thread
constexpr int BG_A_BYTES = BG_BM * 128;         // 16 KB
constexpr int BG_B_BYTES = 256 * 128;           // 32 KB (sized for the widest N tile)
constexpr int BG_STG_LD = 36;
constexpr int BG_STG_BYTES = 32 * BG_STG_LD * 4;
constexpr int BG_LOADERS = 256;
constexpr int BG_THREADS = 32 + 128 + BG_LOADERS;
constexpr size_t BG_SMEM = 1024 + BG_STAGES * (BG_A_BYTES + BG_B_BYTES) + 4 * BG_STG_BYTES + 256;

__device__ __forceinline__ void cp_async16(uint32_t dst, const void* src, uint32_t src_bytes) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst), "l"(src), "r"(src_bytes) : "memory");
}
__device__ __forceinline__ void cp_async4(uint32_t dst, const void* src, uint32_t src_bytes) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4, %2;" ::"r"(dst), "l"(src), "r"(src_bytes) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

// One operand tile [rows x 32 k] into the K-major SWIZZLE_128B layout at `sdst` (shared address):
//   element (r, k) -> r * 128 + (((k >> 2) ^ (r & 7)) << 4) + (k & 3) * 4
// src(r, k) = base + r * s_row + k * s_k (elements); rows >= row_lim and k >= k_lim read as zero.
// lt = loader thread index (0..255).
__device__ __forceinline__ void load_tile(uint32_t sdst, const float* base, int rows, int row_lim, int k_lim, i64 s_row, i64 s_k,
                                          bool k_major, int lt) {
  if (k_major) {
    // thread: chunk c = lt & 7 (k = 4c .. 4c+3), rows (lt >> 3) + 32 i
    const int c = lt & 7;
    const bool k_ok = 4 * c < k_lim;            // K % 4 == 0: a chunk is entirely inside or outside
    const float* src_c = base + 4 * c;
    for (int r = lt >> 3; r < rows; r += 32) {
      const bool ok = k_ok && r < row_lim;
      const float* src = ok ? src_c + (i64)r * s_row : base;
      cp_async16(sdst + r * 128 + ((c ^ (r & 7)) << 4), src, ok ? 16u : 0u);
    }
  } else {
    // warp w: k = 4 w + ki; lane: mi = lane & 7, ki = lane >> 3; row blocks of 8
    const int w = lt >> 5, lane = lt & 31;
    const int mi = lane & 7, ki = lane >> 3;
    const int k = 4 * w + ki;
    const bool k_ok = k < k_lim;
    const float* src_k = base + (i64)k * s_k;
    const uint32_t d0 = sdst + mi * 128 + ((w ^ mi) << 4) + ki * 4;
    for (int rb = 0; rb < rows; rb += 8) {
      const int r = rb + mi;
      const bool ok = k_ok && r < row_lim;
      const float* src = ok ? src_k + (i64)r * s_row : base;
      cp_async4(d0 + rb * 128, src, ok ? 4u : 0u);
    }
  }
}

__global__ void __launch_bounds__(BG_THREADS, 1) bgemm_tc_kernel(const BGemmTc p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  uint8_t* sA = smem;
  uint8_t* sB = sA + BG_STAGES * BG_A_BYTES;
  float* stg = reinterpret_cast<float*>(sB + BG_STAGES * BG_B_BYTES);
  uint64_t* bars = reinterpret_cast<uint64_t*>(reinterpret_cast<uint8_t*>(stg) + 4 * BG_STG_BYTES);
  uint64_t* full = bars;                   // [STAGES] loaders -> MMA
  uint64_t* empty = bars + BG_STAGES;      // [STAGES] MMA -> loaders
  uint64_t* tfull = bars + 2 * BG_STAGES;  // [2] MMA -> epilogue
  uint64_t* tempty = tfull + 2;            // [2] epilogue -> MMA
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty + 2);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int bn = p.bn;
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < BG_STAGES; ++i) { mbar_init(&full[i], BG_LOADERS); mbar_init(&empty[i], 1); }
    for (int i = 0; i < 2; ++i) { mbar_init(&tfull[i], 1); mbar_init(&tempty[i], 4); }
    fence_barrier_init();
  }
  if (warp == 0) tmem_alloc(tmem_slot, 2 * bn);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const int tiles_mn = p.m_tiles * p.n_tiles;

  if (warp == 0) {
    // ---- MMA issuer ------------------------------------------------------------------------------------------------
    if (lane == 0) {
      int stage = 0; uint32_t phase = 0;
      int it = 0;
      for (int tile = blockIdx.x; tile < p.total; tile += gridDim.x, ++it) {
        const int zs = tile / tiles_mn;
        const int split = zs % p.splits;
        const int kb0 = split * p.kb_per_split;
        const int kb1 = min(p.k_blocks, kb0 + p.kb_per_split);
        const int as = it & 1;
        const uint32_t aphase = (it >> 1) & 1;
        mbar_wait(&tempty[as], aphase ^ 1);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + as * bn;
        for (int kb = kb0; kb < kb1; ++kb) {
          mbar_wait(&full[stage], phase);
          tc_fence_after();
          const uint64_t a_desc = umma_desc_k_sw128(smem_u32(sA + stage * BG_A_BYTES));
          const uint64_t b_desc = umma_desc_k_sw128(smem_u32(sB + stage * BG_B_BYTES));
#pragma unroll
          for (int k = 0; k < 4; ++k)
            umma_tf32_ss(d_tmem, a_desc + (uint64_t)(k * 2), b_desc + (uint64_t)(k * 2), p.idesc, (kb > kb0 || k > 0) ? 1u : 0u);
          umma_commit(&empty[stage]);
          if (++stage == BG_STAGES) { stage = 0; phase ^= 1; }
        }
        umma_commit(&tfull[as]);
      }
    }
    __syncwarp();
  } else if (warp >= 5) {
    // ---- loaders -----------------------------------------------------------------------------------------------------
    const int lt = threadIdx.x - 160;
    int stage = 0; uint32_t phase = 0;
    int pub = 0;        // next stage to publish
    int inflight = 0;   // committed groups not yet published
    for (int tile = blockIdx.x; tile < p.total; tile += gridDim.x) {
      const int zs = tile / tiles_mn, rem = tile - zs * tiles_mn;
      const int z = zs / p.splits, split = zs - z * p.splits;
      const int m_blk = rem / p.n_tiles, n_blk = rem - m_blk * p.n_tiles;
      const int kb0 = split * p.kb_per_split;
      const int kb1 = min(p.k_blocks, kb0 + p.kb_per_split);
      const int m0 = m_blk * BG_BM, n0 = n_blk * bn;
      const float* Ab = p.A + (i64)z * p.strA + (i64)m0 * p.sam;
      const float* Bb = p.B + (i64)z * p.strB + (i64)n0 * p.sbn;
      const int m_lim = p.M - m0, n_lim = p.N - n0;
      for (int kb = kb0; kb < kb1; ++kb) {
        const int k0 = kb * BG_BK;
        mbar_wait(&empty[stage], phase ^ 1);
        load_tile(smem_u32(sA + stage * BG_A_BYTES), Ab + (i64)k0 * p.sak, BG_BM, m_lim, p.K - k0, p.sam, p.sak, p.a_km != 0, lt);
        load_tile(smem_u32(sB + stage * BG_B_BYTES), Bb + (i64)k0 * p.sbk, bn, n_lim, p.K - k0, p.sbn, p.sbk, p.b_km != 0, lt);
        cp_async_commit();
        if (++stage == BG_STAGES) { stage = 0; phase ^= 1; }
        if (++inflight == BG_LOOKAHEAD) {
          cp_async_wait<BG_LOOKAHEAD - 1>();
          fence_proxy_async();
          mbar_arrive(&full[pub]);
          if (++pub == BG_STAGES) pub = 0;
          --inflight;
        }
      }
    }
    // drain
    if (inflight == 2) {
      cp_async_wait<1>();
      fence_proxy_async();
      mbar_arrive(&full[pub]);
      if (++pub == BG_STAGES) pub = 0;
      --inflight;
    }
    if (inflight == 1) {
      cp_async_wait<0>();
      fence_proxy_async();
      mbar_arrive(&full[pub]);
    }
  } else {
    // ---- epilogue: TMEM -> registers -> per-warp smem transpose -> coalesced 128-bit global accesses ------------------------
    const int q = warp & 3;   // TMEM lane quarter this warp may access
    float* my_stg = stg + (warp - 1) * (32 * BG_STG_LD);
    const int sub = lane >> 3, cl = (lane & 7) * 4;   // store phase: rows sub, sub+4, ...; 4 columns at cl
    int it = 0;
    for (int tile = blockIdx.x; tile < p.total; tile += gridDim.x, ++it) {
      const int zs = tile / tiles_mn, rem = tile - zs * tiles_mn;
      const int m_blk = rem / p.n_tiles, n_blk = rem - m_blk * p.n_tiles;
      const int as = it & 1;
      const uint32_t aphase = (it >> 1) & 1;
      // with split-K the partials of (z, split) are stored back to back: zs indexes them directly
      const int zl = p.splits > 1 ? zs / p.splits : zs;
      float* Cz = p.C + (i64)zs * p.strC;
      const float* Lz = p.L != nullptr ? p.L + (i64)zl * p.strL : nullptr;
      mbar_wait(&tfull[as], aphase);
      tc_fence_after();
      const uint32_t t_row = tmem_base + ((uint32_t)(q * 32) << 16) + as * bn;
      const int nchunks = bn >> 5;
#pragma unroll 1
      for (int chunk = 0; chunk < nchunks; ++chunk) {
        float v[32];
        tmem_ld_32x32(t_row + chunk * 32, v);
        if (chunk == nchunks - 1) {   // accumulator fully read: hand the TMEM buffer back to the MMA warp
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(&tempty[as]);
        }
#pragma unroll
        for (int i = 0; i < 8; ++i)
          *reinterpret_cast<float4*>(my_stg + lane * BG_STG_LD + i * 4) = make_float4(v[4 * i], v[4 * i + 1], v[4 * i + 2], v[4 * i + 3]);
        __syncwarp();
        const int c = n_blk * bn + chunk * 32 + cl;
        if (c < p.N) {
          float4 acc[8];
#pragma unroll
          for (int i = 0; i < 8; ++i) acc[i] = *reinterpret_cast<const float4*>(my_stg + (i * 4 + sub) * BG_STG_LD + cl);
          const bool vec = (c + 3 < p.N) && ((p.ldc & 3) == 0);
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            const int r = m_blk * BG_BM + q * 32 + i * 4 + sub;
            if (r >= p.M) continue;
            float o[4] = {p.alpha * acc[i].x, p.alpha * acc[i].y, p.alpha * acc[i].z, p.alpha * acc[i].w};
            if (Lz != nullptr) {
#pragma unroll
              for (int e = 0; e < 4; ++e)
                if (c + e < p.N) o[e] = fmaf(p.beta, Lz[(i64)r * p.ldl + c + e], o[e]);
            }
            if (p.diag != 0.f) {
#pragma unroll
              for (int e = 0; e < 4; ++e)
                if (r == c + e) o[e] += p.diag;
            }
            float* dst = Cz + (i64)r * p.ldc + c;
            if (vec) {
              *reinterpret_cast<float4*>(dst) = make_float4(o[0], o[1], o[2], o[3]);
            } else {
#pragma unroll
              for (int e = 0; e < 4; ++e)
                if (c + e < p.N) dst[e] = o[e];
            }
          }
        }
        __syncwarp();
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 2 * bn);
  }
}

// out[z][m,n] = alpha * sum_s P[z][s][m,n] + beta * L[z][m*ldl + n] + (m == n ? diag : 0); fixed summation order
__global__ void __launch_bounds__(256) splitk_reduce_kernel(const float* __restrict__ P, int splits, i64 mn, int N, i64 batch,
                                                            const float* L, i64 ldl, i64 strL, float* C, i64 ldc, i64 strC,
                                                            float alpha, float beta, float diag) {
  const i64 i = (i64)blockIdx.x * 256 + threadIdx.x;
  if (i >= batch * mn) return;
  const i64 z = i / mn, e = i - z * mn;
  const i64 m = e / N, n = e - m * N;
  const float* src = P + z * splits * mn + e;
  float a = 0.f;
  for (int s = 0; s < splits; ++s) a += src[(i64)s * mn];
  float o = alpha * a;
  if (L != nullptr) o = fmaf(beta, L[z * strL + m * ldl + n], o);
  if (m == n) o += diag;
  C[z * strC + m * ldc + n] = o;
}

static int bg_sms() {
  static int sms = 0;
  if (sms == 0) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  }
  return sms;
}

// Split-K plan of one call: number of K slices (1 = none) and k blocks per slice.
static void bg_split_plan(i64 M, i64 N, i64 K, i64 batch, int bn, int* splits, int* kb_per_split) {
  const int k_blocks = (int)ceil_div(K, BG_BK);
  const i64 tiles = batch * ceil_div(M, BG_BM) * ceil_div(N, bn);
  const int sms = bg_sms() > 0 ? bg_sms() : 148;
  int s = 1;
  if (tiles * 2 <= sms && k_blocks >= 16) {
    s = (int)(sms / tiles);
    const int max_s = k_blocks / 8;          // at least 8 k blocks (256 k) per slice
    if (s > max_s) s = max_s;
    if (s < 1) s = 1;
  }
  int per = (int)ceil_div(k_blocks, s);
  s = (int)ceil_div(k_blocks, per);           // no empty slice
  *splits = s;
  *kb_per_split = per;
}
static int bg_pick_bn(i64 N) { return N <= 64 ? 64 : (N <= 128 ? 128 : 256); }

i64 bgemm_tc_scratch_floats(i64 M, i64 N, i64 K, i64 batch) {
  int s, per;
  bg_split_plan(M, N, K, batch, bg_pick_bn(N), &s, &per);
  return s > 1 ? batch * s * M * N : 0;
}

// Same contract as the CUDA-core bgemm of backward.cu.  scratch: bgemm_tc_scratch_floats() floats (may be NULL: then no split-K).
int bgemm_tc(const float* A, const float* B, float* C, const float* L, i64 M, i64 N, i64 K, i64 sam, i64 sak, i64 sbk, i64 sbn, i64 ldc,
             i64 ldl, i64 batch, i64 strA, i64 strB, i64 strC, i64 strL, float alpha, float beta, float diag, float* scratch,
             i64 scratch_floats, cudaStream_t st) {
  TM_CHECK_ARG(M > 0 && N > 0 && K > 0 && batch > 0, TM_ERR_BAD_SHAPE, "bgemm_tc: empty problem");
  TM_CHECK_ARG(M < (1ll << 31) && N < (1ll << 31) && K < (1ll << 31), TM_ERR_BAD_SHAPE, "bgemm_tc: dimension too large");
  TM_CHECK_ARG(tc_available(), TM_ERR_UNSUPPORTED_ARCH, "bgemm_tc needs an sm_100 device");
  static unsigned long long attr_done = 0;
  if (first_use_on_device(&attr_done))
    TM_CUDA(cudaFuncSetAttribute(bgemm_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)BG_SMEM));
  BGemmTc p;
  memset(&p, 0, sizeof(p));
  p.A = A; p.B = B; p.L = L; p.C = C;
  p.M = (int)M; p.N = (int)N; p.K = (int)K;
  p.sam = sam; p.sak = sak; p.sbk = sbk; p.sbn = sbn; p.ldc = ldc; p.ldl = ldl;
  p.strA = strA; p.strB = strB; p.strC = strC; p.strL = strL;
  p.alpha = alpha; p.beta = beta; p.diag = diag;
  // the 16-byte copy path needs k-contiguous, 16-byte aligned rows and K % 4 == 0; anything else takes the element path
  auto vec_ok = [&](const float* ptr, i64 s_row, i64 s_k, i64 str) {
    return s_k == 1 && (K & 3) == 0 && (s_row & 3) == 0 && (str & 3) == 0 && aligned16(ptr);
  };
  p.a_km = vec_ok(A, sam, sak, strA) ? 1 : 0;
  p.b_km = vec_ok(B, sbn, sbk, strB) ? 1 : 0;
  p.bn = bg_pick_bn(N);
  p.idesc = umma_idesc_tf32(BG_BM, p.bn);
  p.m_tiles = (int)ceil_div(M, BG_BM);
  p.n_tiles = (int)ceil_div(N, p.bn);
  p.k_blocks = (int)ceil_div(K, BG_BK);
  bg_split_plan(M, N, K, batch, p.bn, &p.splits, &p.kb_per_split);
  if (p.splits > 1 && (scratch == nullptr || scratch_floats < batch * p.splits * M * N)) {
    p.splits = 1;
    p.kb_per_split = p.k_blocks;
  }
  const i64 total = batch * p.splits * p.m_tiles * p.n_tiles;
  TM_CHECK_ARG(total < (1ll << 31), TM_ERR_BAD_SHAPE, "bgemm_tc: too many tiles");
  p.total = (int)total;
  if (p.splits > 1) {   // partials: [batch][splits][M][N], plain stores
    p.C = scratch; p.ldc = N; p.strC = M * N; p.L = nullptr; p.alpha = 1.f; p.beta = 0.f; p.diag = 0.f;
  }
  const int sms = bg_sms();
  const int grid = total < sms ? (int)total : sms;
  bgemm_tc_kernel<<<grid, BG_THREADS, BG_SMEM, st>>>(p);
  TM_LAUNCH_CHECK();
  if (p.splits > 1) {
    const i64 n_out = batch * M * N;
    splitk_reduce_kernel<<<(unsigned)ceil_div(n_out, 256), 256, 0, st>>>(scratch, p.splits, M * N, (int)N, batch, L, ldl, strL, C, ldc, strC,
                                                                        alpha, beta, diag);
    TM_LAUNCH_CHECK();
  }
  return TM_OK;
}

}  // namespace tmil
